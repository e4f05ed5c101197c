#!/usr/bin/env python
"""Benchmark of the blocked Gibbs sweep (BASELINE.json metric: frames·sweeps/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--config cfg1|cfg2|cfg3|cfg4-shard|cfg5-K20|cfg5-K100|cfg5-K500] [--cold-start]

Workloads (synthetic draws from the model, the example pickles are stripped):
  cfg4-shard (default): BASELINE.json configs[3] — 1000 sessions x 36k frames, 16 keypoints, D=10,
             L=3, K=100 sharded over 8 GPUs, i.e. 125 sessions per GPU.  At --gpus 1 one shard runs (the
             largest slice of the headline config that fits one GPU with its filter workspace); at
             --gpus 8 the job is exactly configs[3] (weak scaling: per-GPU work is fixed).
  cfg1       configs[0] stand-in: 1 session x 36k frames, 8 keypoints, D=7, L=3, K=100 (full sweep)
  cfg2       configs[1] stand-in: AR-HMM on 1 session x 36k frames, D=10, L=3, K=100
  cfg3       configs[2]: 100 sessions x 36k frames, 8 keypoints, D=10, L=3, K=50 (full sweep)
  cfg5-K*    configs[4]: AR-HMM, 32 sessions x 31 250 frames (1 M frames), K = 20 / 100 / 500
A "step" is one sweep: the eight resample_* steps of keypoint_slds.ipynb cell 12 (full model), or
the three steps of depth_ar_hmm.ipynb cell 8 / keypoint_slds.ipynb cell 9 (AR-HMM configs).

`value`   : inputs resident in HBM, CUDA events, max over ranks.
`e2e`     : the same sweep driven from pinned HOST buffers (parallel.gibbs_sweep_host /
            arhmm_sweep_host): every step uploads the data and all states and reads the new states and
            parameters back; the step ends when the host copies do.  `e2e_resident` (reported next to
            it) keeps Y and mask on the device across sweeps and moves only the states.
`roofline`: the dominant entry point timed alone with CUDA events on the launching stream;
            `roofline_per_kernel` the same for every entry point of the sweep, each against the peak
            of the resource that bounds it (HBM copy bandwidth from MEASURED_PEAKS.json; TF32 and FP64
            GEMM peaks measured in this run with cuBLAS).
`cpu_baseline` / `--impl reference`: the float64 NumPy+numba oracle (the reference's JAX path cannot
            run here: JAX/TFP are not installed) on the host cores, on a bounded sample of the same
            workload; the line states the sample that was actually timed.
`--cold-start`: early-fit regime — AR parameters and transitions drawn from the priors, random state
            sequence — where the HMM forgets slowly and the chunk warm-up ladder is exercised.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "frames·sweeps/sec (full Gibbs sweep)"
UNIT = "frames·sweeps/s"
CONFIGS = {
    "cfg1": dict(sessions=1, frames=36000, k=8, dk=2, D=7, L=3, K=100, model="slds"),
    "cfg2": dict(sessions=1, frames=36000, k=8, dk=2, D=10, L=3, K=100, model="arhmm"),
    "cfg3": dict(sessions=100, frames=36000, k=8, dk=2, D=10, L=3, K=50, model="slds"),
    "cfg4-shard": dict(sessions=125, frames=36000, k=16, dk=2, D=10, L=3, K=100, model="slds"),
    "cfg5-K20": dict(sessions=32, frames=31250, k=8, dk=2, D=10, L=3, K=20, model="arhmm"),
    "cfg5-K100": dict(sessions=32, frames=31250, k=8, dk=2, D=10, L=3, K=100, model="arhmm"),
    "cfg5-K500": dict(sessions=32, frames=31250, k=8, dk=2, D=10, L=3, K=500, model="arhmm"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg4-shard", choices=sorted(CONFIGS))
    ap.add_argument("--sessions-per-gpu", type=int, default=0, help="override the config's sessions per GPU")
    ap.add_argument("--frames", type=int, default=0, help="override the config's frames per session")
    ap.add_argument("--cold-start", action="store_true")
    ap.add_argument("--e2e-groups", type=int, default=0,
                    help="session groups of the host-driven sweep (gibbs_sweep_host(groups=)); 0 = its default (2, or 4 with more than 2 ranks)")
    ap.add_argument("--settle", type=int, default=6,
                    help="untimed sweeps before the warm-up sweeps, in which the adaptive chunk warm-up of the filters "
                         "settles (a fit runs hundreds of sweeps; the timed region measures its steady state)")
    ap.add_argument("--chunk-len", default="auto", help="Kalman chunk length: auto, 0 (= one warp per session) or steps")
    ap.add_argument("--burn-in", type=int, default=0, help="Kalman chunk warm-up steps (0 = the package default)")
    ap.add_argument("--hmm-burn-in", type=int, default=0, help="HMM chunk warm-up frames (0 = the package default)")
    ap.add_argument("--cpu-sessions", type=int, default=0, help="0 = one per host core (max 16)")
    ap.add_argument("--cpu-frames", type=int, default=0, help="0 = sized from --steps so that the run takes minutes")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-breakdown", action="store_true")
    return ap.parse_args()


def shape_of(args):
    sh = dict(CONFIGS[args.config])
    if args.sessions_per_gpu:
        sh["sessions"] = args.sessions_per_gpu
    if args.frames:
        sh["frames"] = args.frames
    return sh


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self._stop, self._th = index, [], threading.Event(), None

    def _run(self):
        # NVML in-process (one init, cheap queries); the nvidia-smi subprocess is only the fallback:
        # spawning it several times a second from every rank stalls CUDA calls of all processes.
        nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            hd = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            nv = pynvml
        except Exception:
            nv = None
        while not self._stop.is_set():
            try:
                if nv is not None:
                    sm = nv.nvmlDeviceGetClockInfo(hd, nv.NVML_CLOCK_SM)
                    mx = nv.nvmlDeviceGetMaxClockInfo(hd, nv.NVML_CLOCK_SM)
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(hd) if hasattr(
                        nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(hd)
                    bit = lambda name: "Active" if rs & getattr(nv, name, 0) else "Not Active"
                    self.rows.append([str(sm), str(mx), bit("nvmlClocksThrottleReasonHwSlowdown"),
                                      bit("nvmlClocksThrottleReasonHwThermalSlowdown"),
                                      bit("nvmlClocksThrottleReasonSwThermalSlowdown"),
                                      bit("nvmlClocksThrottleReasonSwPowerCap")])
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                         timeout=5).stdout.strip()
                    if out:
                        self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.05 if nv is not None else 0.5)

    def __enter__(self):
        self._th = threading.Thread(target=self._run, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._th.join(timeout=6)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nm, val in zip(names, r[2:6]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU arm
def cpu_sweep_rate(sh, n_sessions, frames, sweeps=2, warm=1):
    """Time the oracle's sweep (full or AR-HMM) on the host cores (numba prange over sessions).
    Returns (frames·sweeps/s, mean seconds per sweep, per-sweep seconds)."""
    from keypointmoseq_b200.synthetic import simulate, default_hypparams
    from oracle import kpms_oracle as O
    slds = sh["model"] == "slds"
    N, T, k, dk, D, L, K = n_sessions, frames, sh["k"], sh["dk"], sh["D"], sh["L"], sh["K"]
    data, st, pr, Gamma = simulate(7, N, T, k=k, dk=dk, D=D, L=L, K=K)
    d64 = {a: b.astype(np.float64) for a, b in data.items()}
    s64 = {a: (b.astype(np.float64) if a != "z" else b) for a, b in st.items()}
    trans, ar, obs, loc = default_hypparams(D, L, K, full_model=slds)
    hyp = {}
    for h in (trans, ar, obs, loc):
        hyp.update(h)
    rng = np.random.default_rng(3)
    n = D * L

    def draws():
        cnt_max = N * (T - L)
        dr = dict(ar_chi2=rng.chisquare(hyp["nu_0"] + 50.0, size=(K, D)), ar_normals=rng.standard_normal((K, D * (D - 1) // 2)),
                  ar_G=rng.standard_normal((K, D, n + 1)), crp_u=rng.random(cnt_max),
                  betas_g=lambda c: rng.gamma(c), pi_g=lambda c: rng.gamma(c), z_u=rng.random((N, T - L)))
        if slds:
            dr.update(sigmasq_g=rng.gamma((hyp["nu_sigma"] + 3 * N * T) / 2, size=k),
                      x_w=rng.standard_normal((N, T - L, n)), x_w_last=rng.standard_normal((N, n)),
                      h_u=rng.random((N, T, 8, 3)) * 0.998 + 0.001, v_w=rng.standard_normal((N, T - 1, dk)),
                      v_w_last=rng.standard_normal((N, dk)), s_g=rng.gamma((hyp["nu_s"] + 3) / 2, size=(N, T, k)))
        return dr
    times = []
    for i in range(warm + sweeps):
        dr = draws()
        t0 = time.perf_counter()
        if slds:
            s64, pr = O.sweep(d64, s64, pr, hyp, dr, Gamma)
        else:
            s64, pr = O.arhmm_sweep(d64, s64, pr, hyp, dr)
        dt = time.perf_counter() - t0
        if i >= warm:
            times.append(dt)
    per = float(np.mean(times))
    return N * T / per, per, times


def cpu_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_sample_shape(args, sh, total_sweeps, budget_s):
    """Bounded CPU sample of the workload: one session per host thread (never more than the config
    has), frames per session sized so that `total_sweeps` sweeps take about `budget_s` seconds."""
    cores = cpu_cores()
    ns = args.cpu_sessions or max(1, min(cores, 16, sh["sessions"]))
    if args.cpu_frames:
        fr = args.cpu_frames
    else:
        per_frame = 1.0e-3 if sh["model"] == "slds" else 4.5e-4 * max(1.0, sh["K"] / 100.0)   # s per frame per thread, measured
        fr = int(budget_s / max(total_sweeps, 1) / per_frame / max(1, -(-ns // max(cores, 1))))
        fr = max(500, min(sh["frames"], fr // 500 * 500))
    return ns, fr


def run_reference(args, rank):
    """The reference arm: one step = one sweep of the bounded sample; W warm-up sweeps, then exactly K
    timed ones.  The line says what ran (`steps`, `config.sessions/frames` = the sample) and what it
    stands for (`config.extrapolated_from`)."""
    if rank != 0:
        return
    sh = shape_of(args)
    cores = cpu_cores()
    try:
        import numba
        numba.set_num_threads(min(cores, numba.config.NUMBA_NUM_THREADS))
        threads = numba.get_num_threads()
    except Exception:
        threads = 1
    ns, fr = cpu_sample_shape(args, sh, args.steps + args.warmup, 150.0)
    rate, per, times = cpu_sweep_rate(sh, ns, fr, sweeps=args.steps, warm=args.warmup)
    shape = {a: sh[a] for a in ("k", "dk", "D", "L", "K")}
    sample = "%d sessions x %d frames (k=%d,D=%d,L=%d,K=%d), %s sweep, oracle NumPy+numba float64" % (
        ns, fr, sh["k"], sh["D"], sh["L"], sh["K"], sh["model"])
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
            "steps": len(times), "warmup": args.warmup, "ms_per_step": per * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.config, "sweep": sh["model"], "sessions": ns, "frames": fr, **shape,
                       "extrapolated_from": {"sessions_per_gpu": sh["sessions"], "frames": sh["frames"],
                                             "why": "the metric is a per-frame rate and the sweep is linear in "
                                                    "sessions x frames; the full shard is hours per sweep on CPU"},
                       "note": "reference JAX-CPU path unavailable on this box (no jax/tfp); "
                               "oracle port timed on the bounded sample named here"},
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- peaks
def measured_gemm_peaks(torch, dev):
    """cuBLAS GEMM rates measured in this run (best of 5, CUDA events): TF32 (the 3xTF32 emission and
    HMM kernels issue three TF32 MMAs per fp32-equivalent product) and FP64 (DGEMM runs on the FP64
    tensor cores; the DFMA pipe has the same nominal rate)."""
    out = {}
    old = torch.backends.cuda.matmul.allow_tf32
    try:
        for name, dt, n, tf32 in (("tf32_tflops", torch.float32, 8192, True), ("fp64_tflops", torch.float64, 4096, False)):
            torch.backends.cuda.matmul.allow_tf32 = tf32
            a = torch.randn((n, n), dtype=dt, device=dev)
            b = torch.randn((n, n), dtype=dt, device=dev)
            torch.matmul(a, b)
            torch.cuda.synchronize()
            best = 1e30
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                torch.matmul(a, b)
                e1.record()
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            out[name] = 2.0 * n ** 3 / (best * 1e-3) / 1e12
            del a, b
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
    torch.cuda.empty_cache()
    return out


# ----------------------------------------------------------------------------- GPU arm
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    import torch
    import torch.distributed as td
    from keypointmoseq_b200 import _lib, gibbs as G, initialize as I, parallel
    from keypointmoseq_b200.synthetic import simulate, default_hypparams

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        td.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    G.LATENTS_CHUNKING["chunk_len"] = args.chunk_len if args.chunk_len == "auto" else int(args.chunk_len)
    if args.burn_in > 0:
        G.LATENTS_CHUNKING["burn_in"] = args.burn_in
    if args.hmm_burn_in > 0:
        G.STATESEQ_CHUNKING["burn_in"] = args.hmm_burn_in

    sh = shape_of(args)
    slds = sh["model"] == "slds"
    N, T = sh["sessions"], sh["frames"]
    k, dk, D, L, K = sh["k"], sh["dk"], sh["D"], sh["L"], sh["K"]
    off = rank * N          # contiguous shards of N sessions, in rank order
    pr0 = simulate(999, 1, 8, k=k, dk=dk, D=D, L=L, K=K)[2]  # the model all shards are drawn from; replicated params
    data, st, pr, Gamma = simulate(1000 + rank, N, T, k=k, dk=dk, D=D, L=L, K=K, params=pr0)
    trans, ar, obs, loc = default_hypparams(D, L, K, full_model=slds)
    hyp = {}
    for h in (trans, ar, obs, loc):
        hyp.update({a: (torch.as_tensor(b, device=dev) if isinstance(b, np.ndarray) else b) for a, b in h.items()})
    if not slds:
        data = {"mask": data["mask"]}
        st = {"x": st["x"], "z": st["z"]}
        pr0 = {a: pr0[a] for a in ("Ab", "Q", "pi", "betas")}
    host = {a: torch.as_tensor(b).pin_memory() for a, b in data.items()}
    host_st = {a: torch.as_tensor(b.astype(np.int32) if a == "z" else b).pin_memory() for a, b in st.items()}
    d = {a: b.to(dev, non_blocking=True) for a, b in host.items()}
    s = {a: b.to(dev, non_blocking=True) for a, b in host_st.items()}
    p = {a: torch.as_tensor(b, device=dev) for a, b in pr0.items()}
    if args.cold_start:
        # the state of a fit right after initialisation (keypoint_slds.ipynb cell 6): prior-drawn AR
        # parameters and transitions, and a state sequence sampled under them
        p["Ab"], p["Q"] = I.initial_ar_params(0, **hyp, device=dev)
        p["betas"], p["pi"] = I.initial_hdp_transitions(0, num_states=K, alpha=hyp["alpha"], kappa=hyp["kappa"],
                                                        gamma=hyp["gamma"], device=dev)
        s["z"] = torch.randint(0, K, s["z"].shape, device=dev, dtype=torch.int32,
                               generator=torch.Generator(device=dev).manual_seed(5 + rank))
        host_st["z"].copy_(s["z"])
    Gm = torch.as_tensor(Gamma, device=dev)
    frames_total = N * T * world

    def barrier():
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        return float(t.item())

    def sweep(i, dd, ss, pp):
        if slds:
            return parallel.gibbs_sweep((0, i), dd, ss, pp, hyp, Gamma=Gm, session_offset=off, next_key=(0, i + 1))
        return parallel.arhmm_sweep((0, i), dd, ss, pp, hyp, session_offset=off)

    # ---- device-resident arm
    ss, pp = s, p
    if args.cold_start:
        args.settle = 0                         # the cold regime IS the first sweeps
    for i in range(args.settle):                # the adaptive warm-up multipliers settle (not timed)
        ss, pp = sweep(1000 + i, d, ss, pp)
    for i in range(args.warmup):
        ss, pp = sweep(i, d, ss, pp)
    barrier()
    fb0 = {a: (b if not isinstance(b, list) else len(b)) for a, b in G.FALLBACKS.items()}
    l0 = lib.kpms_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        ev0.record()
        for i in range(args.steps):
            ss, pp = sweep(100 + i, d, ss, pp)
        ev1.record()
        barrier()
    launches = int(lib.kpms_launch_count() - l0)
    my_ms = ev0.elapsed_time(ev1) / args.steps
    ms = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
    value = frames_total / (ms * 1e-3)
    fallbacks_timed = {a: (b - fb0.get(a, 0)) for a, b in G.FALLBACKS.items() if not isinstance(b, list)}
    per_rank = None
    if world > 1:   # diagnostics: every rank's own time and chunk retries during the timed sweeps
        per_rank = [None] * world
        td.all_gather_object(per_rank, {"ms": round(my_ms, 2), "fallbacks": fallbacks_timed,
                                        "hmm_burn_boost": G._HMM_ADAPT["boost"]})
    clocks = clk.summary()
    # one more sweep with an event after every phase: where the time of a sweep goes (ms since its start)
    timeline = None
    if slds:
        marks = []

        def _mark(label):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks.append((label, e, time.perf_counter()))
        torch.cuda.synchronize()
        _mark("start")
        parallel.gibbs_sweep((0, 999), d, ss, pp, hyp, Gamma=Gm, session_offset=off, _mark=_mark)
        torch.cuda.synchronize()
        timeline = {lab: {"gpu_ms": round(marks[0][1].elapsed_time(e), 3), "host_ms": round((t - marks[0][2]) * 1e3, 3)}
                    for lab, e, t in marks[1:]}

    # ---- per-step breakdown and rooflines (rank-local)
    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    try:
        mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak, peak_src = float(mp["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        pass
    reps = max(2, min(args.steps, 5))
    kw = dict(Gamma=Gm)
    bd, roofline, per_kernel, boundary_err = {}, None, None, None
    n = D * L
    F_ = N * T
    if not args.no_breakdown:
        gemm = measured_gemm_peaks(torch, dev) if rank == 0 else {}
        bd["ar_stats"] = timed(lambda: G.ar_sufficient_stats(**d, **ss, **hyp), reps)
        bd["hdp_transitions"] = timed(lambda: G.resample_hdp_transitions(1, **d, **ss, **pp, **hyp), reps)
        bd["stateseqs"] = timed(lambda: G.resample_stateseqs(1, **d, **ss, **pp, strict="defer"), reps)
        bd["ar_loglik"] = timed(lambda: G.ar_log_likelihood(**ss, **pp), reps)
        u_pre = torch.rand((N, T - L), device=dev)
        hmm_only = timed(lambda: G.resample_stateseqs(1, **d, **ss, **pp, strict="defer", draws={"u": u_pre}), reps) - bd["ar_loglik"]
        del u_pre
        tf32 = gemm.get("tf32_tflops")
        fp64 = gemm.get("fp64_tflops")

        def row(kernel, bound, work, unit_peak, t_ms, note=""):
            ach = work / (t_ms * 1e-3)
            return {"kernel": kernel, "bound": bound, "ms": t_ms, "achieved": ach / 1e9 if bound == "hbm" else ach / 1e12,
                    "peak": unit_peak, "unit": "GB/s" if bound == "hbm" else "TFLOP/s",
                    "frac": (ach / 1e9 / unit_peak) if bound == "hbm" else ((ach / 1e12 / unit_peak) if unit_peak else None),
                    "note": note}
        per_kernel = [
            row("kpms_ar_loglik (emission GEMM, tcgen05 3xTF32)", "tensor", 840.0 * K * F_, (tf32 / 3.0) if tf32 else None,
                bd["ar_loglik"], "840*K flop/frame fp32-equivalent vs measured cuBLAS TF32 / 3"),
            row("kpms_hmm_ffbs (forward filter + stay bits + sampler)", "hbm", (12.0 * K + 12) * F_, peak, hmm_only,
                "12K+12 B/frame nominal"),
            row("kpms_ar_stats (sort + DMMA SYRK)", "hbm", 48.0 * F_, peak, bd["ar_stats"], "48 B/frame"),
        ]
        if slds:
            pdim = k * dk
            bd["obs_variance_stats"] = timed(lambda: G.obs_variance_stats(**d, **ss, **pp, **kw), reps)
            bd["latents"] = timed(lambda: G.resample_latents(1, **d, **ss, **pp, **kw, strict="defer"), reps)
            bd["heading"] = timed(lambda: G.resample_heading(1, **d, **ss, **pp, **kw), reps)
            bd["location"] = timed(lambda: G.resample_location(1, **d, **ss, **pp, **hyp, **kw), reps)
            bd["scales"] = timed(lambda: G.resample_scales(1, **d, **ss, **pp, **hyp, **kw), reps)
            # latents with pre-drawn normals = the kernels alone (no Philox fill in the timed region)
            w = torch.randn((N, T - L, n), device=dev)
            wl = torch.randn((N, n), device=dev)
            lat_ms = timed(lambda: G.resample_latents(1, **d, **ss, **pp, **kw, strict="defer",
                                                      draws={"w": w, "w_last": wl}), reps)
            _, linfo = G.resample_latents(1, **d, **ss, **pp, **kw, draws={"w": w, "w_last": wl}, return_info=True)
            boundary_err = [float(v) for v in linfo["boundary_err"].cpu().tolist()]
            del w, wl
            bytes_per_frame = 4 * (pdim + k + dk + 3 + D) + 4 * n + 2 * 8 * (n + n * (n + 1) // 2)
            achieved = F_ * bytes_per_frame / (lat_ms * 1e-3) / 1e9
            traffic = None
            try:   # DRAM bytes of the entry point's kernels from an `ncu --set full` capture of this workload
                tj = json.load(open(os.path.join(ROOT, "profiles", "r02_latents_traffic.json")))
                if int(tj["frames"]) == F_ and tj.get("config") == args.config:
                    traffic = float(tj["total"])
            except Exception:
                pass
            roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": traffic, "algorithmic_bytes_per_launch": F_ * bytes_per_frame,
                        "kernel": "kpms_resample_latents (kalman_obs + kalman_forward + kalman_backward)",
                        "algorithmic_bytes_per_frame": bytes_per_frame, "ms_per_launch": lat_ms, "peak_source": peak_src}
            obs_b = 4.0 * (pdim + D + dk + 1 + k)
            flop_step = 2.0 * (n ** 3 / 3 + n ** 3 / 6 + n * n * (D + 1) / 2 + n * (n + 1) * (n + 1) / 2)
            per_kernel += [
                row("kpms_resample_latents (Kalman filter + sampler)", "hbm", F_ * bytes_per_frame, peak, lat_ms,
                    "%d B/frame incl. one write + one read of the filter record" % bytes_per_frame),
                row("kpms_resample_latents, FP64 view", "fp64", flop_step * F_, fp64, lat_ms,
                    "%.0f flop/step (Cholesky + inverse + T + Z, no padding) vs measured cuBLAS DGEMM" % flop_step),
                row("kpms_obs_variance_stats", "hbm", (obs_b + 4) * F_, peak, bd["obs_variance_stats"], "%.0f B/frame" % (obs_b + 4)),
                row("kpms_resample_heading", "hbm", (obs_b + 4) * F_, peak, bd["heading"], "%.0f B/frame" % (obs_b + 4)),
                row("kpms_resample_location", "hbm", (obs_b + 4 * (dk + 1) + 40) * F_, peak, bd["location"], "stats + scalar scan"),
                row("kpms_resample_scales", "hbm", (obs_b + 4 * k) * F_, peak, bd["scales"], "%.0f B/frame" % (obs_b + 4 * k)),
            ]
        else:
            roofline = {"bound": "hbm", "achieved": (12.0 * K + 12) * F_ / (hmm_only * 1e-3) / 1e9, "peak": peak,
                        "unit": "GB/s", "frac": (12.0 * K + 12) * F_ / (hmm_only * 1e-3) / 1e9 / peak, "traffic": None,
                        "algorithmic_bytes_per_launch": (12.0 * K + 12) * F_, "algorithmic_bytes_per_frame": 12 * K + 12,
                        "kernel": "kpms_hmm_ffbs (forward filter + stay bits + backward sampler)",
                        "ms_per_launch": hmm_only, "peak_source": peak_src}
        for r in per_kernel:
            r["peak_source"] = peak_src if r["bound"] == "hbm" else "cuBLAS GEMM measured in this run"
        bd["hmm_only"] = hmm_only
        bd["gemm_peaks_measured"] = gemm

    # ---- end-to-end arm: host buffers in, host buffers out, every step
    e2e = e2e_res = None
    if not args.no_e2e:
        out_host = {a: torch.empty_like(b).pin_memory() for a, b in host_st.items() if (slds or a == "z")}
        out_alt = {a: torch.empty_like(b).pin_memory() for a, b in out_host.items()}
        nbytes = lambda dct: sum(b.numel() * b.element_size() for a, b in dct.items() if not a.startswith("_"))
        h2d = nbytes(host) + nbytes(host_st)
        par_host = None
        cur = {"in": host_st, "out": out_host, "spare": out_alt}

        def e2e_step(i, pp_, resident=None):
            nonlocal par_host
            # public host-side API: uploads ordered by first use, downloads as states become final.  The loop of a
            # fit: every sweep is fed the host tensors the previous sweep wrote (two sets of pinned buffers, used
            # alternately), so from the second sweep on the pose steps run session group by session group.
            if slds:
                o, p_new, _ = parallel.gibbs_sweep_host((0, i), host, cur["in"], pp_, hyp, Gamma=Gm, session_offset=off,
                                                        out=cur["out"], device=dev, resident=resident,
                                                        groups=args.e2e_groups or None)
                cur["in"], cur["out"], cur["spare"] = o, cur["spare"], cur["out"]
            else:
                _, p_new, _ = parallel.arhmm_sweep_host((0, i), host, host_st, pp_, hyp, session_offset=off,
                                                        out=out_host, device=dev)
            par_host = {a: b.cpu() for a, b in p_new.items()}   # small; synchronises the step
            return p_new

        def run_e2e(resident):
            pe = p
            cur.update({"in": host_st, "out": out_host, "spare": out_alt})
            for o_ in (out_host, out_alt):
                o_.pop("_carry", None)
            for i in range(min(args.warmup, 2)):
                pe = e2e_step(200 + i, pe, resident)
            barrier()
            ev0.record()
            for i in range(args.steps):
                pe = e2e_step(300 + i, pe, resident)
            ev1.record()
            barrier()
            return max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
        ems = run_e2e(None)
        d2h = nbytes(out_host) + nbytes(par_host)
        e2e = {"value": frames_total / (ems * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": ems}
        if slds:   # the caller keeps the data (Y, mask) on the device and moves only the states
            rms = run_e2e({"Y": d["Y"], "mask": d["mask"]})
            e2e_res = {"value": frames_total / (rms * 1e-3), "unit": UNIT, "ms_per_step": rms,
                       "h2d_bytes_per_step": int(nbytes(host_st)), "d2h_bytes_per_step": int(d2h),
                       "note": "Y and mask resident on the device across sweeps (gibbs_sweep_host(resident=...))"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            import numba
            threads = numba.get_num_threads()
        except Exception:
            threads = 1
        ns, fr = cpu_sample_shape(args, sh, 3, 24.0)
        rate, per, _ = cpu_sweep_rate(sh, ns, fr, sweeps=2, warm=1)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "%d sessions x %d frames, same k/D/L/K, %s sweep, oracle NumPy+numba float64, %.2f s/sweep" % (
                   ns, fr, sh["model"], per)}

    if rank == 0:
        cfg = {"workload": args.config, "sweep": sh["model"], "sessions_per_gpu": N, "frames": T,
               **{a: sh[a] for a in ("k", "dk", "D", "L", "K")}, "total_sessions": N * world,
               "cold_start": bool(args.cold_start), "settle_sweeps": args.settle, "e2e_session_groups": args.e2e_groups or (2 if world <= 2 else 4),
               "chunk_fallbacks_timed": fallbacks_timed, "chunk_fallbacks_all": dict(G.FALLBACKS),
               "hmm_burn_boost": G._HMM_ADAPT["boost"], "kalman_burn_boost": G._LAT_ADAPT["boost"], "per_rank": per_rank,
               "l2": "per-step working set (filter workspace / ll / messages, GBs) far exceeds the 126 MB L2"}
        if slds:
            cfg.update(kalman_chunk_len=(G.auto_chunk_len(N, T, L, G.LATENTS_CHUNKING["burn_in"], dev)
                                         if G.LATENTS_CHUNKING["chunk_len"] == "auto" else G.LATENTS_CHUNKING["chunk_len"]),
                       kalman_burn_in=G.LATENTS_CHUNKING["burn_in"], kalman_boundary_err=boundary_err)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64",
                "dtype_note": "f64: Kalman filter / sampler, all statistics and parameters; f32 storage of per-frame arrays; "
                              "HMM messages f32, emission GEMM 3xTF32 with fp32 accumulate",
                "data": "synthetic", "config": cfg, "clocks": clocks, "e2e": e2e, "e2e_resident": e2e_res,
                "gpu_launches": launches, "roofline": roofline, "roofline_per_kernel": per_kernel,
                "cpu_baseline": cpu, "breakdown_ms": bd, "sweep_timeline": timeline}
        print(json.dumps(line), flush=True)
    if world > 1:
        td.destroy_process_group()


if __name__ == "__main__":
    main()
