#!/usr/bin/env python
"""Benchmark of the blocked Gibbs sweep (BASELINE.json metric: frames·sweeps/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (config.workload = "cfg4-shard"): BASELINE.json configs[3] — 1000 sessions x 36k
frames, 16 keypoints, D=10, L=3, K=100 sharded over 8 GPUs — i.e. 125 sessions per GPU.
At --gpus 1 one shard is run (the largest slice of the headline config that fits one GPU
together with its filter workspace); at --gpus 8 the job is exactly configs[3] (weak
scaling: per-GPU work is fixed).  Data are synthetic draws from the model.

A "step" is one full sweep (the eight resample_* steps of keypoint_slds.ipynb cell 12).
`value`  : inputs resident in HBM, CUDA events, max over ranks.
`e2e`    : the same sweep driven from pinned HOST buffers through parallel.gibbs_sweep_host —
           every step copies Y, mask and all states to the device (copy stream, ordered by first
           use), runs the sweep, and reads the new states and parameters back (second copy
           stream, each state as soon as it is final); the step ends when the host copies do.
`roofline`: the dominant entry point (kpms_resample_latents: Kalman filter + sampler) timed
           alone with CUDA events on the launching stream.
`cpu_baseline` / `--impl reference`: the float64 NumPy+numba oracle (the reference's JAX
           path cannot run here: JAX/TFP are not installed) on the host cores, on a bounded
           sample of the same workload.
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
SHAPE = dict(k=16, dk=2, D=10, L=3, K=100)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sessions-per-gpu", type=int, default=125)
    ap.add_argument("--frames", type=int, default=36000)
    ap.add_argument("--chunk-len", default="auto", help="Kalman chunk length: auto, 0 (= one warp per session) or steps")
    ap.add_argument("--burn-in", type=int, default=256, help="Kalman chunk warm-up steps")
    ap.add_argument("--cpu-sessions", type=int, default=0, help="0 = one per host core (max 16)")
    ap.add_argument("--cpu-frames", type=int, default=6000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


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
def cpu_sweep_rate(n_sessions, frames, sweeps=2, warm=1):
    """Time the oracle's full sweep on the host cores (numba prange over sessions)."""
    from keypointmoseq_b200.synthetic import simulate, default_hypparams
    from oracle import kpms_oracle as O
    sh = SHAPE
    data, st, pr, Gamma = simulate(7, n_sessions, frames, k=sh["k"], dk=sh["dk"], D=sh["D"], L=sh["L"], K=sh["K"])
    d64 = {a: b.astype(np.float64) for a, b in data.items()}
    s64 = {a: (b.astype(np.float64) if a != "z" else b) for a, b in st.items()}
    trans, ar, obs, loc = default_hypparams(sh["D"], sh["L"], sh["K"])
    hyp = {}
    for h in (trans, ar, obs, loc):
        hyp.update(h)
    rng = np.random.default_rng(3)
    N, T, k, dk, D, L, K = n_sessions, frames, sh["k"], sh["dk"], sh["D"], sh["L"], sh["K"]
    n = D * L

    def draws():
        cnt_max = N * (T - L)
        return dict(
            ar_chi2=rng.chisquare(hyp["nu_0"] + 50.0, size=(K, D)), ar_normals=rng.standard_normal((K, D * (D - 1) // 2)),
            ar_G=rng.standard_normal((K, D, n + 1)), sigmasq_g=rng.gamma((hyp["nu_sigma"] + 3 * N * T) / 2, size=k),
            crp_u=rng.random(cnt_max), thin_u=rng.random((K, 16384)),
            betas_g=lambda c: rng.gamma(c), pi_g=lambda c: rng.gamma(c),
            z_u=rng.random((N, T - L)), x_w=rng.standard_normal((N, T - L, n)), x_w_last=rng.standard_normal((N, n)),
            h_u=rng.random((N, T, 8, 3)) * 0.998 + 0.001, v_w=rng.standard_normal((N, T - 1, dk)),
            v_w_last=rng.standard_normal((N, dk)), s_g=rng.gamma((hyp["nu_s"] + 3) / 2, size=(N, T, k)))
    times = []
    for i in range(warm + sweeps):
        dr = draws()
        t0 = time.perf_counter()
        s64, pr = O.sweep(d64, s64, pr, hyp, dr, Gamma)
        dt = time.perf_counter() - t0
        if i >= warm:
            times.append(dt)
    per = float(np.mean(times))
    return N * T / per, per


def cpu_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference(args, rank):
    if rank != 0:
        return
    cores = cpu_cores()
    try:
        import numba
        numba.set_num_threads(min(cores, numba.config.NUMBA_NUM_THREADS))
        threads = numba.get_num_threads()
    except Exception:
        threads = 1
    ns = args.cpu_sessions or min(cores, 16)
    rate, per = cpu_sweep_rate(ns, args.cpu_frames, sweeps=max(1, min(args.steps, 3)), warm=1)
    sample = "%d sessions x %d frames (k=%d,D=%d,L=%d,K=%d), oracle NumPy+numba float64" % (
        ns, args.cpu_frames, SHAPE["k"], SHAPE["D"], SHAPE["L"], SHAPE["K"])
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": per * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "cfg4-shard", "sessions_per_gpu": args.sessions_per_gpu, "frames": args.frames,
                       **SHAPE, "note": "reference JAX-CPU path unavailable on this box (no jax/tfp); "
                                        "oracle port timed on a bounded sample"},
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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
    from keypointmoseq_b200 import _lib, gibbs as G, parallel
    from keypointmoseq_b200.synthetic import simulate, default_hypparams

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        td.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    if args.chunk_len is not None:
        G.LATENTS_CHUNKING["chunk_len"] = args.chunk_len if args.chunk_len == "auto" else int(args.chunk_len)
    if args.burn_in is not None:
        G.LATENTS_CHUNKING["burn_in"] = args.burn_in

    sh = SHAPE
    N, T = args.sessions_per_gpu, args.frames
    data, st, pr, Gamma = simulate(1000 + rank, N, T, k=sh["k"], dk=sh["dk"], D=sh["D"], L=sh["L"], K=sh["K"])
    pr0 = simulate(999, 1, 8, k=sh["k"], dk=sh["dk"], D=sh["D"], L=sh["L"], K=sh["K"])[2]  # replicated params
    trans, ar, obs, loc = default_hypparams(sh["D"], sh["L"], sh["K"])
    hyp = {}
    for h in (trans, ar, obs, loc):
        hyp.update({a: (torch.as_tensor(b, device=dev) if isinstance(b, np.ndarray) else b) for a, b in h.items()})
    host = {a: torch.as_tensor(b).pin_memory() for a, b in data.items()}
    host_st = {a: torch.as_tensor(b.astype(np.int32) if a == "z" else b).pin_memory() for a, b in st.items()}
    d = {a: b.to(dev, non_blocking=True) for a, b in host.items()}
    s = {a: b.to(dev, non_blocking=True) for a, b in host_st.items()}
    p = {a: torch.as_tensor(b, device=dev) for a, b in pr0.items()}
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
        return parallel.gibbs_sweep((0, i), dd, ss, pp, hyp, Gamma=Gm, n_sessions_global=N * world)

    # ---- device-resident arm
    ss, pp = s, p
    for i in range(args.warmup):
        ss, pp = sweep(i, d, ss, pp)
    barrier()
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
    per_rank = None
    if world > 1:   # diagnostics: every rank's own time and chunk retries during the timed sweeps
        per_rank = [None] * world
        td.all_gather_object(per_rank, {"ms": round(my_ms, 2), "fallbacks": {a: b for a, b in G.FALLBACKS.items()
                                                                             if not a.endswith("_berr")},
                                        "hmm_burn_boost": G._HMM_ADAPT["boost"]})
    clocks = clk.summary()

    # ---- per-step breakdown and the roofline of the dominant entry point (rank-local)
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

    reps = max(2, min(args.steps, 5))
    kw = dict(Gamma=Gm)
    bd = {}
    bd["ar_stats"] = timed(lambda: G.ar_sufficient_stats(**d, **ss, **hyp), reps)
    bd["obs_variance_stats"] = timed(lambda: G.obs_variance_stats(**d, **ss, **pp, **kw), reps)
    bd["hdp_transitions"] = timed(lambda: G.resample_hdp_transitions(1, **d, **ss, **pp, **hyp), reps)
    bd["stateseqs"] = timed(lambda: G.resample_stateseqs(1, **d, **ss, **pp), reps)
    bd["ar_loglik"] = timed(lambda: G.ar_log_likelihood(**ss, **pp), reps)
    bd["latents"] = timed(lambda: G.resample_latents(1, **d, **ss, **pp, **kw), reps)
    bd["heading"] = timed(lambda: G.resample_heading(1, **d, **ss, **pp, **kw), reps)
    bd["location"] = timed(lambda: G.resample_location(1, **d, **ss, **pp, **hyp, **kw), reps)
    bd["scales"] = timed(lambda: G.resample_scales(1, **d, **ss, **pp, **hyp, **kw), reps)
    # latents with pre-drawn normals = the kernels alone (no Philox fill in the timed region)
    n = sh["D"] * sh["L"]
    w = torch.randn((N, T - sh["L"], n), device=dev)
    wl = torch.randn((N, n), device=dev)
    lat_ms = timed(lambda: G.resample_latents(1, **d, **ss, **pp, **kw, draws={"w": w, "w_last": wl}), reps)
    _, linfo = G.resample_latents(1, **d, **ss, **pp, **kw, draws={"w": w, "w_last": wl}, return_info=True)
    boundary_err = [float(v) for v in linfo["boundary_err"].cpu().tolist()]
    del w, wl
    pdim = sh["k"] * sh["dk"]
    bytes_per_frame = 4 * (pdim + sh["k"] + sh["dk"] + 3 + sh["D"]) + 4 * n + 2 * 8 * (n + n * (n + 1) // 2)
    achieved = N * T * bytes_per_frame / (lat_ms * 1e-3) / 1e9
    peak, peak_src = 6650.0, "fallback"
    try:
        mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak, peak_src = float(mp["hbm_gbs"]), "measured"
    except Exception:
        pass
    traffic = None
    try:   # measured DRAM bytes of the entry point's kernels (one ncu --set full capture of this workload)
        tj = json.load(open(os.path.join(ROOT, "profiles", "r01_latents_traffic.json")))
        if int(tj["frames"]) == N * T:
            traffic = float(tj["total"])
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "algorithmic_bytes_per_launch": N * T * bytes_per_frame, "kernel": "kpms_resample_latents (kalman_forward + kalman_backward)",
                "algorithmic_bytes_per_frame": bytes_per_frame, "ms_per_launch": lat_ms, "peak_source": peak_src}

    # ---- end-to-end arm: host buffers in, host buffers out, every step
    e2e = None
    if not args.no_e2e:
        out_host = {a: torch.empty_like(b).pin_memory() for a, b in host_st.items()}
        h2d = sum(b.numel() * b.element_size() for b in host.values()) + \
            sum(b.numel() * b.element_size() for b in host_st.values())
        d2h = sum(b.numel() * b.element_size() for b in out_host.values())
        par_host = None

        def e2e_step(i, pp_):
            nonlocal par_host
            # public host-side API: uploads ordered by first use, downloads as states become final
            _, p_new, _ = parallel.gibbs_sweep_host((0, i), host, host_st, pp_, hyp, Gamma=Gm,
                                                    n_sessions_global=N * world, out=out_host, device=dev)
            par_host = {a: b.cpu() for a, b in p_new.items()}   # small; synchronises the step
            return p_new
        pe = p
        for i in range(min(args.warmup, 2)):
            pe = e2e_step(200 + i, pe)
        barrier()
        ev0.record()
        for i in range(args.steps):
            pe = e2e_step(300 + i, pe)
        ev1.record()
        barrier()
        ems = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
        d2h += sum(b.numel() * b.element_size() for b in par_host.values())
        e2e = {"value": frames_total / (ems * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": ems}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = cpu_cores()
        try:
            import numba
            threads = numba.get_num_threads()
        except Exception:
            threads = 1
        ns = args.cpu_sessions or min(cores, 16)
        rate, per = cpu_sweep_rate(ns, args.cpu_frames, sweeps=2, warm=1)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "%d sessions x %d frames, same k/D/L/K, oracle NumPy+numba float64, %.2f s/sweep" % (
                   ns, args.cpu_frames, per)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64 (Kalman, statistics) / f32 (HMM messages, emission GEMM)",
                "data": "synthetic",
                "config": {"workload": "cfg4-shard", "sessions_per_gpu": N, "frames": T, **sh,
                           "total_sessions": N * world, "kalman_chunk_len": (G.auto_chunk_len(N, T, sh["L"], G.LATENTS_CHUNKING["burn_in"], dev)
                                                if G.LATENTS_CHUNKING["chunk_len"] == "auto" else G.LATENTS_CHUNKING["chunk_len"]),
                           "kalman_burn_in": G.LATENTS_CHUNKING["burn_in"],
                           "kalman_boundary_err": boundary_err,
                           "chunk_fallbacks": dict(G.FALLBACKS), "hmm_burn_boost": G._HMM_ADAPT["boost"],
                           "per_rank": per_rank,
                           "l2": "per-step working set (>= 20 GB of filter workspace + ll/alphan) far exceeds the 126 MB L2"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline,
                "cpu_baseline": cpu, "breakdown_ms": bd}
        print(json.dumps(line), flush=True)
    if world > 1:
        td.destroy_process_group()


if __name__ == "__main__":
    main()
