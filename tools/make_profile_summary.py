"""profiles/r02_summary.md from the committed bench lines and ncu extracts (no GPU needed):
    python tools/make_profile_summary.py"""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")


def line(name):
    path = os.path.join(P, name)
    if not os.path.exists(path):
        return None
    return json.loads(open(path).read().strip().splitlines()[-1])


out = ["# Round 2 — measured on B200 (files in this directory; regenerate with tools/make_profile_summary.py)", ""]
out += ["## bench.py lines, one B200 (`r02_bench_<config>.json`)", "",
        "| config | sweep | sessions x frames | K | ms / sweep | frames·sweeps/s (device) | e2e (host buffers) | CPU oracle, 16 threads | e2e / CPU | chunk fallbacks (timed) |",
        "|---|---|---|---|---|---|---|---|---|---|"]
for c in ("cfg1", "cfg2", "cfg3", "cfg4-shard", "cfg4-shard_cold", "cfg5-K20", "cfg5-K100", "cfg5-K500"):
    d = line("r02_bench_%s.json" % c)
    if d is None:
        continue
    cf = d["config"]
    e2e = (d.get("e2e") or {}).get("value")
    cpu = (d.get("cpu_baseline") or {}).get("value")
    out.append("| %s | %s | %d x %d | %d | %.2f | %.3g | %s | %s | %s | %s |" % (
        c, cf.get("sweep", cf.get("model")), cf["sessions_per_gpu"], cf["frames"], cf["K"], d["ms_per_step"], d["value"],
        "%.3g" % e2e if e2e else "-", "%.3g" % cpu if cpu else "-", "%.0fx" % (e2e / cpu) if e2e and cpu else "-",
        cf.get("chunk_fallbacks_timed")))
d4 = line("r02_bench_cfg4-shard.json")
if d4:
    out += ["", "cfg4-shard is the driver's default (`python bench.py`): BASELINE configs[3] sharded over 8 GPUs, 125 sessions per GPU.",
            "`cfg4-shard_cold` = `--cold-start` (prior-drawn parameters, random z).", "",
            "### cfg4-shard, phases of one sweep (ms; `breakdown_ms`, each entry point timed alone)", ""]
    out += ["| " + " | ".join(k for k in d4["breakdown_ms"] if k != "gemm_peaks_measured") + " |",
            "|" + "---|" * (len(d4["breakdown_ms"]) - 1),
            "| " + " | ".join("%.2f" % v for k, v in d4["breakdown_ms"].items() if k != "gemm_peaks_measured") + " |", ""]
    out += ["`roofline` (dominant entry point): %s" % json.dumps(d4["roofline"]), "",
            "### per-kernel roofline (`roofline_per_kernel`)", "",
            "| entry point | bound | ms | achieved | peak | frac | note |", "|---|---|---|---|---|---|---|"]
    for r in d4.get("roofline_per_kernel", []):
        out.append("| %s | %s | %.2f | %.4g %s | %.4g | %.3f | %s |" % (r["kernel"], r["bound"], r["ms"], r["achieved"], r["unit"],
                                                                       r["peak"], r["frac"], r.get("note", "")))
    out += ["", "clocks during the timed region: %s; gpu_launches per timed run: %s" % (json.dumps(d4.get("clocks")), d4.get("gpu_launches"))]
ref = line("r02_bench_reference_arm.json")
if ref:
    out += ["", "## reference arm (`bench.py --impl reference`, `r02_bench_reference_arm.json`)", "",
            "%.4g frames·sweeps/s on %s — %s" % (ref["value"], ref["cpu_baseline"]["sample"], ref["config"].get("note", ""))]
scale = {n: line("r02_bench_cfg4_%dgpu.json" % n) for n in (2, 4, 8)}
if d4 and any(scale.values()):
    out += ["", "## 1 / 2 / 4 / 8 GPUs, weak scaling (125 sessions x 36k frames per GPU; `r02_bench_cfg4_<N>gpu.json`)", "",
            "| GPUs | ms / sweep | device-resident frames·sweeps/s | efficiency | e2e (host buffers every sweep) | efficiency | e2e, Y and mask resident | session groups |",
            "|---|---|---|---|---|---|---|---|",
            "| 1 | %.2f | %.4g | 1 | %.4g | 1 | %.4g | %s |" % (d4["ms_per_step"], d4["value"], d4["e2e"]["value"], d4["e2e_resident"]["value"],
                                                              d4["config"].get("e2e_session_groups", 1))]
    for n, dn in scale.items():
        if dn is None:
            continue
        out.append("| %d | %.2f | %.4g | %.3f | %.4g | %.3f | %.4g | %s |" % (
            n, dn["ms_per_step"], dn["value"], dn["value"] / (n * d4["value"]), dn["e2e"]["value"],
            dn["e2e"]["value"] / (n * d4["e2e"]["value"]), dn["e2e_resident"]["value"], dn["config"].get("e2e_session_groups", 1)))
    out += ["", "(`r02_e2e_groups.txt` has the comparison of 1 / 2 / 3 / 4 / 6 session groups in the host-driven sweep.)"]
pc = os.path.join(P, "r02_pcie_8ranks.json")
if os.path.exists(pc):
    pj = json.load(open(pc))
    out += ["", "What the host gives 8 ranks copying at once (`r02_pcie_8ranks.json`, pinned buffers): host-to-device %s GiB/s per rank alone, "
            "%s concurrent (sum %.0f); device-to-host %s alone, %s concurrent (sum %.0f)." % (
                pj["h2d"]["alone_GiBs"][0], pj["h2d"]["concurrent_GiBs"], pj["h2d"]["concurrent_sum_GiBs"],
                pj["d2h"]["alone_GiBs"][1], pj["d2h"]["concurrent_GiBs"], pj["d2h"]["concurrent_sum_GiBs"]),
            "A sweep moves 1.13 GB up and 0.54 GB down per rank, i.e. >= 48 ms + 44 ms of copy time on the slower ranks when all eight copy: "
            "end-to-end scaling is bounded by the host, not by NVLink or the kernels."]
kn = os.path.join(P, "r02_sweep_kernels_ncu.json")
if os.path.exists(kn):
    kj = json.load(open(kn))
    out += ["", "## ncu --set full, one sweep on the bench shard (`r02_sweep_kernels_ncu.json`; launch list `r02_ncu_launches.csv`)", "",
            "| kernel | ms | DRAM GB read+written | DRAM % of peak | tensor pipe active % | TF32 ops % of (sparse) peak | FP64 tensor ops % | FP64 pipe % | LSU % | issue % | regs | top stalls (cycles per issue) |",
            "|---|---|---|---|---|---|---|---|---|---|---|---|"]
    for n, r in kj["kernels"].items():
        g = lambda k: r.get(k) if r.get(k) is not None else 0.0
        out.append("| `%s` | %.2f | %.2f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %d | %s |" % (
            n, g("duration_ms"), g("dram_read_GB") + g("dram_write_GB"), g("dram_pct_of_peak"),
            g("tensor_pipe_cycles_active_pct_of_elapsed"), g("tf32_tensor_ops_pct_of_peak_elapsed"),
            g("fp64_tensor_ops_pct_of_peak_elapsed"), g("fp64_pipe_pct_active"), g("lsu_pipe_pct_active"),
            g("issue_active_pct"), int(g("registers")),
            ", ".join("%s %.1f" % kv for kv in list(r["stall_cycles_per_issue"].items())[:3])))
    out += ["", "ncu's TF32 peak (8192 ops/cycle/SM) is the 2:4-sparse figure: 24.6 % of it is 49 % of the dense TF32 peak, the same as the tensor "
            "pipe's active share. Detail pages: `r02_kalman_blk_ncu.txt`, `r02_emission_ncu.txt`, `r02_hmm_forward_ncu.txt`; SASS mnemonics per kernel "
            "(UTCHMMA / LDTM / UTCBAR / DMMA): `r02_sass_evidence.txt`; Kalman kernel variants: `r02_kalman_variants.txt`; FP64 micro-benchmark: "
            "`r02_fp64_probe.txt`; sharded = unsharded check on 2 GPUs: `r02_world_invariance_2gpu.json`."]
open(os.path.join(P, "r02_summary.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
