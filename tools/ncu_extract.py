"""Turn an `ncu --set full` report of one sweep into the summaries kept under profiles/.

    python tools/ncu_extract.py gpurun_out/r02_sweep_full.ncu-rep cfg4-shard 4500000

writes profiles/r02_sweep_kernels_ncu.json (per kernel: duration, DRAM bytes, pipe utilisation, stall mix,
registers, occupancy) and profiles/r02_latents_traffic.json (DRAM bytes of the three kernels of
kpms_resample_latents, which bench.py reports as roofline.traffic for the same config and frame count).
The report itself stays in gpurun_out/ (scratch)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, config, frames = sys.argv[1], sys.argv[2], int(sys.argv[3])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
head, units = rows[0], rows[1]
ix = {n: i for i, n in enumerate(head)}

KEEP = {
    "duration_ms": "gpu__time_duration.sum",
    "dram_read_GB": "dram__bytes_read.sum",
    "dram_write_GB": "dram__bytes_write.sum",
    "dram_pct_of_peak": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm_throughput_pct": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "tensor_pipe_cycles_active_pct_of_elapsed": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "tf32_tensor_ops_pct_of_peak_elapsed": "sm__ops_path_tensor_src_tf32_dst_fp32.sum.pct_of_peak_sustained_elapsed",
    "fp64_tensor_ops_pct_of_peak_elapsed": "sm__ops_path_tensor_src_fp64.sum.pct_of_peak_sustained_elapsed",
    "utcmma_instructions": "smsp__sass_inst_executed_op_utcmma.sum",
    "tmem_load_instructions": "smsp__sass_inst_executed_op_tmem_ldt.sum",
    "tensor_pipe_tc_pct_active": "sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active",
    "tmem_pipe_pct_active": "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
    "dmma_subpipe_pct_active": "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active",
    "fp64_pipe_pct_active": "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "lsu_pipe_pct_active": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "l1_hit_pct": "l1tex__t_sector_hit_rate.pct",
    "l2_hit_pct": "lts__t_sector_hit_rate.pct",
    "registers": "launch__registers_per_thread",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dyn_smem_per_block_KB": "launch__shared_mem_per_block_dynamic",
    "grid": "launch__grid_size",
    "block": "launch__block_size",
}
STALLS = [n for n in head if n.startswith("smsp__average_warp") and n.endswith("_per_issue_active.ratio") or
          n.startswith("smsp__average_warps_issue_stalled") and n.endswith("per_issue_active.ratio")]


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


out = {"source": "ncu --set full --clock-control none --import-source on, python tools/sweep_once.py 125 36000 2, "
                 "first launch of each kernel; " + os.path.basename(rep), "config": config, "frames": frames,
       "units": {}, "kernels": {}}
for r in rows[2:]:
    name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").strip()
    if name in out["kernels"]:
        continue
    rec = {}
    for key, metric in KEEP.items():
        if metric in ix:
            rec[key] = num(r[ix[metric]])
            out["units"][key] = units[ix[metric]]
    stalls = {}
    for n in STALLS:
        v = num(r[ix[n]])
        if v is not None and v >= 0.05:
            stalls[n.replace("smsp__average_warps_issue_stalled_", "").replace("smsp__average_warp_latency_issue_stalled_", "")
                   .replace("_per_issue_active.ratio", "")] = round(v, 2)
    rec["stall_cycles_per_issue"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:6])
    out["kernels"][name] = rec

os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
with open(os.path.join(ROOT, "profiles", "r02_sweep_kernels_ncu.json"), "w") as fh:
    json.dump(out, fh, indent=1)

GB = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
lat = {}
for name, rec in out["kernels"].items():
    if name.startswith(("kalman_obs", "kalman_forward_blk_kernel", "kalman_backward_kernel")):
        sr = GB[out["units"]["dram_read_GB"]]
        sw = GB[out["units"]["dram_write_GB"]]
        lat[name.split("<")[0]] = {"dram_read": rec["dram_read_GB"] * sr, "dram_write": rec["dram_write_GB"] * sw}
total = sum(v["dram_read"] + v["dram_write"] for v in lat.values())
with open(os.path.join(ROOT, "profiles", "r02_latents_traffic.json"), "w") as fh:
    json.dump({"source": out["source"], "unit": "bytes per kpms_resample_latents call", "kernels": lat,
               "frames": frames, "config": config, "total": total}, fh, indent=1)
print(json.dumps({n: {k: v for k, v in r.items() if k != "stall_cycles_per_issue"} for n, r in out["kernels"].items()}, indent=1))
print("latents DRAM bytes per call:", total)
