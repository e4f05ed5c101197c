"""Pinned host <-> device copy bandwidth with ALL ranks copying at the same time (torchrun, one rank
per GPU): the context for the end-to-end scaling numbers, where every rank moves its shard through
the same host every sweep.  Rank 0 prints one JSON line: per-rank and aggregate GiB/s, alone and
concurrent, plus the box topology (nvidia-smi topo -m, NUMA / CPU affinity columns)."""
import json
import os
import subprocess
import sys

import torch
import torch.distributed as td

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    td.init_process_group("nccl", device_id=dev)
MB = 576
h = torch.empty(MB * 1024 * 1024, dtype=torch.uint8).pin_memory()
d = torch.empty_like(h, device=dev)


def bw(fn, reps=4):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return MB / 1024 / (a.elapsed_time(b) / reps * 1e-3)


def gather(v):
    if world == 1:
        return [v]
    out = [None] * world
    td.all_gather_object(out, v)
    return out


res = {}
for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
    alone = []
    for r in range(world):            # one rank at a time
        if world > 1:
            td.barrier()
        v = bw(fn) if r == rank else 0.0
        alone.append(v)
    alone = [max(x) for x in zip(*gather(alone))]
    if world > 1:
        td.barrier()
    together = gather(bw(fn, reps=8))  # all ranks at once
    res[name] = {"alone_GiBs": [round(x, 1) for x in alone], "concurrent_GiBs": [round(x, 1) for x in together],
                 "concurrent_sum_GiBs": round(sum(together), 1)}
if rank == 0:
    try:
        res["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
    except Exception as e:
        res["topo"] = str(e)
    res["world"], res["cpus"] = world, len(os.sched_getaffinity(0))
    print(json.dumps(res), flush=True)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(json.dumps(res, indent=1) + "\n")
if world > 1:
    td.destroy_process_group()
