"""Pinned host <-> device copy bandwidth on this box (context for the e2e number)."""
import torch, time
dev = "cuda"
for mb in (64, 576, 1134):
    h = torch.empty(mb * 1024 * 1024, dtype=torch.uint8).pin_memory()
    d = torch.empty_like(h, device=dev)
    for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            fn()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 3
        print(f"{name} {mb} MB: {ms:.2f} ms  {mb / 1024 / (ms * 1e-3):.1f} GiB/s", flush=True)
# both directions at once
h1 = torch.empty(512 * 1024 * 1024, dtype=torch.uint8).pin_memory(); d1 = torch.empty_like(h1, device=dev)
h2 = torch.empty(512 * 1024 * 1024, dtype=torch.uint8).pin_memory(); d2 = torch.empty_like(h2, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); t0 = time.perf_counter()
with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); print("duplex 512+512 MB: %.2f ms" % ((time.perf_counter() - t0) * 1e3))
