"""Host -> device bandwidth on this box: a flat pinned copy (torch, 1-D) against the library's dense ingest
(twtml_batch_from_dense_async / twtml_batch_from_dense from page-locked rows).  Prints GB/s.  Round 1: a pitched 2-D copy
of 8 KB rows gave 45 GB/s against 55.6 GB/s flat; flat chunks + k_repitch give 53 GB/s."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import twtml_b200 as tw

n, d = 1 << 18, 1004
Xp = torch.empty((n, d), dtype=torch.float64).pin_memory()
Xp.normal_()
yp = torch.zeros(n, dtype=torch.float64).pin_memory()
dev = torch.empty((n, d), dtype=torch.float64, device="cuda")
nbytes = n * d * 8
for rep in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    dev.copy_(Xp, non_blocking=True); torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("flat 1-D pinned copy      %.1f GB/s" % (nbytes / dt / 1e9))
ctx = tw.Context()
Xn, yn = Xp.numpy(), yp.numpy()
for rep in range(3):
    t0 = time.perf_counter()
    b = ctx.batch_from_dense_async(Xn, yn); b.ready(wait=True)
    dt = time.perf_counter() - t0
    print("asynchronous ingest       %.1f GB/s" % (nbytes / dt / 1e9))
    b.free()
for rep in range(2):
    t0 = time.perf_counter()
    b = ctx.batch_from_dense(Xn, yn)
    dt = time.perf_counter() - t0
    print("blocking ingest           %.1f GB/s" % (nbytes / dt / 1e9))
    b.free()
