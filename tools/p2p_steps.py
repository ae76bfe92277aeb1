"""Per-step wall / GPU time of the pix2pix step (diagnostic)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
import bench_extra as B

times = []
orig = B.timed
def timed(fn, steps, warm=3):
    for i in range(14):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        fn()
        t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
        a0 = torch.cuda.memory_allocated() / 2**30
        import gc
        found = gc.collect() if os.environ.get("P2P_GC", "0") == "1" else -1
        times.append(((t1 - t0) * 1e3, (t2 - t0) * 1e3, a0, torch.cuda.memory_allocated() / 2**30, found))
    return 1.0
B.timed = timed
B.pix2pix(8, 1)
for i, t in enumerate(times):
    print("step %2d: host %.1f ms, total %.1f ms, alloc %.2f GiB -> after gc %.2f GiB (unreachable %d)" % (i, *t))
