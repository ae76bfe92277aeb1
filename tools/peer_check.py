"""torchrun --nproc-per-node N tools/peer_check.py : NVLink peer-memory all-reduce (csrc/peer.cu) against NCCL --
values, bit-identity across ranks, CUDA-graph replay, latency per call."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from semantic_segmentation_b200 import ops, parallel

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
world = parallel.init_data_parallel(sync_bn=True)
rank = parallel.rank()
peer = parallel._GROUP["peer"]
assert peer is not None, "peer memory was not set up"
g = torch.Generator().manual_seed(100 + rank)
worst = 0.0
for it in range(40):
    for n in (1, 4, 128, 1000, 2048, 4096):
        v = torch.randn(n, generator=g, dtype=torch.float64).cuda()
        a, b = v.clone(), v.clone()
        parallel._allreduce_sum(a)                       # peer path
        dist.all_reduce(b)                               # NCCL
        worst = max(worst, float((a - b).abs().max() / b.abs().max().clamp_min(1e-300)))
        gathered = [torch.empty_like(a) for _ in range(world)]
        dist.all_gather(gathered, a)
        assert all(torch.equal(gathered[0], t) for t in gathered), "ranks disagree bitwise"
# CUDA-graph replay: the sequence counter lives on the device
x = torch.full((512,), float(rank + 1), dtype=torch.float64, device="cuda")
y = torch.empty_like(x)
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for _ in range(3):
        y.copy_(x); parallel._allreduce_sum(y)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=s):
        for _ in range(10):
            y.copy_(x); parallel._allreduce_sum(y)
    for _ in range(5):
        gr.replay()
torch.cuda.synchronize()
assert torch.equal(y, torch.full_like(y, world * (world + 1) / 2)), y[:4]


def bench(fn, n=200):
    t = torch.zeros(2048, dtype=torch.float64, device="cuda")
    for _ in range(20):
        fn(t)
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn(t)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / n


def bench_graph(fn, n=200):
    t = torch.zeros(2048, dtype=torch.float64, device="cuda")
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(5):
            fn(t)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for _ in range(n):
                fn(t)
        gr.replay()
        dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(5):
            gr.replay()
        b.record(st); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (5 * n)


t_peer = bench(parallel._allreduce_sum)
g_peer = bench_graph(parallel._allreduce_sum)
g_nccl = bench_graph(lambda t: dist.all_reduce(t))
t_nccl = bench(lambda t: dist.all_reduce(t))
if rank == 0:
    print("peer_check world %d: worst rel diff vs NCCL %.2e, bitwise equal across ranks, graph replay ok; "
          "latency per call (2048 fp64): eager peer %.1f us, NCCL %.1f us; graph replay peer %.2f us, NCCL %.2f us"
          % (world, worst, t_peer, t_nccl, g_peer, g_nccl), flush=True)
dist.barrier()
dist.destroy_process_group()
