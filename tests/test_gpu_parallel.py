"""GPU tests of the data-parallel kernels: the fused BatchNorm statistic kernels against the separate
reduce / finalize / param-grad launches, the peer-memory all-reduce kernel at world 1 (same code path, own region
only), and -- when the box has two GPUs -- the 2-rank checks (peer all-reduce == NCCL bit for bit, N-rank == 1-rank)."""
import os
import subprocess
import sys

import pytest
import torch

from semantic_segmentation_b200 import ops
from semantic_segmentation_b200._lib import ACT_NONE, ACT_RELU

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("c,rows", [(64, 148), (8, 3), (512, 37), (1024, 148), (24, 1)])
def test_fused_stats_finalize_matches_separate_launches(c, rows):
    g = torch.Generator().manual_seed(c + rows)
    part = torch.randn(rows, 2 * c, generator=g).abs().cuda()
    part[:, c:] += 5.0                                            # keep the variance positive
    count = 4096.0
    gamma, beta = torch.randn(c, generator=g).cuda(), torch.randn(c, generator=g).cuda()

    def buffers():
        return torch.zeros(c).cuda(), torch.ones(c).cuda(), torch.zeros((), dtype=torch.int64).cuda()
    rm0, rv0, n0 = buffers()
    sums = ops.bn_reduce_partials(part, rows, c)
    ref = ops.bn_finalize(sums, count, gamma, beta, rm0, rv0, n0, 0.1, 1e-5, c, part.device)
    rm1, rv1, n1 = buffers()
    out = ops.bn_stats_finalize(part, rows, c, count, gamma, beta, rm1, rv1, n1, 0.1, 1e-5, None)
    for a, b in zip(out, ref):
        assert torch.equal(a, b)
    assert torch.equal(rm0, rm1) and torch.equal(rv0, rv1) and int(n1) == 1


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("shape", [(2, 24, 40, 64), (3, 7, 9, 128), (1, 16, 16, 8)])
def test_fused_bwd_stats_match_separate_launches(shape, dtype):
    g = torch.Generator().manual_seed(sum(shape))
    n, h, w, c = shape
    y = torch.randn(shape, generator=g).to(dtype).cuda()
    dA = torch.randn(shape, generator=g).to(dtype).cuda()
    scale, shift = (torch.rand(c, generator=g) + 0.5).cuda(), (0.3 * torch.randn(c, generator=g)).cuda()
    mean, invstd = (0.1 * torch.randn(c, generator=g)).cuda(), (torch.rand(c, generator=g) + 0.5).cuda()
    for act in (ACT_RELU, ACT_NONE):
        ref = ops.bn_bwd_reduce(y, dA, None, scale, shift, mean, invstd, act, ACT_NONE)
        rg, rb = ops.bn_param_grad(ref, c)
        sums, dg, db = ops.bn_bwd_stats(y, dA, None, scale, shift, mean, invstd, act, ACT_NONE)
        # bf16 + ReLU takes the dedicated kernel (sum dy*y, normalised after the loop): same sums up to fp32 rounding
        tol = 2e-5 if (dtype == torch.bfloat16 and act == ACT_RELU) else 1e-12
        den = ref.abs().max().clamp_min(1e-30)
        assert float((sums - ref).abs().max() / den) <= tol
        assert torch.allclose(dg, rg, rtol=1e-4, atol=1e-4 * float(rg.abs().max())) and torch.allclose(db, rb, rtol=1e-5, atol=1e-5 * float(rb.abs().max()))


def test_peer_allreduce_world1_and_graph_replay():
    region, handle = ops.peer_create()
    try:
        assert len(handle) == 64 and region != 0
        for n in (1, 7, 1024, 4096):
            v = torch.randn(n, dtype=torch.float64).cuda()
            a = v.clone()
            ops.peer_allreduce_f64(a, [region], 0, 1)
            assert torch.equal(a, v)                               # one rank: the sum is the vector itself
        x = torch.arange(300, dtype=torch.float64).cuda()
        y = torch.empty_like(x)
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            y.copy_(x); ops.peer_allreduce_f64(y, [region], 0, 1)
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                for _ in range(6):                                 # more calls than slots: sequence numbers wrap the ring
                    y.copy_(x); ops.peer_allreduce_f64(y, [region], 0, 1)
            for _ in range(3):
                gr.replay()
        torch.cuda.synchronize()
        assert torch.equal(y, x)
        with pytest.raises(RuntimeError):
            ops.peer_allreduce_f64(torch.zeros(5000, dtype=torch.float64).cuda(), [region], 0, 1)
    finally:
        ops.peer_destroy(region)


def test_sm_reserve_shrinks_persistent_grids():
    import math
    from semantic_segmentation_b200._lib import MODE_CONV, view_nhwc
    x = torch.randn(8, 128, 128, 64).bfloat16().cuda()       # 512 tiles: more than SMs
    wt = (torch.randn(64, 64, 3, 3) / math.sqrt(576)).cuda()
    wp = ops.pack_weight(wt, True, torch.bfloat16)
    y0, y1 = torch.empty_like(x), torch.empty_like(x)
    _, rows0 = ops.conv_gather(view_nhwc(x), wp, None, view_nhwc(y0), MODE_CONV, 3, 1, 1, use_tc=1, want_stats=True)
    assert ops.set_sm_reserve(16) == 0
    try:
        _, rows1 = ops.conv_gather(view_nhwc(x), wp, None, view_nhwc(y1), MODE_CONV, 3, 1, 1, use_tc=1, want_stats=True)
    finally:
        assert ops.set_sm_reserve(0) == 16
    assert rows1 <= rows0 - 16 + 1 and torch.equal(y0, y1)       # one statistics row per CTA: fewer CTAs, same result


def _torchrun(script, nproc, port, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % nproc,
                           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", script)],
                          capture_output=True, text=True, timeout=600, env=e)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one box")
def test_two_ranks_peer_allreduce_and_parity():
    r = _torchrun("peer_check.py", 2, 29561)
    assert r.returncode == 0 and "bitwise equal across ranks" in r.stdout, r.stdout + r.stderr
    # Every kernel is order-independent (no atomics), so a run is bit-reproducible; the 2-rank and the 1-rank run still group
    # their fp32 partial sums differently (tile counts and split factors follow the per-rank batch), which is ordinary
    # rounding noise: <= 2e-6 on the fp32 loss, <= 1e-3 on the bf16 loss.  After ONE RMSprop step (grad_scale = 1: seg_loss is a
    # global-batch mean) the 2-rank parameters equal the 1-rank parameters in the fp32 mode: a step moves a weight by
    # lr * 10 = 1e-3, a wrong 1 / world would show as ~5e-4; rounding-level gradient differences move it by < 2e-5
    # (measured 6.9e-7; gradients agree to 6.4e-8).
    r = _torchrun("dp_check.py", 2, 29562, {"DP_OVERLAP": "1"})
    print([l for l in r.stdout.splitlines() if l.startswith("world 2") or l.startswith("dropout")])
    assert r.returncode == 0, r.stdout + r.stderr
    lines = [l for l in r.stdout.splitlines() if l.startswith("world 2")]
    assert len(lines) == 2
    for l in lines:                                            # loss of the 2-rank run == 1-rank run
        loss, ref = float(l.split("loss")[1].split()[0]), float(l.split("ref")[1].split()[0])
        assert abs(loss - ref) < (2e-6 if "fp32" in l else 1e-3), l
        step = float(l.split("RMSprop step")[1].split()[0])
        if "fp32" in l:       # bf16: two bf16 runs with different tile / split shapes differ by rounding, single gradient elements
            assert step < 2e-5, l   # change sign and RMSprop's first step (lr * g / sqrt(0.01 g^2) = 10 lr) moves them by 2e-3
    drop = [l for l in r.stdout.splitlines() if l.startswith("dropout invariance")]
    assert len(drop) == 1 and all(float(v) < 1e-4 for v in drop[0].split(":")[-1].split(",")), drop
