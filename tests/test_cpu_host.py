"""CPU-side tests: the C-ABI library loads and exports every symbol of include/genseg_b200.h, the drop-in modules
keep the reference's state_dict layout, the product path refuses to run without CUDA, and the data-parallel host
logic works over gloo with world_size 2."""
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "genseg_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(gsb_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_header_symbol():
    import __graft_entry__ as g
    g.build()
    from semantic_segmentation_b200 import _lib
    lib = _lib.load()
    syms = _header_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(lib, s), "missing export " + s
        assert s in _lib.SIGNATURES, "no ctypes signature for " + s
    assert lib.gsb_version() >= 100
    assert set(_lib.SIGNATURES) <= set(syms)


def test_state_dict_layout_matches_reference_spec():
    from oracle import genseg_oracle as O
    from semantic_segmentation_b200.models_pix2pix import networks
    from semantic_segmentation_b200.unet import UNet
    for nch, ncls, bil in ((1, 2, True), (1, 1, False), (3, 3, False)):
        net = UNet(nch, ncls, bil)
        spec = O.unet_state_spec(nch, ncls, bil)
        sd = net.state_dict()
        assert list(sd.keys()) == list(spec.keys())
        assert all(tuple(sd[k].shape) == tuple(spec[k]) for k in spec)
        assert set(net._program.param_names) == set(dict(net.named_parameters()).keys())
        assert (net.n_channels, net.n_classes, net.bilinear) == (nch, ncls, bil)
    G = networks.define_G(1, 1, 64, "unet_256", "batch", True, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic", 3, "batch", "normal", 0.02, [])
    assert list(G.state_dict().keys()) == list(O.generator_state_spec(1, 1, 64).keys())
    assert list(D.state_dict().keys()) == list(O.discriminator_state_spec(2, 64).keys())
    # flag-reachable variants: --netG unet_128 (networks.py:148-151), --netD n_layers --n_layers_D k (networks.py:199-200)
    G7 = networks.define_G(1, 3, 64, "unet_128", "batch", True, "normal", 0.02, [])
    spec7 = O.generator_state_spec(1, 3, 64, 7)
    assert list(G7.state_dict().keys()) == list(spec7.keys())
    assert all(tuple(G7.state_dict()[k].shape) == tuple(spec7[k]) for k in spec7)
    assert set(G7._program.param_names) - {"__upconv_arch__"} == set(dict(G7.named_parameters()).keys())
    assert [L for L in range(1, 7) if G7._program.up[L].dropout] == [4, 5]
    for nl in (1, 2, 4, 5):
        Dn = networks.define_D(4, 64, "n_layers", nl, "batch", "normal", 0.02, [])
        specn = O.discriminator_state_spec(4, 64, nl)
        assert list(Dn.state_dict().keys()) == list(specn.keys())
        assert all(tuple(Dn.state_dict()[k].shape) == tuple(specn[k]) for k in specn)
        assert set(Dn._program.param_names) == set(dict(Dn.named_parameters()).keys())
    assert networks.arch_parameters()[0] is networks.upconv_arch and tuple(networks.upconv_arch.shape) == (8, 3)
    assert networks.upconv_arch.requires_grad
    # init rule of networks.py:84-101: conv weights N(0, 0.02), BN weight N(1, 0.02)
    w = G.state_dict()["model.model.1.model.1.weight"]
    assert abs(w.std().item() - 0.02) < 2e-3
    with pytest.raises(NotImplementedError):
        networks.define_G(1, 1, 64, "resnet_9blocks")


def test_no_cpu_fallback():
    from semantic_segmentation_b200.models_pix2pix import networks
    from semantic_segmentation_b200.unet import UNet
    from semantic_segmentation_b200.util.dice_score import dice_loss, seg_loss
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        UNet(1, 2, True)(torch.zeros(1, 1, 32, 32))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        seg_loss(torch.zeros(2, 2, 8, 8), torch.zeros(2, 8, 8, dtype=torch.long))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        dice_loss(torch.zeros(2, 8, 8), torch.zeros(2, 8, 8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        networks.GANLoss("vanilla")(torch.zeros(1, 1, 30, 30), True)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "semantic_segmentation_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert "oracle" not in src, os.path.join(dp, f)


def test_shard_batch():
    from semantic_segmentation_b200 import parallel
    assert [parallel.shard_batch(64, r, 8) for r in range(8)] == [(8 * r, 8 * r + 8) for r in range(8)]
    with pytest.raises(ValueError):
        parallel.shard_batch(10, 0, 4)


WORKER = r'''
import os, sys, torch
sys.path.insert(0, %r)
from semantic_segmentation_b200 import engine, parallel
world = parallel.init_data_parallel(sync_bn=True, backend="gloo")
r = parallel.rank()
assert world == 2 and parallel.shard_batch(64) == (32 * r, 32 * r + 32)
sums = torch.tensor([1.0 + r, 10.0 * (r + 1)], dtype=torch.float64)
s2, cnt = engine._sync_stats(sums, 100.0)             # BatchNorm statistics: summed over ranks, count scaled
assert s2.tolist() == [3.0, 30.0] and cnt == 200.0
p = torch.nn.Parameter(torch.zeros(3)); p.grad = torch.full((3,), float(r + 1))
parallel.allreduce_gradients([p])                      # gradient sum (1/world folded into the optimiser)
assert p.grad.tolist() == [3.0, 3.0, 3.0]
m = torch.nn.Linear(2, 2)
with torch.no_grad():
    m.weight.fill_(float(r))
parallel.broadcast_parameters(m)
assert float(m.weight.sum()) == 0.0
# evaluation is "replicas only": ranks exchange (sum of per-batch Dice, number of batches)
assert abs(float(parallel.mean_over_ranks(torch.tensor(float(r + 1)), 2 + r)) - 3.0 / 5.0) < 1e-7
parallel.shutdown()
# --- gradient all-reduce inside backward (bucketed, first bucket asynchronous) -- host logic over gloo
world = parallel.init_data_parallel(sync_bn=True, backend="gloo", overlap_grad_allreduce=True)
assert engine._REDUCER["grad_fn"] is not None and engine._REDUCER["sm_reserve"] == 16
bb = engine.BackwardBatch(torch.device("cpu"), {"late": 6, "early": 4})
assert bb.reserve("early") == (0, 4) and bb.offsets["early"] == (0, 4)                 # offsets follow the order of use
assert bb.reserve("unknown") is None
bb.gflat[:4] = float(r + 1)
bb.reduce_async()                                      # bucket 1 = [0, 4): in flight, SM reserve raised
assert bb.flushed == 4 and len(bb.pending) == 1
bb.reserve("late"); bb.gflat[4:10] = 10.0 * (r + 1)
small = [torch.full((3,), float(r)), None, torch.full((2, 2), 2.0 * r)]
assert bb.reduce_finish(small) is True and not bb.pending
assert bb.gflat[:4].tolist() == [3.0] * 4 and bb.gflat[4:10].tolist() == [30.0] * 6
assert small[0].tolist() == [1.0] * 3 and small[2].flatten().tolist() == [2.0] * 4
from semantic_segmentation_b200 import ops
assert ops.set_sm_reserve(0) == 0                      # reduce_finish restored the full SM count
p1 = torch.nn.Parameter(torch.zeros(2)); p1.grad = torch.full((2,), float(r + 1)); p1._gsb_grad_reduced = True
p2 = torch.nn.Parameter(torch.zeros(2)); p2.grad = torch.full((2,), float(r + 1))
parallel.allreduce_gradients([p1, p2])                 # p1 was summed inside backward: skipped, flag cleared
assert p1.grad.tolist() == [r + 1.0] * 2 and p2.grad.tolist() == [3.0, 3.0] and p1._gsb_grad_reduced is False
parallel.shutdown()
assert engine._REDUCER["grad_fn"] is None
print("rank", r, "ok")
'''


def test_data_parallel_host_logic_gloo_world2(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("ok") == 2


def test_fused_optimisers_are_torch_optimizers():
    """what the reference hangs on its optimisers -- ReduceLROnPlateau (train_end2end_jsrt.py:73), state_dict
    round trips with torch.optim (checkpoints, Betty roll_back) -- works on the fused classes; step() has no CPU path."""
    from semantic_segmentation_b200.optim import Adam, RMSprop
    p = torch.nn.Parameter(torch.ones(4))
    p.grad = torch.ones(4)
    o = RMSprop([p], lr=1e-2, weight_decay=1e-8, momentum=0.9)
    assert isinstance(o, torch.optim.Optimizer) and o.param_groups[0]["momentum"] == 0.9
    sched = torch.optim.lr_scheduler.ReduceLROnPlateau(o, "max", patience=0)
    sched.step(1.0)
    sched.step(0.5)
    assert abs(o.param_groups[0]["lr"] - 1e-3) < 1e-12
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        o.step()
    assert len(o.state) == 0                                  # nothing was touched before the refusal
    q = torch.nn.Parameter(torch.ones(4))
    q.grad = torch.full((4,), 0.5)
    ref = torch.optim.Adam([q], lr=2e-5, betas=(0.5, 0.999))
    ref.step()
    ref.step()
    a = Adam([p], lr=1.0)
    a.load_state_dict(ref.state_dict())                       # an optimiser checkpoint of the reference loads
    assert int(a.state[p]["step"]) == 2 and a.param_groups[0]["betas"] == (0.5, 0.999) and a.param_groups[0]["lr"] == 2e-5
    assert torch.equal(a.state[p]["exp_avg"], ref.state[q]["exp_avg"]) and torch.equal(a.state[p]["exp_avg_sq"], ref.state[q]["exp_avg_sq"])
    back = torch.optim.Adam([q], lr=1.0)
    back.load_state_dict(a.state_dict())                      # and the other way round
    back.step()
    assert int(back.state[q]["step"]) == 3
    o.zero_grad()
    assert p.grad is None
    with pytest.raises(NotImplementedError):
        RMSprop([p], centered=True)
