"""Betty-compatible bilevel driver (semantic_segmentation_b200/betty) -- host logic, CPU only.

betty-ml is absent and the reference pins none of its numerics (PARITY UNPINNED): the driver is checked against
closed-form bilevel problems and against the way running_files/train_end2end_jsrt.py:286-342 constructs and runs it."""
import os
import subprocess
import sys

import pytest
import torch
from torch import nn

from semantic_segmentation_b200.betty import hypergradient
from semantic_segmentation_b200.betty.configs import Config, EngineConfig
from semantic_segmentation_b200.betty.engine import Engine
from semantic_segmentation_b200.betty.problems import ImplicitProblem

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
torch.set_default_dtype(torch.float32)


class Vec(nn.Module):
    def __init__(self, init):
        super().__init__()
        self.v = nn.Parameter(torch.as_tensor(init, dtype=torch.float64).clone())

    def forward(self):
        return self.v


def quadratic_problems(unroll=1, lr_w=0.25, lr_lam=0.1, c=0.3, seed=0):
    """Lower: L_low(w, lam) = 1/2 |w - A lam|^2  (w* = A lam, dw*/dlam = A, Hessian in w = I).
    Upper: L_up(w, lam) = 1/2 |w - t|^2 + c/2 |lam|^2  ->  dL_up/dlam = c lam + A^T (w - t)  exactly."""
    g = torch.Generator().manual_seed(seed)
    A = torch.randn(5, 3, generator=g, dtype=torch.float64)
    t = torch.randn(5, generator=g, dtype=torch.float64)
    wm, lm = Vec(torch.randn(5, generator=g, dtype=torch.float64)), Vec(torch.randn(3, generator=g, dtype=torch.float64))

    class Lower(ImplicitProblem):
        def training_step(self, batch):
            return 0.5 * ((self.module() - A @ self.upper.module()) ** 2).sum()

    class Upper(ImplicitProblem):
        def training_step(self, batch):
            return 0.5 * ((self.lower.module() - t) ** 2).sum() + 0.5 * c * (self.module() ** 2).sum()

    lower = Lower(name="lower", module=wm, optimizer=torch.optim.SGD(wm.parameters(), lr=lr_w),
                  config=Config(type="darts", unroll_steps=unroll))
    upper = Upper(name="upper", module=lm, optimizer=torch.optim.SGD(lm.parameters(), lr=lr_lam), config=Config(retain_graph=True))
    return lower, upper, A, t, c, lr_w, lr_lam


def make_engine(lower, upper, **engine_kw):
    deps = {"l2u": {lower: [upper]}, "u2l": {upper: [lower]}}
    return Engine(config=EngineConfig(**engine_kw), problems=[lower, upper], dependencies=deps)


def test_darts_hypergradient_matches_closed_form():
    lower, upper, A, t, c, lr_w, lr_lam = quadratic_problems()
    eng = make_engine(lower, upper, train_iters=1, valid_step=0)
    assert upper.paths == [[upper, lower, upper]] and eng.leaves == [lower] and lower.parents == [upper]
    w0, lam0 = lower.module.v.detach().clone(), upper.module.v.detach().clone()
    eng.run()
    w1 = w0 - lr_w * (w0 - A @ lam0)                       # the lower step
    hyper = c * lam0 + A.T @ (w1 - t)                      # direct + implicit (exact for this quadratic)
    assert torch.allclose(lower.module.v.detach(), w1, atol=1e-12)
    assert torch.allclose(upper.module.v.detach(), lam0 - lr_lam * hyper, atol=1e-8)
    assert lower.module.v.grad is None and upper.module.v.grad is None      # zero_grad after every step


def test_darts_function_alone_and_zero_direction():
    lower, upper, A, t, c, _, _ = quadratic_problems(seed=3)
    make_engine(lower, upper, train_iters=1, valid_step=0)
    lower.cur_batch = None
    v = torch.randn(5, dtype=torch.float64, generator=torch.Generator().manual_seed(9))
    w_before = lower.module.v.detach().clone()
    out = hypergradient.darts([v], lower, upper)
    assert torch.allclose(out[0], A.T @ v, atol=1e-8)
    assert torch.equal(lower.module.v.detach(), w_before) or torch.allclose(lower.module.v.detach(), w_before, atol=1e-15)
    assert hypergradient.darts([None], lower, upper) == [None]                       # upper loss does not reach this level
    assert hypergradient.darts([torch.zeros(5, dtype=torch.float64)], lower, upper) == [None]


def test_first_order_drops_the_implicit_term():
    lower, upper, A, t, c, lr_w, lr_lam = quadratic_problems()
    upper.config.first_order = True
    eng = make_engine(lower, upper, train_iters=1, valid_step=0)
    lam0 = upper.module.v.detach().clone()
    eng.run()
    assert torch.allclose(upper.module.v.detach(), lam0 - lr_lam * c * lam0, atol=1e-12)


def test_unroll_steps_and_validation_schedule():
    lower, upper, *_ = quadratic_problems(unroll=3)
    calls = []

    class E(Engine):
        def validation(self):
            calls.append((self.global_step, self.lower.count, self.upper.count, self.lower._training))
            return {"score": 1.0}

    deps = {"l2u": {lower: [upper]}, "u2l": {upper: [lower]}}
    eng = E(config=EngineConfig(train_iters=12, valid_step=6), problems=[lower, upper], dependencies=deps)
    eng.run()
    assert lower.count == 12 and upper.count == 4            # one upper step per 3 lower steps
    assert calls == [(6, 6, 2, False), (12, 12, 4, False)] and lower._training is True
    assert eng.global_step == 12 and not Engine.is_implemented(Engine(problems=[]), "validation")


def test_roll_back_returns_to_the_unroll_start_and_takes_one_step():
    unroll = 2
    lower, upper, A, t, c, lr_w, lr_lam = quadratic_problems(unroll=unroll, seed=5)
    eng = make_engine(lower, upper, train_iters=unroll, valid_step=0, roll_back=True)
    w0, lam0 = lower.module.v.detach().clone(), upper.module.v.detach().clone()
    eng.run()
    w = w0.clone()
    for _ in range(unroll):                                   # look-ahead steps
        w = w - lr_w * (w - A @ lam0)
    lam1 = lam0 - lr_lam * (c * lam0 + A.T @ (w - t))         # upper step at the look-ahead point
    w_expected = w0 - lr_w * (w0 - A @ lam1)                  # snapshot + ONE step under the new lam
    assert torch.allclose(upper.module.v.detach(), lam1, atol=1e-8)
    assert torch.allclose(lower.module.v.detach(), w_expected, atol=1e-8)


def test_roll_back_restores_buffers_and_optimizer_state():
    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.bn = nn.BatchNorm1d(2)
            self.lin = nn.Linear(2, 1)

        def forward(self, x):
            return self.lin(self.bn(x))

    torch.manual_seed(0)
    m = M().double()
    opt = torch.optim.RMSprop(m.parameters(), lr=1e-2, momentum=0.9)
    data = [torch.randn(4, 2, dtype=torch.float64) for _ in range(8)]

    class P(ImplicitProblem):
        def training_step(self, batch):
            return (self.module(batch) ** 2).mean()

    p = P(name="p", module=m, optimizer=opt, train_data_loader=data, config=Config())
    p.initialize()
    p._count = 1
    p.one_step_descent()
    p.cache_states()
    snap = {k: v.clone() for k, v in m.state_dict().items()}
    mom = opt.state_dict()["state"][0]["momentum_buffer"].clone()
    p.one_step_descent()
    p.one_step_descent()
    assert int(m.bn.num_batches_tracked) == 3 and not torch.equal(m.lin.weight.detach(), snap["lin.weight"])
    ids = [id(q) for q in m.parameters()]
    p.recover_states()
    for k, v in m.state_dict().items():
        assert torch.equal(v, snap[k]), k
    assert torch.equal(opt.state_dict()["state"][0]["momentum_buffer"], mom)
    assert ids == [id(q) for q in m.parameters()]              # nn.Parameter objects survive (the optimiser holds them)


def test_reference_script_graph():
    """train_end2end_jsrt.py:295-339: four problems, ``l2u = {netG: [unet], netG: [arch]}`` (repeated key -> {netG: [arch]}),
    ``u2l = {arch: [netG]}``, inner Config(type='darts', unroll_steps=k), outer Config(retain_graph=True), roll_back=True."""
    torch.manual_seed(1)
    G, D, U = nn.Linear(4, 4), nn.Linear(8, 1), nn.Linear(4, 1)
    alpha = nn.Parameter(torch.zeros(3))
    loader = [{"image": torch.randn(2, 4), "mask": torch.rand(2, 1)} for _ in range(3)]     # 3 batches: exercises the epoch wrap
    seen = {"D_via_name": 0, "val": []}

    class Generator(ImplicitProblem):
        def training_step(self, batch):
            fake = self.module(batch["mask"].expand(-1, 4)) * torch.softmax(alpha, 0).sum()
            seen["D_via_name"] += 1
            return self.netD(torch.cat([batch["mask"].expand(-1, 4), fake], 1)).mean() + (fake - batch["image"]).abs().mean()

    class Discriminator(ImplicitProblem):
        def training_step(self, batch):
            fake = self.netG(batch["mask"].expand(-1, 4))
            return self.module(torch.cat([batch["image"], fake.detach()], 1)).pow(2).mean()

    class Unet(ImplicitProblem):
        def training_step(self, batch):
            return (self.module(batch["image"]) - batch["mask"]).pow(2).mean()

    class Arch(ImplicitProblem):
        def training_step(self, batch):
            return (self.unet(batch["image"]) - batch["mask"]).pow(2).mean()

    class SSEngine(Engine):
        @torch.no_grad()
        def validation(self):
            seen["val"].append((self.global_step, type(self.unet.module).__name__))

    inner, outer = Config(type="darts", unroll_steps=2), Config(retain_graph=True)
    netG = Generator(name="netG", module=G, optimizer=torch.optim.Adam(G.parameters(), lr=1e-3, betas=(0.5, 0.999)),
                     train_data_loader=loader, config=inner, device=torch.device("cpu"))
    netD = Discriminator(name="netD", module=D, optimizer=torch.optim.Adam(D.parameters(), lr=1e-3), train_data_loader=loader,
                         config=inner, device=torch.device("cpu"))
    unet = Unet(name="unet", module=U, optimizer=torch.optim.RMSprop(U.parameters(), lr=1e-3, momentum=0.9),
                train_data_loader=loader, config=inner, device=torch.device("cpu"))
    arch = Arch(name="arch", module=U, optimizer=torch.optim.Adam([alpha], lr=1e-2, betas=(0.5, 0.999), weight_decay=1e-5),
                train_data_loader=loader, config=outer, device=torch.device("cpu"))
    l2u = {netG: [unet], netG: [arch]}     # noqa: F601 -- verbatim from the reference script
    u2l = {arch: [netG]}
    eng = SSEngine(config=EngineConfig(valid_step=4, train_iters=8, roll_back=True), problems=[netG, netD, unet, arch],
                   dependencies={"l2u": l2u, "u2l": u2l})
    assert eng.leaves == [netG, netD, unet] and netG.parents == [arch] and arch.children == [netG]
    assert arch.paths == [[arch, netG, arch]] and unet.parents == [] and eng.netD is netD
    g0 = [p.detach().clone() for p in G.parameters()]
    eng.run()
    assert (netG.count, netD.count, unet.count, arch.count) == (8, 8, 8, 4)
    assert seen["val"] == [(4, "Linear"), (8, "Linear")]
    assert netG.epoch_counter >= 2 and any(not torch.equal(a, b.detach()) for a, b in zip(g0, G.parameters()))
    # arch's loss does not reach netG's parameters: no implicit term, and alpha (not a parameter of arch.module) keeps its value
    assert torch.equal(alpha.detach(), torch.zeros(3))
    assert all(p.grad is None for p in U.parameters())


def test_three_level_paths_and_errors():
    def prob(name):
        m = Vec([0.0])
        return ImplicitProblem(name=name, module=m, optimizer=torch.optim.SGD(m.parameters(), lr=0.1), config=Config())

    a, b, c = prob("a"), prob("b"), prob("c")
    eng = Engine(problems=[a, b, c], dependencies={"l2u": {a: [b, c], b: [c]}, "u2l": {c: [a]}})
    assert sorted([[p.name for p in path] for path in c.paths]) == [["c", "a", "c"], ["c", "b", "a", "c"]]
    assert eng.leaves == [a] and eng.check_leaf(a) and not eng.check_leaf(c)
    with pytest.raises(ValueError):
        Engine(problems=[prob("x"), prob("x")])
    x, y = prob("x"), prob("y")
    with pytest.raises(ValueError):
        Engine(problems=[x, y], dependencies={"l2u": {}, "u2l": {y: [x]}})      # no route back up
    with pytest.raises(NotImplementedError):
        Config(type="neumann")
    with pytest.raises(AttributeError):
        Engine(problems=[prob("run")])


WORKER = r'''
import os, sys, torch
sys.path.insert(0, %r)
import torch.distributed as dist
sys.path.insert(0, os.path.join(%r, "tests"))
from test_betty_driver import quadratic_problems, make_engine
from semantic_segmentation_b200.betty.problems import ImplicitProblem
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("gloo")
r, world = dist.get_rank(), dist.get_world_size()
lower, upper, A, t, c, lr_w, lr_lam = quadratic_problems()
# rank-dependent data term on the lower loss: the averaged problem is the mean over ranks
shift = torch.full((5,), float(r), dtype=torch.float64)
base = type(lower).training_step
type(lower).training_step = lambda self, batch: base(self, batch) + (self.module() * shift).sum()
eng = make_engine(lower, upper, train_iters=1, valid_step=0, distributed=True)
w0, lam0 = lower.module.v.detach().clone(), upper.module.v.detach().clone()
eng.run()
mean_shift = torch.full((5,), sum(range(world)) / world, dtype=torch.float64)
w1 = w0 - lr_w * (w0 - A @ lam0 + mean_shift)
lam1 = lam0 - lr_lam * (c * lam0 + A.T @ (w1 - t))
assert torch.allclose(lower.module.v.detach(), w1, atol=1e-10), (lower.module.v, w1)
assert torch.allclose(upper.module.v.detach(), lam1, atol=1e-7), (upper.module.v, lam1)
both = [torch.zeros(3, dtype=torch.float64) for _ in range(world)]
dist.all_gather(both, upper.module.v.detach())
assert torch.equal(both[0], both[1])           # replicas stay identical
print("rank", r, "ok")
'''


def test_distributed_gradient_average_gloo_world2(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER % (ROOT, ROOT))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)],
                       capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("ok") == 2


def test_gradient_accumulation_warmup_and_scheduler():
    """gradient_accumulation g: the optimiser steps every g-th training_step on the mean of the g losses; warmup_steps: no
    upper-level step before the lower level has taken them; a scheduler attached to a problem steps with its optimiser."""
    torch.manual_seed(0)
    w = Vec([1.0, -2.0, 0.5])
    data = [torch.tensor([float(i + 1)] * 3, dtype=torch.float64) for i in range(4)]

    class P(ImplicitProblem):
        def training_step(self, batch):
            return 0.5 * ((self.module() - batch) ** 2).sum()

    opt = torch.optim.SGD(w.parameters(), lr=0.5)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.5)
    p = P(name="p", module=w, optimizer=opt, scheduler=sched, train_data_loader=data, config=Config(gradient_accumulation=2))
    eng = Engine(problems=[p], config=EngineConfig(train_iters=4, valid_step=0))
    w0 = w.v.detach().clone()
    eng.run()
    g1 = 0.5 * ((w0 - data[0]) + (w0 - data[1]))                 # mean of two gradients at the same point
    w1 = w0 - 0.5 * g1
    g2 = 0.5 * ((w1 - data[2]) + (w1 - data[3]))
    w2 = w1 - 0.25 * g2                                           # StepLR halved the rate after the first optimiser step
    assert torch.allclose(w.v.detach(), w2, atol=1e-12) and p.count == 4
    assert abs(opt.param_groups[0]["lr"] - 0.125) < 1e-12

    lower, upper, *_ = quadratic_problems(unroll=1)
    lower.warmup_steps = 3
    eng = make_engine(lower, upper, train_iters=5, valid_step=0)
    eng.run()
    assert lower.count == 5 and upper.count == 2                  # upper steps only after lower steps 4 and 5


def test_problem_state_dict_round_trip_and_batch_forms():
    lower, upper, *_ = quadratic_problems()
    eng = make_engine(lower, upper, train_iters=3, valid_step=0)
    eng.run()
    state = lower.state_dict()
    assert state["count"] == 3 and "v" in state["module"]
    fresh_l, fresh_u, *_ = quadratic_problems(seed=9)
    make_engine(fresh_l, fresh_u, train_iters=1, valid_step=0)
    fresh_l.load_state_dict(state)
    assert torch.equal(fresh_l.module.v.detach(), lower.module.v.detach()) and fresh_l.count == 3

    # training_step may return a tensor, a dict with "loss", or a tuple whose first entry is the loss; batches may be
    # tensors, tuples or dicts (moved to the problem's device)
    seen = []

    class P(ImplicitProblem):
        def training_step(self, batch):
            seen.append(type(batch).__name__)
            x = batch["x"] if isinstance(batch, dict) else (batch[0] if isinstance(batch, (tuple, list)) else batch)
            loss = ((self.module() - x) ** 2).sum()
            return {"loss": loss, "aux": 1} if isinstance(batch, dict) else ((loss, 2) if isinstance(batch, (tuple, list)) else loss)

    m = Vec([0.0])
    one = torch.ones(1, dtype=torch.float64)
    p = P(name="p", module=m, optimizer=torch.optim.SGD(m.parameters(), lr=0.25),
          train_data_loader=[one, (one, 0), {"x": one}], config=Config(), device=torch.device("cpu"))
    Engine(problems=[p], config=EngineConfig(train_iters=3, valid_step=0)).run()
    assert seen == ["Tensor", "tuple", "dict"]
    assert torch.allclose(m.v.detach(), torch.tensor([1.0 - 0.5 ** 3], dtype=torch.float64))
