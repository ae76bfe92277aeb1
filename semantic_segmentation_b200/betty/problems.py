"""``Problem`` / ``ImplicitProblem``: one optimisation level of the bilevel program (module + optimiser + data
loader + ``training_step``), as train_end2end_jsrt.py:141-236,295-330 subclasses and constructs it.

Step semantics (restated from Betty's documentation; PARITY UNPINNED, see the package docstring):
  * a leaf problem (no children) is stepped once per Engine.train_step; after every ``unroll_steps`` of its own steps it
    marks itself ready at each parent, and a parent steps when all its children are ready;
  * a step = next batch -> ``training_step`` -> gradients w.r.t. ``trainable_parameters()`` -> optimiser step -> zero_grad;
  * a problem with paths (an upper level) adds, for every path [upper, lower_k .. lower_1, upper], the implicit term
    obtained by chaining the lower levels' hyper-gradient functions (``hypergradient.darts``);
  * roll_back: a lower level snapshots parameters, buffers and optimiser state when an unroll starts; after its parents
    have stepped it returns to the snapshot and takes ONE step under the new upper-level variables (the unrolled steps
    were a look-ahead, DARTS eq. 7's w').
"""
from __future__ import annotations

import copy
import logging
from typing import Any, List, Optional

import torch
import torch.distributed as dist

from .configs import Config
from . import hypergradient


def _to_device(batch: Any, device) -> Any:
    if device is None:
        return batch
    if torch.is_tensor(batch):
        return batch.to(device, non_blocking=True)
    if isinstance(batch, dict):
        return {k: _to_device(v, device) for k, v in batch.items()}
    if isinstance(batch, (list, tuple)):
        return type(batch)(_to_device(v, device) for v in batch)
    return batch


class Problem:
    def __init__(self, name: str, config: Optional[Config] = None, module=None, optimizer=None, scheduler=None,
                 train_data_loader=None, device=None, extra_config=None):
        self._name = name
        self._config = config if config is not None else Config()
        self.module = module
        self.optimizer = optimizer
        self.scheduler = scheduler
        self.train_data_loader = train_data_loader
        self.device = device
        self.extra_config = extra_config
        self.train_data_iterator = None
        self.epoch_counter = 0
        self.cur_batch = None
        self._parents: List["Problem"] = []
        self._children: List["Problem"] = []
        self._paths: List[List["Problem"]] = []
        self.ready: List[bool] = []
        self._count = 0
        self._global_step = 0
        self._training = True
        self._inner_loop_start = True
        self._roll_back = False
        self._distributed = False
        self._unroll_steps = self._config.unroll_steps
        self.gas = self._config.gradient_accumulation
        self.warmup_steps = self._config.warmup_steps
        self._cache = None
        self._env = None
        self._last_loss = None

    # ---- wiring (Engine.parse_dependency) ---------------------------------------------------------------------------
    @property
    def name(self) -> str:
        return self._name

    @property
    def config(self) -> Config:
        return self._config

    @property
    def count(self) -> int:
        return self._count

    @property
    def parents(self):
        return self._parents

    @property
    def children(self):
        return self._children

    @property
    def paths(self):
        return self._paths

    @property
    def leaf(self) -> bool:
        return len(self._children) == 0

    def add_parent(self, problem: "Problem") -> None:
        if problem not in self._parents:
            self._parents.append(problem)

    def add_child(self, problem: "Problem") -> None:
        if problem not in self._children:
            self._children.append(problem)
            self.ready.append(False)

    def add_paths(self, paths) -> None:
        self._paths.extend(paths)

    def set_problem_attr(self, problem: "Problem") -> None:
        """Other problems are reachable by name: ``self.netD(x)``, ``self.unet.module`` (train_end2end_jsrt.py:148,241)."""
        if hasattr(self, problem.name) and getattr(self, problem.name) is not problem:
            raise AttributeError("problem name %r collides with an attribute of problem %r" % (problem.name, self._name))
        setattr(self, problem.name, problem)

    def set_env(self, env) -> None:
        self._env = env

    def initialize(self, engine_config=None) -> None:
        if engine_config is not None:
            self._roll_back = bool(engine_config.roll_back or self._config.roll_back)
            self._distributed = bool(engine_config.distributed)
        else:
            self._roll_back = bool(self._config.roll_back)
        for attr in ("module", "optimizer", "scheduler", "train_data_loader"):
            hook = getattr(self, "configure_" + attr, None)
            if getattr(self, attr) is None and callable(hook):
                setattr(self, attr, hook())
        if self.device is not None and isinstance(self.module, torch.nn.Module):
            self.module.to(self.device)
        if self.train_data_loader is not None:
            self.train_data_iterator = iter(self.train_data_loader)

    def is_implemented(self, fn_name: str) -> bool:
        return callable(getattr(self, fn_name, None))

    # ---- the module ------------------------------------------------------------------------------------------------------
    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def __call__(self, *args, **kwargs):
        return self.forward(*args, **kwargs)

    def training_step(self, batch):
        raise NotImplementedError("%s.training_step" % type(self).__name__)

    def training_step_exec(self, batch):
        return self.training_step(batch)

    def parameters(self):
        raise NotImplementedError

    def trainable_parameters(self):
        raise NotImplementedError

    def train(self) -> None:
        self._training = True

    def eval(self) -> None:
        self._training = False

    # ---- one step ----------------------------------------------------------------------------------------------------------
    def get_batch(self):
        if self.train_data_iterator is None:
            if self.train_data_loader is None:
                return None
            self.train_data_iterator = iter(self.train_data_loader)
        try:
            batch = next(self.train_data_iterator)
        except StopIteration:
            self.epoch_counter += 1
            sampler = getattr(self.train_data_loader, "sampler", None)
            if hasattr(sampler, "set_epoch"):
                sampler.set_epoch(self.epoch_counter)
            self.train_data_iterator = iter(self.train_data_loader)
            batch = next(self.train_data_iterator)
        return _to_device(batch, self.device)

    def get_loss(self, batch):
        out = self.training_step_exec(batch)
        if isinstance(out, dict):
            loss = out["loss"]
        elif isinstance(out, (tuple, list)):
            loss = out[0]
        else:
            loss = out
        self._last_loss = loss.detach()
        return loss / self.gas if self.gas > 1 else loss

    def backward(self, loss, params, paths) -> None:
        """Direct gradient, plus one implicit term per path; accumulated into ``p.grad``."""
        cfg = self._config
        implicit = bool(paths) and not cfg.first_order
        grads = torch.autograd.grad(loss, params, retain_graph=cfg.retain_graph or implicit, allow_unused=cfg.allow_unused)
        grads = list(grads)
        if implicit:
            for i, path in enumerate(paths):
                keep = cfg.retain_graph or i + 1 < len(paths)
                extra = hypergradient.get_grads(loss, path, retain_graph=keep, sync=self._distributed)
                for j, e in enumerate(extra):
                    if e is not None:
                        grads[j] = e if grads[j] is None else grads[j] + e
        for p, g in zip(params, grads):
            if g is None:
                continue
            if p.grad is None:
                p.grad = g.detach().clone()
            else:
                p.grad.add_(g.detach())

    def sync_grads(self, params) -> None:
        """Data-parallel hook: average the gradients over the ranks (every rank holds a shard of the batch)."""
        if not (self._distributed and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
            return
        hypergradient.allreduce_mean([p.grad for p in params])

    def optimizer_step(self) -> None:
        raise NotImplementedError

    def zero_grad(self) -> None:
        for p in self.trainable_parameters():
            p.grad = None
        if self.optimizer is not None:
            self.optimizer.zero_grad(set_to_none=True)

    def one_step_descent(self, batch=None):
        if batch is None:
            self.cur_batch = self.get_batch()
            batch = self.cur_batch
        loss = self.get_loss(batch)
        params = list(self.trainable_parameters())
        self.backward(loss, params, self._paths)
        if self._count % self.gas == 0:
            self.sync_grads(params)
            self.optimizer_step()
            if self.is_implemented("param_callback"):
                self.param_callback()
            self.zero_grad()
        if self._config.log_step > 0 and self._count % self._config.log_step == 0:
            logging.info("[%s] step %d (global %d) loss %.6f", self._name, self._count, self._global_step, float(self._last_loss))
        return loss

    def check_ready(self) -> bool:
        return all(self.ready)

    def step(self, global_step: Optional[int] = None) -> None:
        """Called by the Engine on leaves; parents are stepped from here when an unroll completes."""
        if global_step is not None:
            self._global_step = global_step
        if not self.check_ready():
            return
        if self._inner_loop_start:
            if self.is_implemented("on_inner_loop_start"):
                self.on_inner_loop_start()
            self._inner_loop_start = False
            if self._roll_back and self._parents:
                self.cache_states()
        if self._training:
            self._count += 1
        self.one_step_descent()
        boundary = self._count % (self._unroll_steps * self.gas) == 0 and self._count > self.warmup_steps
        if boundary and self._parents:
            stepped = False
            for parent in self._parents:
                parent.ready[parent._children.index(self)] = True
                if parent.check_ready():
                    parent.step(global_step=self._global_step)
                    stepped = True
            self._inner_loop_start = True
            if stepped and self._roll_back and self._training:
                self.recover_states()
                self.one_step_descent()
        self.ready = [False] * len(self._children)

    # ---- roll-back snapshot ----------------------------------------------------------------------------------------------
    def cache_states(self) -> None:
        raise NotImplementedError

    def recover_states(self) -> None:
        raise NotImplementedError

    def state_dict(self) -> dict:
        return {"module": self.module.state_dict() if isinstance(self.module, torch.nn.Module) else None,
                "optimizer": self.optimizer.state_dict() if self.optimizer is not None else None,
                "scheduler": self.scheduler.state_dict() if self.scheduler is not None else None,
                "count": self._count}

    def load_state_dict(self, state: dict) -> None:
        if state.get("module") is not None:
            self.module.load_state_dict(state["module"])
        if state.get("optimizer") is not None and self.optimizer is not None:
            self.optimizer.load_state_dict(state["optimizer"])
        if state.get("scheduler") is not None and self.scheduler is not None:
            self.scheduler.load_state_dict(state["scheduler"])
        self._count = int(state.get("count", self._count))


class ImplicitProblem(Problem):
    """A level whose variables are the parameters of ``module`` and whose update is ``optimizer.step()``."""

    def parameters(self):
        return list(self.module.parameters())

    def trainable_parameters(self):
        return [p for p in self.module.parameters() if p.requires_grad]

    def optimizer_step(self) -> None:
        if self.is_implemented("custom_optimizer_step"):
            self.custom_optimizer_step()
        else:
            if self._config.gradient_clipping > 0.0:
                torch.nn.utils.clip_grad_norm_(self.trainable_parameters(), self._config.gradient_clipping)
            self.optimizer.step()
        if self.scheduler is not None:
            self.scheduler.step()

    def cache_states(self) -> None:
        # parameters and buffers are copied on their device (one D2D pass); the optimiser state is deep-copied
        self._cache = ({k: v.detach().clone() for k, v in self.module.state_dict().items()},
                       copy.deepcopy(self.optimizer.state_dict()) if self.optimizer is not None else None)

    def recover_states(self) -> None:
        if self._cache is None:
            return
        module_state, opt_state = self._cache
        with torch.no_grad():
            for k, v in self.module.state_dict().items():
                v.copy_(module_state[k])
        if opt_state is not None:
            self.optimizer.load_state_dict(opt_state)
        self._cache = None
