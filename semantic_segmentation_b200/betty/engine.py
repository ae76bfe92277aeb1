"""``Engine``: wires the problems by their dependencies and runs the iteration loop, as
train_end2end_jsrt.py:239-342 subclasses (``validation``), constructs and ``run()``s it.  PARITY UNPINNED (package docstring).

Dependencies: ``{"l2u": {lower: [upper, ...]}, "u2l": {upper: [lower, ...]}}``.  An l2u edge makes ``upper`` a parent of
``lower`` (it steps after every ``lower.config.unroll_steps`` lower steps); a u2l edge says ``upper``'s variables enter
``lower``'s loss, so ``upper`` receives one hyper-gradient path [upper, lower, ..., upper] for every l2u route from that
lower level back up to it.  (The reference's ``l2u = {netG: [unet], netG: [arch]}`` is a dict literal with a repeated key:
Python keeps ``{netG: [arch]}`` -- the engine sees exactly that.)
"""
from __future__ import annotations

import logging
from typing import Dict, List, Optional

from .configs import EngineConfig
from .problems import Problem


class Engine:
    def __init__(self, problems: List[Problem], config: Optional[EngineConfig] = None,
                 dependencies: Optional[Dict[str, dict]] = None, env=None):
        self.config = config if config is not None else EngineConfig()
        self.train_iters = self.config.train_iters
        self.valid_step = self.config.valid_step
        self._roll_back = self.config.roll_back
        self.problems = list(problems)
        self.dependencies = dependencies if dependencies is not None else {"l2u": {}, "u2l": {}}
        self.env = env
        self.leaves: List[Problem] = []
        self.global_step = 0
        self._training = True
        self.initialize()

    # ---- wiring ----------------------------------------------------------------------------------------------------------
    def initialize(self) -> None:
        names = [p.name for p in self.problems]
        if len(set(names)) != len(names):
            raise ValueError("problem names must be unique: %s" % names)
        self.parse_dependency()
        self.set_problem_attr()
        for problem in self.problems:
            problem.initialize(engine_config=self.config)
            if self.env is not None:
                problem.set_env(self.env)

    def parse_dependency(self) -> None:
        l2u = self.dependencies.get("l2u", {}) or {}
        u2l = self.dependencies.get("u2l", {}) or {}
        for lower, uppers in l2u.items():
            for upper in uppers:
                lower.add_parent(upper)
                upper.add_child(lower)
        for upper, lowers in u2l.items():
            for lower in lowers:
                upper.add_paths(self.find_paths(src=lower, dst=upper))
        self.leaves = [p for p in self.problems if p.leaf]

    def find_paths(self, src: Problem, dst: Problem) -> List[List[Problem]]:
        """Every l2u route src -> ... -> dst, returned in back-propagation order [dst, ..., src, dst]."""
        l2u = self.dependencies.get("l2u", {}) or {}
        routes: List[List[Problem]] = []

        def dfs(node, route):
            if node is dst:
                routes.append(list(route))
                return
            for nxt in l2u.get(node, []):
                if nxt not in route:
                    dfs(nxt, route + [nxt])

        dfs(src, [src])
        if not routes:
            raise ValueError("u2l edge %s -> %s has no l2u route back up" % (dst.name, src.name))
        return [list(reversed(r)) + [dst] for r in routes]

    def set_problem_attr(self) -> None:
        for problem in self.problems:
            if hasattr(self, problem.name):
                raise AttributeError("problem name %r collides with an Engine attribute" % problem.name)
            setattr(self, problem.name, problem)
            for other in self.problems:
                if other is not problem:
                    problem.set_problem_attr(other)

    def check_leaf(self, problem: Problem) -> bool:
        return problem in self.leaves

    def is_implemented(self, fn_name: str) -> bool:
        return getattr(type(self), fn_name, None) is not getattr(Engine, fn_name, None)

    # ---- loop ------------------------------------------------------------------------------------------------------------
    def train(self) -> None:
        self._training = True
        for problem in self.problems:
            problem.train()

    def eval(self) -> None:
        self._training = False
        for problem in self.problems:
            problem.eval()

    def train_step(self) -> None:
        for leaf in self.leaves:
            leaf.step(global_step=self.global_step)

    def validation(self):
        """Overridden by the caller (train_end2end_jsrt.py:239-253); may return a dict of statistics."""
        return None

    def run(self) -> None:
        self.train()
        for it in range(1, self.train_iters + 1):
            self.global_step += 1
            self.train_step()
            if self.valid_step > 0 and it % self.valid_step == 0 and self.is_implemented("validation"):
                self.eval()
                stats = self.validation()
                if isinstance(stats, dict) and stats:
                    logging.info("[validation] step %d %s", self.global_step,
                                 " ".join("%s: %s" % (k, v) for k, v in stats.items()))
                self.train()
