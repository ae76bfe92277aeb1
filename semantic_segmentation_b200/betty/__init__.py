"""Betty-compatible bilevel driver -- the part of ``betty-ml`` that
running_files/train_end2end_jsrt.py:35-37,141-342 uses, so the GenSeg end-to-end loop runs where betty-ml is absent:

    from semantic_segmentation_b200.betty.engine import Engine
    from semantic_segmentation_b200.betty.configs import Config, EngineConfig
    from semantic_segmentation_b200.betty.problems import ImplicitProblem

PARITY UNPINNED: betty-ml (pinned 0.2.0 by env.sh:12) is an external dependency with no source under the reference
tree and is not installed here; the reference holds no test of its numerics.  This is a behavioural restatement of its
documented API (problems, l2u / u2l dependencies, unroll_steps, roll_back, valid_step) and of the published DARTS
finite-difference hyper-gradient (Liu et al. 2019, eq. 7-8), validated against closed-form bilevel problems in
tests/test_betty_driver.py.  Host logic only: every tensor operation is the caller's (our CUDA modules or any nn.Module).
"""
from .configs import Config, EngineConfig  # noqa: F401
from .engine import Engine  # noqa: F401
from .problems import ImplicitProblem, Problem  # noqa: F401
