"""``Config`` (per problem) and ``EngineConfig`` -- the fields train_end2end_jsrt.py:287-293 sets plus the documented
ones that change the driver's arithmetic.  Unknown keyword arguments raise TypeError (dataclass), like the original."""
from dataclasses import dataclass


@dataclass
class Config:
    type: str = "darts"              # hyper-gradient method of THIS problem when it is a lower level on a path
    unroll_steps: int = 1            # lower-level steps between two upper-level steps
    retain_graph: bool = False       # keep the loss graph after backward (train_end2end_jsrt.py:286)
    allow_unused: bool = True        # parameters the loss does not reach get no gradient instead of an error
    gradient_accumulation: int = 1   # optimiser step every g-th training_step
    gradient_clipping: float = 0.0   # > 0: clip_grad_norm_ before the optimiser step
    warmup_steps: int = 0            # lower-level steps before the first upper-level step
    roll_back: bool = False          # per-problem override of EngineConfig.roll_back
    darts_alpha: float = 0.01        # finite-difference radius: eps = darts_alpha / ||v||
    first_order: bool = False        # drop the implicit term (first-order DARTS)
    log_step: int = -1               # > 0: logging.info the loss every log_step steps

    def __post_init__(self):
        if self.type not in ("darts",):
            raise NotImplementedError("hyper-gradient type %r: only 'darts' (the one train_end2end_jsrt.py:287 uses) is built" % self.type)
        if self.unroll_steps < 1 or self.gradient_accumulation < 1:
            raise ValueError("unroll_steps and gradient_accumulation must be >= 1")


@dataclass
class EngineConfig:
    train_iters: int = 50000         # calls of Engine.train_step
    valid_step: int = 500            # Engine.validation() every valid_step iterations
    roll_back: bool = False          # lower levels return to their pre-unroll state after the upper-level step
    distributed: bool = False        # average gradients (and the DARTS direction) over torch.distributed ranks
    logger_type: str = "none"        # accepted for signature compatibility; logging goes through ``logging``
