"""Optimizers, LR schedulers and loss scaling."""
from hetseq_b200.optim.tensor_optimizers import Adam, Adadelta  # noqa: F401
from hetseq_b200.optim.flat_optimizer import _Optimizer, _Adam, _Adadelta, DynamicLossScaler  # noqa: F401
from hetseq_b200.optim.lr_scheduler import _LRScheduler, PolynomialDecayScheduler, FixedScheduler  # noqa: F401
