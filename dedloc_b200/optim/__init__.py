from .lamb import Lamb, lamb_bias_correction  # noqa: F401
from .lars import LARC, build_lars_sgd  # noqa: F401
from .schedules import LinearWarmupCosineAnnealingLR, NoOpScheduler, get_linear_schedule_with_warmup  # noqa: F401
