from .lamb import Lamb, lamb_bias_correction  # noqa: F401
from .lars import LARC, build_lars_sgd  # noqa: F401
