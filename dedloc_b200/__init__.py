"""dedloc_b200 -- a Blackwell-native (B200 / sm_100a) collaborative-training framework with the capabilities
of yhn112/DeDLOC: DHT peer discovery, CollaborativeOptimizer (accumulate to a collaboration-wide batch, then a
weighted butterfly all-reduce fused with LAMB/LARS over NVLink peer memory), progress tracking, state download
from peers, auxiliary / first-peer roles and the albert / sahajbert / swav trainers. See SURVEY.md / DESIGN.md.
"""
__version__ = "0.1.0"

from .utils.compression import CompressionType  # noqa: F401


def __getattr__(name):
    # hivemind-shaped top-level API, imported lazily so that `import dedloc_b200` stays light
    if name in ("DHT", "get_dht_time"):
        from . import dht as _dht
        return getattr(_dht, name)
    if name in ("CollaborativeOptimizer",):
        from .collab.optimizer import CollaborativeOptimizer
        return CollaborativeOptimizer
    raise AttributeError(name)
