"""stgcn_b200 -- B200-native ST-GCN block stack (drop-in for nagyrajmund/st-gcn's src/network)."""
from . import graph  # noqa: F401
from .modules import SpatialConv, SpatialTemporalConv  # noqa: F401
from .model import L_STGCN, STGCN  # noqa: F401
from . import optim  # noqa: F401
from .functional import cross_entropy  # noqa: F401
from .training import CapturedForward, CapturedTrainStep  # noqa: F401

__all__ = ['SpatialConv', 'SpatialTemporalConv', 'L_STGCN', 'STGCN', 'graph', 'optim', 'cross_entropy', 'CapturedTrainStep', 'CapturedForward']
