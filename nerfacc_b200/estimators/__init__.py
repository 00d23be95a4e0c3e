from .base import AbstractEstimator
from .occ_grid import OccGridEstimator
from .prop_net import PropNetEstimator

__all__ = ["AbstractEstimator", "OccGridEstimator", "PropNetEstimator"]
