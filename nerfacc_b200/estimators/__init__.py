from .base import AbstractEstimator
from .occ_grid import OccGridEstimator

__all__ = ["AbstractEstimator", "OccGridEstimator"]
