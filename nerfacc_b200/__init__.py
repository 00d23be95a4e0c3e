"""nerfacc_b200 — B200-native (sm_100a) implementation of nerfacc's rendering hot path behind
nerfacc's own Python API (reference nerfacc/__init__.py:23-46).  No CPU fallback for packed ops."""
from .chunked import render_image_with_occgrid_test  # additive: the test-time loop
from .data_specs import RayIntervals, RaySamples
from .estimators.occ_grid import OccGridEstimator
from .estimators.prop_net import PropNetEstimator
from .grid import ray_aabb_intersect, traverse_grids
from .pack import pack_info, sample_positions  # sample_positions: additive
from .pdf import importance_sampling, searchsorted
from .scan import exclusive_prod, exclusive_sum, inclusive_prod, inclusive_sum
from .version import __version__
from .volrend import (
    accumulate_along_rays,
    accumulate_along_rays_,
    render_transmittance_from_alpha,
    render_transmittance_from_density,
    render_visibility_from_alpha,
    render_visibility_from_density,
    render_weight_from_alpha,
    render_weight_from_density,
    rendering,
)

# the 22 public names of the reference (nerfacc/__init__.py:23-46): `from nerfacc import *` works
# the same when this package is imported under the alias `nerfacc`
__all__ = [
    "__version__",
    "inclusive_prod",
    "exclusive_prod",
    "inclusive_sum",
    "exclusive_sum",
    "pack_info",
    "render_visibility_from_alpha",
    "render_visibility_from_density",
    "render_weight_from_alpha",
    "render_weight_from_density",
    "render_transmittance_from_alpha",
    "render_transmittance_from_density",
    "accumulate_along_rays",
    "rendering",
    "importance_sampling",
    "searchsorted",
    "RayIntervals",
    "RaySamples",
    "ray_aabb_intersect",
    "traverse_grids",
    "OccGridEstimator",
    "PropNetEstimator",
]
