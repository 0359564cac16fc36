"""The reference's `vcm.cubedsphere` names for the coarse-graining path
(external/vcm/vcm/cubedsphere/__init__.py:1-21), backed by CUDA."""
from .coarsen import (  # noqa: F401
    add_coordinates,
    block_coarsen,
    block_edge_coarsen,
    block_edge_sum,
    block_median,
    block_upsample,
    block_upsample_like,
    coarsen_coords,
    coarsen_coords_coord_func,
    edge_weighted_block_average,
    horizontal_block_reduce,
    shift_edge_var_to_center,
    weighted_block_average,
    xarray_block_reduce,
)
from .regridz import (  # noqa: F401
    pressure_level_average,
    regrid_to_area_weighted_pressure,
    regrid_to_edge_weighted_pressure,
    regrid_vertical,
)
from .coarsen_restarts import (  # noqa: F401
    coarsen_restarts_on_pressure,
    coarsen_restarts_on_sigma,
    coarsen_restarts_via_blended_method,
)
from .constants import *  # noqa: F401,F403
