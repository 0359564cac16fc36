"""The fine_res_budget composition on device-resident data.

Restates what workflows/fine_res_budget/budget/budgets.py:217-297
(`compute_recoarsened_budget_inputs`) computes for ONE timestep — the unit of work the
reference hands to one Beam worker (budget/pipeline.py:35-53) — as a handful of CUDA
launches with no fine-grid intermediate:

  delp_coarse  = weighted_block_average(delp, area)                    (budgets.py:275)
  3-D fields   = Grid.pressure_level_average(...)  for 22 first moments,
                 2 second moments (T*omega, sphum*omega) and 2 storage terms
                 ((end - begin)/dt of T, sphum)                         (:49-57, :80, :151)
  2-D fields   = weighted_block_average(field, area)                    (:121)
  exposed_area = Grid.area_above_fine_surface(delp, delp_coarse, exposed_area_c384) (:62-77)
  area         = block_sum(area)                                        (:59, :288-292)

Variable lists are those of budget/config.py:7-58.
"""
from collections import OrderedDict

from . import ops

FACTOR = 8          # budget/config.py:57
DT = 15 * 60        # budget/config.py:58

PHYSICS_3D = [
    "t_dt_fv_sat_adj_coarse", "t_dt_nudge_coarse", "t_dt_phys_coarse", "qv_dt_fv_sat_adj_coarse",
    "qv_dt_phys_coarse", "eddy_flux_vulcan_omega_sphum", "eddy_flux_vulcan_omega_temp",
    "vulcan_omega_coarse",
]
RESTART_3D = ["sphum", "T", "DZ"]
GFSPHYSICS_3D = [
    "dq3dt_deep_conv_coarse", "dq3dt_mp_coarse", "dq3dt_pbl_coarse", "dq3dt_shal_conv_coarse",
    "dt3dt_deep_conv_coarse", "dt3dt_lw_coarse", "dt3dt_mp_coarse", "dt3dt_ogwd_coarse",
    "dt3dt_pbl_coarse", "dt3dt_shal_conv_coarse", "dt3dt_sw_coarse",
]
GFSPHYSICS_2D = [
    "DLWRFsfc_coarse", "DSWRFsfc_coarse", "DSWRFtoa_coarse", "ULWRFsfc_coarse", "ULWRFtoa_coarse",
    "USWRFsfc_coarse", "USWRFtoa_coarse", "LHTFLsfc_coarse", "SHTFLsfc_coarse", "PRATEsfc_coarse",
]
FIRST_MOMENTS_3D = RESTART_3D + PHYSICS_3D + GFSPHYSICS_3D          # 22 fields
SECOND_MOMENTS = (("T", "vulcan_omega_coarse"), ("sphum", "vulcan_omega_coarse"))
STORAGE_TERMS = ("T", "sphum")


def recoarsen_budget_inputs(middle3d, middle2d, delp, area, exposed_area_fine, begin, end,
                            dt=DT, factor=FACTOR, second_moments=SECOND_MOMENTS,
                            storage_terms=STORAGE_TERMS):
    """One timestep.  All arguments are CUDA tensors (float32):
      middle3d: {name: (tile, z, y, x)} first moments at step="middle"
      middle2d: {name: (tile, y, x)}
      delp (tile, z, y, x), area / exposed_area_fine (tile, y, x)
      begin / end: {name: (tile, z, y, x)} for the storage terms
    Returns an OrderedDict of coarse tensors keyed like the reference's output Dataset."""
    names3d = list(middle3d)
    fields = [middle3d[n] for n in names3d]
    out_names = list(names3d)
    for a, b in second_moments:                      # budgets.py:151-158, name f"{a}_{b}"
        fields.append(ops.product(middle3d[a], middle3d[b]))
        out_names.append(f"{a}_{b}")
    for n in storage_terms:                          # budgets.py:80-85, name f"{n}_storage"
        fields.append(ops.storage(end[n], begin[n], dt))
        out_names.append(f"{n}_storage")

    coarse3d, delp_coarse = ops.pressure_level_average(fields, delp, area, factor, mask_below=False)
    out = OrderedDict()
    if middle2d:
        names2d = list(middle2d)
        c2 = ops.weighted_block_average([middle2d[n].unsqueeze(1) for n in names2d], w2d=area,
                                        fy=factor, fx=factor)
        for n, t in zip(names2d, c2):
            out[n] = t.squeeze(1)
    for n, t in zip(out_names, coarse3d):
        out[n] = t
    out["delp"] = delp_coarse
    out["exposed_area"] = ops.area_above_fine_surface(delp, delp_coarse, exposed_area_fine, factor)
    out["area"] = ops.block_reduce(area, "sum", factor, factor)
    return out


def algorithmic_bytes(n_tile, nz, ny, nx, n3d_reads, n3d_out, n2d, factor=FACTOR, itemsize=4):
    """Every input element read once + every output element written once."""
    h, hc = ny * nx, (ny // factor) * (nx // factor)
    fine3, coarse3 = n_tile * nz * h * itemsize, n_tile * nz * hc * itemsize
    fine2, coarse2 = n_tile * h * itemsize, n_tile * hc * itemsize
    reads = (n3d_reads + 1) * fine3 + (n2d + 2) * fine2          # fields + delp; 2-D + area + exposed
    writes = (n3d_out + 2) * coarse3 + (n2d + 1) * coarse2       # + delp_coarse + exposed_area; + area
    return reads + writes
