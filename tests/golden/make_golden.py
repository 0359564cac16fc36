"""Regenerate tests/golden/*.npz from the reference checkout (run in the build
container only; /root/reference does not exist on the GPU box).

    python tests/golden/make_golden.py [/root/reference]

What is extracted (data only, no reference source):
  * coarsen_restarts_reference.npz — the expected outputs of the reference's
    regression test external/vcm/tests/test_coarsen_restarts.py, i.e. every
    data variable of the 24 JSON files under
    external/vcm/tests/_coarsen_restarts_regression_tests/reference/
    (key "<tag>/<category>/<var>", plus "<...>@dims" with the dim names).
  * coarsen_restarts_meta.json — per golden file the coordinates (dims, values) and the
    attrs of every output variable, for checking the labelled results' metadata.
  * restart_schemas.json — name/dims/shape/dtype/attrs of every input variable from
    .../schemas/*.json, enough to regenerate the synthetic inputs the way
    external/synth/synth/core.py:63-67 does
    (np.random.seed(0); np.random.uniform(lo, hi, shape).astype(dtype)).
  * skimage_block_reduce.npz — outputs of the reference's own
    vcm/cubedsphere/_skimage.py:block_reduce (loaded by file path; it imports
    only numpy) for mean/median on the regtest input of
    test_cubedsphere.py:275-290 and on random blocks with NaNs.
"""
import glob
import importlib.util
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def main(ref="/root/reference"):
    tests = os.path.join(ref, "external/vcm/tests")
    reg = os.path.join(tests, "_coarsen_restarts_regression_tests")

    out, meta = {}, {}
    for path in sorted(glob.glob(os.path.join(reg, "reference", "*.json"))):
        base = os.path.basename(path)[: -len(".json")]
        for cat in ("fv_core.res", "fv_srf_wnd.res", "fv_tracer.res", "sfc_data"):
            if base.endswith("-" + cat):
                tag = base[: -len(cat) - 1]
                break
        else:
            raise RuntimeError(base)
        with open(path) as f:
            d = json.load(f)
        for var, v in d["data_vars"].items():
            arr = np.array(v["data"], dtype=np.float64)
            out[f"{tag}/{cat}/{var}"] = arr
            out[f"{tag}/{cat}/{var}@dims"] = np.array(v["dims"])
        meta[f"{tag}/{cat}"] = {
            "coords": {k: {"dims": c["dims"], "data": c["data"]} for k, c in d["coords"].items()},
            "attrs": {k: v["attrs"] for k, v in d["data_vars"].items()},
        }
    with open(os.path.join(HERE, "coarsen_restarts_meta.json"), "w") as f:
        json.dump(meta, f, indent=0, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "coarsen_restarts_reference.npz"), **out)

    schemas = {}
    for path in sorted(glob.glob(os.path.join(reg, "schemas", "*-schema.json"))):
        cat = os.path.basename(path)[: -len("-schema.json")]
        with open(path) as f:
            s = json.load(f)["schema"]
        schemas[cat] = {
            "variables": {
                k: {"dims": v["dims"], "shape": v["array"]["shape"], "dtype": v["array"]["dtype"],
                    "attrs": v.get("attrs", {})}
                for k, v in s["variables"].items()
            },
            "coords": {k: {"dims": v["dims"], "value": v["value"], "attrs": v.get("attrs", {})}
                       for k, v in s["coords"].items() if k != "time"},
        }
    with open(os.path.join(HERE, "restart_schemas.json"), "w") as f:
        json.dump(schemas, f, indent=0, sort_keys=True)

    # the reference's own block_reduce, imported by path (numpy only)
    spec = importlib.util.spec_from_file_location(
        "_ref_skimage", os.path.join(ref, "external/vcm/vcm/cubedsphere/_skimage.py"))
    sk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sk)
    g = {}
    a = np.arange(32, dtype=np.float32).reshape(4, 4, 2)       # test_cubedsphere.py:262-272
    g["regtest_in"] = a
    g["regtest_mean"] = sk.block_reduce(a, (2, 2, 1), np.mean, cval=np.nan)
    g["regtest_median"] = sk.block_reduce(a, (2, 2, 1), np.median, cval=np.nan)
    rng = np.random.RandomState(7)
    for f, dt in ((8, np.float32), (16, np.float32), (2, np.float64), (4, np.float32)):
        x = rng.uniform(-5, 5, size=(3, 4 * f, 2 * f)).astype(dt)
        x[0, 1, 1] = np.nan                                     # NaN poisons one block only
        x[1, :f, :f] = 1.0                                      # ties
        x[2, f:2 * f, :f] = np.round(x[2, f:2 * f, :f])         # repeated values
        g[f"median_in_f{f}_{np.dtype(dt).name}"] = x
        g[f"median_out_f{f}_{np.dtype(dt).name}"] = sk.block_reduce(
            x, (1, f, f), np.median, cval=np.nan)
    np.savez_compressed(os.path.join(HERE, "skimage_block_reduce.npz"), **g)
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main(*sys.argv[1:])
