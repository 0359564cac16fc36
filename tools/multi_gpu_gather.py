"""One C384 timestep sharded over the ranks of one box (regime "fewer timesteps than GPUs"):
every rank coarse-grains its (tile, row band) shards with the CUDA kernels, ONE NCCL gather brings the coarse
outputs (1/64 of the input bytes) to rank 0, which checks the assembled fields bit for bit against its own
unsharded run.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/multi_gpu_gather.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fv3net_b200 import ops, sharding as S, synthetic as syn  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    n_tile, n, f, nfields = 6, int(os.environ.get("CG_N", 384)), 8, 4
    nz = syn.NZ
    # the same synthetic timestep on every rank (seeded), so each rank can cut its own shards
    delp = syn.torch_delp(n_tile, n, n, 11, dev)
    area = syn.torch_area((n_tile, n, n), 12, dev)
    fields = [syn.torch_field((n_tile, nz, n, n), 20 + i, dev, scale=10.0 ** i) for i in range(nfields)]
    shards = S.plan(1, world, n_tile=n_tile, ny=n, factor=f)
    names = [f"q{i}" for i in range(nfields)] + ["delp"]
    local = {k: [] for k in names}
    for s in shards[rank]:
        r1 = n if s.row1 is None else s.row1
        cut = lambda t: t[s.tile0:s.tile1, ..., s.row0:r1, :].contiguous()   # noqa: E731
        outs, dc = ops.pressure_level_average([cut(t) for t in fields], cut(delp), cut(area), f)
        for k, o in zip(names, outs + [dc]):
            local[k].append(o)
    shapes = {k: ((nz,), n, n, f) for k in names}
    S.gather_coarse(local, shards, shapes, dst=0)        # warm-up: the first gather sets up NCCL's point-to-point channels
    torch.cuda.synchronize()
    dist.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    gathered = S.gather_coarse(local, shards, shapes, dst=0)
    t1.record()
    torch.cuda.synchronize()
    ok = True
    if rank == 0:
        ref, ref_dc = ops.pressure_level_average(fields, delp, area, f)
        for k, e in zip(names, ref + [ref_dc]):
            got = S.assemble(gathered[k], 1, n_tile, (nz,), n, n, f)[0]
            ok = ok and torch.equal(torch.nan_to_num(got), torch.nan_to_num(e))
        nbytes = sum(t.numel() * 4 for k in names for t in gathered[k].values())
        print(json.dumps({"world": world, "shards_per_rank": [len(r) for r in shards], "bit_identical_to_unsharded": bool(ok),
                          "gather_ms": t0.elapsed_time(t1), "gathered_bytes": nbytes, "backend": "nccl"}))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
