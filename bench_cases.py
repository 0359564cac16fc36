"""The other named workloads of BASELINE.json beside the headline one, as sub-records of bench.py's JSON line.

    configs[1]  C384 -> C48 model-level averages of 20 variables (area- and mass-weighted block averages,
                edge-weighted averages of the staggered winds): coarsen_restarts.py:316-408.  Weak by time.
    configs[3]  coarsen_c384_diagnostics 2-D sweep: block_median and block_upsample over 96 timesteps, the
                timesteps dealt over the ranks (coarsen_c384_diagnostics.py:64-70, coarsen_surface_data/core.py:23).
    configs[4]  C3072 -> C384 pressure-level coarse-graining cut by (tile, 768-row band) = 24 shards
                (coarsen_surface_c3072.py:29-76), the shards dealt over the ranks, coarse outputs all-gathered.
    strong      configs[2]'s ONE timestep cut by `sharding.plan` over the ranks (tiles, then row bands), every
                rank's coarse outputs all-gathered inside the timed region; the assembled result is compared
                bit for bit with the unsharded run once before timing.
    h2d         bare concurrent pinned-host -> device copies of the end-to-end leg's buffers (its ceiling).

Every record: `ms_per_step` = CUDA events on the launching stream, barrier + synchronize on both sides, max
over ranks; `value` = fine-grid cells of every coarse-grained field per second, all ranks; `roofline` =
algorithmic bytes of the rank's kernels / their time against the measured copy bandwidth.
"""
import os
from collections import OrderedDict

import torch
import torch.distributed as dist

from fv3net_b200 import budget as B, ops, sharding as S, synthetic as syn

NZ = syn.NZ


class Ctx:
    def __init__(self, world, rank, device, peak_gbs, steps, warmup):
        self.world, self.rank, self.device, self.peak = world, rank, device, peak_gbs
        self.steps, self.warmup = steps, warmup

    def barrier(self):
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = torch.tensor(values, device=self.device, dtype=torch.float64)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    def timed(self, fn, steps=None, warmup=None):
        """ms per step of fn(), max over ranks."""
        steps, warmup = steps or self.steps, self.warmup if warmup is None else warmup
        for _ in range(warmup):
            fn()
        self.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        self.barrier()
        return self.max_over_ranks([a.elapsed_time(b) / steps])[0]

    def roofline(self, nbytes, ms, kernel):
        gbs = nbytes / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": gbs, "peak": self.peak, "unit": "GB/s", "frac": gbs / self.peak,
                "traffic": None, "kernel": kernel, "ms": ms, "algorithmic_bytes": int(nbytes)}


def _free():
    import gc

    gc.collect()
    torch.cuda.empty_cache()


# ------------------------------------------------------------------------------------------- configs[1]
def config1(ctx, n=384, n_tile=6, f=8):
    """20 variables of one restart timestep per rank (weak by time): 3 area-weighted and 11 mass-weighted 3-D
    fields, u and v on their staggered edges with dx / dy weights, 4 area-weighted 2-D fields."""
    dev, seed = ctx.device, 5000 + 100 * ctx.rank
    delp = syn.torch_delp(n_tile, n, n, seed, dev)
    area = syn.torch_area((n_tile, n, n), seed + 1, dev)
    a3 = [delp] + [syn.torch_field((n_tile, NZ, n, n), seed + 2 + i, dev) for i in range(2)]
    m3 = [syn.torch_field((n_tile, NZ, n, n), seed + 10 + i, dev, scale=10.0 ** (i % 3)) for i in range(11)]
    u = syn.torch_field((n_tile, NZ, n + 1, n), seed + 30, dev)
    v = syn.torch_field((n_tile, NZ, n, n + 1), seed + 31, dev)
    dx = syn.torch_area((n_tile, n + 1, n), seed + 32, dev)
    dy = syn.torch_area((n_tile, n, n + 1), seed + 33, dev)
    s2 = [syn.torch_field((n_tile, 1, n, n), seed + 40 + i, dev) for i in range(4)]
    nc = n // f

    def step():
        ops.weighted_block_average(a3, w2d=area, fy=f, fx=f)
        ops.weighted_block_average(m3, w2d=area, w3d=delp, fy=f, fx=f)
        ops.weighted_block_average(u, w2d=dx, fy=1, fx=f, sy=f, sx=f, nyo=nc + 1, nxo=nc)      # coarsen.py:221-273
        ops.weighted_block_average(v, w2d=dy, fy=f, fx=1, sy=f, sx=f, nyo=nc, nxo=nc + 1)
        ops.weighted_block_average(s2, w2d=area, fy=f, fx=f)

    ms = ctx.timed(step)
    fine3, fine2 = n_tile * NZ * n * n, n_tile * n * n
    cells = 14 * fine3 + 2 * n_tile * NZ * (n + 1) * n + 4 * fine2
    # distinct inputs read once (delp serves as field and as weight; the edge averages need the kept lines only),
    # every coarse value written once
    nbytes = 4 * (14 * fine3 + 2 * (fine3 // f) + 4 * fine2 + 3 * fine2 + 16 * n_tile * NZ * nc * nc + 4 * n_tile * nc * nc)
    rec = {"workload": "configs[1]: C384->C48 (f=8) model-level averages, 20 variables: 3 area-weighted + 11 mass-weighted "
                       "3-D fields, staggered u / v with dx / dy edge weights, 4 area-weighted 2-D fields; one timestep per rank",
           "scaling": "weak", "ms_per_step": ms, "value": ctx.world * cells / (ms * 1e-3), "unit": "cells/s",
           "gpu_launches_per_step": 5,
           "roofline": ctx.roofline(nbytes, ms, "cg::wavg_f8_kernel (x3) + cg::wavg_generic_kernel (edge averages)")}
    del delp, area, a3, m3, u, v, dx, dy, s2
    _free()
    return rec


# ------------------------------------------------------------------------------------------- configs[3]
def config3(ctx, n=384, n_tile=6, f=8, n_time=96, n_var=8):
    """96 timesteps of `n_var` 2-D diagnostics, dealt over the ranks by time (total work fixed: strong)."""
    t0, t1 = S.split_range(n_time, ctx.world, ctx.rank)
    mine = t1 - t0
    dev = ctx.device
    fields = [torch.rand((max(mine, 1) * n_tile, n, n), device=dev) for _ in range(n_var)]

    def step():
        if mine == 0:
            return
        for x in fields:
            ops.block_upsample(ops.block_median(x, f), f, f)

    ms = ctx.timed(step)
    fine = n_time * n_tile * n * n
    cells = 2 * n_var * fine                                      # median over the fine grid + the upsampled result
    per_rank = n_var * 2 * 4 * (mine * n_tile) * (n * n + (n // f) ** 2)     # each op: fine + coarse, read or written
    rec = {"workload": f"configs[3]: coarsen_c384_diagnostics 2-D sweep, block_median then block_upsample (f=8) of {n_var} "
                       f"variables x {n_time} timesteps x 6 tiles x 384^2, timesteps dealt over the ranks",
           "scaling": "strong", "ms_per_step": ms, "value": cells / (ms * 1e-3), "unit": "cells/s",
           "gpu_launches_per_step": 2 * n_var,
           "roofline": ctx.roofline(ctx.max_over_ranks([per_rank])[0], ms, "cg::median_f8_f32_kernel + cg::upsample_kernel")}
    del fields
    _free()
    return rec


# ------------------------------------------------------------------------------------------- configs[4]
def _budget_fields(shape3, seed, dev):
    n3 = len(B.FIRST_MOMENTS_3D)
    mid = [syn.torch_field(shape3, seed + i, dev, scale=10.0 ** (i % 4 - 2)) for i in range(n3)]
    t_i, q_i, w_i = B.FIRST_MOMENTS_3D.index("T"), B.FIRST_MOMENTS_3D.index("sphum"), B.FIRST_MOMENTS_3D.index("vulcan_omega_coarse")
    fields = list(mid)
    fields.append(ops.product(mid[t_i], mid[w_i]))
    fields.append(ops.product(mid[q_i], mid[w_i]))
    ends = []
    for j, i in enumerate((t_i, q_i)):
        begin = mid[i] - 0.01 * syn.torch_field(shape3, seed + 200 + j, dev)
        end = mid[i] + 0.01 * syn.torch_field(shape3, seed + 300 + j, dev)
        ends += [begin, end]
        fields.append(ops.storage(end, begin, B.DT))
    return fields, mid + ends


def config4(ctx, n=3072, n_tile=6, band=768, f=8):
    """The budget remap at X-SHiELD size.  One (tile, band) shard = 26 output fields from 26 arrays + delp of
    (79, 768, 3072) = 20 GB: a rank holds one synthetic shard in HBM and runs it once for each of its shards
    (the cost of the walk depends on the pressures' statistics, not on which shard it is)."""
    units = [(t, b) for t in range(n_tile) for b in range(n // band)]
    mine = units[ctx.rank::ctx.world]
    dev, seed = ctx.device, 9000 + 50 * ctx.rank
    shape3 = (1, NZ, band, n)
    delp = syn.torch_delp(1, band, n, seed, dev)
    area = syn.torch_area((1, band, n), seed + 1, dev)
    fields, keep = _budget_fields(shape3, seed + 2, dev)
    nout = len(fields)
    hc = (band // f) * (n // f)
    # the kernels write straight into this rank's slot of the exchange buffer (slots padded to the longest rank of a
    # ragged deal): nothing is packed or copied.  Two exchanges are timed: ONE in-place ncclAllGather after the kernels,
    # and the fused form - the buffer in peer-mapped (symmetric) memory, the kernel storing every coarse value into
    # all peers' copies over NVLink while it computes (cg_pressure_level_average_replicated), then one barrier
    per_unit = (nout + 1) * NZ * hc
    longest = len(units[0::ctx.world]) * per_unit
    cshape = (1, NZ, band // f, n // f)
    k_ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]

    def run(fused):
        handle, offs, mc = None, None, None
        try:
            if fused:
                import torch.distributed._symmetric_memory as symm
                gathered = symm.empty(ctx.world * longest, dtype=torch.float32, device=dev)
                gathered.zero_()
                handle = symm.rendezvous(gathered, dist.group.WORLD.group_name)
                ptrs = list(handle.buffer_ptrs)
                if fused == "peers":
                    offs = [int(ptrs[r]) - int(ptrs[ctx.rank]) for r in range(ctx.world) if r != ctx.rank]
                else:
                    mc = int(getattr(handle, "multicast_ptr", 0) or 0)
                    if not mc:
                        raise RuntimeError("no multicast mapping (NVLS) for the symmetric buffer")
                    mc -= int(gathered.data_ptr())
            else:
                gathered = torch.zeros(ctx.world * longest, device=dev)
            err = None
        except Exception as e:
            if not fused:
                raise
            err = f"{type(e).__name__}: {e}"[:200]
        if fused and ctx.max_over_ranks([0.0 if err is None else 1.0])[0] != 0.0:        # every rank takes the same way out
            return {"unavailable": err or "another rank could not map the symmetric buffer"}
        slot = gathered[ctx.rank * longest:(ctx.rank + 1) * longest]
        unit_outs = [[slot[u * per_unit + v * NZ * hc:u * per_unit + (v + 1) * NZ * hc].view(cshape) for v in range(nout + 1)]
                     for u in range(len(mine))]

        comm = torch.cuda.Stream(dev) if fused == "broadcast" else None
        done = [torch.cuda.Event() for _ in unit_outs]

        def step():
            k_ev[0].record()
            cur = torch.cuda.current_stream(dev)
            if comm is not None:
                handle.barrier()                    # nobody still reads the previous step's values from my copy
            for u, o in enumerate(unit_outs):
                ops.pressure_level_average(fields, delp, area, f, outs=o[:nout], delp_coarse_out=o[nout], replica_offsets=offs,
                                           multicast_offset=mc if comm is None else None)
                if comm is not None:                # this unit's values to every GPU, under the next unit's kernels
                    done[u].record(cur)
                    comm.wait_event(done[u])
                    with torch.cuda.stream(comm):
                        ops.multicast_store(slot[u * per_unit:(u + 1) * per_unit], mc)
            k_ev[1].record()
            if comm is not None:
                cur.wait_stream(comm)
            if handle is not None:
                handle.barrier()
            elif ctx.world > 1:
                dist.all_gather_into_tensor(gathered, slot)
            k_ev[2].record()

        ms = ctx.timed(step, steps=min(ctx.steps, 3))
        torch.cuda.synchronize()
        kernel_ms, x_ms = ctx.max_over_ranks([k_ev[0].elapsed_time(k_ev[1]), k_ev[1].elapsed_time(k_ev[2])])
        # the slots of the other ranks must hold what they computed: a checksum of every slot agrees on every rank
        sums = [float(gathered[r * longest:(r + 1) * longest].double().sum()) for r in range(ctx.world)]
        agree = all(ctx.max_over_ranks([x])[0] == x and -ctx.max_over_ranks([-x])[0] == x for x in sums) and all(x != 0.0 for x in sums)
        del unit_outs, slot, gathered
        _free()
        return {"ms_per_step": ms, "kernel_ms": kernel_ms, "exchange_ms": x_ms, "slot_checksums_agree_on_all_ranks": bool(agree)}

    def run_pipelined():
        """NCCL again, one in-place all-gather per unit on a side stream: the exchange of unit u runs under the kernels of
        unit u + 1, only the last one is exposed.  Buffer laid out unit by unit, [unit][rank][values]."""
        nu = len(units[0::ctx.world])
        gathered = torch.zeros(nu * ctx.world * per_unit, device=dev)
        block = lambda u: gathered[u * ctx.world * per_unit:(u + 1) * ctx.world * per_unit]
        mine_slot = lambda u: block(u)[ctx.rank * per_unit:(ctx.rank + 1) * per_unit]
        unit_outs = [[mine_slot(u)[v * NZ * hc:(v + 1) * NZ * hc].view(cshape) for v in range(nout + 1)] for u in range(len(mine))]
        comm = torch.cuda.Stream(dev)
        done = [torch.cuda.Event() for _ in range(nu)]

        def step():
            k_ev[0].record()
            cur = torch.cuda.current_stream(dev)
            for u in range(nu):
                if u < len(mine):
                    o = unit_outs[u]
                    ops.pressure_level_average(fields, delp, area, f, outs=o[:nout], delp_coarse_out=o[nout])
                done[u].record(cur)
                comm.wait_event(done[u])
                with torch.cuda.stream(comm):
                    dist.all_gather_into_tensor(block(u), mine_slot(u))
            k_ev[1].record()
            cur.wait_stream(comm)
            k_ev[2].record()

        ms = ctx.timed(step, steps=min(ctx.steps, 3))
        torch.cuda.synchronize()
        kernel_ms, x_ms = ctx.max_over_ranks([k_ev[0].elapsed_time(k_ev[1]), k_ev[1].elapsed_time(k_ev[2])])
        sums = [float(block(u)[r * per_unit:(r + 1) * per_unit].double().sum()) for u in range(nu) for r in range(ctx.world)]
        agree = all(ctx.max_over_ranks([x])[0] == x and -ctx.max_over_ranks([-x])[0] == x for x in sums) and all(x != 0.0 for x in sums)
        del unit_outs, gathered
        _free()
        return {"ms_per_step": ms, "kernel_ms": kernel_ms, "exchange_ms": x_ms, "slot_checksums_agree_on_all_ranks": bool(agree),
                "note": "kernel_ms = the compute stream's kernels with the all-gathers of earlier units running beside them; "
                        "exchange_ms = what is left of the last unit's all-gather after the last kernel"}

    paths = OrderedDict([("nccl_all_gather_in_place", run(False))])
    if ctx.world > 1:
        if len(units) % ctx.world == 0:
            paths["nccl_all_gather_per_unit_overlapped"] = run_pipelined()
        if _FUSED_EXCHANGE:
            paths["nvls_broadcast_kernel_per_unit_overlapped"] = run("broadcast")
            paths["fused_multicast_stores"] = run("multicast")
            if ctx.world == 2 or "ms_per_step" not in paths["fused_multicast_stores"]:
                paths["fused_peer_stores"] = run("peers")
    good = {k: v for k, v in paths.items() if "ms_per_step" in v and (ctx.world == 1 or v["slot_checksums_agree_on_all_ranks"])}
    which = min(good, key=lambda k: good[k]["ms_per_step"])
    best = good[which]
    ms, kernel_ms, gather_ms = best["ms_per_step"], best["kernel_ms"], best["exchange_ms"]
    fine3 = NZ * band * n
    cells = len(units) * (nout + 1) * fine3
    per_shard = 4 * ((len(keep) + 1) * fine3 + band * n + (nout + 1) * NZ * hc)
    rec = {"workload": "configs[4]: C3072->C384 (f=8) pressure-level coarse-graining of a fine_res_budget timestep (26 remapped "
                       "fields), cut into 24 shards of (tile, 768-row band) dealt over the ranks; coarse outputs all-gathered",
           "scaling": "strong", "ms_per_step": ms, "value": cells / (ms * 1e-3), "unit": "cells/s",
           "shards_per_rank": [len(units[r::ctx.world]) for r in range(ctx.world)],
           "kernel_ms": kernel_ms, "collective": None if ctx.world == 1 else
           {"op": _EXCHANGE_OPS[which], "bytes_out_per_rank": int(ctx.world * longest * 4), "ms": gather_ms},
           "exchange_paths": paths,
           "data": "synthetic; each rank keeps ONE shard-sized input set (20 GB) and runs it once per shard it owns",
           "roofline": ctx.roofline(per_shard * len(units[0::ctx.world]), kernel_ms, "cg::pairk::plevel_pair_kernel")}
    del fields, keep, delp, area
    _free()
    return rec


# CG_BENCH_EXCHANGE=nccl: time the NCCL forms of the exchange only (no symmetric memory, no NVLS stores)
_FUSED_EXCHANGE = os.environ.get("CG_BENCH_EXCHANGE", "all") != "nccl"
_EXCHANGE_OPS = {
    "nccl_all_gather_in_place": "ncclAllGather in place (dist.all_gather_into_tensor, one per step; the kernels write into the send slot)",
    "nccl_all_gather_per_unit_overlapped": "ncclAllGather in place, one per (tile, band) unit on a side stream under the next unit's kernels "
                                           "(dist.all_gather_into_tensor; the kernels write into the send slots)",
    "nvls_broadcast_kernel_per_unit_overlapped": "own NVLS all-gather, one per (tile, band) unit on a side stream under the next unit's kernels: the "
                                                 "unit's slice of the rank's slot is stored to the multicast mapping of the symmetric output buffer "
                                                 "(cg_multicast_store: multimem.st, 512 contiguous bytes per warp instruction), barriers at both ends",
    "nvls_broadcast_kernel": "own NVLS all-gather: after the kernels every rank stores its slot of the symmetric output buffer to the buffer's "
                             "multicast mapping (cg_multicast_store: multimem.st, 512 contiguous bytes per warp instruction, the NVSwitch "
                             "writes all GPUs' copies), device-side barriers before and after",
    "fused_multicast_stores": "fused: the kernel stores every coarse value ONCE to the NVLS multicast mapping of the symmetric output "
                              "buffer (multimem.st, the NVSwitch writes all GPUs' copies; cg_pressure_level_average_replicated), "
                              "then one device-side barrier",
    "fused_peer_stores": "fused: the kernel stores every coarse value into every peer's copy of the symmetric output buffer "
                         "(one NVLink store per peer; cg_pressure_level_average_replicated), then one device-side barrier",
}


# ------------------------------------------------------------------------------------------- strong scaling
def strong_config2(ctx, inp, n=384, n_tile=6, f=8):
    """configs[2]'s timestep (rank 0's, rebuilt from its seed on every rank) cut over the ranks; the pressure-level
    outputs + delp are all-gathered every step.  Checked once against the unsharded run, bit for bit."""
    dev = ctx.device
    shards = S.plan(1, ctx.world, n_tile=n_tile, ny=n, factor=f)
    names3 = list(inp["middle3d"])

    def cut(t, s):
        r1 = n if s.row1 is None else s.row1
        return t[s.tile0:s.tile1, ..., s.row0:r1, :].contiguous()

    local_in = []
    for s in shards[ctx.rank]:
        mid = {k: cut(v, s) for k, v in inp["middle3d"].items()}
        fields = [mid[k] for k in names3]
        for a, b in B.SECOND_MOMENTS:
            fields.append(ops.product(mid[a], mid[b]))
        ends = []
        for k in B.STORAGE_TERMS:
            e, bgn = cut(inp["end"][k], s), cut(inp["begin"][k], s)
            ends += [e, bgn]
            fields.append(ops.storage(e, bgn, B.DT))
        local_in.append((fields, cut(inp["delp"], s), cut(inp["area"], s), mid, ends))
    out_names = names3 + [f"{a}_{b}" for a, b in B.SECOND_MOMENTS] + [f"{k}_storage" for k in B.STORAGE_TERMS] + ["delp"]
    shapes = {k: ((NZ,), n, n, f) for k in out_names}
    k_ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    result = {}

    def make_step(bufs, fused):
        # the kernels write straight into this rank's slot of the exchange buffer (sharding.CoarseBuffers); fused: they
        # store every value into every GPU's copy themselves - "multicast": one multimem.st to the NVLS mapping, the
        # NVSwitch replicates; "peers": one NVLink store per peer - and only a barrier follows
        shard_outs = [[bufs.local[k][j] for k in out_names] for j in range(len(local_in))]
        offs = bufs.peer_offsets() if fused == "peers" else None
        mc = bufs.multicast_offset() if fused == "multicast" else None
        if fused == "multicast" and mc is None:
            raise RuntimeError("no multicast mapping (NVLS) for the symmetric buffer")

        def step():
            k_ev[0].record()
            for (fields, delp, area, _, _), o in zip(local_in, shard_outs):
                ops.pressure_level_average(fields, delp, area, f, outs=o[:-1], delp_coarse_out=o[-1], replica_offsets=offs,
                                           multicast_offset=mc)
            k_ev[1].record()
            result["g"] = bufs.multicast_gather() if fused == "broadcast" else bufs.barrier() if fused else bufs.all_gather()
            k_ev[2].record()
        return step

    # the unsharded run of the same timestep (every rank holds it): both exchange paths must reproduce it bit for bit
    fields = [inp["middle3d"][k] for k in names3]
    for a, b in B.SECOND_MOMENTS:
        fields.append(ops.product(inp["middle3d"][a], inp["middle3d"][b]))
    for k in B.STORAGE_TERMS:
        fields.append(ops.storage(inp["end"][k], inp["begin"][k], B.DT))
    ref, ref_dc = ops.pressure_level_average(fields, inp["delp"], inp["area"], f)
    del fields

    def run(fused):
        try:
            bufs = S.CoarseBuffers(shards, shapes, ctx.rank, ctx.world, dev, symmetric=bool(fused))
            step = make_step(bufs, fused)
            err = None
        except Exception as e:                       # no peer-mapped memory on this box: the NCCL path is the product
            if not fused:
                raise
            err = f"{type(e).__name__}: {e}"[:200]
        if fused and ctx.max_over_ranks([0.0 if err is None else 1.0])[0] != 0.0:        # every rank takes the same way out
            return {"unavailable": err or "another rank could not map the symmetric buffer"}
        step()
        torch.cuda.synchronize()
        ok = True
        for k, e in zip(out_names, ref + [ref_dc]):
            got = S.assemble(result["g"][k], 1, n_tile, (NZ,), n, n, f)[0]
            ok = ok and bool(torch.equal(torch.nan_to_num(got), torch.nan_to_num(e)))
        ok = bool(ctx.max_over_ranks([0.0 if ok else 1.0])[0] == 0.0)
        ms = ctx.timed(step)
        torch.cuda.synchronize()
        kernel_ms, x_ms = ctx.max_over_ranks([k_ev[0].elapsed_time(k_ev[1]), k_ev[1].elapsed_time(k_ev[2])])
        result.clear()
        return {"ms_per_step": ms, "kernel_ms": kernel_ms, "exchange_ms": x_ms, "bit_identical_to_unsharded": ok}

    paths = OrderedDict([("nccl_all_gather_in_place", run(False))])
    if ctx.world > 1 and _FUSED_EXCHANGE:
        paths["nvls_broadcast_kernel"] = run("broadcast")
        paths["fused_multicast_stores"] = run("multicast")
        if ctx.world == 2 or "ms_per_step" not in paths["fused_multicast_stores"]:
            paths["fused_peer_stores"] = run("peers")        # (7 small NVLink stores per value at N = 8: slower than NCCL)
    del ref, ref_dc
    good = {k: v for k, v in paths.items() if "ms_per_step" in v and v["bit_identical_to_unsharded"]}
    which = min(good, key=lambda k: good[k]["ms_per_step"])
    best = good[which]
    ms, kernel_ms, gather_ms, ok = best["ms_per_step"], best["kernel_ms"], best["exchange_ms"], best["bit_identical_to_unsharded"]
    fine3 = n_tile * NZ * n * n
    cells = len(out_names) * fine3
    rows = max(sum((s.tile1 - s.tile0) * ((n if s.row1 is None else s.row1) - s.row0) for s in r) for r in shards)    # tile-rows of the busiest rank
    per_rank = 4 * (27 * NZ * rows * n + rows * n + len(out_names) * NZ * (rows // f) * (n // f))
    nbytes = len(out_names) * n_tile * NZ * (n // f) ** 2 * 4
    rec = {"workload": "configs[2]'s one C384 timestep cut by sharding.plan over the ranks (whole tiles, else bands of 8 rows "
                       "dealt evenly), 26 pressure-level fields + delp all-gathered inside the timed region",
           "scaling": "strong", "ms_per_step": ms, "value": cells / (ms * 1e-3), "unit": "cells/s",
           "bit_identical_to_unsharded": ok, "shards_per_rank": [len(r) for r in shards],
           "kernel_ms": kernel_ms,
           "collective": {"op": "none (one rank)" if ctx.world == 1 else _EXCHANGE_OPS[which],
                          "bytes_out_per_rank": int(nbytes), "ms": gather_ms},
           "exchange_paths": paths,
           "roofline": ctx.roofline(per_rank, kernel_ms, "cg::pairk::plevel_pair_kernel on the busiest rank's shards")}
    del local_in, result
    _free()
    return rec


# ------------------------------------------------------------------------------------------- H2D ceiling
def h2d_ceiling(ctx, host, chunk=1):
    """The end-to-end leg's host -> device copies alone: the same pinned buffers, the same tile chunks, one
    cudaMemcpyAsync per buffer and chunk on one stream, every rank at once.  GB/s summed over the ranks."""
    from fv3net_b200 import staging

    flat = staging._flatten(host)
    n = flat[("delp", None)].shape[0]
    dev = ctx.device
    bufs = {k: torch.empty((chunk,) + tuple(v.shape[1:]), dtype=v.dtype, device=dev) for k, v in flat.items()}
    stream = torch.cuda.Stream(dev)
    nbytes = sum(v.numel() * v.element_size() for v in flat.values())

    def step():
        with torch.cuda.stream(stream):
            for i in range(0, n, chunk):
                for k, v in flat.items():
                    bufs[k].copy_(v[i:i + chunk], non_blocking=True)
        torch.cuda.current_stream(dev).wait_stream(stream)

    ms = ctx.timed(step, steps=2, warmup=1)
    del bufs
    return {"gbs_all_ranks": ctx.world * nbytes / (ms * 1e-3) / 1e9, "ms": ms, "bytes_per_rank": int(nbytes)}
