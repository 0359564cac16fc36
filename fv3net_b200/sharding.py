"""Sharding of the coarse-graining path over the GPUs of one box (one process per GPU).

Every coarse cell depends only on its own f x f block of fine columns, so the path
partitions with no data-path exchange.  Work is split the way the reference
splits it - by time first (fine_res_budget/budget/pipeline.py:46-53 hands one timestep to
one worker) - then by tile, then by bands of rows that are multiples of the coarsening
factor.  The only collective is an optional gather of the (64x smaller) coarse outputs onto
one rank for a consumer that wants the global coarse field; ranks that write their own
region, as the reference's workers do, never communicate.
"""
from collections import OrderedDict, namedtuple

import torch
import torch.distributed as dist

Shard = namedtuple("Shard", "time tile0 tile1 row0 row1")


def split_range(n, parts, index, multiple=1):
    """[start, stop) of part `index` of `n` items in `parts` contiguous, balanced pieces whose
    boundaries are multiples of `multiple` (the remainder goes to the leading pieces)."""
    if n % multiple:
        raise ValueError(f"{n} is not a multiple of {multiple}")
    units = n // multiple
    base, extra = divmod(units, parts)
    start = index * base + min(index, extra)
    stop = start + base + (1 if index < extra else 0)
    return start * multiple, stop * multiple


def plan(n_time, world, n_tile=6, ny=None, factor=1):
    """Shards for every rank: a list (one entry per rank) of lists of `Shard`.

    n_time >= world: whole timesteps, dealt in contiguous runs.  Fewer timesteps than ranks:
    the ranks are divided among the timesteps and each group splits its timestep by whole tiles
    when they divide evenly, else by row bands that are multiples of `factor` (coarsen_surface_c3072.py:29-76
    cuts a tile into sub-tiles the same way)."""
    if world < 1 or n_time < 1:
        raise ValueError("need at least one rank and one timestep")
    shards = [[] for _ in range(world)]
    if n_time >= world:
        for r in range(world):
            t0, t1 = split_range(n_time, world, r)
            shards[r] = [Shard(t, 0, n_tile, 0, ny) for t in range(t0, t1)]
        return shards
    for t in range(n_time):
        r0, r1 = split_range(world, n_time, t)                 # ranks working on timestep t
        group = r1 - r0
        if group <= n_tile and n_tile % group == 0:
            for g in range(group):
                a, b = split_range(n_tile, group, g)
                shards[r0 + g].append(Shard(t, a, b, 0, ny))
            continue
        # Otherwise the timestep's n_tile * (ny / factor) row bands of `factor` rows, taken tile after tile, are
        # dealt evenly: a rank owns a contiguous run of bands, which may continue into the next tile (one Shard
        # per tile it touches).  Every rank then has the same work to within one band.
        if ny is None:
            raise ValueError("ny is needed to split tiles into row bands")
        if ny % factor:
            raise ValueError(f"{ny} is not a multiple of {factor}")
        per_tile = ny // factor
        if n_tile * per_tile < group:
            raise ValueError(f"{n_tile} tiles of {ny} rows give fewer than {group} bands of {factor} rows")
        for g in range(group):
            u0, u1 = split_range(n_tile * per_tile, group, g)
            while u0 < u1:
                tile, first = divmod(u0, per_tile)
                last = min(per_tile, first + (u1 - u0))
                shards[r0 + g].append(Shard(t, tile, tile + 1, first * factor, last * factor))
                u0 += last - first
    return shards


def take(shard, array, time_axis=0, tile_axis=1, row_axis=-2):
    """The part of a (time, tile, ..., y, x) array a shard owns (a view).  `time_axis=None`:
    the array has no time dim (grid data such as area: (tile, y, x), tile_axis=0)."""
    if time_axis is not None:
        idx = [slice(None)] * array.ndim
        idx[time_axis] = shard.time
        array = array[tuple(idx)]
        tile_axis = tile_axis - 1 if tile_axis > time_axis else tile_axis
    idx = [slice(None)] * array.ndim
    idx[tile_axis] = slice(shard.tile0, shard.tile1)
    if shard.row1 is not None:
        idx[row_axis] = slice(shard.row0, shard.row1)
    return array[tuple(idx)]


def coarse_shape(shard, lead, ny, nx, factor):
    """Shape of a shard's coarse output: (tiles, *lead, rows / f, nx / f)."""
    rows = (shard.row1 if shard.row1 is not None else ny) - shard.row0
    return (shard.tile1 - shard.tile0,) + tuple(lead) + (rows // factor, nx // factor)


def gather_coarse(local, shards, shapes, dst=0, group=None):
    """One collective per step: every rank's coarse outputs onto rank `dst` (`dst=None`: onto every rank,
    one all-gather).

    local:  {name: [tensor per shard of this rank]} in the order of `shards[rank]`
    shapes: {name: (lead, ny, nx, factor)} describing every output (same on all ranks)
    Returns on `dst` {name: {shard: tensor}} for every shard of every rank, elsewhere None.
    The buffers are padded to the largest rank's size so a plain gather (NCCL or gloo) works
    for ragged plans; the plan is known everywhere, so no sizes are exchanged."""
    if dist.is_available() and dist.is_initialized():
        rank, world = dist.get_rank(group), dist.get_world_size(group)
    else:
        rank, world = 0, 1                                   # one process: the plan has one rank
    names = list(shapes)

    def rank_numel(r):
        total = 0
        for name in names:
            lead, ny, nx, f = shapes[name]
            for s in shards[r]:
                n = 1
                for d in coarse_shape(s, lead, ny, nx, f):
                    n *= d
                total += n
        return total

    longest = max(rank_numel(r) for r in range(world))
    first = next(iter(local.values()))[0] if local and shards[rank] else None
    device = first.device if first is not None else torch.device("cpu")
    dtype = first.dtype if first is not None else torch.float32
    buf = torch.zeros(max(longest, 1), dtype=dtype, device=device)
    pos = 0
    for name in names:
        for t in local.get(name, []):
            flat = t.reshape(-1)
            buf[pos:pos + flat.numel()] = flat
            pos += flat.numel()
    if pos != rank_numel(rank):
        raise ValueError(f"rank {rank} packed {pos} values, the plan says {rank_numel(rank)}")
    if world > 1 and dst is None:
        # every rank gets every shard: ONE ncclAllGather (ring / NVLS over NVSwitch), no per-peer send / recv
        whole = torch.empty(world * buf.numel(), dtype=dtype, device=device)
        dist.all_gather_into_tensor(whole, buf, group=group)
        parts = list(whole.split(buf.numel()))
    elif world > 1:
        parts = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
        dist.gather(buf, parts, dst=dst, group=group)
    else:
        parts = [buf]
    if dst is not None and rank != dst:
        return None
    out = OrderedDict((name, OrderedDict()) for name in names)
    for r in range(world):
        pos = 0
        for name in names:
            lead, ny, nx, f = shapes[name]
            for s in shards[r]:
                shape = coarse_shape(s, lead, ny, nx, f)
                n = 1
                for d in shape:
                    n *= d
                out[name][s] = parts[r][pos:pos + n].reshape(shape)
                pos += n
    return out


class CoarseBuffers:
    """The coarse outputs of one step, of every rank, in ONE buffer laid out for an in-place all-gather.

    Slot r = [r * longest, (r + 1) * longest) holds rank r's outputs name by name, shard by shard (the order of
    `gather_coarse`; slots are padded to the longest rank of a ragged plan).  `local[name]` are views into this
    rank's own slot: hand them to the kernels as their output arrays (`ops.pressure_level_average(..., outs=...)`),
    so nothing is packed or copied before the collective, and `all_gather()` is one `ncclAllGather` whose send
    buffer is the rank's slot of the receive buffer.  The plan is known everywhere, so no sizes are exchanged.

    `symmetric=True`: the buffer lives in peer-mapped (symmetric) memory; the kernels can then store each output
    into every peer's copy themselves (`peer_offsets()` -> `ops.pressure_level_average(replica_offsets=...)`,
    NVLink stores from inside the compute kernel) and the collective reduces to `barrier()`."""

    def __init__(self, shards, shapes, rank, world, device, dtype=torch.float32, whole=None, symmetric=False,
                 group=None):
        self.shards, self.shapes, self.rank, self.world = shards, shapes, rank, world
        self.handle, self.group = None, group
        self.names = list(shapes)

        def numel(shape):
            n = 1
            for d in shape:
                n *= d
            return n

        self._layout = []                                    # per rank: [(name, shard, offset in the slot, shape)]
        sizes = []
        for r in range(world):
            pos, items = 0, []
            for name in self.names:
                lead, ny, nx, f = shapes[name]
                for s in shards[r]:
                    shape = coarse_shape(s, lead, ny, nx, f)
                    items.append((name, s, pos, shape))
                    pos += numel(shape)
            self._layout.append(items)
            sizes.append(pos)
        # slots start on 16-byte boundaries whatever the dtype (vector stores, NCCL's preferred alignment)
        q = max(1, 16 // torch.empty((), dtype=dtype).element_size())
        self.longest = max(1, (max(sizes) + q - 1) // q * q)
        if symmetric and world > 1:
            # peer-mapped memory (CUDA VMM handles exchanged over the process group): every rank's buffer has the
            # same layout and is addressable from every GPU of the box through NVLink
            import torch.distributed._symmetric_memory as symm

            whole = symm.empty(world * self.longest, dtype=dtype, device=device)
            whole.zero_()
            self.handle = symm.rendezvous(whole, (group or dist.group.WORLD).group_name)
        elif whole is None:
            whole = torch.zeros(world * self.longest, dtype=dtype, device=device)
        elif whole.numel() != world * self.longest or whole.dtype != dtype:
            raise ValueError(f"the buffer must hold {world * self.longest} values of {dtype}")
        self.whole = whole
        self.local = OrderedDict((name, []) for name in self.names)
        for name, s, pos, shape in self._layout[rank]:
            self.local[name].append(self._view(rank, pos, shape))

    def _view(self, r, pos, shape):
        n = 1
        for d in shape:
            n *= d
        return self.whole[r * self.longest + pos:r * self.longest + pos + n].view(shape)

    def slot(self, r=None):
        r = self.rank if r is None else r
        return self.whole[r * self.longest:(r + 1) * self.longest]

    def peer_offsets(self):
        """Byte offsets from this rank's buffer to every other rank's copy (symmetric=True), for the kernels that
        store their outputs into the peers themselves (ops.pressure_level_average(replica_offsets=...))."""
        if self.handle is None:
            return []
        ptrs = list(self.handle.buffer_ptrs)
        return [int(ptrs[r]) - int(ptrs[self.rank]) for r in range(self.world) if r != self.rank]

    def multicast_offset(self):
        """Byte offset from this rank's buffer to the NVLS multicast mapping of the symmetric buffer (a store there is
        written into every rank's copy by the NVSwitch), or None where the box has no multicast support."""
        mc = int(getattr(self.handle, "multicast_ptr", 0) or 0) if self.handle is not None else 0
        return mc - int(self.whole.data_ptr()) if mc else None

    def barrier(self):
        """After kernels that stored into the peers: every rank's stores have landed everywhere when this returns on
        the stream (a device-side barrier over the symmetric memory's signal pads, no data moves).  Returns the views
        of `all_gather`.  Call it once more before the buffers are overwritten if a peer may still be reading."""
        if self.handle is not None:
            self.handle.barrier()
        return self.views()

    def multicast_gather(self):
        """The all-gather as ONE small kernel per rank (symmetric=True on a box with NVLS): this rank's slot is stored
        to the multicast mapping of the buffer (ops.multicast_store: 512 contiguous bytes per warp instruction, the
        NVSwitch writes every GPU's copy), then the device-side barrier.  Returns the views of `all_gather`."""
        from . import ops

        mc = self.multicast_offset()
        if mc is None:
            raise RuntimeError("no multicast mapping (NVLS) for the symmetric buffer")
        self.handle.barrier()                  # nobody is still reading the previous step's values from my copy
        ops.multicast_store(self.slot(), mc)
        self.handle.barrier()
        return self.views()

    def all_gather(self, group=None):
        """In place: afterwards every rank's `whole` holds every slot.  Returns {name: {shard: view}}."""
        if self.world > 1:
            dist.all_gather_into_tensor(self.whole, self.slot(), group=group)
        return self.views()

    def views(self):
        out = OrderedDict((name, OrderedDict()) for name in self.names)
        for r in range(self.world):
            for name, s, pos, shape in self._layout[r]:
                out[name][s] = self._view(r, pos, shape)
        return out


def assemble(gathered, n_time, n_tile, lead, ny, nx, factor):
    """{shard: tensor} of one output -> the global (time, tile, *lead, ny/f, nx/f) tensor."""
    first = next(iter(gathered.values()))
    full = torch.empty((n_time, n_tile) + tuple(lead) + (ny // factor, nx // factor),
                       dtype=first.dtype, device=first.device)
    for s, t in gathered.items():
        r0 = s.row0 // factor
        r1 = (s.row1 if s.row1 is not None else ny) // factor
        full[s.time, s.tile0:s.tile1, ..., r0:r1, :] = t
    return full
