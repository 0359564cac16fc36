"""Device staging under the fine_res_budget step: host arrays in, host arrays out.

The reference hands one timestep to a worker as host (numpy / zarr-chunk) arrays
(workflows/fine_res_budget/budget/pipeline.py:35-53, data.py:62-131) and gets coarse host arrays
back.  On a B200 the kernels need 7 ms for a C384 timestep while its 7.6 GB of inputs need
~140 ms on the PCIe link, so what matters end to end is that the link never waits:

  * the timestep is cut into chunks along the leading (tile) dimension - a coarse cell needs only
    its own fine columns, so chunks are independent;
  * two device buffer sets; a copy stream fills set b with chunk i + 1 (pinned host -> device,
    back to back, no synchronisation with the host) while the compute stream runs the step on set
    1 - b with chunk i and sends its coarse outputs (1/64 of the input bytes) device -> host;
  * events order the two streams: `filled[b]` (copy -> compute), `drained[b]` (compute -> copy).

The host thread only enqueues; it blocks once, at the end (`synchronize=True`), so several
timesteps can be queued behind each other.
"""
from collections import OrderedDict

import torch

from . import budget as B

_GROUPS = ("middle3d", "middle2d", "begin", "end")
_SINGLES = ("delp", "area", "exposed_area_fine")


def _flatten(inputs):
    flat = OrderedDict()
    for g in _GROUPS:
        for k, v in inputs.get(g, {}).items():
            flat[(g, k)] = v
    for k in _SINGLES:
        flat[(k, None)] = inputs[k]
    return flat


def _unflatten(flat):
    out = {g: OrderedDict() for g in _GROUPS}
    for (g, k), v in flat.items():
        if k is None:
            out[g] = v
        else:
            out[g][k] = v
    return out


def bind_host_thread(device=None):
    """Run the calling process on the CPUs next to `device` (NVML's affinity mask) so that the pinned buffers it
    allocates afterwards, and the copies it enqueues, stay on the GPU's own NUMA node and PCIe root.  With one
    process per GPU on an 8-GPU box the host -> device streams otherwise meet in one socket's memory.  Returns the
    CPU set, or None when NVML or the affinity call is not available (nothing is changed then)."""
    import os
    import warnings

    try:
        import pynvml

        index = torch.cuda.current_device() if device is None else torch.device(device).index
        pynvml.nvmlInit()
        # NVML enumerates physical devices, CUDA the visible ones: match on the UUID, else on the full PCI id
        # (domain:bus:device), never on the bus number alone
        props = torch.cuda.get_device_properties(index)
        uuid = str(getattr(props, "uuid", "") or "").lower().replace("gpu-", "")
        pci = tuple(getattr(props, k, None) for k in ("pci_domain_id", "pci_bus_id", "pci_device_id"))
        handle = None
        for i in range(pynvml.nvmlDeviceGetCount()):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            u = pynvml.nvmlDeviceGetUUID(h)
            u = (u.decode() if isinstance(u, bytes) else str(u)).lower().replace("gpu-", "")
            info = pynvml.nvmlDeviceGetPciInfo(h)
            if (uuid and u == uuid) or (None not in pci and (info.domain, info.bus, info.device) == pci):
                handle = h
                break
        if handle is None:
            warnings.warn(f"bind_host_thread: cuda:{index} not found among NVML's devices; affinity left unchanged")
            return None
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            warnings.warn("bind_host_thread: NVML's CPU set for the device does not meet this process's; affinity left unchanged")
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception as e:                  # no NVML, no permission, not Linux: leave the process where it is, and say so
        warnings.warn(f"bind_host_thread: skipped ({type(e).__name__}: {e})")
        return None


def pin(inputs):
    """The same nested dict with every array in pinned host memory (numpy or CPU tensors in)."""
    flat = _flatten(inputs)
    pinned = OrderedDict()
    for key, v in flat.items():
        t = v if isinstance(v, torch.Tensor) else torch.from_numpy(v)
        if t.is_cuda:
            raise ValueError("pin() expects host arrays")
        p = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        p.copy_(t)
        pinned[key] = p
    return _unflatten(pinned)


class BudgetPipeline:
    """Double-buffered host -> device -> host pipeline around `budget.recoarsen_budget_inputs`."""

    def __init__(self, device=None, chunk=1, **step_kwargs):
        if not torch.cuda.is_available():
            raise RuntimeError("fv3net_b200.staging needs a CUDA device; there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.chunk = int(chunk)                      # leading-dim entries (tiles) per chunk
        self.step_kwargs = step_kwargs
        self.copy_stream = torch.cuda.Stream(self.device)
        self.compute_stream = torch.cuda.Stream(self.device)
        self._buffers = None
        self._signature = None
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _ensure_buffers(self, flat):
        sig = tuple((k, tuple(v.shape[1:]), v.dtype) for k, v in flat.items())
        if sig == self._signature:
            return
        self._signature = sig
        self._buffers = []
        for _ in range(2):
            bufs = OrderedDict((k, torch.empty((self.chunk,) + tuple(v.shape[1:]), dtype=v.dtype, device=self.device))
                               for k, v in flat.items())
            self._buffers.append(bufs)
        self._filled = [torch.cuda.Event() for _ in range(2)]
        self._drained = [torch.cuda.Event() for _ in range(2)]

    def run(self, host_inputs, out_host=None, synchronize=True):
        """One timestep.  `host_inputs`: the keyword arguments of `recoarsen_budget_inputs` as
        HOST tensors (pinned for full link speed, see `pin`), leading dim = tile.  Returns an
        OrderedDict of pinned host tensors keyed like the reference's output Dataset
        (`out_host` from a previous call is reused)."""
        flat = _flatten(host_inputs)
        n = flat[("delp", None)].shape[0]
        for k, v in flat.items():
            if v.is_cuda:
                raise ValueError(f"{k}: staging takes host arrays (use budget.recoarsen_budget_inputs on the device)")
            if v.shape[0] != n:
                raise ValueError(f"{k}: leading (tile) dimension {v.shape[0]} != {n}")
        if n % self.chunk:
            raise ValueError(f"{n} tiles do not split into chunks of {self.chunk}")
        self._ensure_buffers(flat)
        nchunks = n // self.chunk
        caller = torch.cuda.current_stream(self.device)
        self.copy_stream.wait_stream(caller)
        self.compute_stream.wait_stream(caller)
        h2d = d2h = 0
        for i in range(nchunks):
            b = i & 1
            lo, hi = i * self.chunk, (i + 1) * self.chunk
            bufs = self._buffers[b]
            with torch.cuda.stream(self.copy_stream):
                if i >= 2:
                    self.copy_stream.wait_event(self._drained[b])      # set b is free again
                for k, v in flat.items():
                    bufs[k].copy_(v[lo:hi], non_blocking=True)
                    h2d += bufs[k].numel() * bufs[k].element_size()
                self._filled[b].record(self.copy_stream)
            with torch.cuda.stream(self.compute_stream):
                self.compute_stream.wait_event(self._filled[b])
                out = B.recoarsen_budget_inputs(**_unflatten(bufs), **self.step_kwargs)
                if out_host is None:
                    out_host = OrderedDict((k, torch.empty((n,) + tuple(v.shape[1:]), dtype=v.dtype, pin_memory=True))
                                           for k, v in out.items())
                for k, v in out.items():
                    out_host[k][lo:hi].copy_(v, non_blocking=True)
                    d2h += v.numel() * v.element_size()
                self._drained[b].record(self.compute_stream)
        caller.wait_stream(self.compute_stream)
        caller.wait_stream(self.copy_stream)
        self.h2d_bytes, self.d2h_bytes = h2d, d2h
        if synchronize:
            torch.cuda.current_stream(self.device).synchronize()
        return out_host
