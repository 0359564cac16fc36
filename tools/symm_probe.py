"""Is peer (symmetric) memory usable on this box?  torchrun --nproc-per-node N tools/symm_probe.py"""
import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
import torch.distributed._symmetric_memory as symm_mem
t = symm_mem.empty(1 << 20, dtype=torch.float32, device=f"cuda:{torch.cuda.current_device()}")
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
ptrs = list(hdl.buffer_ptrs)
mc = getattr(hdl, "multicast_ptr", None)
print(rank, "buffer_ptrs", [hex(p) for p in ptrs], "multicast", hex(mc) if mc else mc, "local", hex(t.data_ptr()), flush=True)
t.fill_(float(rank))
hdl.barrier()
# read a peer's buffer through its mapped address
peer = (rank + 1) % world
pt = hdl.get_buffer(peer, (16,), torch.float32)
print(rank, "peer", peer, "value", pt[:2].tolist(), flush=True)
hdl.barrier()
dist.destroy_process_group()
