"""Probe: does torch's symmetric memory give this box a multicast (NVLS) mapping and peer pointers?  (plumbing only)"""
import os
import torch
import torch.distributed as dist

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
    h = symm.rendezvous(t, dist.group.WORLD.group_name)
    t.fill_(rank + 1)
    h.barrier()
    print(f"rank {rank}: world {h.world_size} multicast_ptr {getattr(h, 'multicast_ptr', None)} "
          f"buffer_ptrs {[hex(p) for p in h.buffer_ptrs][:3]} signal_pads {len(h.signal_pad_ptrs)}", flush=True)
    other = h.get_buffer((rank + 1) % h.world_size, (16,), torch.float32)
    print(f"rank {rank}: peer value {float(other[0])}", flush=True)
    h.barrier()
except Exception as exc:
    print(f"rank {rank}: symmetric memory unavailable: {type(exc).__name__}: {exc}", flush=True)
dist.barrier()
dist.destroy_process_group()
