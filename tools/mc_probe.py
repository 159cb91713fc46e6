"""Does torch symmetric memory expose an NVSwitch multicast mapping on this box?  torchrun --nproc-per-node N tools/mc_probe.py"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
buf = symm_mem.empty((1024,), dtype=torch.float32, device=torch.device("cuda", local))
hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
mc = getattr(hdl, "multicast_ptr", None)
print(rank, "multicast_ptr", hex(int(mc)) if mc else mc, "buffer_ptrs", [hex(int(p)) for p in hdl.buffer_ptrs][:2], flush=True)
dist.barrier()
dist.destroy_process_group()
