import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as sm
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
t = sm.empty(1 << 20, dtype=torch.uint8, device=torch.device("cuda", lr))
hdl = sm.rendezvous(t, dist.group.WORLD)
print(rank, "world", hdl.world_size, "mc_ptr", hex(hdl.multicast_ptr), "bufs", [hex(p) for p in hdl.buffer_ptrs][:3],
      "sig", len(hdl.signal_pad_ptrs), "has_mc", sm._SymmetricMemory.has_multicast_support(torch._C._autograd.DeviceType.CUDA if False else __import__('torch').device('cuda').type and sm.DeviceType.CUDA, lr), flush=True)
dist.barrier(); dist.destroy_process_group()
