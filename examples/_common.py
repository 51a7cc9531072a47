"""Shared bootstrap for the on-box example / benchmark scripts (one process per GPU via torchrun)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def init():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(lr)
        dev = torch.device("cuda", lr)
        if world > 1:
            dist.init_process_group("nccl", device_id=dev)
    else:
        dev = torch.device("cpu")
        if world > 1:
            dist.init_process_group("gloo")
    return rank, world, dev


def device_time_ms(fn, dev, world):
    """Run ``fn`` once bracketed by barrier + synchronize; CUDA-event time, max over ranks."""
    if dev.type != "cuda":
        import time
        t0 = time.perf_counter(); fn(); ms = (time.perf_counter() - t0) * 1e3
    else:
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def finish(rank, world, result, engine=None):
    if rank == 0:
        print(json.dumps(result))
    if engine is not None:
        engine.close()  # unmaps peer memory; needs the process group for its barriers
    if world > 1:
        dist.destroy_process_group()
