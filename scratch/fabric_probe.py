#!/usr/bin/env python
"""Host fabric probe for the N>1 end-to-end numbers: every rank copies pinned host memory to its own GPU
at the same time (plain cudaMemcpyAsync through torch, no library code) and reports its GB/s, alone and
with a concurrent device->host stream.  Run under torchrun with one rank per GPU:
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scratch/fabric_probe.py [--bind]"""
import os, sys, time, json
import torch, torch.distributed as dist

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
bound = None
if "--bind" in sys.argv:
    import pynvml
    pynvml.nvmlInit()
    pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(lr))
    bound = sorted(os.sched_getaffinity(0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
nbytes = 128 * 307200
host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(3)]
for h in host: h.fill_(7)
dev = [torch.empty(nbytes, dtype=torch.uint8, device="cuda") for _ in range(3)]
back_d = torch.empty(10_765_328, dtype=torch.uint8, device="cuda")
back_h = torch.empty(10_765_328, dtype=torch.uint8).pin_memory()
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

def run(iters, with_d2h):
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(iters):
        with torch.cuda.stream(s_up):
            dev[i % 3].copy_(host[i % 3], non_blocking=True)
        if with_d2h:
            with torch.cuda.stream(s_dn):
                back_h.copy_(back_d, non_blocking=True)
    torch.cuda.synchronize()
    return nbytes * iters / (time.perf_counter() - t0) / 1e9

out = {}
for tag, d2h in (("h2d", False), ("h2d+d2h", True)):
    run(5, d2h)
    g = torch.tensor([run(40, d2h)], device="cuda")
    if world > 1:
        allg = [torch.zeros_like(g) for _ in range(world)]
        dist.all_gather(allg, g)
        out[tag] = [round(float(x), 1) for x in allg]
    else:
        out[tag] = [round(float(g), 1)]
if world > 1:
    cpus = [None] * world
    dist.all_gather_object(cpus, (bound[0], bound[-1], len(bound)) if bound else None)
    out["cpus"] = cpus
if rank == 0:
    out["sum_h2d"] = round(sum(out["h2d"]), 1); out["sum_h2d_with_d2h"] = round(sum(out["h2d+d2h"]), 1)
    print(json.dumps(out))
if world > 1: dist.destroy_process_group()
