"""Host<->device copy rates of this box with pinned memory: each direction alone and both at once,
on ONE GPU (plain `python`) or on all N ranks CONCURRENTLY (under torchrun): the end-to-end leg of
bench.py cannot move bytes faster than this, so it is the ceiling its `e2e` figure is held against.

    python tools/probes/pcie_probe.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29530 tools/probes/pcie_probe.py [--bind]

--bind pins every rank's host threads to the CPU cores the driver reports as local to its GPU before the
buffers are allocated and pinned (first touch then lands on that NUMA node).  Prints one JSON line per run.
"""
import json
import os
import sys
import time

import torch

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
bind_note = "no binding"
if "--bind" in sys.argv:
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 1) + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1} & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        bind_note = f"bound to {len(cpus)} GPU-local cores"
    except Exception as e:  # noqa: BLE001
        bind_note = f"binding failed: {type(e).__name__}"
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

n = 1 << 30  # 1 GiB per buffer
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
h_in.fill_(1); h_out.fill_(2)  # first touch under the current CPU affinity
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def run(h2d, d2h, reps=8):
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    barrier()
    gbs = torch.tensor([reps * n / (time.perf_counter() - t0) / 1e9], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(gbs, op=dist.ReduceOp.MIN)  # the slowest rank is what a lock-step job sees
    return float(gbs.item())


run(True, True, 2)
res = {"n_gpus": world, "binding": bind_note, "cpu_affinity_cores": len(os.sched_getaffinity(0)),
       "h2d_alone_gbs_per_gpu": run(True, False), "d2h_alone_gbs_per_gpu": run(False, True),
       "both_gbs_per_gpu_each_way": run(True, True), "note": "minimum over ranks, all ranks copying at once"}
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
