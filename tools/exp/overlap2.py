"""Experiment (torchrun, N ranks): event trace of one overlapped step, rank 0."""
import os, sys
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import helpers
import kubeflow_tfx_workshop_b200 as ktw
from kubeflow_tfx_workshop_b200 import synth
R = 100_000_000
fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
t = synth.taxi_table(R, row0=rank * R)
eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA), comm=ktw.Comm())
rpr = [R] * world
out, state = eng.analyze_and_transform(t, rows_per_rank=rpr)

def loop(overlap, steps=8):
    eng.overlap_transform = overlap
    o = out
    for _ in range(2):
        o, _ = eng.analyze_and_transform(t, o, rows_per_rank=rpr)
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        o, _ = eng.analyze_and_transform(t, o, rows_per_rank=rpr)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / steps

for ov in (True, False):
    ms = loop(ov)
    if rank == 0:
        print(f"world {world} overlap {ov}: {ms:.3f} ms per step", flush=True)
# event timeline of one overlapped step (TXF_TRACE=1), every rank's view of the slowest points
if os.environ.get("TXF_TRACE") == "1":
    eng.overlap_transform = True
    for _ in range(3):
        eng.trace_events = []
        dist.barrier(); torch.cuda.synchronize()
        o2, _ = eng.analyze_and_transform(t, out, rows_per_rank=rpr)
        torch.cuda.synchronize()
        rep = eng.trace_report()
    allrep = [None] * world
    dist.all_gather_object(allrep, rep)
    if rank == 0:
        for k in rep:
            vals = [r.get(k, float("nan")) for r in allrep]
            print(f"trace {k:22s} min {min(vals):7.3f}  max {max(vals):7.3f}", flush=True)
eng._profile = True
for _ in range(2):
    eng.timeline.clear()
    eng.analyze(t, rpr)
    torch.cuda.synchronize()
if rank == 0:
    print("profiled analyze segments (wall ms, syncs inside):", {k: round(v, 3) for k, v in eng.timeline.items()}, flush=True)
dist.barrier(); dist.destroy_process_group()
