"""Experiment (torchrun): per-repetition time of Engine.analyze alone after an overlapped loop, multi-rank."""
import os, sys, time
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
out = None
for _ in range(5):
    out, st = eng.analyze_and_transform(t, out, rows_per_rank=rpr)
torch.cuda.synchronize()
for rep in range(6):
    a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    h0 = time.perf_counter()
    a.record()
    st2 = eng.analyze(t, rpr)
    b.record()
    h1 = time.perf_counter()
    out = eng.transform(t, st2, out)
    c.record()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"rep {rep}: analyze {a.elapsed_time(b):.3f} ms (host {1e3 * (h1 - h0):.3f}), transform {b.elapsed_time(c):.3f} ms", flush=True)
dist.barrier(); dist.destroy_process_group()
