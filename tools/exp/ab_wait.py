"""Experiment (torchrun, N ranks): back-to-back steps with the side streams at high / normal priority and the host
waiting on the state event / on the main stream; GPU ms per step and host ms per step (enqueue + waits)."""
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
out, state = eng.analyze_and_transform(t, rows_per_rank=rpr)


def loop(steps=12):
    o = out
    for _ in range(3):
        o, _ = eng.analyze_and_transform(t, o, rows_per_rank=rpr)
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    h0 = time.perf_counter()
    evs = [a]
    for _ in range(steps):
        o, _ = eng.analyze_and_transform(t, o, rows_per_rank=rpr)
        evs.append(torch.cuda.Event(enable_timing=True)); evs[-1].record()
    host = (time.perf_counter() - h0) / steps * 1e3
    b.record(); torch.cuda.synchronize()
    per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(steps))
    v = torch.tensor([a.elapsed_time(b) / steps, host, per[-1], -per[0], per[steps // 2]], dtype=torch.float64, device="cuda")
    dist.all_reduce(v, op=dist.ReduceOp.MAX)
    return v.tolist()


for prio, ev in ((-1, True), (0, True), (-1, False), (0, False), (-1, True)):
    eng._side, eng._side_priority, eng._host_waits_on_event = None, prio, ev
    gpu, host, pmax, pmin, pmed = loop()
    if rank == 0:
        print(f"world {world} side priority {prio:2d} host waits on {'event' if ev else 'main stream'}: {gpu:.3f} ms per step (host {host:.3f}; single steps min {-pmin:.2f} median {pmed:.2f} max {pmax:.2f})", flush=True)
dist.barrier(); dist.destroy_process_group()
