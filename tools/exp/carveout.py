"""Experiment: what makes the back-to-back overlapped step slower than a single traced one?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, helpers
import kubeflow_tfx_workshop_b200 as ktw
from kubeflow_tfx_workshop_b200 import synth
mode = sys.argv[1]
fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
t = synth.taxi_table(100_000_000)
eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
out, state = eng.analyze_and_transform(t)
for rep in range(3):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        if mode == "events":
            e = torch.cuda.Event(enable_timing=True); e.record()
        if mode == "sync":
            torch.cuda.synchronize()
        out, state = eng.analyze_and_transform(t, out)
        if mode == "events":
            e2 = torch.cuda.Event(enable_timing=True); e2.record()
    b.record(); torch.cuda.synchronize()
    print("mode", mode, os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS"), "overlapped ms/step", round(a.elapsed_time(b) / 10, 3), flush=True)
