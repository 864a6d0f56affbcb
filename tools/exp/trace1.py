"""Experiment: CUDA-event timeline of one overlapped step on one GPU (TXF_TRACE=1)."""
import os, sys
os.environ["TXF_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, helpers
import kubeflow_tfx_workshop_b200 as ktw
from kubeflow_tfx_workshop_b200 import synth
fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
t = synth.taxi_table(100_000_000)
eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
out = None
for i in range(5):
    eng.trace_events = []
    torch.cuda.synchronize()
    out, st = eng.analyze_and_transform(t, out)
    torch.cuda.synchronize()
    rep = eng.trace_report()
print({k: v for k, v in rep.items()})
# steady state: 10 steps back to back
torch.cuda.synchronize()
eng._trace_on = False
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    out, st = eng.analyze_and_transform(t, out)
b.record(); torch.cuda.synchronize()
print("back to back ms/step", a.elapsed_time(b) / 10)
