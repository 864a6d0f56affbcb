"""Experiment: back-to-back overlapped steps when torch is imported (and CUDA initialised) BEFORE the package,
with and without a timing event recorded on the main stream at the start of every step."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
if sys.argv[1] == "init-first":
    torch.zeros(1, device="cuda")  # the CUDA context exists before the package sets CUDA_DEVICE_MAX_CONNECTIONS
import helpers
import kubeflow_tfx_workshop_b200 as ktw
from kubeflow_tfx_workshop_b200 import synth
fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
t = synth.taxi_table(100_000_000)
eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
out = None
for mode in ("plain", "entry-event", "plain"):
    for _ in range(3):
        out, st = eng.analyze_and_transform(t, out)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        if mode == "entry-event":
            e = torch.cuda.Event(enable_timing=True); e.record()
        out, st = eng.analyze_and_transform(t, out)
    b.record(); torch.cuda.synchronize()
    print(sys.argv[1], os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS"), mode, round(a.elapsed_time(b) / 10, 3), flush=True)
