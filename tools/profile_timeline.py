import os, sys, torch
os.environ["TXF_PROFILE"]="1"
sys.path.insert(0,"/root/repo"); sys.path.insert(0,"/root/repo/tests")
import helpers, kubeflow_tfx_workshop_b200 as ktw
from kubeflow_tfx_workshop_b200 import synth
fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
t = synth.taxi_table(100_000_000)
eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
for i in range(4):
    eng.timeline.clear()
    st = eng.analyze(t)
    torch.cuda.synchronize()
    print({k: round(v,3) for k,v in eng.timeline.items()})
