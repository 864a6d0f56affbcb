import os, sys, torch, torch.distributed as dist
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0,ROOT); sys.path.insert(0,os.path.join(ROOT,'tests'))
os.environ["TXF_PROFILE"]="1"
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda",local))
import helpers, kubeflow_tfx_workshop_b200 as ktw
from kubeflow_tfx_workshop_b200 import synth
fn=helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
eng=ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA), device=f"cuda:{local}", comm=ktw.Comm())
R=50_000_000
t=synth.taxi_table(R,row0=rank*R,backend="torch",device=f"cuda:{local}")
for i in range(4): st=eng.analyze(t)
eng.timeline.clear()
for i in range(5): st=eng.analyze(t)
if rank==0: print({k:round(v/5,3) for k,v in eng.timeline.items()})
dist.barrier(); dist.destroy_process_group()
