"""Experiment: vocab_count_kernel time against the number of warps of its CTA (variants of the
library built with -DTXF_VC_WARPS=N under build_variants/).  Run on the GPU box:
    python tools/exp/vocab_warps.py            # parent: one subprocess per variant
"""
import os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, "kubeflow-tfx-workshop_b200", "libtaxi_xf.so")

def child(rows):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch, helpers
    import kubeflow_tfx_workshop_b200 as ktw
    from kubeflow_tfx_workshop_b200 import synth
    fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
    t = synth.taxi_table(rows)
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    res = {}
    for vk, spec in eng.vocab_specs.items():
        for _ in range(2):
            eng._vocab_count(vk, spec, [t])
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            eng._vocab_count(vk, spec, [t])
        b.record(); torch.cuda.synchronize()
        res[vk[0]] = a.elapsed_time(b) / 5
    print("RESULT", os.environ.get("VARIANT"), rows, {k: round(v, 3) for k, v in res.items()}, flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        child(int(sys.argv[2]))
    else:
        rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
        keep = LIB + ".keep"
        shutil.copy(LIB, keep)
        try:
            for v in sorted(os.listdir(os.path.join(ROOT, "build_variants"))):
                shutil.copy(os.path.join(ROOT, "build_variants", v), LIB)
                env = dict(os.environ, VARIANT=v)
                subprocess.run([sys.executable, __file__, "child", str(rows)], env=env)
        finally:
            shutil.copy(keep, LIB); os.remove(keep)
