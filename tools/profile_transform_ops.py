"""Achieved DRAM rate of the fused transform kernel, one column op at a time (100 M taxi rows).

    python tools/profile_transform_ops.py [rows]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
import kubeflow_tfx_workshop_b200 as ktw  # noqa: E402
from kubeflow_tfx_workshop_b200 import plan as P, synth  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
    fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
    table = synth.taxi_table(n)
    full = ktw.trace(fn, synth.TAXI_SCHEMA)
    eng = ktw.Engine(full)
    state = eng.analyze(table)
    out = eng.transform(table, state)
    torch.cuda.synchronize()

    def timed(e, st, reps=5):
        o = e.transform(table, st)
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            e.transform(table, st, out=o)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return sorted(ts)[len(ts) // 2]

    print(f"all {len(full.specs)} ops: {timed(eng, state):.3f} ms")
    for dt, name in ((torch.int64, "int64"), (torch.float32, "float32")):  # torch's own copy at the same sizes, for scale
        a_, b_ = torch.empty(n, dtype=dt, device="cuda"), torch.empty(n, dtype=dt, device="cuda")
        ts = []
        for _ in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            b_.copy_(a_)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[len(ts) // 2]
        print(f"torch copy_ {name:8s} {ms:7.3f} ms -> {2 * a_.nbytes / ms / 1e6:7.1f} GB/s")
    total = 0.0
    for spec in full.specs:
        sub = P.Plan(schema=full.schema, specs=[spec])
        e = ktw.Engine(sub)
        e._ws = eng._ws
        ms = timed(e, state)
        col_in = [table[k] for k in P.spec_inputs(spec)]
        coded = isinstance(spec, P.VocabSpec)
        rd = sum((4 * n if coded else c.nbytes()) for c in col_in)
        wr = out[spec.out_key].nbytes()
        total += ms
        print(f"{type(spec).__name__:16s} {spec.out_key:28s} {ms:7.3f} ms  read {rd / 1e9:5.2f} GB write {wr / 1e9:5.2f} GB  -> {(rd + wr) / ms / 1e6:7.1f} GB/s")
    print(f"sum of single-op launches: {total:.3f} ms")


if __name__ == "__main__":
    main()
