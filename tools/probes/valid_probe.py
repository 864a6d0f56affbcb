import sys, torch
sys.path.insert(0, "/root/repo")
import kubeflow_tfx_workshop_b200 as ktw
from kubeflow_tfx_workshop_b200 import plan as P
n = 100_000_000
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.rand(n, device="cuda", generator=g)
bits_all = torch.full(((n + 7) // 8,), 255, dtype=torch.uint8, device="cuda")
bits_rand = torch.randint(0, 256, ((n + 7) // 8,), dtype=torch.uint8, device="cuda", generator=g) | 0x55
tf, tft = ktw.tf, ktw.tft
def dense(x):
    return tf.squeeze(tf.sparse.to_dense(tf.SparseTensor(x.indices, x.values, [x.dense_shape[0], 1]), 0), axis=1)
for name, valid in (("no bitmap", None), ("bitmap all ones", bits_all), ("bitmap ~37% missing", bits_rand)):
    table = ktw.Table({"x": ktw.NumericColumn(x, valid)})
    for what, fn in (("fill", lambda i: {"o": dense(i["x"])}), ("bucketize", lambda i: {"o": tft.bucketize(dense(i["x"]), 10)}),
                     ("zscore", lambda i: {"o": tft.scale_to_z_score(dense(i["x"]))})):
        eng = ktw.Engine(ktw.trace(fn, {"x": "float32"}))
        st = eng.analyze(table)
        o = eng.transform(table, st)
        ts = []
        for _ in range(7):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); eng.transform(table, st, out=o); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        print(f"{name:22s} {what:10s} {sorted(ts)[3]:.3f} ms")
