"""Experiment: pieces of the overlapped analyze+transform step, timed alone and together (100 M taxi rows)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, helpers
import kubeflow_tfx_workshop_b200 as ktw
from kubeflow_tfx_workshop_b200 import synth

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
t = synth.taxi_table(rows)
eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
out, state = eng.analyze_and_transform(t)

def timeit(name, f, reps=5):
    for _ in range(2):
        f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        f()
    b.record(); torch.cuda.synchronize()
    print(f"{name:50s} {a.elapsed_time(b) / reps:8.3f} ms", flush=True)

timeit("state-free ops, stream kernel alone", lambda: eng.transform_state_free(t, out))
def dep():
    done = {k: out[k] for k in out.keys() if k in free_keys}
    eng.transform(t, state, out, done)
free_keys = set(eng.transform_state_free(t, out).keys())
timeit("dependent ops alone (tiles kernel)", dep)
timeit("all ops, tiles kernel", lambda: eng.transform(t, state, out))
for w in (0, 24):
    def an():
        eng._count_warps = w
        eng.analyze(t)
        eng._count_warps = 0
    timeit(f"analyze alone, count warps {w or 32}", an)
for vk, spec in eng.vocab_specs.items():
    for w in (0, 24):
        def vc():
            eng._count_warps = w
            eng._vocab_count(vk, spec, [t])
            eng._count_warps = 0
        timeit(f"vocab_count {vk[0]} warps {w or 32}", vc)
# vocab count with the stream kernel next to it
xs = torch.cuda.Stream()
for vk, spec in eng.vocab_specs.items():
    def both():
        main = torch.cuda.current_stream()
        xs.wait_stream(main)
        with torch.cuda.stream(xs):
            eng.transform_state_free(t, out)
        eng._count_warps = 24
        eng._vocab_count(vk, spec, [t])
        eng._count_warps = 0
        main.wait_stream(xs)
    timeit(f"vocab_count {vk[0]} (24 warps) + stream kernel", both)
# how long does the count pass itself take while the stream kernel is running? (events on the main stream only)
for vk, spec in eng.vocab_specs.items():
    tot = 0.0
    for _ in range(5):
        main = torch.cuda.current_stream()
        torch.cuda.synchronize()
        xs.wait_stream(main)
        with torch.cuda.stream(xs):
            eng.transform_state_free(t, out)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eng._count_warps = 24
        a.record()
        eng._vocab_count(vk, spec, [t])
        b.record()
        eng._count_warps = 0
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    print(f"{'vocab_count ' + vk[0] + ' (24 warps) WHILE the stream kernel runs':50s} {tot / 5:8.3f} ms", flush=True)
timeit("analyze_and_transform (overlapped)", lambda: eng.analyze_and_transform(t, out))
eng.overlap_transform = False
timeit("analyze_and_transform (not overlapped)", lambda: eng.analyze_and_transform(t, out))

# where does the overlapped step spend its time?  events on each stream, no syncs inside
eng.overlap_transform = True
import types
def traced_step(count_warps):
    main = torch.cuda.current_stream()
    xs = eng._xstream or torch.cuda.Stream()
    eng._xstream = xs
    e = {k: torch.cuda.Event(enable_timing=True) for k in ("start", "x_end", "a_end", "end")}
    torch.cuda.synchronize()
    e["start"].record(main)
    xs.wait_stream(main)
    with torch.cuda.stream(xs):
        done = eng.transform_state_free(t, out)
        e["x_end"].record(xs)
    eng._count_warps = count_warps
    st = eng.analyze(t)
    eng._count_warps = 0
    e["a_end"].record(main)
    eng.transform(t, st, out, done)
    main.wait_stream(xs)
    e["end"].record(main)
    torch.cuda.synchronize()
    return {k: round(e["start"].elapsed_time(e[k]), 3) for k in ("x_end", "a_end", "end")}
for w in (24, 0):
    for _ in range(3):
        r = traced_step(w)
    print(f"traced overlapped step, count warps {w or 32}: stream kernel ends {r['x_end']}, analyze ends {r['a_end']}, step ends {r['end']}", flush=True)

# back-to-back steps: per-step GPU span and the gap between steps, from events (no syncs)
import time
evs = []
torch.cuda.synchronize()
t0 = time.perf_counter()
host = []
for i in range(6):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0 = time.perf_counter()
    a.record()
    eng.analyze_and_transform(t, out)
    b.record()
    host.append((time.perf_counter() - h0) * 1e3)
    evs.append((a, b))
torch.cuda.synchronize()
print("host ms per call", [round(h, 2) for h in host])
print("GPU span per step", [round(a.elapsed_time(b), 3) for a, b in evs])
print("gap to next step", [round(evs[i][1].elapsed_time(evs[i + 1][0]), 3) for i in range(5)])
print("start-to-start", [round(evs[i][0].elapsed_time(evs[i + 1][0]), 3) for i in range(5)])
