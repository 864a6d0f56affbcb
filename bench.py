#!/usr/bin/env python
"""Benchmark of the hot path: preprocessing_fn analyze + transform rows/s (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W                      # this repo's CUDA path (configs[1] per GPU)
    python bench.py --gpus N --global-rows 1000000000 ...              # configs[2] as written: 1 B rows over N GPUs
    python bench.py --workload vocab_stress | bucket_stress            # configs[3] / configs[4]
    python bench.py --impl reference --gpus N --steps K --warmup W     # CPU arm (oracle port, host cores)

For N > 1 launch with `python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py
--gpus N ...` (one rank per GPU, NCCL).  A step is one full analyze+transform pass
(`Engine.analyze_and_transform`, what tft_beam.AnalyzeAndTransformDataset does) over the synthetic
rows already resident in HBM; rank 0 prints ONE JSON line.  After the timed region the last step is
verified (`parity_check`): analyzer state identical on every rank, equal to a single-rank
recomputation over all shards, outputs equal to the oracle's on a deterministic 1/64 of the rows; a
mismatch makes the run exit non-zero.
"""
import argparse
import importlib
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# The pass keeps four streams busy at once (main, two analyzer streams, the state-free transform stream).  With the
# default of 8 host->device work queues two of them can share a queue and serialise falsely (measured: 7.55 ms per
# step against 6.65 ms); the variable must be set before the CUDA context exists.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

METRIC = "preprocessing_fn analyze+transform rows/sec"
UNIT = "rows/s"
TAXI_WORKLOAD = "taxi preprocessing_fn analyze+transform on synthetic Chicago-taxi-schema rows "
ROWS_PER_GPU = 100_000_000  # BASELINE.json configs[1]: 100M synthetic taxi rows on one B200
CHUNK_ROWS = 100_000_000    # string pools hold < 2^31 bytes: larger shards are lists of chunks


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="taxi", choices=["taxi", "vocab_stress", "bucket_stress"])
    ap.add_argument("--rows-per-gpu", type=int, default=ROWS_PER_GPU)
    ap.add_argument("--global-rows", type=int, default=0, help="total rows sharded over the GPUs (strong scaling; configs[2] = 1e9)")
    ap.add_argument("--chunk-rows", type=int, default=CHUNK_ROWS)
    ap.add_argument("--num-buckets", type=int, default=100, help="bucket_stress: num_buckets (10 / 100 / 1000)")
    ap.add_argument("--e2e-rows", type=int, default=0, help="rows per GPU of the end-to-end leg (0 = same as the shard, bounded by host RAM)")
    ap.add_argument("--e2e-steps", type=int, default=16)
    ap.add_argument("--cpu-rows", type=int, default=6_000_000, help="sample of the workload timed on the host cores")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity-check", action="store_true")
    ap.add_argument("--out-ring", type=int, default=0, help="write the outputs of a multi-chunk shard into this many reused output sets (0 = one set per chunk)")
    ap.add_argument("--no-overlap", action="store_true", help="time Engine.analyze + Engine.transform back to back instead")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores (reported baseline / --impl reference).  It never
# imports the product package (and so never maps libtaxi_xf.so): the synthetic-data module is loaded
# under a private package name whose __init__ is not executed.
# ------------------------------------------------------------------------------------------------
def _data_modules():
    """(synth, columns) of the package directory WITHOUT running the package's __init__ (which loads the CUDA library)."""
    name = "_txf_data"
    if name not in sys.modules:
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(ROOT, "kubeflow-tfx-workshop_b200")]
        sys.modules[name] = pkg
    return importlib.import_module(name + ".synth"), importlib.import_module(name + ".columns")


def _oracle():
    from oracle import tft_oracle as O
    return O


def _oracle_cols(table):
    """Table (host) -> the oracle's column objects."""
    O = _oracle()
    cols = {}
    for k, c in table.items():
        if c.kind == "string":
            off, pool, valid = c.numpy()
            cols[k] = O.StrCol(off, pool, valid)
        else:
            v, valid = c.numpy()
            cols[k] = O.NumCol(v, valid)
    return cols


def _cpu_partial(args):
    lo, hi = args
    synth, _ = _data_modules()
    O = _oracle()
    cols = _oracle_cols(synth.taxi_table(hi - lo, lo, backend="numpy"))
    t0 = time.perf_counter()
    part = O.analyze_partial(cols)
    return part, time.perf_counter() - t0


def _cpu_transform(args):
    lo, hi, state = args
    synth, _ = _data_modules()
    O = _oracle()
    cols = _oracle_cols(synth.taxi_table(hi - lo, lo, backend="numpy"))
    t0 = time.perf_counter()
    out = O.transform(cols, state)
    return sum(len(v) for v in out.values()), time.perf_counter() - t0


def cpu_pass(rows: int, cores: int, pool=None):
    """One analyze+transform pass of the oracle over `rows` synthetic rows, row-sharded over
    `cores` processes, partial analyzer states merged like the GPU ranks do.  Returns seconds of
    the slowest worker per phase + the merge (input generation is not timed, as on the GPU)."""
    O = _oracle()
    cuts = [rows * i // cores for i in range(cores + 1)]
    spans = [(a, b) for a, b in zip(cuts, cuts[1:])]
    if pool is None:
        parts = [_cpu_partial(s) for s in spans]
    else:
        parts = pool.map(_cpu_partial, spans)
    t0 = time.perf_counter()
    state = O.finalize_state(O.merge_partials([p for p, _ in parts]))
    t_merge = time.perf_counter() - t0
    if pool is None:
        outs = [_cpu_transform(s + (state,)) for s in spans]
    else:
        outs = pool.map(_cpu_transform, [s + (state,) for s in spans])
    return max(t for _, t in parts) + t_merge + max(t for _, t in outs)


def run_reference(args):
    """`--impl reference`: the reference's own implementation (tft/Beam) is not installable here,
    so this arm times the oracle port on all host cores (cpu_baseline.kind = "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    _oracle().build_c()  # gcc: the oracle's bulk FarmHash helper; the CUDA library is neither built nor loaded here
    cores = len(os.sched_getaffinity(0))
    rows = args.cpu_rows * max(1, min(cores, 16) // 2)
    times = []
    with mp.get_context("spawn").Pool(cores) as pool:
        for i in range(args.warmup + args.steps):
            t = cpu_pass(rows, cores, pool)
            if i >= args.warmup:
                times.append(t)
    sec = sum(times) / len(times)
    value = rows / sec
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32/i64/u8", "data": "synthetic",
        "config": {"workload": TAXI_WORKLOAD + "(BASELINE configs[1] per GPU)", "rows_per_step": rows,
                   "note": f"bounded sample ({rows} rows per step) of the 100M-row workload so that the arm ends within minutes; the port's "
                           "time is linear in rows (row-sharded numpy passes + one merge), so rows/s carries over to 100M rows. "
                           "tft/Beam itself is not installable on this box"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{rows} synthetic taxi rows per step, row-sharded over {cores} processes"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def _load_example_fn(ktw, name="taxi_preprocessing"):
    """examples/<name>.py imported with the tf / tft stand-ins bound (the GPU box has no /root/reference)."""
    with ktw.install_shims():
        spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "examples", name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    return mod


def table_bytes(table, keys):
    return sum(table[k].nbytes() for k in keys)


def _peak():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def _rows_per_rank(args, ktw, world):
    if args.global_rows:
        return [hi - lo for lo, hi in (ktw.shard_range(args.global_rows, r, world) for r in range(world))]
    return [args.rows_per_gpu] * world


def _workload(args, ktw, synth, dev, rank, world):
    """-> (plan, chunks, rows_local, description).  Shards are lists of chunks of <= chunk_rows rows."""
    import torch
    if args.workload == "taxi":
        fn = _load_example_fn(ktw).preprocessing_fn
        plan = ktw.trace(fn, synth.TAXI_SCHEMA)
        if args.global_rows:
            lo, hi = ktw.shard_range(args.global_rows, rank, world)
        else:
            lo, hi = rank * args.rows_per_gpu, (rank + 1) * args.rows_per_gpu
        chunks = [synth.taxi_table(min(args.chunk_rows, hi - s), row0=s, backend="torch", device=dev) for s in range(lo, hi, args.chunk_rows)]
        desc = (TAXI_WORKLOAD
                + ("(BASELINE configs[2]: global rows sharded over the GPUs)" if args.global_rows else "(BASELINE configs[1] per GPU)"))
        return plan, chunks, hi - lo, desc
    mod = _load_example_fn(ktw, "stress_preprocessing")
    R = args.rows_per_gpu
    if args.workload == "vocab_stress":
        table = synth.vocab_stress_table(R, row0=rank * R, device=dev)
        plan = ktw.trace(mod.vocab_stress_fn, table.schema())
        return plan, [table], R, "vocabulary stress (BASELINE configs[3]): 2 string columns, 10M distinct terms, top_k=1000, 10 OOV buckets"
    table = synth.bucket_stress_table(R, row0=rank * R, device=dev)
    plan = ktw.trace(mod.bucket_stress_fn(args.num_buckets), table.schema(), dense_keys=set(table.keys()))
    return plan, [table], R, f"bucketize stress (BASELINE configs[4]): 8 float columns, num_buckets={args.num_buckets}"


def run_b200(args):
    import torch
    import torch.distributed as dist
    from __graft_entry__ import build

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank == 0:
        build()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()
    import kubeflow_tfx_workshop_b200 as ktw
    from kubeflow_tfx_workshop_b200 import synth
    nat = ktw._native

    plan, chunks, R, desc = _workload(args, ktw, synth, dev, rank, world)
    comm = ktw.Comm()
    big_vocab = args.workload == "vocab_stress"
    eng = ktw.Engine(plan, device=dev, comm=comm, **({"vocab_capacity": 1 << 25, "arena_bytes": 1 << 29} if big_vocab else {}))
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    rows_per_rank = _rows_per_rank(args, ktw, world)  # the benchmark sharded the rows itself: no gather of row counts

    def step(out):
        if args.no_overlap:
            state = eng.analyze(chunks, rows_per_rank)
            return [eng.transform(c, state, o) for c, o in zip(chunks, out)], state
        return eng.analyze_and_transform(chunks, out, rows_per_rank)

    # warm-up (also allocates every workspace and the output buffers)
    ring = args.out_ring if 0 < args.out_ring < len(chunks) else 0
    if ring:  # a shard larger than a chunk writes its outputs into a reused ring of output sets
        sets, _ = eng.analyze_and_transform(chunks[:ring])
        out = [sets[i % ring] for i in range(len(chunks))]
    else:
        out = [None] * len(chunks)
    for _ in range(max(args.warmup, 3)):
        out, state = step(out)
    barrier()

    sampler = ClockSampler(local_rank)
    if rank == 0 and os.environ.get("TXF_NO_CLOCK_SAMPLER") != "1":
        sampler.start()
    launches0 = nat.launch_count()
    barrier()
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for i in range(args.steps):
        out, state = step(out)
    t_end.record()
    barrier()
    launches = nat.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    total_ms = t_start.elapsed_time(t_end)

    # the two phases on their own (not overlapped), for the roofline of the dominant kernel
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    reps = 3
    a_ms = x_ms = 0.0
    for _ in range(reps):
        barrier()  # ranks enter each repetition together: otherwise one rank's lateness is booked as the others' analyze time
        ev[0].record()
        st2 = eng.analyze(chunks, rows_per_rank)
        ev[1].record()
        out = [eng.transform(c, st2, o) for c, o in zip(chunks, out)]
        ev[2].record()
        torch.cuda.synchronize()
        a_ms += ev[0].elapsed_time(ev[1]) / reps
        x_ms += ev[1].elapsed_time(ev[2]) / reps
    out, state = step(out)  # leave the outputs of the timed entry point in place for the parity check
    torch.cuda.synchronize()
    t = torch.tensor([total_ms, a_ms, x_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, analyze_ms, transform_ms = t.tolist()
    ms_per_step = total_ms / args.steps
    rows_all = torch.tensor([R], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(rows_all)
    global_rows = int(rows_all.item())
    value = global_rows / (ms_per_step * 1e-3)

    # roofline of the dominant kernel (the fused transform): algorithmic bytes / measured time
    in_keys = plan.input_keys()
    read_b = sum(table_bytes(c, in_keys) for c in chunks)
    write_b = sum(o[k].nbytes() for o in out for k in o.keys())  # a ring is written once per chunk that maps to it
    peak, peak_src = _peak()
    xf_gbs = (read_b + write_b) / (transform_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r1_transform_traffic.json")
    if os.path.exists(tpath) and args.workload == "taxi":
        tj = json.load(open(tpath))
        traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) * (R / tj["rows"])
    analyze_keys = eng.moment_keys + [k for ks in eng.bucket_groups.values() for k in ks] + [s.in_key for s in eng.vocab_specs.values()]
    analyze_b = sum(table_bytes(c, analyze_keys) for c in chunks)
    whole_gbs = (analyze_b + read_b + write_b) / (ms_per_step * 1e-3) / 1e9
    import ctypes
    g, b, s = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    nat._lib.txf_transform_launch_info(ctypes.byref(g), ctypes.byref(b), ctypes.byref(s))

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if args.global_rows else "weak", "vs_baseline": None,
        "dtype": "f32/i64/u8", "data": "synthetic",
        "config": {"workload": desc, "rows_per_gpu": R, "global_rows": global_rows, "chunks_per_gpu": len(chunks),
                   "l2": "inputs (%.1f GB per GPU) larger than L2" % (read_b / 1e9),
                   "quantile_path": {str(k): v for k, v in eng.last_quantile_path.items()},
                   "entry_point": "Engine.analyze + Engine.transform" if args.no_overlap else
                                  "Engine.analyze_and_transform (state-free ops streamed under the analyzers)",
                   "parallelism": f"rows sharded over {world} GPU(s)",
                   "inter_step": ("the host lets a step drain before it queues the next (Engine.drain_between_steps_world)"
                                  if (comm.active and world >= eng.drain_between_steps_world and not args.no_overlap)
                                  else "steps queued back to back")},
        "phases_ms": {"analyze_alone": analyze_ms, "transform_alone": transform_ms, "sum_not_overlapped": analyze_ms + transform_ms},
        "roofline": {"bound": "hbm", "kernel": "transform_tiles_kernel", "achieved": xf_gbs, "peak": peak, "unit": "GB/s",
                     "frac": xf_gbs / peak, "traffic": traffic, "peak_source": peak_src,
                     "traffic_source": None if traffic is None else "constant from one committed ncu --set full capture (profiles/r1_transform_traffic.json), scaled to this row count; not measured by this run",
                     "traffic_frac": None if traffic is None else traffic / (transform_ms * 1e-3) / 1e9 / peak,
                     "timing": "CUDA events around Engine.transform run alone (all ops in one launch), average of 3",
                     "note": "achieved/frac use SURVEY 8(d) algorithmic bytes, which count the string columns' term bytes; the kernel "
                             "reads 4-byte dictionary codes written by the count pass instead, so its DRAM traffic (traffic, "
                             "traffic_frac) is lower and frac can exceed 1",
                     "algorithmic_bytes": read_b + write_b,
                     "algorithmic_bytes_per_row": (read_b + write_b) / R, "launch": {"grid": g.value, "block": b.value, "smem": s.value},
                     "whole_pass": {"achieved": whole_gbs, "frac": whole_gbs / peak,
                                    "algorithmic_bytes_per_row": (analyze_b + read_b + write_b) / R}},
        "gpu_launches": launches,
        "clocks": clocks,
    }

    ok = True
    if not args.no_parity_check:
        pc = parity_check(args, ktw, synth, eng, plan, chunks, out, state, dev, rank, world)
        ok = bool(pc["ok"])
        line["parity_check"] = pc
    # end-to-end: pinned host buffers -> H2D -> analyze+transform -> D2H of every output column
    if not args.no_e2e and args.workload == "taxi":
        line["e2e"] = run_e2e(args, eng, plan, chunks[0], out[0], dev, world, rank, barrier)
    del chunks, out
    torch.cuda.empty_cache()
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.workload == "taxi":
        t0 = time.perf_counter()
        sec = cpu_pass(args.cpu_rows, 1)
        line["cpu_baseline"] = {"value": args.cpu_rows / sec, "unit": UNIT, "cores": 1, "kind": "port",
                                "sample": f"{args.cpu_rows} synthetic taxi rows, oracle port (numpy + C FarmHash), 1 process; "
                                          f"tft/Beam itself is not installable here ({time.perf_counter() - t0:.0f}s incl. generation)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        sys.exit(3)


# ------------------------------------------------------------------------------------------------
# parity check of the last step (after the timed region)
# ------------------------------------------------------------------------------------------------
def _state_signature(state):
    """Everything the transform depends on, as bytes: moments, boundaries, vocabulary terms in id order."""
    import numpy as np
    parts = []
    if state.moments is not None:
        parts.append(state.moments.cpu().numpy().tobytes())
        parts.append(state.mean_var.cpu().numpy().tobytes())
    for nb in sorted(state.buckets):
        res = state.buckets[nb]
        for k in res.keys:
            parts.append(res.boundaries_of(k).astype(np.float32).tobytes())
    for key in sorted(state.vocabs, key=str):
        terms = state.vocabs[key].terms()
        parts.append(repr((key, len(terms))).encode())
        parts.append(b"\x00".join(terms))
    return parts


def _host_rows(table, lo, hi):
    """Rows [lo, hi) of a device table as a host table (lo % 8 == 0)."""
    import kubeflow_tfx_workshop_b200 as ktw
    cols = {}
    for k, c in table.items():
        valid = None if c.valid is None else c.valid[lo // 8:(hi + 7) // 8].cpu()
        if c.kind == "numeric":
            cols[k] = ktw.NumericColumn(c.values[lo:hi].cpu(), valid)
        else:
            off = c.offsets[lo:hi + 1].cpu()
            b0, b1 = int(off[0]), int(off[-1])
            pool = c.pool[b0:b1 + 64].cpu().clone()
            cols[k] = ktw.StringColumn((off - b0).to(off.dtype), pool, valid, b1 - b0)
    return ktw.Table(cols)


def parity_check(args, ktw, synth, eng, plan, chunks, out, state, dev, rank, world):
    """(1) the analyzer state is bit-identical on every rank; (2) it equals a single-rank recomputation
    over ALL shards done on rank 0 (vocabulary and boundaries bit-exact, moments to 1e-12) when those
    rows fit next to the benchmark's buffers; (3) taxi workload: the outputs of this rank equal the
    oracle's transform (given that state) on a deterministic 1/64 of the rows -- integers and strings
    bit-exact, z-scores within 1e-6 relative.  Returns a dict for the JSON line."""
    import hashlib
    import numpy as np
    import torch
    import torch.distributed as dist
    res = {"ok": True}
    sig = _state_signature(state)
    digest = hashlib.sha256(b"".join(sig)).digest()
    d = torch.tensor(list(digest), dtype=torch.uint8, device=dev)
    if world > 1:
        g = [torch.empty_like(d) for _ in range(world)]
        dist.all_gather(g, d)
        same = all(bool((x == g[0]).all()) for x in g)
    else:
        same = True
    res["state_identical_on_all_ranks"] = same
    res["state_sha256"] = digest.hex()[:16]
    res["ok"] &= same

    # (2) single-rank recomputation on rank 0
    if world > 1 and args.workload == "taxi":
        total_rows = args.global_rows or world * args.rows_per_gpu
        need = total_rows * 135  # input bytes + dictionary codes per row
        free = torch.cuda.mem_get_info(dev)[0]
        verdict = None
        if rank == 0:
            if need < 0.8 * free:
                spans = [ktw.shard_range(args.global_rows, r, world) if args.global_rows else (r * args.rows_per_gpu, (r + 1) * args.rows_per_gpu)
                         for r in range(world)]
                allc = []
                for lo, hi in spans:
                    allc += [synth.taxi_table(min(args.chunk_rows, hi - s), row0=s, backend="torch", device=dev) for s in range(lo, hi, args.chunk_rows)]
                solo = ktw.Engine(plan, device=dev, comm=_NoComm())
                st1 = solo.analyze(allc)
                same_vb = all(a == b for a, b in zip(_state_signature(st1)[2:], sig[2:])) and len(_state_signature(st1)) == len(sig)
                m1, mN = st1.moments.cpu().numpy(), state.moments.cpu().numpy()
                rel = float(np.max(np.abs(m1 - mN) / np.maximum(np.abs(m1), 1e-300)))
                verdict = {"rows": total_rows, "vocab_and_boundaries_bit_exact": bool(same_vb), "moments_max_rel_diff": rel,
                           "ok": bool(same_vb and rel <= 1e-12)}
                del allc, solo, st1
                torch.cuda.empty_cache()
            else:
                verdict = {"skipped": f"{total_rows} rows need {need / 1e9:.0f} GB on one GPU next to the benchmark's buffers ({free / 1e9:.0f} GB free)", "ok": True}
        flag = torch.tensor([1 if (verdict is None or verdict["ok"]) else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        res["single_rank_recomputation"] = verdict
        res["ok"] &= bool(flag.item())

    # (3) outputs against the oracle on 1/64 of this rank's rows
    if args.workload == "taxi":
        O = _oracle()
        O.build_c()
        ostate = {"zscore": {k: tuple(np.float32(v) for v in state.mean_var_of(k)) for k in state.moment_keys},
                  "vocab": {key[0]: state.vocabs[key].terms() for key in state.vocabs},
                  "boundaries": {k: res_.boundaries_of(k) for res_ in state.buckets.values() for k in res_.keys}}
        bad, checked = [], 0
        live = range(len(chunks)) if not args.out_ring or args.out_ring >= len(chunks) else range(len(chunks) - args.out_ring, len(chunks))
        for c, o in [(chunks[i], out[i]) for i in live]:  # with an output ring only the last sets still hold their chunk
            n = len(c)
            seg = max(8, (n // (64 * 16)) // 8 * 8)
            for i in range(16):
                lo = (n * i // 16) // 8 * 8
                hi = min(lo + seg, n)
                if hi <= lo:
                    continue
                want = O.transform(_oracle_cols(_host_rows(c, lo, hi)), ostate)
                for key, w in want.items():
                    col = o[key]
                    if isinstance(w, list):
                        off = col.offsets[lo:hi + 1].cpu().numpy()
                        pool = col.pool[int(off[0]):int(off[-1])].cpu().numpy().tobytes()
                        got = [pool[a - off[0]:b - off[0]] for a, b in zip(off[:-1], off[1:])]
                        good = got == w
                    elif w.dtype == np.float32:
                        mean, var = ostate["zscore"][key[:-3]]
                        gotv = col.values[lo:hi].cpu().numpy()
                        std = float(np.sqrt(np.float32(var))) or 1.0
                        atol = 4 * float(np.spacing(np.float32(abs(mean)) + np.float32(1))) / std  # x ~ mean is ill-conditioned
                        good = bool(np.allclose(gotv, w, rtol=1e-6, atol=atol, equal_nan=True))
                    else:
                        good = bool(np.array_equal(col.values[lo:hi].cpu().numpy(), w))
                    if not good:
                        bad.append(key)
                checked += hi - lo
        flag = torch.tensor([0 if bad else 1, checked], dtype=torch.int64, device=dev)
        if world > 1:
            g = [torch.empty_like(flag) for _ in range(world)]
            dist.all_gather(g, flag)
            allok, checked = all(int(x[0]) for x in g), sum(int(x[1]) for x in g)
        else:
            allok = not bad
        res["outputs_vs_oracle"] = {"rows_checked": int(checked), "mismatching_outputs_rank0": sorted(set(bad)), "ok": bool(allok),
                                    "tolerance": "integers/strings bit-exact; z-scores rtol 1e-6 + 4 ulp(mean)/std"}
        res["ok"] &= bool(allok)
    res["ok"] = bool(res["ok"])
    return res


class _NoComm:
    """A Comm that is never active: the single-rank recomputation must not talk to the other ranks."""
    active, enabled, rank, world = False, False, 0, 1


def run_e2e(args, eng, plan, table, out, dev, world, rank, barrier):
    """End-to-end leg: every step copies ALL inputs from pinned host memory to the GPU, runs
    Engine.analyze_and_transform (the public API), and copies ALL outputs back to pinned host
    memory.  Steps are software-pipelined the way a streaming ingest would run them: two device
    input sets and two device output sets, so step i+1's host->device copy overlaps step i's compute
    and device->host copy (PCIe is full duplex).  The serial (un-pipelined) step time is reported too."""
    import psutil
    import torch
    import torch.distributed as dist
    import kubeflow_tfx_workshop_b200 as ktw

    R = args.e2e_rows or len(table)
    per_row = (table.nbytes() + sum(out[k].nbytes() for k in out.keys())) / len(table)
    budget = psutil.virtual_memory().available * 0.5 / max(world, 1)
    free_dev = torch.cuda.mem_get_info(dev)[0]
    R = int(min(R, len(table), budget // per_row, (free_dev * 0.8) // (2 * per_row)))
    R -= R % 8
    in_keys = plan.input_keys()
    sub = ktw.Table({k: _slice_col(table[k], R) for k in in_keys})
    _bind_to_gpu_numa_node(dev)
    host_in = ktw.Table({k: _pin(c) for k, c in sub.items()})
    # device staging buffers and pinned output buffers are allocated once, outside the timed region
    dev_in = [ktw.Table({k: _empty_like(c, dev) for k, c in host_in.items()}) for _ in range(2)]
    out_dev = [eng.analyze_and_transform(sub)[0], eng.analyze_and_transform(sub)[0]]
    host_out = {k: _pinned_like(c) for k, c in out_dev[0].items()}
    h2d = host_in.nbytes()
    d2h = sum(c.nbytes() for c in out_dev[0].columns.values())
    torch.cuda.synchronize()
    s_h2d, s_cmp, s_d2h = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(steps, pipelined):
        ev_in = [None, None]      # H2D of buffer b finished
        ev_cmp = [None, None]     # compute finished reading input buffer b / writing output buffer b
        ev_out = [None, None]     # D2H of output buffer b finished

        def upload(i):
            b = i & 1
            with torch.cuda.stream(s_h2d):
                if ev_cmp[b] is not None:
                    s_h2d.wait_event(ev_cmp[b])       # the step that last used this input set is done with it
                for k, c in host_in.items():
                    _copy_col(dev_in[b][k], c)
                ev_in[b] = torch.cuda.Event()
                ev_in[b].record(s_h2d)

        upload(0)
        for i in range(steps):
            b = i & 1
            if pipelined and i + 1 < steps:
                upload(i + 1)
            with torch.cuda.stream(s_cmp):
                s_cmp.wait_event(ev_in[b])
                if ev_out[b] is not None:
                    s_cmp.wait_event(ev_out[b])       # this output set has been read back
                o, _ = eng.analyze_and_transform(dev_in[b], out_dev[b])
                ev_cmp[b] = torch.cuda.Event()
                ev_cmp[b].record(s_cmp)
            with torch.cuda.stream(s_d2h):
                s_d2h.wait_event(ev_cmp[b])
                for k, c in o.items():
                    _copy_col_to_host(host_out[k], c)
                ev_out[b] = torch.cuda.Event()
                ev_out[b].record(s_d2h)
            if not pipelined:
                torch.cuda.synchronize()
                if i + 1 < steps:
                    upload(i + 1)
        torch.cuda.synchronize()

    def timed(steps, pipelined):
        barrier()
        t0 = time.perf_counter()
        run(steps, pipelined)
        barrier()
        ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def copies_only(reps):
        """The same bytes with no compute in between: every input column up, every output column down, both directions at
        once on all ranks -- the ceiling this box's host links put on the end-to-end step (measured, not assumed)."""
        o = out_dev[0]
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            with torch.cuda.stream(s_h2d):
                for k, c in host_in.items():
                    _copy_col(dev_in[0][k], c)
            with torch.cuda.stream(s_d2h):
                for k, c in o.items():
                    _copy_col_to_host(host_out[k], c)
        torch.cuda.synchronize()
        barrier()
        ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    run(2, True)  # warm-up (allocates every workspace)
    serial_ms = timed(max(args.e2e_steps // 2, 1), False)
    piped_ms = timed(args.e2e_steps, True)
    copies_only(1)
    ceiling_ms = copies_only(3)
    return {"value": world * R / (piped_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "rows_per_gpu": R, "ms_per_step": piped_ms, "steps": args.e2e_steps, "pipelined": True,
            "serial_ms_per_step": serial_ms, "serial_value": world * R / (serial_ms * 1e-3),
            "copy_gbs_per_gpu_each_way": {"h2d": h2d / (piped_ms * 1e-3) / 1e9, "d2h": d2h / (piped_ms * 1e-3) / 1e9},
            "pcie_ceiling": {"ms_per_step": ceiling_ms, "gbs_per_gpu_each_way": max(h2d, d2h) / (ceiling_ms * 1e-3) / 1e9,
                             "how": "the step's H2D and D2H copies alone, both directions at once on all ranks, 3 repetitions, max over ranks"},
            "frac_of_ceiling": ceiling_ms / piped_ms,
            "host_numa": _NUMA_NOTE.get("note"),
            "timing": "host wall clock around barrier + synchronize (three streams); max over ranks",
            "api": "Engine.analyze_and_transform on tables copied from pinned host memory; every output column copied back; "
                   "step i+1's upload overlaps step i's compute and download"}


_NUMA_NOTE = {}


def _bind_to_gpu_numa_node(dev):
    """Pin this rank's host threads (and so the pages it first-touches and pins) to the CPU cores of the
    GPU's NUMA node, when the driver reports one: with all ranks on one node the host<->device copies of
    8 GPUs share that node's memory controllers and inter-socket link."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(dev.index or 0)
        ncpu = os.cpu_count() or 1
        words = (ncpu + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if use and use != allowed:
            os.sched_setaffinity(0, use)
            _NUMA_NOTE["note"] = f"host threads bound to the GPU's {len(use)} local cores before pinning"
        else:
            _NUMA_NOTE["note"] = f"GPU's ideal CPU set equals the allowed set ({len(allowed)} cores): nothing to bind"
    except Exception as e:  # noqa: BLE001 -- optional placement; the copy still works without it
        _NUMA_NOTE["note"] = f"no NUMA placement ({type(e).__name__})"


def _slice_col(c, n):
    import kubeflow_tfx_workshop_b200 as ktw
    if c.kind == "numeric":
        return ktw.NumericColumn(c.values[:n], None if c.valid is None else c.valid[: (n + 7) // 8])
    nbytes = int(c.offsets[n].item())
    return ktw.StringColumn(c.offsets[: n + 1], c.pool[: nbytes + 64], None if c.valid is None else c.valid[: (n + 7) // 8], nbytes)


def _pin(c):
    return c.to("cpu").pin()


def _empty_like(c, dev):
    import torch
    import kubeflow_tfx_workshop_b200 as ktw
    e = lambda t: None if t is None else torch.empty_like(t, device=dev)
    if c.kind == "numeric":
        return ktw.NumericColumn(e(c.values), e(c.valid))
    return ktw.StringColumn(e(c.offsets), e(c.pool), e(c.valid), c.pool_bytes)


def _pinned_like(c):
    import torch
    p = lambda t: None if t is None else torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    if c.kind == "numeric":
        return (p(c.values),)
    return (p(c.offsets), p(c.pool))


def _copy_col(dst, src):
    if dst.kind == "numeric":
        dst.values.copy_(src.values, non_blocking=True)
    else:
        dst.offsets.copy_(src.offsets, non_blocking=True)
        dst.pool.copy_(src.pool, non_blocking=True)
    if src.valid is not None:
        dst.valid.copy_(src.valid, non_blocking=True)


def _copy_col_to_host(dst, c):
    if c.kind == "numeric":
        dst[0].copy_(c.values, non_blocking=True)
    else:
        dst[0].copy_(c.offsets, non_blocking=True)
        dst[1].copy_(c.pool, non_blocking=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
