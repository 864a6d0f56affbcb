#!/usr/bin/env python
"""Benchmark of the hot path: preprocessing_fn analyze + transform rows/s (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # CPU arm (oracle port, host cores)

For N > 1 launch with `python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py
--gpus N ...` (one rank per GPU, NCCL).  A step is one full analyze+transform pass over the
synthetic taxi rows already resident in HBM; rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "preprocessing_fn analyze+transform rows/sec"
UNIT = "rows/s"
ROWS_PER_GPU = 100_000_000  # BASELINE.json configs[1]: 100M synthetic taxi rows on one B200


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows-per-gpu", type=int, default=ROWS_PER_GPU)
    ap.add_argument("--e2e-rows", type=int, default=0, help="rows per GPU of the end-to-end leg (0 = same as --rows-per-gpu, bounded by host RAM)")
    ap.add_argument("--e2e-steps", type=int, default=16)
    ap.add_argument("--cpu-rows", type=int, default=6_000_000, help="sample of the workload timed on the host cores")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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
                                          "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
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
# CPU arm: the oracle port on the host cores (reported baseline / --impl reference)
# ------------------------------------------------------------------------------------------------
def _cpu_partial(args):
    lo, hi = args
    import helpers
    from helpers import O
    from kubeflow_tfx_workshop_b200 import synth
    cols = helpers.oracle_from_table(synth.taxi_table(hi - lo, lo, backend="numpy"))
    t0 = time.perf_counter()
    part = O.analyze_partial(cols)
    return part, time.perf_counter() - t0


def _cpu_transform(args):
    lo, hi, state = args
    import helpers
    from helpers import O
    from kubeflow_tfx_workshop_b200 import synth
    cols = helpers.oracle_from_table(synth.taxi_table(hi - lo, lo, backend="numpy"))
    t0 = time.perf_counter()
    out = O.transform(cols, state)
    return sum(len(v) for v in out.values()), time.perf_counter() - t0


def cpu_pass(rows: int, cores: int, pool=None):
    """One analyze+transform pass of the oracle over `rows` synthetic rows, row-sharded over
    `cores` processes, partial analyzer states merged like the GPU ranks do.  Returns seconds of
    the slowest worker per phase + the merge (input generation is not timed, as on the GPU)."""
    import helpers
    from helpers import O
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
    from __graft_entry__ import build
    build()
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
        "config": {"workload": "taxi preprocessing_fn analyze+transform, synthetic Chicago-taxi schema", "rows_per_step": rows,
                   "note": "bounded sample of the 100M-row workload; tft/Beam itself is not installable on this box"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{rows} synthetic taxi rows per step, row-sharded over {cores} processes"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def table_bytes(table, keys):
    return sum(table[k].nbytes() for k in keys)


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
    import helpers
    nat = ktw._native

    R = args.rows_per_gpu
    fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
    plan = ktw.trace(fn, synth.TAXI_SCHEMA)
    comm = ktw.Comm()
    eng = ktw.Engine(plan, device=dev, comm=comm)
    table = synth.taxi_table(R, row0=rank * R, backend="torch", device=dev)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # warm-up (also allocates every workspace and the output buffers)
    out = None
    for _ in range(max(args.warmup, 3)):
        state = eng.analyze(table)
        out = eng.transform(table, state, out)
    barrier()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    launches0 = nat.launch_count()
    barrier()
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for i in range(args.steps):
        ev[i][0].record()
        state = eng.analyze(table)
        ev[i][1].record()
        out = eng.transform(table, state, out)
        ev[i][2].record()
    t_end.record()
    barrier()
    launches = nat.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    total_ms = t_start.elapsed_time(t_end)
    analyze_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / args.steps
    transform_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps
    t = torch.tensor([total_ms, analyze_ms, transform_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, analyze_ms, transform_ms = t.tolist()
    ms_per_step = total_ms / args.steps
    value = world * R / (ms_per_step * 1e-3)

    # roofline of the dominant kernel (the fused transform): algorithmic bytes / measured time
    in_keys = plan.input_keys()
    read_b = table_bytes(table, in_keys)
    write_b = sum(out[k].nbytes() for k in out.keys())
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    xf_gbs = (read_b + write_b) / (transform_ms * 1e-3) / 1e9
    traffic = None  # dram__bytes_read.sum + dram__bytes_write.sum of one transform launch (ncu --set full capture)
    tpath = os.path.join(ROOT, "profiles", "r1_transform_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) * (R / tj["rows"])
    analyze_keys = eng.moment_keys + [k for ks in eng.bucket_groups.values() for k in ks] + [s.in_key for s in eng.vocab_specs.values()]
    analyze_b = table_bytes(table, analyze_keys)
    whole_gbs = (analyze_b + read_b + write_b) / (ms_per_step * 1e-3) / 1e9
    import ctypes
    g, b, s = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    nat._lib.txf_transform_launch_info(ctypes.byref(g), ctypes.byref(b), ctypes.byref(s))

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32/i64/u8", "data": "synthetic",
        "config": {"workload": "taxi preprocessing_fn analyze+transform on synthetic Chicago-taxi-schema rows (BASELINE configs[1])",
                   "rows_per_gpu": R, "global_rows": world * R, "l2": "inputs (%.1f GB per GPU) larger than L2" % (read_b / 1e9),
                   "quantile_path": eng.last_quantile_path.get(10), "parallelism": f"rows sharded over {world} GPU(s)"},
        "phases_ms": {"analyze": analyze_ms, "transform": transform_ms},
        "roofline": {"bound": "hbm", "kernel": "transform_tiles_kernel", "achieved": xf_gbs, "peak": peak, "unit": "GB/s",
                     "frac": xf_gbs / peak, "traffic": traffic, "peak_source": peak_src,
                     "traffic_frac": None if traffic is None else traffic / (transform_ms * 1e-3) / 1e9 / peak,
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

    # end-to-end: pinned host buffers -> H2D -> analyze+transform -> D2H of every output column
    if not args.no_e2e:
        line["e2e"] = run_e2e(args, eng, plan, table, out, dev, world, rank, barrier)
    del table, out
    torch.cuda.empty_cache()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
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


def run_e2e(args, eng, plan, table, out, dev, world, rank, barrier):
    """End-to-end leg: every step copies ALL inputs from pinned host memory to the GPU, runs
    Engine.analyze + Engine.transform (the public API), and copies ALL outputs back to pinned host
    memory.  Steps are software-pipelined the way a streaming ingest would run them: two device
    input sets and two device output sets, so step i+1's host->device copy overlaps step i's compute
    and device->host copy (PCIe is full duplex).  The serial (un-pipelined) step time is reported too."""
    import psutil
    import torch
    import torch.distributed as dist
    import kubeflow_tfx_workshop_b200 as ktw

    R = args.e2e_rows or args.rows_per_gpu
    per_row = (table.nbytes() + sum(out[k].nbytes() for k in out.keys())) / len(table)
    budget = psutil.virtual_memory().available * 0.5 / max(world, 1)
    free_dev = torch.cuda.mem_get_info(dev)[0]
    R = int(min(R, len(table), budget // per_row, (free_dev * 0.8) // (2 * per_row)))
    R -= R % 8
    in_keys = plan.input_keys()
    sub = ktw.Table({k: _slice_col(table[k], R) for k in in_keys})
    host_in = ktw.Table({k: _pin(c) for k, c in sub.items()})
    # device staging buffers and pinned output buffers are allocated once, outside the timed region
    dev_in = [ktw.Table({k: _empty_like(c, dev) for k, c in host_in.items()}) for _ in range(2)]
    state = eng.analyze(sub)
    out_dev = [eng.transform(sub, state), eng.transform(sub, state)]
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
                st = eng.analyze(dev_in[b])
                o = eng.transform(dev_in[b], st, out_dev[b])
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

    run(2, True)  # warm-up (allocates every workspace)
    serial_ms = timed(max(args.e2e_steps // 2, 1), False)
    piped_ms = timed(args.e2e_steps, True)
    return {"value": world * R / (piped_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "rows_per_gpu": R, "ms_per_step": piped_ms, "steps": args.e2e_steps, "pipelined": True,
            "serial_ms_per_step": serial_ms, "serial_value": world * R / (serial_ms * 1e-3),
            "timing": "host wall clock around barrier + synchronize (three streams); max over ranks",
            "api": "Engine.analyze + Engine.transform on tables copied from pinned host memory; every output column copied back; "
                   "step i+1's upload overlaps step i's compute and download"}


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
