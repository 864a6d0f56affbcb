"""Throughput of the GPU CSV ingest and tf.Example/TFRecord encode next to the host readers, on taxi-shaped text.

    python tools/bench_ingest.py [--mb 512] [--reps 5]

The text is the body of tests/golden/taxi_head200.csv repeated to the requested size (same field
shapes as the bundled file).  Prints one JSON line: device-resident parse rate, end-to-end rate
with the host->device copy from pinned memory, and the host baselines (pyarrow's C++ reader on all
cores; the reference-style csv + int()/float() row loop on a bounded sample)."""
import argparse
import csv
import io
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kubeflow_tfx_workshop_b200 as ktw  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=512)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    with open(os.path.join(ROOT, "tests", "golden", "taxi_head200.csv"), "rb") as f:
        head = f.read()
    header, body = head.split(b"\n", 1)
    k = max(args.mb * (1 << 20) // len(body), 1)
    text = header + b"\n" + body * k
    nbytes, nrows = len(text), 200 * k
    host = torch.from_numpy(np.frombuffer(text, np.uint8).copy()).pin_memory()
    dev = host.to("cuda")
    schema = None
    launches0 = ktw._native.launch_count()
    table = ktw.read_csv(dev)  # warm-up (+ type inference)
    launches = ktw._native.launch_count() - launches0
    schema = table.schema()
    assert len(table) == nrows

    def timed(fn):
        best = []
        for _ in range(args.reps):
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            best.append(a.elapsed_time(b))
        return float(np.median(best))

    ms_infer = timed(lambda: ktw.read_csv(dev))
    ms_typed = timed(lambda: ktw.read_csv(dev, schema=schema))
    ms_e2e = timed(lambda: ktw.read_csv(host, schema=schema))
    out_bytes = table.nbytes()

    sub = ktw.Table({k: c.slice(0, 1 << 20) if c.kind == "numeric" else c for k, c in table.items() if c.kind == "numeric"})
    ms_enc = timed(lambda: ktw.encode_tfrecord(sub))
    enc_bytes = int(ktw.encode_tfrecord(sub)[1][-1].item())

    # decode of those records back into columns (record index + protobuf parse on the device)
    enc_data = ktw.encode_tfrecord(sub)[0]
    dec_schema = sub.schema()
    assert len(ktw.read_tfrecord(enc_data, dec_schema)) == len(sub)
    ms_dec = timed(lambda: ktw.read_tfrecord(enc_data, dec_schema))
    ms_dec_nocrc = timed(lambda: ktw.read_tfrecord(enc_data, dec_schema, verify_crc=False))

    import pyarrow.csv as pacsv
    t0 = time.perf_counter()
    pacsv.read_csv(io.BytesIO(text))
    t_arrow = time.perf_counter() - t0

    sample_rows = min(nrows, 200_000)
    sample = body * (sample_rows // 200)
    kinds = list(schema.values())
    t0 = time.perf_counter()
    for row in csv.reader(io.StringIO(sample.decode())):  # reference-style typed loop (notebooks/04_...:222-240)
        for v, kd in zip(row, kinds):
            if v != "" and kd != "bytes":
                (int if kd == "int64" else float)(v)
    t_py = time.perf_counter() - t0
    print(json.dumps({
        "metric": "taxi CSV ingest", "text_mb": nbytes / 1e6, "rows": nrows, "out_mb": out_bytes / 1e6, "gpu_launches_per_call": launches,
        "gpu_parse_inferred": {"ms": ms_infer, "GBps_text": nbytes / ms_infer / 1e6, "rows_per_s": nrows / ms_infer * 1e3},
        "gpu_parse_typed": {"ms": ms_typed, "GBps_text": nbytes / ms_typed / 1e6, "rows_per_s": nrows / ms_typed * 1e3},
        "gpu_e2e_from_pinned_host": {"ms": ms_e2e, "GBps_text": nbytes / ms_e2e / 1e6, "rows_per_s": nrows / ms_e2e * 1e3},
        "gpu_tfrecord_encode": {"rows": len(sub), "features": len(list(sub.keys())), "ms": ms_enc, "out_mb": enc_bytes / 1e6,
                                "rows_per_s": len(sub) / ms_enc * 1e3, "GBps_out": enc_bytes / ms_enc / 1e6},
        "gpu_tfrecord_decode": {"rows": len(sub), "features": len(list(sub.keys())), "in_mb": enc_bytes / 1e6, "ms": ms_dec,
                                "rows_per_s": len(sub) / ms_dec * 1e3, "GBps_in": enc_bytes / ms_dec / 1e6, "ms_without_data_crc": ms_dec_nocrc},
        "pyarrow_host": {"s": t_arrow, "GBps_text": nbytes / t_arrow / 1e9, "rows_per_s": nrows / t_arrow, "threads": os.cpu_count()},
        "python_row_loop": {"rows": sample_rows, "s": t_py, "rows_per_s": sample_rows / t_py, "threads": 1},
    }))


if __name__ == "__main__":
    main()
