"""Generate tests/golden/taxi_15000.npz from the reference's bundled CSV (BASELINE config 1).

Run in the authoring container only (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_taxi_golden.py

The fixture holds (a) the 15 000 parsed input columns in the device layout (values + validity
for numerics; int32 offsets + byte pool + validity for strings) and (b) the oracle's analyzer
state and all 17 transformed outputs.  The real tft/Beam output cannot be produced here (tft,
TF and Beam are not installable), so (b) is the oracle's output, anchored against the
independently computed facts in SURVEY.md Appendix B by tests/test_oracle_taxi.py.

Column typing follows what CsvExampleGen + SchemaGen infer: a column whose
present fields are all integers -> int64, any decimal -> float32, otherwise bytes; an
all-missing column -> bytes; an empty CSV field is a missing value.
"""
import csv
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import tft_oracle as o  # noqa: E402

CSV = "/root/reference/airflow-dags/data/taxi_data/data.csv"


def _is_int(s):
    try:
        int(s)
        return True
    except ValueError:
        return False


def _is_float(s):
    try:
        float(s)
        return True
    except ValueError:
        return False


def load_csv(path):
    with open(path, newline="") as f:
        rows = list(csv.reader(f))
    header, rows = rows[0], rows[1:]
    cols = {}
    for j, name in enumerate(header):
        raw = [r[j] for r in rows]
        present = [s for s in raw if s != ""]
        valid = np.array([s != "" for s in raw], bool)
        if present and all(_is_int(s) for s in present):
            cols[name] = o.NumCol(np.array([int(s) if s else 0 for s in raw], np.int64), valid)
        elif present and all(_is_float(s) for s in present):
            cols[name] = o.NumCol(np.array([float(s) if s else 0.0 for s in raw], np.float32), valid)
        else:
            cols[name] = o.StrCol.from_list([s.encode() if s else None for s in raw])
    return cols


def main():
    cols = load_csv(CSV)
    out, state = o.analyze_and_transform(cols)
    blob = {}
    for k, c in cols.items():
        if isinstance(c, o.StrCol):
            blob[f"in/{k}/offsets"] = c.offsets
            blob[f"in/{k}/pool"] = c.pool
        else:
            blob[f"in/{k}/values"] = c.values
        blob[f"in/{k}/valid"] = np.packbits(c.valid, bitorder="little")
    blob["n"] = np.int64(len(cols["fare"]))
    for k, v in out.items():
        if isinstance(v, list):  # filled bytes column
            sc = o.StrCol.from_list(v)
            blob[f"out/{k}/offsets"] = sc.offsets
            blob[f"out/{k}/pool"] = sc.pool
        else:
            blob[f"out/{k}"] = v
    for k, (m, v) in state["zscore"].items():
        blob[f"state/zscore/{k}"] = np.array([m, v], np.float32)
    for k, vocab in state["vocab"].items():
        sc = o.StrCol.from_list(vocab)
        blob[f"state/vocab/{k}/offsets"] = sc.offsets
        blob[f"state/vocab/{k}/pool"] = sc.pool
    for k, b in state["boundaries"].items():
        blob[f"state/boundaries/{k}"] = b
    np.savez_compressed(os.path.join(HERE, "taxi_15000.npz"), **blob)
    print("rows", int(blob["n"]), "tips positives", int(out["tips_xf"].sum()))


if __name__ == "__main__":
    main()


# tests/golden/taxi_head200.csv: header + first 200 data lines of the same file, byte for byte
# (`head -201 data.csv`), for the GPU CSV-ingest test: parsing it must reproduce rows [0, 200) of
# the columns stored in taxi_15000.npz.
