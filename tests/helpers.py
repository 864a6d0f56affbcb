"""Shared test utilities: golden fixture access and Table <-> oracle column conversion."""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import tft_oracle as O  # noqa: E402


def load_taxi_golden():
    """-> (oracle columns of the 15 000 bundled rows, expected outputs, expected state)."""
    z = np.load(os.path.join(GOLDEN, "taxi_15000.npz"))
    n = int(z["n"])
    names = sorted({k.split("/")[1] for k in z.files if k.startswith("in/")})
    cols = {}
    for k in names:
        valid = np.unpackbits(z[f"in/{k}/valid"], bitorder="little")[:n].astype(bool)
        if f"in/{k}/offsets" in z.files:
            cols[k] = O.StrCol(z[f"in/{k}/offsets"], z[f"in/{k}/pool"], valid)
        else:
            cols[k] = O.NumCol(z[f"in/{k}/values"], valid)
    out = {}
    for k in z.files:
        if k.startswith("out/"):
            parts = k.split("/")
            if len(parts) == 2:
                out[parts[1]] = z[k]
            elif parts[2] == "offsets":
                sc = O.StrCol(z[k], z[f"out/{parts[1]}/pool"])
                out[parts[1]] = sc.to_list()
    state = {"zscore": {}, "vocab": {}, "boundaries": {}}
    for k in z.files:
        p = k.split("/")
        if p[0] != "state":
            continue
        if p[1] == "zscore":
            state["zscore"][p[2]] = (np.float32(z[k][0]), np.float32(z[k][1]))
        elif p[1] == "boundaries":
            state["boundaries"][p[2]] = z[k]
        elif p[1] == "vocab" and p[3] == "offsets":
            state["vocab"][p[2]] = O.StrCol(z[k], z[f"state/vocab/{p[2]}/pool"]).to_list()
    return cols, out, state


def table_from_oracle(cols, ktw):
    t = {}
    for k, c in cols.items():
        if isinstance(c, O.StrCol):
            t[k] = ktw.StringColumn.from_numpy(c.offsets, c.pool, c.valid)
        else:
            t[k] = ktw.NumericColumn.from_numpy(c.values, c.valid)
    return ktw.Table(t)


def oracle_from_table(table):
    cols = {}
    for k, c in table.items():
        if c.kind == "string":
            off, pool, valid = c.numpy()
            cols[k] = O.StrCol(off, pool, valid)
        else:
            v, valid = c.numpy()
            cols[k] = O.NumCol(v, valid)
    return cols


def load_example_preprocessing_fn(ktw):
    """examples/taxi_preprocessing.py imported with the tf / tft stand-ins bound."""
    with ktw.install_shims():
        spec = importlib.util.spec_from_file_location("taxi_preprocessing", os.path.join(ROOT, "examples", "taxi_preprocessing.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    return mod


def zscore_close(got, want, mean, var):
    """north_star tolerance: 1e-6 relative, plus a few ulp(mean)/std absolute (the
    quotient is ill-conditioned where x ~ mean)."""
    got = np.asarray(got, np.float32)
    want = np.asarray(want, np.float32)
    std = float(np.sqrt(np.float32(var))) or 1.0
    atol = 4 * float(np.spacing(np.float32(abs(mean)) + np.float32(1))) / std
    return np.allclose(got, want, rtol=1e-6, atol=atol, equal_nan=True)
