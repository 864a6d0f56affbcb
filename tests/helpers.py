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


# ---- Python model of the long-term fingerprint mixer (csrc/vocab.cuh KeyMixer), used to construct
# adversarial inputs: a window that zeroes one factor of each lane.
_M32 = 0xFFFFFFFF


class KeyMixerModel:
    """`fold=True`: the current mixer (x*y + (x<<32|y)); `fold=False`: the round-1 mixer (bare product),
    whose lanes collapse to zero when a factor is zero."""

    def __init__(self, n, fold=True):
        self.a1 = (n * 0x9E3779B1 + 0x7F4A7C15) & _M32
        self.b1 = (n ^ 0x632BE5AB) & _M32
        self.a2 = ((n ^ 0xD6E8FEB8) * 0xC2B2AE3D) & _M32
        self.b2 = (n + 0x27D4EB2F) & _M32
        self.fold = fold

    def _mul(self, x, y):
        p = x * y + (((x << 32) | y) if self.fold else 0)
        return p & 0xFFFFFFFFFFFFFFFF

    def zeroing_window(self):
        """8 bytes that make x == 0 in both lanes for the current state."""
        lo = self.a1 ^ 0x53C5CA59
        hi = self.a2 ^ 0xA0761D65
        return (lo | (hi << 32)).to_bytes(8, "little")

    def add(self, w8: bytes):
        w = int.from_bytes(w8, "little")
        lo, hi = w & _M32, w >> 32
        p = self._mul(self.a1 ^ lo ^ 0x53C5CA59, self.b1 ^ hi ^ 0x74743C1B)
        q = self._mul(self.a2 ^ hi ^ 0xA0761D65, self.b2 ^ lo ^ 0xE7037ED1)
        self.a1, self.b1, self.a2, self.b2 = p & _M32, p >> 32, q & _M32, q >> 32

    def state(self):
        return (self.a1, self.b1, self.a2, self.b2)


def lane_zeroing_pair(n=32):
    """Two different n-byte terms (n = 32) whose second 8-byte window zeroes one factor of each mixer
    lane: under a bare-product mixer both reach the all-zero state and share every later window."""
    assert n == 32
    terms = []
    for k in range(256):  # first windows whose zeroing window holds no line break (such terms never enter a vocabulary)
        first = b"TERM%04d" % k
        m = KeyMixerModel(n, fold=False)
        m.add(first)
        second = m.zeroing_window()
        if b"\n" in second or b"\r" in second:
            continue
        terms.append(first + second + b"tail-16-bytes-xx")
        if len(terms) == 2:
            break
    states = []  # construction check: the round-1 mixer ends in the same state for both
    for t in terms:
        m = KeyMixerModel(n, fold=False)
        for i in range(0, 32, 8):
            m.add(t[i:i + 8])
        states.append(m.state())
    assert states[0] == states[1] and terms[0] != terms[1]
    return terms
