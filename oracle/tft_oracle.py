"""oracle/tft_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (numpy + one small C helper) of the analyze-and-transform pass that
tf.Transform runs for ``preprocessing_fn`` in the reference
(``/root/reference/airflow-dags/taxi_utils.py:86-145``).

The arithmetic of that path does not live in the reference repo: it is delegated to
tensorflow-transform 0.13.0 / TensorFlow 1.13 / Apache Beam (pin:
``notebooks/03_TensorFlow_Transform.ipynb:47``), none of which is installed or installable in
this image.  This file restates their published behaviour and follows
the reference's own call sites for structure.

PARITY STATUS (see DESIGN.md "Oracle"):
  * pinned by the reference itself: most-frequent-first, 0-based vocabulary ids and
    mean-centering (prose known-answers, ``notebooks/03_TensorFlow_Transform.ipynb:102-105,
    227-233``);
  * pinned by independent implementations/test expectations: FarmHash Fingerprint64 for
    lengths 0..32 (TensorFlow string_to_hash_bucket_fast expectations for <=16; abseil
    CityHash64 compiled in pyarrow for <=32, ``tests/golden/make_farmhash_golden.py``);
  * PARITY UNPINNED: z-score arithmetic order, the ``''``/newline vocabulary filter, the
    (count, bytes)-descending tie-break, quantile boundary rule, Fingerprint64 for len > 32.
    No real tft/TF is available to pin these; each rule is isolated in one function below so
    it can be flipped if a real tft ever shows up.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package never does.
"""
from __future__ import annotations

import collections
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_MASK = (1 << 64) - 1

# ----------------------------------------------------------------------------------------------
# Feature keys of the taxi pipeline (reference constants, taxi_utils.py:32-63)
# ----------------------------------------------------------------------------------------------
DENSE_FLOAT_KEYS = ["trip_miles", "fare", "trip_seconds"]  # taxi_utils.py:40
BUCKET_KEYS = ["pickup_latitude", "pickup_longitude", "dropoff_latitude", "dropoff_longitude"]  # :45-48
VOCAB_KEYS = ["payment_type", "company"]  # :56-59
CATEGORICAL_KEYS = [
    "trip_start_hour", "trip_start_day", "trip_start_month", "pickup_census_tract",
    "dropoff_census_tract", "pickup_community_area", "dropoff_community_area",
]  # :34-38
LABEL_KEY, FARE_KEY = "tips", "fare"  # :62-63
FEATURE_BUCKET_COUNT, VOCAB_SIZE, OOV_SIZE = 10, 1000, 10  # :43, :51, :54
EMPTY_VOCAB_SENTINEL = b"49d0cd50-04bb-48c0-bc6f-5b575dce351a"  # Appendix A.2 rule 6


def xf(key):
    """Output naming rule, taxi_utils.py:66-67."""
    return key + "_xf"


# ----------------------------------------------------------------------------------------------
# Column containers used by the oracle (host twins of the device layout)
# ----------------------------------------------------------------------------------------------
class NumCol:
    """values (float32 or int64) + validity (bool per row)."""

    def __init__(self, values, valid=None):
        self.values = np.ascontiguousarray(values)
        self.valid = np.ones(len(self.values), bool) if valid is None else np.asarray(valid, bool)

    def __len__(self):
        return len(self.values)

    def slice(self, lo, hi):
        return NumCol(self.values[lo:hi], self.valid[lo:hi])


class StrCol:
    """Arrow-style string column: int32 offsets[n+1] + uint8 pool + validity."""

    def __init__(self, offsets, pool, valid=None):
        self.offsets = np.ascontiguousarray(offsets, np.int32)
        self.pool = np.ascontiguousarray(pool, np.uint8)
        n = len(self.offsets) - 1
        self.valid = np.ones(n, bool) if valid is None else np.asarray(valid, bool)

    @staticmethod
    def from_list(items):
        """items: list of bytes or None."""
        valid = np.array([it is not None for it in items], bool)
        lens = np.array([0 if it is None else len(it) for it in items], np.int64)
        offsets = np.zeros(len(items) + 1, np.int32)
        np.cumsum(lens, out=offsets[1:])
        pool = np.frombuffer(b"".join(it for it in items if it is not None), np.uint8)
        return StrCol(offsets, pool, valid)

    def __len__(self):
        return len(self.offsets) - 1

    def to_list(self):
        """Filled byte strings (missing -> b'')."""
        buf = self.pool.tobytes()
        o = self.offsets
        v = self.valid
        return [buf[o[i]:o[i + 1]] if v[i] else b"" for i in range(len(self))]

    def slice(self, lo, hi):
        o = self.offsets[lo:hi + 1]
        return StrCol(o - o[0], self.pool[o[0]:o[-1]], self.valid[lo:hi])


# ----------------------------------------------------------------------------------------------
# a1  _fill_in_missing  (taxi_utils.py:86-103)
# ----------------------------------------------------------------------------------------------
def fill_in_missing(col):
    """Missing -> '' for strings, 0 for numerics; dtype preserved (taxi_utils.py:98-103)."""
    if isinstance(col, StrCol):
        return col.to_list()
    out = col.values.copy()
    out[~col.valid] = 0
    return out


# ----------------------------------------------------------------------------------------------
# a2  tft.scale_to_z_score  (call site taxi_utils.py:116-119; semantics Appendix A.1)
# ----------------------------------------------------------------------------------------------
def moments(x_filled):
    """Analyzer state (count, mean, M2) in fp64 over the fp32-rounded filled column."""
    x = np.asarray(x_filled).astype(np.float32).astype(np.float64)
    n = x.size
    if n == 0:
        return (0.0, 0.0, 0.0)
    mean = float(np.sum(x) / n)
    m2 = float(np.sum((x - mean) ** 2))
    return (float(n), mean, m2)


def merge_moments(a, b):
    """Chan et al. pairwise merge -- how partial states from shards/ranks are combined."""
    na, ma, sa = a
    nb, mb, sb = b
    if na == 0:
        return b
    if nb == 0:
        return a
    n = na + nb
    d = mb - ma
    return (n, ma + d * (nb / n), sa + sb + d * d * (na * nb / n))


def finalize_moments(state):
    """-> (mean_f32, var_f32): population variance (divide by N), rounded to fp32."""
    n, mean, m2 = state
    var = m2 / n if n > 0 else 0.0
    return np.float32(mean), np.float32(var)


def z_score(x_filled, mean_f32, var_f32):
    """(f32(x) - mean) / sqrt(var), all fp32; var == 0 -> x - mean (Appendix A.1 rule 3)."""
    x = np.asarray(x_filled).astype(np.float32)
    num = (x - np.float32(mean_f32)).astype(np.float32)
    den = np.sqrt(np.float32(var_f32), dtype=np.float32)
    if den != 0:
        return (num / den).astype(np.float32)
    return num


# ----------------------------------------------------------------------------------------------
# FarmHash Fingerprint64 (Appendix A.3) -- pure Python (small cases) + C helper (bulk)
# ----------------------------------------------------------------------------------------------
_K0, _K1, _K2 = 0xC3A5C85C97CB3127, 0xB492B66FBE98F273, 0x9AE16A3B2F90404F


def _f64(s, i):
    return int.from_bytes(s[i:i + 8], "little")


def _f32(s, i):
    return int.from_bytes(s[i:i + 4], "little")


def _rot(v, s):
    return v if s == 0 else ((v >> s) | (v << (64 - s))) & _MASK


def _smix(v):
    return v ^ (v >> 47)


def _h16(u, v, m):
    a = ((u ^ v) * m) & _MASK
    a ^= a >> 47
    b = ((v ^ a) * m) & _MASK
    b ^= b >> 47
    return (b * m) & _MASK


def _weak(s, p, a, b):
    w, x, y, z = _f64(s, p), _f64(s, p + 8), _f64(s, p + 16), _f64(s, p + 24)
    a = (a + w) & _MASK
    b = _rot((b + a + z) & _MASK, 21)
    c = a
    a = (a + x + y) & _MASK
    b = (b + _rot(a, 44)) & _MASK
    return (a + z) & _MASK, (b + c) & _MASK


def fingerprint64(s: bytes) -> int:
    """farmhashna::Hash64, little-endian fetches, arithmetic mod 2**64 (Appendix A.3)."""
    n = len(s)
    if n == 0:
        return _K2
    if n <= 3:
        y = (s[0] + (s[n >> 1] << 8)) & 0xFFFFFFFF
        z = (n + (s[n - 1] << 2)) & 0xFFFFFFFF
        return (_smix(((y * _K2) ^ (z * _K0)) & _MASK) * _K2) & _MASK
    if n <= 7:
        m = (_K2 + 2 * n) & _MASK
        return _h16((n + (_f32(s, 0) << 3)) & _MASK, _f32(s, n - 4), m)
    if n <= 16:
        m = (_K2 + 2 * n) & _MASK
        a = (_f64(s, 0) + _K2) & _MASK
        b = _f64(s, n - 8)
        c = (_rot(b, 37) * m + a) & _MASK
        d = ((_rot(a, 25) + b) * m) & _MASK
        return _h16(c, d, m)
    if n <= 32:
        m = (_K2 + 2 * n) & _MASK
        a = (_f64(s, 0) * _K1) & _MASK
        b = _f64(s, 8)
        c = (_f64(s, n - 8) * m) & _MASK
        d = (_f64(s, n - 16) * _K2) & _MASK
        return _h16((_rot((a + b) & _MASK, 43) + _rot(c, 30) + d) & _MASK,
                    (a + _rot((b + _K2) & _MASK, 18) + c) & _MASK, m)
    if n <= 64:
        m = (_K2 + 2 * n) & _MASK
        a = (_f64(s, 0) * _K2) & _MASK
        b = _f64(s, 8)
        c = (_f64(s, n - 8) * m) & _MASK
        d = (_f64(s, n - 16) * _K2) & _MASK
        y = (_rot((a + b) & _MASK, 43) + _rot(c, 30) + d) & _MASK
        z = _h16(y, (a + _rot((b + _K2) & _MASK, 18) + c) & _MASK, m)
        e = (_f64(s, 16) * m) & _MASK
        f = _f64(s, 24)
        g = ((y + _f64(s, n - 32)) * m) & _MASK
        h = ((z + _f64(s, n - 24)) * m) & _MASK
        return _h16((_rot((e + f) & _MASK, 43) + _rot(g, 30) + h) & _MASK,
                    (e + _rot((f + a) & _MASK, 18) + g) & _MASK, m)
    x = 81
    y = (81 * _K1 + 113) & _MASK
    z = (_smix((y * _K2 + 113) & _MASK) * _K2) & _MASK
    v0 = v1 = w0 = w1 = 0
    x = (x * _K2 + _f64(s, 0)) & _MASK
    end = ((n - 1) // 64) * 64
    p = 0
    while True:
        x = (_rot((x + y + v0 + _f64(s, p + 8)) & _MASK, 37) * _K1) & _MASK
        y = (_rot((y + v1 + _f64(s, p + 48)) & _MASK, 42) * _K1) & _MASK
        x ^= w1
        y = (y + v0 + _f64(s, p + 40)) & _MASK
        z = (_rot((z + w0) & _MASK, 33) * _K1) & _MASK
        v0, v1 = _weak(s, p, (v1 * _K1) & _MASK, (x + w0) & _MASK)
        w0, w1 = _weak(s, p + 32, (z + w1) & _MASK, (y + _f64(s, p + 16)) & _MASK)
        z, x = x, z
        p += 64
        if p == end:
            break
    m = (_K1 + ((z & 0xFF) << 1)) & _MASK
    p = n - 64
    w0 = (w0 + ((n - 1) & 63)) & _MASK
    v0 = (v0 + w0) & _MASK
    w0 = (w0 + v0) & _MASK
    x = (_rot((x + y + v0 + _f64(s, p + 8)) & _MASK, 37) * m) & _MASK
    y = (_rot((y + v1 + _f64(s, p + 48)) & _MASK, 42) * m) & _MASK
    x ^= (w1 * 9) & _MASK
    y = (y + v0 * 9 + _f64(s, p + 40)) & _MASK
    z = (_rot((z + w0) & _MASK, 33) * m) & _MASK
    v0, v1 = _weak(s, p, (v1 * m) & _MASK, (x + w0) & _MASK)
    w0, w1 = _weak(s, p + 32, (z + w1) & _MASK, (y + _f64(s, p + 16)) & _MASK)
    z, x = x, z
    return _h16((_h16(v0, w0, m) + (_smix(y) * _K0) + z) & _MASK, (_h16(v1, w1, m) + x) & _MASK, m)


_CLIB = None


def build_c(force=False):
    """Compile oracle/farmhash_oracle.c -> oracle/_build/liboracle.so (gcc). Returns path."""
    out_dir = os.path.join(_HERE, "_build")
    so = os.path.join(out_dir, "liboracle.so")
    src = os.path.join(_HERE, "farmhash_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        os.makedirs(out_dir, exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", so, src])
    return so


def _clib():
    global _CLIB
    if _CLIB is None:
        lib = ctypes.CDLL(build_c())
        lib.oracle_fingerprint64.restype = ctypes.c_uint64
        lib.oracle_fingerprint64.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
        lib.oracle_fingerprint64_bulk.restype = None
        lib.oracle_fingerprint64_bulk.argtypes = [ctypes.c_void_p] * 2 + [ctypes.c_int64, ctypes.c_void_p]
        _CLIB = lib
    return _CLIB


def fingerprint64_c(s: bytes) -> int:
    return int(_clib().oracle_fingerprint64(s, len(s)))


def fingerprint64_bulk(offsets, pool):
    """uint64[n] fingerprints of an offsets/pool string column (C helper)."""
    offsets = np.ascontiguousarray(offsets, np.int32)
    pool = np.ascontiguousarray(pool, np.uint8)
    if pool.size == 0:
        pool = np.zeros(1, np.uint8)
    n = len(offsets) - 1
    out = np.empty(n, np.uint64)
    _clib().oracle_fingerprint64_bulk(offsets.ctypes.data, pool.ctypes.data, n, out.ctypes.data)
    return out


# ----------------------------------------------------------------------------------------------
# a3  tft.compute_and_apply_vocabulary (call site taxi_utils.py:121-126; Appendix A.2)
# ----------------------------------------------------------------------------------------------
def vocab_counts(strings):
    """Analyzer partial state: exact occurrence count per distinct byte string (rule 1)."""
    return collections.Counter(strings)


def merge_vocab_counts(a, b):
    out = collections.Counter(a)
    out.update(b)
    return out


def vocab_keep(term: bytes) -> bool:
    """Rule 2: the empty string and any term containing \\n or \\r never enter a vocabulary."""
    return len(term) > 0 and b"\n" not in term and b"\r" not in term


def finalize_vocabulary(counts, top_k=None, frequency_threshold=None):
    """Rules 2-6: filter, order by (count, bytes) DESCENDING, keep the first top_k."""
    items = [(c, t) for t, c in counts.items() if vocab_keep(t)]
    if frequency_threshold is not None:
        items = [it for it in items if it[0] >= frequency_threshold]
    items.sort(reverse=True)
    if top_k is not None:
        items = items[:top_k]
    vocab = [t for _, t in items]
    if not vocab:
        vocab = [EMPTY_VOCAB_SENTINEL]
    return vocab


def apply_vocabulary(strings, vocab, num_oov_buckets=0, default_value=-1, fingerprints=None):
    """Rule 7: id in vocab, else len(vocab) + Fingerprint64 % num_oov_buckets, int64."""
    table = {t: i for i, t in enumerate(vocab)}
    out = np.empty(len(strings), np.int64)
    for i, s in enumerate(strings):
        j = table.get(s)
        if j is None:
            if num_oov_buckets > 0:
                fp = int(fingerprints[i]) if fingerprints is not None else fingerprint64(s)
                j = len(vocab) + fp % num_oov_buckets
            else:
                j = default_value
        out[i] = j
    return out


# ----------------------------------------------------------------------------------------------
# a4  tft.bucketize (call site taxi_utils.py:128-130; Appendix A.4)
# ----------------------------------------------------------------------------------------------
def canonical_f32(x):
    """-0.0 -> +0.0 and every NaN -> the positive quiet NaN, so order statistics are unique."""
    x = np.asarray(x, np.float32) + np.float32(0.0)
    return np.where(np.isnan(x), np.float32(np.nan), x).astype(np.float32)


def order_key(x):
    """Order-preserving uint32 key of canonical fp32 (NaN sorts last)."""
    u = canonical_f32(x).view(np.uint32)
    nan = np.isnan(x)
    k = np.where(u & 0x80000000, ~u, u | 0x80000000).astype(np.uint32)
    return np.where(nan, np.uint32(0xFFFFFFFF), k)


def quantile_ranks(n, num_buckets):
    """1-based target ranks ceil(i*n/num_buckets), i = 1..num_buckets-1 (Appendix A.4 rule 5)."""
    i = np.arange(1, num_buckets, dtype=np.int64)
    r = (i * n + num_buckets - 1) // num_buckets
    return np.clip(r, 1, max(n, 1))


def quantile_boundaries(x_filled, num_buckets):
    """Exact order statistics at quantile_ranks, consecutive duplicates dropped (rules 3, 5)."""
    if not isinstance(num_buckets, (int, np.integer)) or isinstance(num_buckets, bool):
        raise TypeError("num_buckets must be an int")
    if num_buckets < 1:
        raise ValueError("num_buckets must be >= 1")
    x = canonical_f32(np.asarray(x_filled).astype(np.float32))
    n = x.size
    if n == 0 or num_buckets == 1:
        return np.zeros(0, np.float32)
    ranks = quantile_ranks(n, num_buckets)
    # np.partition on keys gives the same order statistics as a full sort
    keys = order_key(x)
    part = np.partition(keys, np.unique(ranks - 1))
    sel = part[ranks - 1]
    sel = sel[np.concatenate([[True], sel[1:] != sel[:-1]])]
    return key_to_f32(sel)


def key_to_f32(k):
    k = np.asarray(k, np.uint32)
    u = np.where(k & 0x80000000, k & 0x7FFFFFFF, ~k).astype(np.uint32)
    return u.view(np.float32)


def boundary_rank_ok(x_filled, boundaries, num_buckets, epsilon=None):
    """Acceptance test of rule 5 for an APPROXIMATE boundary list of full length: each b_i
    must be a data value whose rank interval meets [(i/nb - eps) n, (i/nb + eps) n]."""
    eps = min(1.0 / num_buckets, 0.01) if epsilon is None else epsilon
    keys = np.sort(order_key(np.asarray(x_filled).astype(np.float32)))
    n = keys.size
    for i, b in enumerate(np.asarray(boundaries, np.float32), start=1):
        kb = order_key(np.array([b], np.float32))[0]
        lo = np.searchsorted(keys, kb, "left")
        hi = np.searchsorted(keys, kb, "right")
        if hi == lo:
            return False
        if hi < (i / num_buckets - eps) * n or lo + 1 > (i / num_buckets + eps) * n:
            return False
    return True


def bucketize(x_filled, boundaries):
    """std::upper_bound: id = number of leading boundaries b with not (x < b); NaN -> last."""
    x = np.asarray(x_filled).astype(np.float32)
    b = np.asarray(boundaries, np.float32)
    out = np.searchsorted(b, x, side="right").astype(np.int64)
    out[np.isnan(x)] = len(b)
    return out


# ----------------------------------------------------------------------------------------------
# a6  label  (taxi_utils.py:135-143; Appendix A.5)
# ----------------------------------------------------------------------------------------------
def tips_label(fare_filled, tips_filled, ratio=0.2):
    """isnan(fare) ? 0 : int64(tips > fare * f32(ratio)); the product is rounded to fp32."""
    fare = np.asarray(fare_filled, np.float32)
    tips = np.asarray(tips_filled, np.float32)
    thr = (fare * np.float32(ratio)).astype(np.float32)
    return np.where(np.isnan(fare), 0, (tips > thr)).astype(np.int64)


# ----------------------------------------------------------------------------------------------
# Whole path: analyze (shardable) + transform, following preprocessing_fn line by line
# ----------------------------------------------------------------------------------------------
def analyze_partial(cols):
    """Per-shard analyzer state for preprocessing_fn (phase 0 of AnalyzeDataset)."""
    st = {"n": len(next(iter(cols.values())))}
    st["moments"] = {k: moments(fill_in_missing(cols[k])) for k in DENSE_FLOAT_KEYS}
    st["vocab"] = {k: vocab_counts(fill_in_missing(cols[k])) for k in VOCAB_KEYS}
    # exact quantiles are merged by concatenating keys (the CPU twin of summing histograms)
    st["qkeys"] = {k: np.sort(order_key(fill_in_missing(cols[k]).astype(np.float32))) for k in BUCKET_KEYS}
    return st


def merge_partials(parts):
    out = {"n": sum(p["n"] for p in parts), "moments": {}, "vocab": {}, "qkeys": {}}
    for k in DENSE_FLOAT_KEYS:
        acc = (0.0, 0.0, 0.0)
        for p in parts:  # rank order, like the device allgather+merge
            acc = merge_moments(acc, p["moments"][k])
        out["moments"][k] = acc
    for k in VOCAB_KEYS:
        acc = collections.Counter()
        for p in parts:
            acc.update(p["vocab"][k])
        out["vocab"][k] = acc
    for k in BUCKET_KEYS:
        out["qkeys"][k] = np.sort(np.concatenate([p["qkeys"][k] for p in parts]))
    return out


def finalize_state(merged, num_buckets=FEATURE_BUCKET_COUNT, top_k=VOCAB_SIZE):
    state = {"zscore": {}, "vocab": {}, "boundaries": {}}
    for k in DENSE_FLOAT_KEYS:
        state["zscore"][k] = finalize_moments(merged["moments"][k])
    for k in VOCAB_KEYS:
        state["vocab"][k] = finalize_vocabulary(merged["vocab"][k], top_k=top_k)
    for k in BUCKET_KEYS:
        keys = merged["qkeys"][k]
        n = keys.size
        if n == 0 or num_buckets == 1:
            state["boundaries"][k] = np.zeros(0, np.float32)
            continue
        sel = keys[quantile_ranks(n, num_buckets) - 1]
        sel = sel[np.concatenate([[True], sel[1:] != sel[:-1]])]
        state["boundaries"][k] = key_to_f32(sel)
    return state


def analyze(cols, num_buckets=FEATURE_BUCKET_COUNT, top_k=VOCAB_SIZE):
    return finalize_state(merge_partials([analyze_partial(cols)]), num_buckets, top_k)


def transform(cols, state, num_oov_buckets=OOV_SIZE, use_c_hash=True):
    """The per-row phase of preprocessing_fn (taxi_utils.py:115-145)."""
    out = {}
    for k in DENSE_FLOAT_KEYS:  # :116-119
        mean, var = state["zscore"][k]
        out[xf(k)] = z_score(fill_in_missing(cols[k]), mean, var)
    for k in VOCAB_KEYS:  # :121-126
        strings = fill_in_missing(cols[k])
        fps = fingerprint64_bulk(cols[k].offsets, cols[k].pool) if use_c_hash else None
        if fps is not None:  # missing rows hash as '' regardless of what the pool holds
            fps = np.where(cols[k].valid, fps, np.uint64(_K2))
        out[xf(k)] = apply_vocabulary(strings, state["vocab"][k], num_oov_buckets, fingerprints=fps)
    for k in BUCKET_KEYS:  # :128-130
        out[xf(k)] = bucketize(fill_in_missing(cols[k]), state["boundaries"][k])
    for k in CATEGORICAL_KEYS:  # :132-133
        out[xf(k)] = fill_in_missing(cols[k])
    fare = fill_in_missing(cols[FARE_KEY])  # :136
    tips = fill_in_missing(cols[LABEL_KEY])  # :137
    out[xf(LABEL_KEY)] = tips_label(fare, tips, 0.2)  # :138-143
    return out


def analyze_and_transform(cols):
    state = analyze(cols)
    return transform(cols, state), state


# =================================================================================================
# Sibling analyzers / mappers of the workshop notebooks.  TEST INFRASTRUCTURE like
# everything above.  Semantics [UPSTREAM-RECALLED, tft 0.13 mappers.py / analyzers.py]:
#   tft.mean(x)             -> mean of all values, in x's dtype (float32 here: fp64 sum, one rounding)
#   tft.scale_to_0_1(x)     == scale_by_min_max(x, 0, 1): where(min < max, (x - min) / (max - min), 0.5),
#                              then * (output_max - output_min) + output_min; every step in float32
#   tft.vocabulary(x, ...)  == the analyzer half of compute_and_apply_vocabulary
# Pinned by the reference: notebooks/03_TensorFlow_Transform.ipynb:227-231 (x=[1,2,3], y=[1,2,3],
# s=[hello,world,hello] -> [-1,0,1], [0,.5,1], [0,1,0], [-0,0,1]); see notebook03_example().
# Unpinned: behaviour for NaN input (here: NaN cells never become min or max).
# =================================================================================================
def min_max(x_filled):
    """(min, max) as float32 of the filled values; (+inf, -inf) for an empty column."""
    x = np.asarray(x_filled).astype(np.float32)
    x = x[~np.isnan(x)]
    if x.size == 0:
        return np.float32(np.inf), np.float32(-np.inf)
    return np.float32(x.min()), np.float32(x.max())


def merge_min_max(a, b):
    return np.float32(min(a[0], b[0])), np.float32(max(a[1], b[1]))


def mean_f32(x_filled):
    """tft.mean: the same fp64-accumulated mean the z-score uses, rounded once to float32."""
    return finalize_moments(moments(x_filled))[0]


def scale_by_min_max(x_filled, mn, mx, output_min=0.0, output_max=1.0):
    x = np.asarray(x_filled).astype(np.float32)
    mn, mx = np.float32(mn), np.float32(mx)
    omin, omax = np.float32(output_min), np.float32(output_max)
    with np.errstate(all="ignore"):
        if mn < mx:
            r = ((x - mn).astype(np.float32) / np.float32(mx - mn)).astype(np.float32)
        else:
            r = np.full(x.shape, 0.5, np.float32)
        return ((r * np.float32(omax - omin)).astype(np.float32) + omin).astype(np.float32)


_ARITH = {"add": np.add, "sub": np.subtract, "mul": np.multiply, "div": np.divide}


def arith_f32(op, a, b):
    """One float32 operation per row; scalars broadcast."""
    with np.errstate(all="ignore"):
        return _ARITH[op](np.asarray(a, np.float32), np.asarray(b, np.float32)).astype(np.float32)


def notebook03_example(x, y, s):
    """The preprocessing_fn of notebooks/03_TensorFlow_Transform.ipynb:141-157 on dense columns."""
    x = np.asarray(x, np.float32)
    y = np.asarray(y, np.float32)
    x_centered = arith_f32("sub", x, mean_f32(x))
    y_normalized = scale_by_min_max(y, *min_max(y))
    vocab = finalize_vocabulary(vocab_counts(s))
    return {
        "x_centered": x_centered,
        "y_normalized": y_normalized,
        "s_integerized": apply_vocabulary(s, vocab, num_oov_buckets=0, default_value=-1),
        "x_centered_times_y_normalized": arith_f32("mul", x_centered, y_normalized),
    }
