"""Synthetic Chicago-taxi-schema data (BASELINE.json configs 2-5), bit-reproducible on CPU and GPU.

Every cell is a pure function of (seed, column, global row): a splitmix64-style integer hash,
integer arithmetic, int->float conversions and IEEE fp32 multiplies only -- so the numpy twin
(`backend="numpy"`, used by the tests and the CPU baseline) and the torch-on-GPU generator
(`backend="torch"`, used to fill HBM for the benchmark) produce identical bytes, and a rank can
generate just its own row range of one global dataset.  Distributions follow the bundled CSV
(SURVEY.md Appendix B): heavy-tailed fare / miles / seconds, 65 % zero tips, ~800 lat/long
centroids with a skewed popularity and 3.2 % missing drop-offs, 7 payment types, 34 % missing
company over ~1000 names of 17-32 bytes.  ``trip_start_timestamp`` is not generated:
``preprocessing_fn`` never reads it (taxi_utils.py:106-145).

This is data plumbing (PyTorch ops), not part of the measured path.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .columns import NumericColumn, StringColumn, Table, POOL_SLACK

SEED = 20190501
_M64 = (1 << 64) - 1

TAXI_SCHEMA: Dict[str, str] = {
    "pickup_community_area": "int64", "fare": "float32", "trip_start_month": "int64", "trip_start_hour": "int64",
    "trip_start_day": "int64", "pickup_latitude": "float32", "pickup_longitude": "float32",
    "dropoff_latitude": "float32", "dropoff_longitude": "float32", "trip_miles": "float32",
    "pickup_census_tract": "bytes", "dropoff_census_tract": "int64", "payment_type": "bytes", "company": "bytes",
    "trip_seconds": "int64", "dropoff_community_area": "int64", "tips": "float32",
}
_COL_ID = {k: i for i, k in enumerate(TAXI_SCHEMA)}

PAYMENT_TYPES = [b"Cash", b"Credit Card", b"No Charge", b"Unknown", b"Dispute", b"Pcard", b"Prcard"]
_PAYMENT_THRESHOLDS = [6600, 9920, 9975, 9993, 9996, 9999]  # of 10000
N_CENTROIDS = 800
N_COMPANIES = 1000


def _s64(c: int) -> int:
    """Two's-complement view of a 64-bit constant (torch has no uint64 arithmetic)."""
    c &= _M64
    return c - (1 << 64) if c >= (1 << 63) else c


class _Np:
    name = "numpy"

    def __init__(self, device=None):
        pass

    def rows(self, lo, hi):
        return np.arange(lo, hi, dtype=np.uint64)

    def const(self, c):
        return np.uint64(c & _M64)

    def shr(self, a, s):
        return a >> np.uint64(s)

    def small(self, a):  # 31-bit draw as int64
        return (a >> np.uint64(33)).astype(np.int64)

    def f32(self, a):
        return a.astype(np.float32)

    def mulf(self, a, c):
        return (a * np.float32(c)).astype(np.float32)

    def where(self, c, a, b):
        return np.where(c, a, b)

    def take(self, table, idx):
        return table[idx]

    def table(self, arr):
        return np.asarray(arr)

    def i64(self, a):
        return np.asarray(a, np.int64)

    def cumsum(self, a):
        return np.cumsum(a)

    def repeat_rows(self, n, lens):
        return np.repeat(np.arange(n, dtype=np.int64), lens)

    def arange(self, n):
        return np.arange(n, dtype=np.int64)

    def zeros_u8(self, n):
        return np.zeros(n, np.uint8)

    def to_numpy(self, a):
        return np.asarray(a)


class _Th:
    name = "torch"

    def __init__(self, device):
        self.device = torch.device(device)

    def rows(self, lo, hi):
        return torch.arange(lo, hi, dtype=torch.int64, device=self.device)

    def const(self, c):
        return _s64(c)

    def shr(self, a, s):
        return (a >> s) & ((1 << (64 - s)) - 1)

    def small(self, a):
        return (a >> 33) & ((1 << 31) - 1)

    def f32(self, a):
        return a.to(torch.float32)

    def mulf(self, a, c):
        return a * float(np.float32(c))

    def where(self, c, a, b):
        if not torch.is_tensor(a):
            a = torch.full_like(b, a) if torch.is_tensor(b) else torch.tensor(a, device=self.device)
        if not torch.is_tensor(b):
            b = torch.full_like(a, b)
        return torch.where(c, a, b)

    def take(self, table, idx):
        return table[idx]

    def table(self, arr):
        return torch.from_numpy(np.array(arr)).to(self.device)

    def i64(self, a):
        return a.to(torch.int64)

    def cumsum(self, a):
        return torch.cumsum(a, 0)

    def repeat_rows(self, n, lens):
        return torch.repeat_interleave(torch.arange(n, dtype=torch.int64, device=self.device), lens)

    def arange(self, n):
        return torch.arange(n, dtype=torch.int64, device=self.device)

    def zeros_u8(self, n):
        return torch.zeros(n, dtype=torch.uint8, device=self.device)

    def to_numpy(self, a):
        return a.cpu().numpy()


def _mix(xp, rows, seed: int, stream: int):
    """splitmix64 of (seed, stream, row) -> 64-bit hash (wrapping arithmetic on both backends)."""
    z = rows * xp.const(0x9E3779B97F4A7C15) + xp.const(seed * 0xD1B54A32D192ED03 + stream * 0x8CB92BA72F3D8DD7 + 0x2545F4914F6CDD1D)
    z = (z ^ xp.shr(z, 30)) * xp.const(0xBF58476D1CE4E5B9)
    z = (z ^ xp.shr(z, 27)) * xp.const(0x94D049BB133111EB)
    return z ^ xp.shr(z, 31)


def _draw(xp, rows, seed, col: str, k: int = 0):
    """31-bit non-negative draw number k of a column."""
    return xp.small(_mix(xp, rows, seed, _COL_ID[col] * 16 + k))


def centroid_tables(seed: int = SEED) -> Tuple[np.ndarray, np.ndarray]:
    k = np.arange(N_CENTROIDS, dtype=np.uint64)
    a = _Np().small(_mix(_Np(), k, seed, 1001))
    b = _Np().small(_mix(_Np(), k, seed, 1002))
    lat = (np.float32(41.65) + (a % 40000).astype(np.float32) * np.float32(1e-5)).astype(np.float32)
    lon = (np.float32(-87.95) + (b % 45000).astype(np.float32) * np.float32(1e-5)).astype(np.float32)
    return lat, lon


def company_names() -> List[bytes]:
    words = [b"Taxi", b"Cab", b"Affiliation", b"Association", b"Management", b"Leasing", b"Medallion", b"Services",
             b"Dispatch", b"Chicago", b"Blue", b"Ribbon", b"Choice", b"Northwest", b"Elite", b"Star", b"Royal", b"Corp"]
    names = []
    for k in range(N_COMPANIES):
        length = 17 + (k * 13) % 16
        base = b"%04d - %05d " % ((k * 7919) % 10000, (k * 104729) % 100000)
        i = k
        while len(base) < length + 8:
            base += words[i % len(words)] + b" "
            i = i * 5 + 3
        names.append(base[:length])
    assert len(set(names)) == N_COMPANIES
    return names


def _string_table(xp, terms: List[bytes]):
    lens = np.array([len(t) for t in terms], np.int64)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    pool = np.frombuffer(b"".join(terms), np.uint8)
    return xp.table(lens), xp.table(offs), xp.table(pool)


def _build_strings(xp, idx, missing, tab):
    """idx: int64 term index per row; missing: bool per row -> (offsets int64 [n+1], pool uint8)."""
    tab_len, tab_off, tab_pool = tab
    n = int(idx.shape[0])
    lens = xp.where(missing, 0, xp.take(tab_len, idx))
    ends = xp.cumsum(lens)
    total = int(ends[-1]) if n else 0
    if total == 0:
        return ends, xp.zeros_u8(0)
    row_of_byte = xp.repeat_rows(n, lens)
    starts = ends - lens
    pos = xp.arange(total) - xp.take(starts, row_of_byte)
    src = xp.take(tab_off, xp.take(idx, row_of_byte)) + pos
    return ends, xp.take(tab_pool, src)


def _pack_bits(present_np: np.ndarray) -> Optional[np.ndarray]:
    if present_np.all():
        return None
    return np.packbits(present_np, bitorder="little")


def _skewed_index(xp, rows, seed, col, k, n_items):
    a = _draw(xp, rows, seed, col, k) % n_items
    b = _draw(xp, rows, seed, col, k + 1) % n_items
    return (a * b) // n_items


def taxi_columns(xp, lo: int, hi: int, seed: int = SEED):
    """Raw per-column arrays of rows [lo, hi) on backend xp: dict name -> (values|('str', ends, pool), present|None)."""
    rows = xp.rows(lo, hi)
    out = {}
    lat_np, lon_np = centroid_tables(seed)
    lat_t, lon_t = xp.table(lat_np), xp.table(lon_np)

    def tail(col, k, common_mod, tail_mod, scale):
        u, v, w = _draw(xp, rows, seed, col, k), _draw(xp, rows, seed, col, k + 1), _draw(xp, rows, seed, col, k + 2)
        return xp.where(u % 100 == 0, v % tail_mod, w % common_mod) * scale

    fare_cents = xp.where(_draw(xp, rows, seed, "fare", 0) % 100 == 0, _draw(xp, rows, seed, "fare", 1) % 70008,
                          325 + _draw(xp, rows, seed, "fare", 2) % 4000)
    out["fare"] = (xp.mulf(xp.f32(fare_cents), 0.01), None)
    out["trip_miles"] = (xp.mulf(xp.f32(tail("trip_miles", 0, 300, 17101, 1)), 0.1), None)
    out["trip_seconds"] = (xp.i64(tail("trip_seconds", 0, 60, 1441, 60)), _draw(xp, rows, seed, "trip_seconds", 3) % 2500 != 0)
    tips_cents = xp.where(_draw(xp, rows, seed, "tips", 0) % 100 < 65, 0, _draw(xp, rows, seed, "tips", 1) % 4701)
    out["tips"] = (xp.mulf(xp.f32(tips_cents), 0.01), None)

    ip = _skewed_index(xp, rows, seed, "pickup_latitude", 0, N_CENTROIDS)
    idp = _skewed_index(xp, rows, seed, "dropoff_latitude", 0, N_CENTROIDS)
    drop_present = _draw(xp, rows, seed, "dropoff_latitude", 2) % 1000 >= 32
    out["pickup_latitude"] = (xp.take(lat_t, ip), None)
    out["pickup_longitude"] = (xp.take(lon_t, ip), None)
    out["dropoff_latitude"] = (xp.take(lat_t, idp), drop_present)
    out["dropoff_longitude"] = (xp.take(lon_t, idp), drop_present)

    out["trip_start_hour"] = (xp.i64(_draw(xp, rows, seed, "trip_start_hour") % 24), None)
    out["trip_start_day"] = (xp.i64(_draw(xp, rows, seed, "trip_start_day") % 7 + 1), None)
    out["trip_start_month"] = (xp.i64(_draw(xp, rows, seed, "trip_start_month") % 12 + 1), None)
    out["pickup_community_area"] = (xp.i64(_draw(xp, rows, seed, "pickup_community_area") % 77 + 1), None)
    out["dropoff_community_area"] = (xp.i64(_draw(xp, rows, seed, "dropoff_community_area") % 77 + 1),
                                     _draw(xp, rows, seed, "dropoff_community_area", 1) % 1000 >= 34)
    out["dropoff_census_tract"] = (xp.i64(17031000000 + _draw(xp, rows, seed, "dropoff_census_tract") % 1000000),
                                   _draw(xp, rows, seed, "dropoff_census_tract", 1) % 100 >= 28)

    r = _draw(xp, rows, seed, "payment_type") % 10000
    pidx = (r >= _PAYMENT_THRESHOLDS[0]) * 1
    for t in _PAYMENT_THRESHOLDS[1:]:
        pidx = pidx + (r >= t) * 1
    pidx = xp.i64(pidx)
    never = pidx < 0
    out["payment_type"] = (("str",) + _build_strings(xp, pidx, never, _string_table(xp, PAYMENT_TYPES)), None)
    cidx = _skewed_index(xp, rows, seed, "company", 0, N_COMPANIES)
    cmiss = _draw(xp, rows, seed, "company", 2) % 100 < 34
    out["company"] = (("str",) + _build_strings(xp, cidx, cmiss, _string_table(xp, company_names())), ~cmiss)
    always = pidx >= 0
    out["pickup_census_tract"] = (("str",) + _build_strings(xp, pidx, always, _string_table(xp, PAYMENT_TYPES)), never)
    return out


def _assemble(xp, cols, n: int, device) -> Table:
    """Raw arrays -> Table (torch tensors on `device`; numpy backend -> host tensors)."""
    table = {}
    for name in TAXI_SCHEMA:
        val, present = cols[name]
        valid = None
        if present is not None:
            bits = _pack_bits(xp.to_numpy(present).astype(bool)) if n else None
            valid = None if bits is None else torch.from_numpy(bits).to(device)
        if isinstance(val, tuple):
            _, ends, pool = val
            total = int(ends[-1]) if n else 0
            if total >= 2 ** 31:
                raise ValueError("string pool exceeds 2^31-1 bytes; generate fewer rows per chunk")
            if xp.name == "torch":
                offsets = torch.zeros(n + 1, dtype=torch.int32, device=device)
                offsets[1:] = ends.to(torch.int32)
                buf = torch.zeros(total + POOL_SLACK, dtype=torch.uint8, device=device)
                buf[:total] = pool
            else:
                offsets = torch.zeros(n + 1, dtype=torch.int32)
                offsets[1:] = torch.from_numpy(np.asarray(ends, np.int32))
                buf = torch.zeros(total + POOL_SLACK, dtype=torch.uint8)
                buf[:total] = torch.from_numpy(np.ascontiguousarray(pool))
                offsets, buf = offsets.to(device), buf.to(device)
            table[name] = StringColumn(offsets, buf, valid, total)
        else:
            t = val if xp.name == "torch" else torch.from_numpy(np.ascontiguousarray(val)).to(device)
            table[name] = NumericColumn(t.contiguous(), valid)
    return Table(table)


def taxi_table(n: int, row0: int = 0, seed: int = SEED, backend: str = "torch", device="cuda", piece: int = 1 << 23) -> Table:
    """Rows [row0, row0+n) of the synthetic taxi dataset as a Table on `device`.

    Large tables are generated in pieces of `piece` rows and concatenated, to bound temporaries."""
    xp = _Th(device) if backend == "torch" else _Np()
    dev = device if backend == "torch" else "cpu"
    if n <= piece:
        return _assemble(xp, taxi_columns(xp, row0, row0 + n, seed), n, dev)
    parts = [taxi_table(min(piece, n - s), row0 + s, seed, backend, device, piece) for s in range(0, n, piece)]
    return concat_tables(parts)


def concat_tables(parts: List[Table]) -> Table:
    """Row-wise concatenation (piece sizes must be multiples of 8 rows except the last)."""
    out = {}
    n_before = np.cumsum([0] + [len(p) for p in parts])
    for name in parts[0].keys():
        cols = [p[name] for p in parts]
        any_valid = any(c.valid is not None for c in cols)
        valid = None
        if any_valid:
            for c, p in zip(cols[:-1], parts[:-1]):
                if len(p) % 8:
                    raise ValueError("only the last piece may have a row count that is not a multiple of 8")
            vs = [c.valid if c.valid is not None else torch.full(((len(c) + 7) // 8,), 255, dtype=torch.uint8, device=c.device)
                  for c in cols]
            valid = torch.cat(vs)
        if isinstance(cols[0], StringColumn):
            total = sum(c.pool_bytes for c in cols)
            if total >= 2 ** 31:
                raise ValueError("concatenated string pool exceeds 2^31-1 bytes; keep the pieces as separate chunks")
            dev = cols[0].device
            offsets = torch.zeros(int(n_before[-1]) + 1, dtype=torch.int32, device=dev)
            pool = torch.zeros(total + POOL_SLACK, dtype=torch.uint8, device=dev)
            base = 0
            for c, r0 in zip(cols, n_before[:-1]):
                offsets[int(r0) + 1: int(r0) + len(c) + 1] = c.offsets[1:] + base
                pool[base: base + c.pool_bytes] = c.pool[: c.pool_bytes]
                base += c.pool_bytes
            out[name] = StringColumn(offsets, pool, valid, total)
        else:
            out[name] = NumericColumn(torch.cat([c.values for c in cols]), valid)
    return Table(out)


# ------------------------------------------------------------------------------------------------
# stress workloads (BASELINE.json configs[3] and configs[4]); torch only, generated on the device
# ------------------------------------------------------------------------------------------------
def decimal_string_column(ids: torch.Tensor) -> StringColumn:
    """The decimal text of every int64 as a string column (what tf.as_string gives), built with the
    library's own formatting kernels.  Data plumbing for the stress workloads, outside any timed region."""
    from . import _native as nat
    n = ids.numel()
    dev = ids.device
    with torch.cuda.device(dev):
        s = torch.cuda.current_stream().cuda_stream
        lens = torch.empty(n, dtype=torch.int32, device=dev)
        nat.call("txf_i64_format_lengths", ids.data_ptr(), None, n, lens.data_ptr(), s)
        ends = torch.cumsum(lens, 0, dtype=torch.int64)
        total = int(ends[-1].item())
        if total >= 1 << 31:
            raise ValueError("string pool exceeds 2^31-1 bytes; generate fewer rows per chunk")
        offsets = torch.zeros(n + 1, dtype=torch.int32, device=dev)
        offsets[1:] = ends.to(torch.int32)
        pool = torch.zeros(total + POOL_SLACK, dtype=torch.uint8, device=dev)
        nat.call("txf_i64_format_write", ids.data_ptr(), None, n, offsets.data_ptr(), pool.data_ptr(), s)
    return StringColumn(offsets, pool, None, total)


def vocab_stress_table(n: int, row0: int = 0, n_distinct: int = 10_000_000, seed: int = SEED, device="cuda") -> Table:
    """Two string columns over `n_distinct` distinct terms each with a skewed popularity (term index =
    floor(D u^3)): `code` holds terms of at most 13 bytes (their own key), `name` terms of 18-20 bytes
    (fingerprint class).  Rows are a pure function of (seed, row), so ranks generate disjoint shards."""
    xp = _Th(device)
    rows = xp.rows(row0, row0 + n)
    cols = {}
    for name, stream, shift in (("code", 2001, 24), ("name", 2002, 1)):
        u = _mix(xp, rows, seed, stream)
        a = ((u >> 11) & ((1 << 53) - 1)).to(torch.float64) * (1.0 / (1 << 53))
        idx = (a * a * a * n_distinct).to(torch.int64).clamp_(max=n_distinct - 1)
        ids = _mix(xp, idx, seed, stream + 10)  # scrambled 64-bit id of the term
        ids = (ids >> shift) & ((1 << (63 - shift)) - 1) if shift > 1 else ids & ((1 << 62) - 1) | (1 << 61)
        cols[name] = decimal_string_column(ids.contiguous())
    return Table(cols)


def bucket_stress_table(n: int, row0: int = 0, seed: int = SEED, device="cuda") -> Table:
    """Eight float32 lat/long-style columns: four drawn from ~800 centroids with a skewed popularity and
    3.2 % missing cells (heavy duplicates, an atom at 0.0 after fill-missing), four fine-grained."""
    xp = _Th(device)
    rows = xp.rows(row0, row0 + n)
    lat_np, lon_np = centroid_tables(seed)
    cent = [xp.table(lat_np), xp.table(lon_np)]
    cols = {}
    for c in range(8):
        if c < 4:
            idx = _skewed_index(xp, rows, seed, "pickup_latitude", 4 * c + 3, N_CENTROIDS)
            x = (xp.take(cent[c & 1], idx) + float(c)).to(torch.float32)
            present = _draw(xp, rows, seed, "dropoff_latitude", 8 + c) % 1000 >= 32
            bits = torch.from_numpy(np.packbits(present.cpu().numpy(), bitorder="little")).to(device)
            cols[f"c{c}"] = NumericColumn(x.contiguous(), bits)
        else:
            u = _draw(xp, rows, seed, "pickup_longitude", 8 + c)  # 31-bit draw
            x = (u.to(torch.float32) * float(np.float32(2.0 ** -31)) * 0.5 - 87.9 + float(c)).to(torch.float32)
            cols[f"c{c}"] = NumericColumn(x.contiguous(), None)
    return Table(cols)
