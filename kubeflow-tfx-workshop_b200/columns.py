"""Columnar containers for the Chicago-taxi schema (north_star subsystem 1: ingest).

The reference feeds tf.Transform row-major Python objects -- a list of dicts whose values are
0/1-element lists (``notebooks/06_Airflow_Feature_Analysis.ipynb:237-256``), typed by the schema
as int64 / float32 / bytes (``notebooks/04_TensorFlow_Model_Analysis.ipynb:222-240``), where an
empty CSV field is a missing value.  Here a dataset is a ``Table`` of Arrow-layout columns:

* ``NumericColumn``: ``values`` (float32 or int64, 16-byte aligned) + optional validity bitmap
  (uint8, LSB first, 1 = present);
* ``StringColumn``: ``offsets`` (int32, n+1) + ``pool`` (uint8, with TXF_POOL_SLACK readable
  bytes past the end) + optional validity bitmap.  Invariant: a missing row has zero length.

Tensors may live on the host (optionally pinned) or on the GPU; kernels only take GPU tables.
PyTorch is used purely as the buffer owner.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np
import torch

POOL_SLACK = 64
FLOAT, INT, BYTES = "float32", "int64", "bytes"


def _pack_valid(valid: Optional[np.ndarray]) -> Optional[np.ndarray]:
    if valid is None or bool(np.all(valid)):
        return None
    return np.packbits(np.asarray(valid, bool), bitorder="little")


def _unpack_valid(bits: Optional[torch.Tensor], n: int) -> np.ndarray:
    if bits is None:
        return np.ones(n, bool)
    return np.unpackbits(bits.cpu().numpy(), bitorder="little")[:n].astype(bool)


class NumericColumn:
    kind = "numeric"

    def __init__(self, values: torch.Tensor, valid: Optional[torch.Tensor] = None):
        if values.dtype not in (torch.float32, torch.int64):
            raise TypeError(f"numeric columns are float32 or int64, got {values.dtype}")
        if values.dim() != 1:
            raise ValueError("numeric column values must be rank 1")
        if valid is not None and (valid.dtype != torch.uint8 or valid.numel() < (values.numel() + 7) // 8):
            raise ValueError("validity must be a uint8 bitmap of ceil(n/8) bytes")
        self.values = values.contiguous()
        self.valid = valid
        self.dtype = FLOAT if values.dtype == torch.float32 else INT

    def __len__(self):
        return self.values.numel()

    @property
    def device(self):
        return self.values.device

    @staticmethod
    def from_numpy(values, valid=None) -> "NumericColumn":
        values = np.ascontiguousarray(values)
        if values.dtype == np.float64:
            values = values.astype(np.float32)
        elif values.dtype.kind in "iu" and values.dtype != np.int64:
            values = values.astype(np.int64)
        v = _pack_valid(valid)
        return NumericColumn(torch.from_numpy(values.copy()), None if v is None else torch.from_numpy(v.copy()))

    def to(self, device, non_blocking=False) -> "NumericColumn":
        return NumericColumn(self.values.to(device, non_blocking=non_blocking),
                             None if self.valid is None else self.valid.to(device, non_blocking=non_blocking))

    def pin(self) -> "NumericColumn":
        return NumericColumn(self.values.pin_memory(), None if self.valid is None else self.valid.pin_memory())

    def numpy(self):
        """(values, valid bool[n]) on the host."""
        return self.values.cpu().numpy(), _unpack_valid(self.valid, len(self))

    def slice(self, lo: int, hi: int) -> "NumericColumn":
        """Rows [lo, hi); lo must be a multiple of 8 so the bitmap can be sliced by bytes."""
        if lo % 8:
            raise ValueError("slice start must be a multiple of 8 rows")
        valid = None if self.valid is None else self.valid[lo // 8:(hi + 7) // 8]
        return NumericColumn(self.values[lo:hi], valid)

    def nbytes(self):
        return self.values.numel() * self.values.element_size() + (0 if self.valid is None else self.valid.numel())


class StringColumn:
    kind = "string"
    dtype = BYTES

    def __init__(self, offsets: torch.Tensor, pool: torch.Tensor, valid: Optional[torch.Tensor] = None,
                 pool_bytes: Optional[int] = None):
        if offsets.dtype != torch.int32 or offsets.dim() != 1 or offsets.numel() < 1:
            raise TypeError("offsets must be a rank-1 int32 tensor of n+1 entries")
        if pool.dtype != torch.uint8:
            raise TypeError("pool must be uint8")
        self.offsets = offsets.contiguous()
        self.pool = pool.contiguous()  # capacity >= pool_bytes + POOL_SLACK
        self.valid = valid
        self.pool_bytes = int(pool_bytes if pool_bytes is not None else max(pool.numel() - POOL_SLACK, 0))
        if self.pool.numel() < self.pool_bytes + POOL_SLACK:
            raise ValueError(f"pool needs {POOL_SLACK} readable bytes of slack past its last byte")

    def __len__(self):
        return self.offsets.numel() - 1

    @property
    def device(self):
        return self.offsets.device

    @staticmethod
    def from_numpy(offsets, pool, valid=None) -> "StringColumn":
        offsets = np.ascontiguousarray(offsets, np.int32)
        pool = np.ascontiguousarray(pool, np.uint8)
        n = len(offsets) - 1
        if valid is not None:  # enforce the invariant: a missing row has zero length
            lens = np.diff(offsets)
            if np.any(lens[~np.asarray(valid, bool)] != 0):
                items = [bytes(pool[offsets[i]:offsets[i + 1]]) if valid[i] else None for i in range(n)]
                return StringColumn.from_list(items)
        nbytes = int(offsets[-1] - offsets[0])
        buf = np.zeros(nbytes + POOL_SLACK, np.uint8)
        buf[:nbytes] = pool[offsets[0]:offsets[-1]]
        v = _pack_valid(valid)
        return StringColumn(torch.from_numpy((offsets - offsets[0]).astype(np.int32)), torch.from_numpy(buf),
                            None if v is None else torch.from_numpy(v.copy()), nbytes)

    @staticmethod
    def from_list(items: Sequence[Optional[bytes]]) -> "StringColumn":
        valid = np.array([it is not None for it in items], bool)
        lens = np.array([0 if it is None else len(it) for it in items], np.int64)
        if lens.sum() >= 2 ** 31:
            raise ValueError("a string chunk holds at most 2^31-1 bytes; split the rows into chunks")
        offsets = np.zeros(len(items) + 1, np.int32)
        np.cumsum(lens, out=offsets[1:])
        pool = np.frombuffer(b"".join(it for it in items if it is not None), np.uint8)
        return StringColumn.from_numpy(offsets, pool, valid)

    def to(self, device, non_blocking=False) -> "StringColumn":
        return StringColumn(self.offsets.to(device, non_blocking=non_blocking), self.pool.to(device, non_blocking=non_blocking),
                            None if self.valid is None else self.valid.to(device, non_blocking=non_blocking), self.pool_bytes)

    def pin(self) -> "StringColumn":
        return StringColumn(self.offsets.pin_memory(), self.pool.pin_memory(),
                            None if self.valid is None else self.valid.pin_memory(), self.pool_bytes)

    def to_list(self) -> List[bytes]:
        """Filled byte strings on the host (missing -> b'')."""
        o = self.offsets.cpu().numpy()
        buf = self.pool.cpu().numpy().tobytes()
        return [buf[o[i]:o[i + 1]] for i in range(len(self))]

    def numpy(self):
        return self.offsets.cpu().numpy(), self.pool.cpu().numpy()[:self.pool_bytes], _unpack_valid(self.valid, len(self))

    def nbytes(self):
        return self.offsets.numel() * 4 + self.pool_bytes + (0 if self.valid is None else self.valid.numel())


Column = NumericColumn  # typing alias root


class Table:
    """Ordered mapping feature name -> column; all columns have the same number of rows."""

    def __init__(self, columns: Dict[str, object]):
        if not columns:
            raise ValueError("a table needs at least one column")
        lens = {len(c) for c in columns.values()}
        if len(lens) != 1:
            raise ValueError(f"columns disagree on the number of rows: {sorted(lens)}")
        self.columns = dict(columns)
        self.num_rows = lens.pop()

    def __len__(self):
        return self.num_rows

    def __getitem__(self, key):
        return self.columns[key]

    def __contains__(self, key):
        return key in self.columns

    def keys(self):
        return self.columns.keys()

    def items(self):
        return self.columns.items()

    @property
    def device(self):
        return next(iter(self.columns.values())).device

    def schema(self) -> Dict[str, str]:
        return {k: c.dtype for k, c in self.columns.items()}

    def to(self, device, non_blocking=False) -> "Table":
        return Table({k: c.to(device, non_blocking=non_blocking) for k, c in self.columns.items()})

    def pin(self) -> "Table":
        return Table({k: c.pin() for k, c in self.columns.items()})

    def nbytes(self) -> int:
        return sum(c.nbytes() for c in self.columns.values())

    def select(self, keys: Iterable[str]) -> "Table":
        return Table({k: self.columns[k] for k in keys})

    # ---- ingest ---------------------------------------------------------------------------
    @staticmethod
    def from_rows(rows: Sequence[dict], schema: Dict[str, str]) -> "Table":
        """Row-major ingest in the shape tft_beam receives (notebooks/06_...:237-256): every value is
        a scalar, a 0/1-element list, or None; missing keys are missing values."""
        cols = {}
        for name, dtype in schema.items():
            cells = []
            for r in rows:
                v = r.get(name)
                if isinstance(v, (list, tuple, np.ndarray)):
                    if len(v) > 1:
                        raise ValueError(f"feature {name!r}: at most one value per row (taxi_utils.py:92-93)")
                    v = v[0] if len(v) else None
                cells.append(v)
            valid = np.array([c is not None for c in cells], bool)
            if dtype == BYTES:
                cols[name] = StringColumn.from_list(
                    [None if c is None else (c if isinstance(c, bytes) else str(c).encode()) for c in cells])
            elif dtype == FLOAT:
                cols[name] = NumericColumn.from_numpy(np.array([0.0 if c is None else c for c in cells], np.float32), valid)
            elif dtype == INT:
                cols[name] = NumericColumn.from_numpy(np.array([0 if c is None else c for c in cells], np.int64), valid)
            else:
                raise TypeError(f"feature {name!r}: unknown dtype {dtype!r}")
        return Table(cols)

    @staticmethod
    def from_arrow(tbl) -> "Table":
        """Zero-copy-ish ingest of a pyarrow Table (float32/float64/ints/binary/string columns)."""
        import pyarrow as pa

        cols = {}
        for name in tbl.column_names:
            arr = tbl.column(name).combine_chunks()
            if isinstance(arr, pa.ChunkedArray):
                arr = arr.chunk(0) if arr.num_chunks else pa.array([], arr.type)
            valid = None if arr.null_count == 0 else np.asarray(arr.is_valid())
            t = arr.type
            if pa.types.is_null(t):
                cols[name] = StringColumn.from_list([None] * len(arr))
            elif pa.types.is_floating(t):
                cols[name] = NumericColumn.from_numpy(np.asarray(arr.fill_null(0).cast(pa.float32())), valid)
            elif pa.types.is_integer(t):
                cols[name] = NumericColumn.from_numpy(np.asarray(arr.fill_null(0).cast(pa.int64())), valid)
            elif pa.types.is_binary(t) or pa.types.is_string(t) or pa.types.is_large_binary(t) or pa.types.is_large_string(t):
                arr = arr.cast(pa.binary())
                bufs = arr.buffers()
                off = np.frombuffer(bufs[1], np.int32, len(arr) + 1, arr.offset * 4)
                pool = np.frombuffer(bufs[2], np.uint8) if bufs[2] is not None else np.zeros(0, np.uint8)
                cols[name] = StringColumn.from_numpy(off, pool, valid)
            else:
                raise TypeError(f"column {name!r}: unsupported arrow type {t}")
        return Table(cols)

    @staticmethod
    def from_csv(path: str, schema: Optional[Dict[str, str]] = None) -> "Table":
        """CSV ingest with the typing CsvExampleGen + SchemaGen apply (taxi_pipeline.py:70-79): all
        present fields integers -> int64, any decimal -> float32, otherwise bytes; an all-missing
        column -> bytes; an empty field is a missing value."""
        import pyarrow as pa
        import pyarrow.csv as pacsv

        raw = pacsv.read_csv(path, convert_options=pacsv.ConvertOptions(strings_can_be_null=True, null_values=[""]))
        fields = []
        for name in raw.column_names:
            t = raw.schema.field(name).type
            want = (schema or {}).get(name)
            if want == BYTES or (want is None and not (pa.types.is_integer(t) or pa.types.is_floating(t) or pa.types.is_null(t))):
                fields.append((name, pa.binary()))
            elif want == FLOAT or (want is None and pa.types.is_floating(t)):
                fields.append((name, pa.float32()))
            elif want == INT or (want is None and pa.types.is_integer(t)):
                fields.append((name, pa.int64()))
            else:
                fields.append((name, pa.null()))
        cols = {}
        for name, t in fields:
            col = raw.column(name)
            cols[name] = col if pa.types.is_null(t) else col.cast(t)
        return Table.from_arrow(pa.table(cols))
