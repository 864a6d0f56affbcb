"""CSV ingest on the GPU (north_star subsystem 1).

The reference gets its rows from TFX ``CsvExampleGen`` (``airflow-dags/taxi_pipeline.py:70-73``)
or from the schema-typed CSV loop of ``notebooks/04_TensorFlow_Model_Analysis.ipynb:222-240``:
every field becomes ``int(value)`` / ``float(value)`` / bytes, an empty field is a missing value,
and column types are: all present fields integers -> int64, any decimal -> float32, otherwise
bytes, nothing present -> bytes.  ``read_csv`` does the same with the text copied to the device
once and parsed there (``csrc/csv.cu``); the host only reads the header line and sizes buffers.
"""
from __future__ import annotations

import csv as _csv
from typing import Dict, List, Optional, Union

import numpy as np
import torch

from . import _native as nat
from .columns import BYTES, FLOAT, INT, NumericColumn, StringColumn, Table

_UINT64_MAX = (1 << 64) - 1


def _device_text(data: Union[bytes, bytearray, memoryview, np.ndarray, torch.Tensor], device) -> torch.Tensor:
    """Text on `device`, 16-byte aligned, zero-padded with at least 16 readable bytes past its end."""
    if isinstance(data, torch.Tensor):
        src = data.reshape(-1).view(torch.uint8)
    else:
        src = torch.from_numpy(np.frombuffer(data, dtype=np.uint8).copy() if not isinstance(data, np.ndarray) else data.view(np.uint8))
    n = src.numel()
    buf = torch.zeros(((n + 15) // 16 + 2) * 16, dtype=torch.uint8, device=device)
    buf[:n].copy_(src, non_blocking=True)
    return buf


def read_csv(source, schema: Optional[Dict[str, str]] = None, device="cuda", header: bool = True,
             column_names: Optional[List[str]] = None) -> Table:
    """CSV file (path), bytes, or a uint8 tensor -> GPU-resident Table.

    `schema` fixes column types ('int64' | 'float32' | 'bytes'); columns it does not name are
    inferred on the device with the rule above.  Raises NativeError naming the first bad line."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("read_csv parses on the GPU: it needs a CUDA device (Table.from_csv is the host-side reader)")
    if isinstance(source, str):
        source = np.fromfile(source, dtype=np.uint8)
    nbytes = source.numel() if isinstance(source, torch.Tensor) else len(source)
    if nbytes >= 1 << 31:
        raise ValueError("read_csv takes at most 2 GiB of text per call: split the file at a line boundary")
    text = _device_text(source, device)
    stream = torch.cuda.current_stream().cuda_stream

    # ---- lines ------------------------------------------------------------------------------
    bb = nat.query("txf_csv_block_bytes")
    nblocks = max((nbytes + bb - 1) // bb, 1)
    block_lines = torch.zeros(nblocks, dtype=torch.int32, device=device)
    nat.call("txf_csv_count_lines", text.data_ptr(), nbytes, block_lines.data_ptr(), stream)
    incl = torch.cumsum(block_lines, 0, dtype=torch.int64)
    block_base = incl - block_lines
    newlines = int(incl[-1].item())
    ends_with_newline = nbytes > 0 and int(text[nbytes - 1].item()) == 0x0A
    nlines = newlines + (0 if ends_with_newline or nbytes == 0 else 1)
    line_start = torch.zeros(nlines + 2, dtype=torch.int64, device=device)
    nat.call("txf_csv_line_starts", text.data_ptr(), nbytes, block_base.data_ptr(), line_start.data_ptr(), stream)
    line_start[nlines] = nbytes

    # ---- header (host: one line) ---------------------------------------------------------------
    first = 0
    if header:
        if nlines == 0:
            raise ValueError("empty CSV: no header line")
        end = int(line_start[1].item())
        names = next(_csv.reader([bytes(text[:end].cpu().numpy()).decode("utf-8").rstrip("\r\n")]))
        first = 1
    else:
        if not column_names:
            raise ValueError("header=False needs column_names")
        names = list(column_names)
    ncols, nrows = len(names), nlines - first
    if ncols > nat.TXF_CSV_MAX_COLS:
        raise NotImplementedError(f"at most {nat.TXF_CSV_MAX_COLS} CSV columns")
    status = torch.tensor([0, -1], dtype=torch.int64, device=device)  # (error bits, first bad line = UINT64_MAX)

    def check():
        st = status.cpu().tolist()
        bits = st[0]
        if bits:
            why = "; ".join(msg for b, msg in nat.CSV_ERRORS.items() if bits & b)
            raise nat.NativeError(f"CSV line {(st[1] & _UINT64_MAX) + 1}: {why}")

    # ---- column types ------------------------------------------------------------------------
    kinds = {}
    schema = dict(schema or {})
    if any(n not in schema for n in names):
        flags = torch.zeros(ncols, dtype=torch.int32, device=device)
        nat.call("txf_csv_classify", text.data_ptr(), line_start.data_ptr(), first, nrows, ncols, flags.data_ptr(), status.data_ptr(),
                 stream)
        check()
        for name, fl in zip(names, flags.cpu().tolist()):
            kinds[name] = BYTES if (fl & nat.CSV_SEEN_TEXT or not fl & nat.CSV_SEEN_VALUE) else FLOAT if fl & nat.CSV_SEEN_DECIMAL else INT
    kinds.update({k: v for k, v in schema.items() if k in names})

    # ---- parse ---------------------------------------------------------------------------------
    cols = (nat.TxfCsvColumn * ncols)()
    held = {}
    nwords = (nrows + 31) // 32
    for i, name in enumerate(names):
        kind = kinds[name]
        valid = torch.zeros(max(nwords, 1) * 4, dtype=torch.uint8, device=device)
        if kind == BYTES:
            values = torch.empty(max(nrows, 1), dtype=torch.int32, device=device)
            lens = torch.empty(max(nrows, 1), dtype=torch.int32, device=device)
            cols[i].kind, cols[i].lens = nat.CSV_BYTES, lens.data_ptr()
        else:
            values = torch.empty(max(nrows, 1), dtype=torch.int64 if kind == INT else torch.float32, device=device)
            lens = None
            cols[i].kind = nat.CSV_I64 if kind == INT else nat.CSV_F32
        cols[i].values, cols[i].valid = values.data_ptr(), valid.data_ptr()
        held[name] = (values, lens, valid)
    nat.call("txf_csv_parse", text.data_ptr(), line_start.data_ptr(), first, nrows, ncols, cols, status.data_ptr(), stream)
    check()

    # ---- assemble ------------------------------------------------------------------------------
    out = {}
    nvalid_bytes = (nrows + 7) // 8
    for name in names:
        values, lens, valid = held[name]
        valid = valid[:nvalid_bytes]
        if kinds[name] != BYTES:
            out[name] = NumericColumn(values[:nrows], valid)
            continue
        offsets = torch.zeros(nrows + 1, dtype=torch.int32, device=device)
        total = 0
        if nrows:
            ends = torch.cumsum(lens[:nrows], 0, dtype=torch.int64)
            total = int(ends[-1].item())
            offsets[1:] = ends.to(torch.int32)
        pool = torch.zeros(total + nat.TXF_POOL_SLACK, dtype=torch.uint8, device=device)
        nat.call("txf_csv_copy_strings", text.data_ptr(), values.data_ptr(), offsets.data_ptr(), nrows, pool.data_ptr(), stream)
        out[name] = StringColumn(offsets, pool, valid, total)
    return Table(out)


# ------------------------------------------------------------------------------------------------
# TFRecord / tf.Example decode on the GPU (csrc/example_decode.cu)
# ------------------------------------------------------------------------------------------------
def read_tfrecord(source, schema: Dict[str, str], device="cuda", compressed: Optional[bool] = None, verify_crc: bool = True) -> Table:
    """TFRecord file of tf.Example protos (path, bytes, or a uint8 tensor of the UNCOMPRESSED stream) ->
    GPU-resident Table with the features `schema` names ('int64' | 'float32' | 'bytes').

    This is what a TFX Transform executor is fed (``airflow-dags/taxi_pipeline.py:70-89``; the
    reference's reader is the gzip TFRecord reader of ``airflow-dags/taxi_utils.py:79-83``).  gunzip
    runs on the host (`compressed`: None = by the gzip magic); record framing, protobuf parsing and
    column building run on the device.  A feature that is absent from a record, or present with an
    empty value list, is a missing cell -- the same "empty string -> empty value list" convention as
    the reference's typed CSV loop (``notebooks/04_TensorFlow_Model_Analysis.ipynb:222-240``); more
    than one value in a cell raises (the reference requires at most one, ``taxi_utils.py:92-93``)."""
    import gzip
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("read_tfrecord decodes on the GPU: it needs a CUDA device")
    if not schema:
        raise ValueError("read_tfrecord needs the features to extract (name -> 'int64' | 'float32' | 'bytes')")
    if len(schema) > nat.TXF_EX_MAX_FEATURES:
        raise NotImplementedError(f"at most {nat.TXF_EX_MAX_FEATURES} features per Example")
    if isinstance(source, str):
        with open(source, "rb") as f:
            source = f.read()
    if not isinstance(source, torch.Tensor):
        raw = bytes(source)
        if compressed or (compressed is None and raw[:2] == b"\x1f\x8b"):
            raw = gzip.decompress(raw)
        source = raw
    nbytes = source.numel() if isinstance(source, torch.Tensor) else len(source)
    if nbytes >= 1 << 31:
        raise ValueError("read_tfrecord takes at most 2 GiB of records per call: split the file at a record boundary")
    data = _device_text(source, device)
    stream = torch.cuda.current_stream().cuda_stream

    # ---- record index (no sequential walk: every byte position is tested as a record start) ------
    bb = nat.query("txf_tfrecord_block_bytes")
    nblocks = max((nbytes + bb - 1) // bb, 1)
    counts = torch.zeros(nblocks, dtype=torch.int32, device=device)
    nat.call("txf_tfrecord_count_candidates", data.data_ptr(), nbytes, counts.data_ptr(), stream)
    incl = torch.cumsum(counts, 0, dtype=torch.int64)
    ncand = int(incl[-1].item())
    cand = torch.empty(max(ncand, 1), dtype=torch.int64, device=device)
    nat.call("txf_tfrecord_candidates", data.data_ptr(), nbytes, (incl - counts).contiguous().data_ptr(), cand.data_ptr(), stream)
    mark = torch.zeros(max(ncand, 1), dtype=torch.uint8, device=device)
    nat.call("txf_tfrecord_link", data.data_ptr(), cand.data_ptr(), ncand, mark.data_ptr(), stream)
    starts = cand[:ncand][mark[:ncand].bool()]
    n = int(starts.numel())
    starts = torch.cat([starts, torch.tensor([nbytes], dtype=torch.int64, device=device)])
    if nbytes and (n == 0 or int(starts[0].item()) != 0):
        raise nat.NativeError("TFRecord stream does not start with a record header")
    if n:  # every record must end where the next begins, the last one at the end of the stream
        lens = torch.zeros(n, dtype=torch.int64, device=device)
        for k in range(4):
            lens |= data[starts[:n] + k].to(torch.int64) << (8 * k)
        bad = torch.nonzero(starts[:n] + 16 + lens != starts[1:]).flatten()
        if bad.numel():
            raise nat.NativeError(f"TFRecord stream is corrupt after record {int(bad[0].item())} (length or header CRC)")

    # ---- parse -----------------------------------------------------------------------------------
    names = list(schema)
    feats = (nat.TxfExampleColumn * len(names))()
    blob = b""
    held = {}
    nwords = max((n + 31) // 32, 1)
    for i, k in enumerate(names):
        kind, nb = schema[k], k.encode("utf-8")
        f = feats[i]
        f.name_off, f.name_len = len(blob), len(nb)
        blob += nb
        valid = torch.zeros(nwords * 4, dtype=torch.uint8, device=device)
        if kind == BYTES:
            values = torch.zeros(max(n, 1), dtype=torch.int32, device=device)
            lens_t = torch.zeros(max(n, 1), dtype=torch.int32, device=device)
            f.kind, f.lens = nat.TXF_EX_BYTES, lens_t.data_ptr()
        elif kind in (INT, FLOAT):
            values = torch.zeros(max(n, 1), dtype=torch.int64 if kind == INT else torch.float32, device=device)
            lens_t = None
            f.kind = nat.TXF_I64 if kind == INT else nat.TXF_F32
        else:
            raise TypeError(f"feature {k!r}: unknown dtype {kind!r}")
        f.values, f.valid = values.data_ptr(), valid.data_ptr()
        held[k] = (values, lens_t, valid)
    names_dev = torch.tensor(list(blob) or [0], dtype=torch.uint8, device=device)
    status = torch.tensor([0, -1], dtype=torch.int64, device=device)
    nat.call("txf_example_decode", feats, len(names), names_dev.data_ptr(), data.data_ptr(), starts.data_ptr(), n, 1 if verify_crc else 0,
             status.data_ptr(), stream)
    st = status.cpu().tolist()
    if st[0]:
        why = "; ".join(msg for b, msg in nat.EX_ERRORS.items() if st[0] & b)
        raise nat.NativeError(f"TFRecord record {st[1] & _UINT64_MAX}: {why}")

    # ---- assemble --------------------------------------------------------------------------------
    out = {}
    nvalid_bytes = (n + 7) // 8
    for k in names:
        values, lens_t, valid = held[k]
        valid = valid[:nvalid_bytes]
        if schema[k] != BYTES:
            out[k] = NumericColumn(values[:n], valid)
            continue
        offsets = torch.zeros(n + 1, dtype=torch.int32, device=device)
        total = 0
        if n:
            ends = torch.cumsum(lens_t[:n], 0, dtype=torch.int64)
            total = int(ends[-1].item())
            if total >= 1 << 31:
                raise ValueError(f"feature {k!r}: more than 2 GiB of bytes in one call; split the file")
            offsets[1:] = ends.to(torch.int32)
        pool = torch.zeros(total + nat.TXF_POOL_SLACK, dtype=torch.uint8, device=device)
        nat.call("txf_gather_bytes", data.data_ptr(), values.data_ptr(), offsets.data_ptr(), n, pool.data_ptr(), stream)
        out[k] = StringColumn(offsets, pool, valid, total)
    return Table(out)
