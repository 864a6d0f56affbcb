"""tf.Example / TFRecord encode of a GPU table (egress half).

The reference materialises transformed rows with ``ExampleProtoCoder.encode`` +
``WriteToTFRecord`` (``notebooks/03a_TensorFlow_Transform_Advanced.ipynb:445-455``) and reads the
gzip-compressed TFRecord files back in the trainer (``airflow-dags/taxi_utils.py:79-83,260-281``).
``encode_tfrecord`` serialises every row of a GPU-resident ``Table`` into one framed record on the
device (``csrc/example.cu``); ``write_tfrecord`` copies the stream to a file, optionally through
the host's gzip (compression is host-side plumbing, the records are produced on the GPU).
"""
from __future__ import annotations

import gzip
from typing import Optional, Sequence, Tuple

import torch

from . import _native as nat
from .columns import Table


def encode_tfrecord(table: Table, keys: Optional[Sequence[str]] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (uint8 device tensor with the framed records back to back, int64 device offsets [n+1]).
    Features are written in name order (protobuf's deterministic map order)."""
    if table.device.type != "cuda":
        raise RuntimeError("encode_tfrecord runs on the GPU: call table.to('cuda') first")
    names = sorted(keys if keys is not None else table.keys(), key=lambda k: k.encode("utf-8"))
    if not names:
        raise ValueError("no features to encode")
    if len(names) > nat.TXF_EX_MAX_FEATURES:
        raise NotImplementedError(f"at most {nat.TXF_EX_MAX_FEATURES} features per Example")
    n, dev = len(table), table.device
    stream = torch.cuda.current_stream().cuda_stream
    feats = (nat.TxfExampleFeature * len(names))()
    blob = b""
    for i, k in enumerate(names):
        col, nb = table[k], k.encode("utf-8")
        f = feats[i]
        f.name_off, f.name_len = len(blob), len(nb)
        blob += nb
        f.valid = None if col.valid is None else col.valid.data_ptr()
        if col.kind == "string":
            f.kind, f.values, f.pool = nat.TXF_EX_BYTES, col.offsets.data_ptr(), col.pool.data_ptr()
        else:
            f.kind, f.values = (nat.TXF_F32 if col.dtype == "float32" else nat.TXF_I64), col.values.data_ptr()
    names_dev = torch.tensor(list(blob), dtype=torch.uint8, device=dev)
    offsets = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    if n == 0:
        return torch.empty(0, dtype=torch.uint8, device=dev), offsets
    sizes = torch.empty(n, dtype=torch.int64, device=dev)
    nat.call("txf_example_sizes", feats, len(names), n, sizes.data_ptr(), stream)
    torch.cumsum(sizes, 0, out=offsets[1:])
    total = int(offsets[-1].item())
    out = torch.empty(total, dtype=torch.uint8, device=dev)
    nat.call("txf_example_encode", feats, len(names), names_dev.data_ptr(), n, offsets.data_ptr(), out.data_ptr(), stream)
    return out, offsets


def write_tfrecord(table: Table, path: str, keys: Optional[Sequence[str]] = None, compress: bool = False) -> int:
    """Encode on the GPU and write one TFRecord file (gzip-compressed like the reference's when
    `compress`); returns the number of uncompressed bytes."""
    data, _ = encode_tfrecord(table, keys)
    host = data.cpu().numpy().tobytes()
    opener = gzip.open if compress else open
    with opener(path, "wb") as f:
        f.write(host)
    return len(host)
