"""GPU tests (-m gpu) of the tf.Example / TFRecord egress: records produced on the
device must be byte for byte what the protobuf runtime's deterministic serialisation plus the
TFRecord framing give for the same rows (oracle/example_oracle.py)."""
import gzip
import os
import sys

import numpy as np
import pytest
import torch

import helpers
from helpers import O

sys.path.insert(0, helpers.ROOT)
from oracle import example_oracle as E  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ktw():
    import kubeflow_tfx_workshop_b200 as k
    assert torch.cuda.is_available()
    return k


def rows_of(cols):
    """oracle columns -> list of {feature: (kind, value | None)}"""
    n = len(next(iter(cols.values())))
    per = {}
    for k, c in cols.items():
        if isinstance(c, O.StrCol):
            per[k] = ("bytes", c.to_list(), c.valid)
        else:
            per[k] = ("float32" if c.values.dtype == np.float32 else "int64", c.values.tolist(), c.valid)
    return [{k: (kind, vals[i] if valid[i] else None) for k, (kind, vals, valid) in per.items()} for i in range(n)]


def test_crc32c_known_answer():
    assert E.crc32c(b"123456789") == 0xE3069283  # published check value of CRC-32C (Castagnoli)


@pytest.mark.parametrize("n", [1, 2, 255, 256, 3001])
def test_records_equal_protobuf_deterministic_serialisation(ktw, n):
    rng = np.random.default_rng(n)
    words = [b"", b"Cash", b"Credit Card", bytes(range(256)), b"x" * 300, "ünï".encode()]
    ints = rng.integers(-2 ** 63, 2 ** 63 - 1, n, dtype=np.int64)
    ints[: min(n, 8)] = np.array([0, 1, -1, 127, 128, 2 ** 63 - 1, -2 ** 63, 300])[: min(n, 8)]
    cols = {
        "zeta": O.NumCol(rng.standard_normal(n).astype(np.float32), rng.random(n) > 0.2),
        "alpha_int": O.NumCol(ints, rng.random(n) > 0.1),
        "text": O.StrCol.from_list([words[i] for i in rng.integers(0, len(words), n)]),
        "Beta": O.NumCol(np.array([np.nan, np.inf, -0.0, 1e-45][: min(n, 4)] + [1.0] * max(n - 4, 0), np.float32), None),
        "small": O.NumCol(rng.integers(0, 1010, n).astype(np.int64), None),
        "a-very-long-feature-name-" + "n" * 120: O.NumCol(rng.integers(0, 10, n).astype(np.int64), None),
    }
    keep = rng.random(n) > 0.15  # some missing strings (a missing row has zero length)
    cols["text"] = O.StrCol.from_list([w if k else None for w, k in zip(cols["text"].to_list(), keep)])
    table = helpers.table_from_oracle(cols, ktw).to("cuda")
    data, offsets = ktw.encode_tfrecord(table)
    torch.cuda.synchronize()
    got = data.cpu().numpy().tobytes()
    want = E.encode_rows(rows_of(cols))
    assert len(got) == len(want) == int(offsets[-1])
    assert got == want
    decoded = E.decode_tfrecord(got)  # also re-verifies both CRCs of every record
    assert len(decoded) == n
    ex = decoded[0].features.feature
    assert sorted(ex.keys()) == sorted(cols.keys())


def test_transformed_taxi_rows_round_trip(ktw, tmp_path):
    """The transformed taxi table (the thing the reference writes for the trainer) -> gzip TFRecord
    file -> parsed back with protobuf: every value equals the device column."""
    from kubeflow_tfx_workshop_b200 import synth
    n = 2000
    fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
    table = synth.taxi_table(n, seed=3, backend="numpy").to("cuda")
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    out, _ = eng.analyze_and_transform(table)
    path = str(tmp_path / "transformed-00000-of-00001.gz")
    nbytes = ktw.write_tfrecord(out, path, compress=True)
    with gzip.open(path, "rb") as f:
        buf = f.read()
    assert len(buf) == nbytes
    decoded = E.decode_tfrecord(buf)
    assert len(decoded) == n
    for key, col in out.items():
        if col.kind == "string":
            want, _, valid = col.to_list(), None, col.numpy()[2]
            for i in (0, 1, n // 2, n - 1):
                got = list(decoded[i].features.feature[key].bytes_list.value)
                assert got == ([want[i]] if valid[i] else []), (key, i)
            continue
        vals, valid = col.numpy()
        assert valid.all()
        if vals.dtype == np.float32:
            got = np.array([d.features.feature[key].float_list.value[0] for d in decoded], np.float32)
            assert np.array_equal(got.view(np.uint32), vals.view(np.uint32)), key
        else:
            got = np.array([d.features.feature[key].int64_list.value[0] for d in decoded], np.int64)
            assert np.array_equal(got, vals), key
    # subset of features, uncompressed
    data, offsets = ktw.encode_tfrecord(out, keys=["tips_xf", "fare_xf"])
    sub = E.decode_tfrecord(data.cpu().numpy().tobytes())
    assert sorted(sub[5].features.feature.keys()) == ["fare_xf", "tips_xf"]


def test_egress_edge_cases(ktw):
    empty = ktw.Table({"a": ktw.NumericColumn(torch.empty(0, dtype=torch.float32, device="cuda"))})
    data, offsets = ktw.encode_tfrecord(empty)
    assert data.numel() == 0 and offsets.tolist() == [0]
    with pytest.raises(RuntimeError, match="runs on the GPU"):
        ktw.encode_tfrecord(ktw.Table({"a": ktw.NumericColumn(torch.zeros(3))}))
    many = ktw.Table({f"f{i}": ktw.NumericColumn(torch.zeros(2, device="cuda")) for i in range(65)})
    with pytest.raises(NotImplementedError, match="at most 64"):
        ktw.encode_tfrecord(many)
