"""GPU tests (-m gpu) of the CSV ingest: text parsed on the device must equal
what the reference's typed CSV loop produces field by field (int(value) / float32(float(value)) /
bytes, empty -> missing; notebooks/04_TensorFlow_Model_Analysis.ipynb:222-240)."""
import csv
import io
import os

import numpy as np
import pytest
import torch

import helpers
from helpers import O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ktw():
    import kubeflow_tfx_workshop_b200 as k
    assert torch.cuda.is_available()
    return k


def python_parse(text: bytes, kinds, header=True):
    """The reference-style row loop: csv module + int()/float() per field."""
    rows = list(csv.reader(io.StringIO(text.decode("utf-8"), newline="")))
    names = rows[0] if header else list(kinds)
    body = rows[1:] if header else rows
    out = {}
    for j, name in enumerate(names):
        vals = [r[j] for r in body]
        present = np.array([v != "" for v in vals], bool)
        if kinds[name] == "int64":
            out[name] = (np.array([int(v) if v != "" else 0 for v in vals], np.int64), present)
        elif kinds[name] == "float32":
            out[name] = (np.array([float(v) if v != "" else 0.0 for v in vals], np.float64).astype(np.float32), present)
        else:
            out[name] = ([v.encode("utf-8") for v in vals], present)
    return out


def check_table(table, want):
    assert list(table.keys()) == list(want.keys())
    for name, (vals, present) in want.items():
        col = table[name]
        if isinstance(vals, list):
            assert col.kind == "string" and col.to_list() == vals, name
            _, _, valid = col.numpy()
        else:
            got, valid = col.numpy()
            assert got.dtype == vals.dtype, (name, got.dtype)
            same = got.view(np.uint32) == vals.view(np.uint32) if vals.dtype == np.float32 else got == vals
            assert same.all(), (name, np.flatnonzero(~same)[:5], got[~same][:5], vals[~same][:5])
        assert np.array_equal(valid, present), name


def test_bundled_head_matches_golden_columns(ktw):
    """First 200 rows of the bundled taxi CSV: types inferred on the device, values equal to the
    golden columns (which were produced by the host-side reader from the same file)."""
    path = os.path.join(helpers.GOLDEN, "taxi_head200.csv")
    table = ktw.read_csv(path)
    cols, _, _ = helpers.load_taxi_golden()
    assert len(table) == 200
    for name, c in cols.items():
        got = table[name]
        if isinstance(c, O.StrCol):
            assert got.kind == "string" and got.to_list() == c.to_list()[:200], name
            assert np.array_equal(got.numpy()[2], c.valid[:200]), name
        else:
            v, valid = got.numpy()
            assert v.dtype == c.values.dtype, (name, v.dtype, c.values.dtype)
            assert np.array_equal(valid, c.valid[:200]), name
            assert np.array_equal(v[valid], c.values[:200][valid]), name
    with open(path, "rb") as f:
        text = f.read()
    check_table(table, python_parse(text, table.schema()))


def make_csv(n, seed, crlf=False, trailing_newline=True):
    rng = np.random.default_rng(seed)
    f32 = (rng.standard_normal(n) * 10.0 ** rng.integers(-6, 7, n)).astype(np.float32)
    f64 = rng.standard_normal(n) * 10.0 ** rng.integers(-30, 30, n)
    words = ["Cash", "Credit Card", 'say "hi"', "a,b", "", "Taxi Affiliation Services", "ünïcode", "x" * 70]
    lines = ["i,short,f32repr,f64repr,sci,s,allmissing,ints_with_gap"]
    for k in range(n):
        i = int(rng.integers(-2 ** 62, 2 ** 62)) if k % 7 else [0, -1, 2 ** 63 - 1, -2 ** 63, 7][k % 5]
        short = "%.2f" % (rng.random() * 700) if k % 11 else ""
        s = words[int(rng.integers(0, len(words)))]
        q = '"' + s.replace('"', '""') + '"' if (',' in s or '"' in s or k % 13 == 0) and s != "" else s
        sci = "%de%d" % (rng.integers(1, 10 ** 9), rng.integers(-20, 20)) if k % 3 else "%.3E" % (rng.random() * 1e5)
        gap = str(int(rng.integers(0, 78))) if k % 4 else ""
        lines.append(",".join([str(i), short, repr(float(f32[k])), repr(float(f64[k])), sci, q, "", gap]))
    eol = "\r\n" if crlf else "\n"
    text = eol.join(lines) + (eol if trailing_newline else "")
    return text.encode("utf-8")


@pytest.mark.parametrize("n,crlf,trailing", [(1, False, True), (255, False, False), (256, True, True), (257, False, True), (20_003, True, False),
                                              (300_000, False, True)])
def test_synthetic_text_matches_python_parse(ktw, n, crlf, trailing):
    """Every numeric spelling (17-digit reprs, fixed point, exponents, int64 extremes), quoted
    fields with commas and doubled quotes, empty fields, CRLF, missing final newline."""
    text = make_csv(n, seed=n, crlf=crlf, trailing_newline=trailing)
    table = ktw.read_csv(text)
    kinds = table.schema()
    one = n == 1  # row 0 leaves `short` and `ints_with_gap` empty: a column with no value at all is bytes
    assert kinds == {"i": "int64", "short": "bytes" if one else "float32", "f32repr": "float32", "f64repr": "float32", "sci": "float32",
                     "s": "bytes", "allmissing": "bytes", "ints_with_gap": "bytes" if one else "int64"}
    assert len(table) == n
    check_table(table, python_parse(text, kinds))


def test_schema_override_headerless_and_errors(ktw):
    nat = ktw._native
    text = b"1,2.5,x\n3,,y\n"
    t = ktw.read_csv(text, schema={"a": "float32", "c": "bytes"}, header=False, column_names=["a", "b", "c"])
    assert t.schema() == {"a": "float32", "b": "float32", "c": "bytes"}
    assert t["a"].values.cpu().tolist() == [1.0, 3.0] and t["b"].numpy()[1].tolist() == [True, False]
    with pytest.raises(nat.NativeError, match=r"line 3: a line has more or fewer fields"):
        ktw.read_csv(b"a,b\n1,2\n3\n4,5\n")
    with pytest.raises(nat.NativeError, match=r"line 2: a line has more or fewer fields"):
        ktw.read_csv(b"a,b\n1,2,3\n")
    with pytest.raises(nat.NativeError, match=r"line 3: a field of a numeric column is not a number"):
        ktw.read_csv(b"a\n1\nx\n", schema={"a": "int64"})
    with pytest.raises(nat.NativeError, match="integer outside int64"):
        ktw.read_csv(b"a\n9223372036854775808\n", schema={"a": "int64"})
    with pytest.raises(nat.NativeError, match="unterminated quote"):
        ktw.read_csv(b'a,b\n"oops,1\n')
    with pytest.raises(nat.NativeError, match="too long to convert exactly"):
        ktw.read_csv(b"a\n1.00000000000000000001\n", schema={"a": "float32"})
    empty = ktw.read_csv(b"a,b\n")
    assert len(empty) == 0 and empty.schema() == {"a": "bytes", "b": "bytes"}


def test_ingested_table_feeds_the_transform(ktw):
    """read_csv -> trace -> analyze -> transform on the bundled head: same outputs as the host-read table."""
    path = os.path.join(helpers.GOLDEN, "taxi_head200.csv")
    fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
    gpu_table = ktw.read_csv(path)
    host_table = ktw.Table.from_csv(path).to("cuda")
    assert gpu_table.schema() == host_table.schema()
    plan = ktw.trace(fn, gpu_table.schema())
    eng = ktw.Engine(plan)
    a, _ = eng.analyze_and_transform(gpu_table)
    a = {k: (c.to_list() if c.kind == "string" else c.values.clone()) for k, c in a.items()}
    b, _ = eng.analyze_and_transform(host_table)
    for k, c in b.items():
        if c.kind == "string":
            assert a[k] == c.to_list(), k
        else:
            assert torch.equal(a[k], c.values), k


def test_fuzz_small_files_against_csv_module(ktw):
    """300 small random files written by Python's csv.writer (quoting of commas / quotes / spaces,
    empty and quoted-empty fields, one to twelve columns, with and without a final newline, CRLF)
    parsed with a given schema: every cell must equal the csv module + int()/float()/bytes."""
    rng = np.random.default_rng(2024)
    texts = ["", "a", "a,b", 'say "x"', "  lead", "trail  ", "ünï", "7", "-3.5e2", ",", '"', "x" * 40]

    def cell(kind):
        if rng.random() < 0.2:
            return ""
        if kind == "int64":
            return str(int(rng.integers(-10 ** int(rng.integers(1, 19)), 10 ** int(rng.integers(1, 19)))))
        if kind == "float32":
            v = float(rng.standard_normal() * 10.0 ** int(rng.integers(-12, 12)))
            return [repr(v), "%.3f" % v, "%e" % v, str(int(v)) if abs(v) < 1e15 else repr(v), "%.10g" % v][int(rng.integers(0, 5))]
        return texts[int(rng.integers(0, len(texts)))]

    for trial in range(300):
        ncols, nrows = int(rng.integers(1, 13)), int(rng.integers(0, 41))
        kinds = {f"c{j}": ["int64", "float32", "bytes"][int(rng.integers(0, 3))] for j in range(ncols)}
        buf = io.StringIO()
        w = csv.writer(buf, lineterminator="\r\n" if trial % 3 == 0 else "\n")
        w.writerow(list(kinds))
        for _ in range(nrows):
            row = [cell(k) for k in kinds.values()]
            if ncols == 1 and row[0] == "":
                row[0] = "1" if kinds["c0"] != "bytes" else "z"  # a lone empty field is a blank line
            w.writerow(row)
        text = buf.getvalue()
        if trial % 2:
            text = text.rstrip("\r\n")
        data = text.encode("utf-8")
        table = ktw.read_csv(data, schema=kinds)
        assert len(table) == nrows, (trial, text)
        check_table(table, python_parse(data, kinds))


# ------------------------------------------------------------------------------------------------
# TFRecord / tf.Example decode on the GPU (csrc/example_decode.cu, ingest.read_tfrecord)
# ------------------------------------------------------------------------------------------------
def _assert_tables_equal(got, want_cols):
    for k, c in want_cols.items():
        g = got[k]
        if isinstance(c, O.StrCol):
            assert g.to_list() == c.to_list(), k
            off, pool, valid = g.numpy()
            assert np.array_equal(valid, c.valid), k
        else:
            v, valid = g.numpy()
            assert v.dtype == c.values.dtype and np.array_equal(valid, c.valid), k
            a, b = np.where(valid, v, 0), np.where(c.valid, c.values, 0)
            assert np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a, b.view(np.uint32) if b.dtype == np.float32 else b), k


def _example_cols(n, rng):
    words = [b"", b"Cash", b"Credit Card", bytes(range(256)), b"x" * 300, "ünï".encode(), b"\x0a\x12\x1a"]
    ints = rng.integers(-2 ** 63, 2 ** 63 - 1, n, dtype=np.int64)
    ints[: min(n, 8)] = np.array([0, 1, -1, 127, 128, 2 ** 63 - 1, -2 ** 63, 300])[: min(n, 8)]
    floats = rng.standard_normal(n).astype(np.float32)
    floats[: min(n, 4)] = np.array([np.inf, -0.0, 1e-45, 3.4e38], np.float32)[: min(n, 4)]
    text = [words[i] for i in rng.integers(0, len(words), n)]
    keep = rng.random(n) > 0.15
    return {
        "zeta": O.NumCol(floats, rng.random(n) > 0.2),
        "alpha_int": O.NumCol(ints, rng.random(n) > 0.1),
        "text": O.StrCol.from_list([w if k else None for w, k in zip(text, keep)]),
        "small": O.NumCol(rng.integers(0, 1010, n).astype(np.int64), None),
        "a-very-long-feature-name-" + "n" * 120: O.NumCol(rng.integers(0, 10, n).astype(np.int64), None),
    }


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 127, 128, 129, 3001, 50_000])
def test_tfrecord_decode_round_trip(ktw, n):
    """Records produced by the GPU encoder (byte-identical to protobuf's, tests/test_gpu_egress.py) decode back
    to the same columns: values, validity (an empty value list is a missing cell), string bytes."""
    rng = np.random.default_rng(n)
    cols = _example_cols(n, rng)
    # a present-but-empty string is written as a one-value list holding b"": it must come back present
    table = helpers.table_from_oracle(cols, ktw).to("cuda")
    data, _ = ktw.encode_tfrecord(table)
    schema = {k: ("bytes" if isinstance(c, O.StrCol) else "float32" if c.values.dtype == np.float32 else "int64") for k, c in cols.items()}
    got = ktw.read_tfrecord(data, schema)
    assert len(got) == n
    _assert_tables_equal(got, cols)
    # a subset of the features, in another order; a feature no record has is all-missing
    sub = {"text": "bytes", "never_there": "int64", "zeta": "float32"}
    got2 = ktw.read_tfrecord(data.cpu().numpy().tobytes(), sub)
    _assert_tables_equal(got2, {"text": cols["text"], "zeta": cols["zeta"]})
    assert got2["never_there"].numpy()[1].sum() == 0


def test_tfrecord_decode_of_protobuf_runtime_bytes_gzip_and_unpacked_lists(ktw, tmp_path):
    """Records written by the protobuf runtime (the oracle's encoder), through a gzip file like the reference's;
    plus hand-built Examples with UNPACKED repeated fields, unknown features, unknown fields and the map entries
    in reverse order -- all legal wire forms of the same message."""
    sys_path = helpers.ROOT
    import sys
    sys.path.insert(0, sys_path)
    from oracle import example_oracle as E
    import gzip
    rng = np.random.default_rng(9)
    n = 777
    cols = _example_cols(n, rng)
    per = {}
    for k, c in cols.items():
        if isinstance(c, O.StrCol):
            per[k] = ("bytes", c.to_list(), c.valid)
        else:
            per[k] = ("float32" if c.values.dtype == np.float32 else "int64", c.values.tolist(), c.valid)
    rows = [{k: (kind, vals[i] if valid[i] else None) for k, (kind, vals, valid) in per.items()} for i in range(n)]
    path = tmp_path / "part-00000.gz"
    with gzip.open(path, "wb") as f:
        f.write(E.encode_rows(rows))
    schema = {k: v[0] for k, v in per.items()}
    _assert_tables_equal(ktw.read_tfrecord(str(path), schema), cols)

    def varint(v):
        out = b""
        v &= (1 << 64) - 1
        while v >= 0x80:
            out += bytes([(v & 0x7F) | 0x80]); v >>= 7
        return out + bytes([v])

    def ld(tag, payload):  # length-delimited field
        return bytes([tag]) + varint(len(payload)) + payload

    def example(entries):
        feats = b"".join(ld(0x0A, ld(0x0A, k) + ld(0x12, feat)) for k, feat in entries)
        return ld(0x0A, feats)

    f_unpacked = ld(0x12, bytes([0x0D]) + np.float32(2.5).tobytes())                 # FloatList, value as fixed32 field
    i_unpacked = ld(0x1A, bytes([0x08]) + varint(-7))                                  # Int64List, value as varint field
    b_val = ld(0x0A, ld(0x0A, b"hello"))
    recs = [
        example([(b"s", b_val), (b"i", i_unpacked), (b"f", f_unpacked)]),              # reverse name order, unpacked
        example([(b"unknown", ld(0x1A, ld(0x0A, varint(5)))), (b"f", ld(0x12, b""))]) + bytes([0x18, 0x05]),  # unknown feature / field, empty list
        example([(b"i", ld(0x1A, ld(0x0A, varint(2 ** 63 - 1))))]),
        b"",                                                                            # an empty Example
    ]
    stream = b"".join(E.frame(r) for r in recs)
    got = ktw.read_tfrecord(stream, {"f": "float32", "i": "int64", "s": "bytes"})
    assert got["f"].numpy()[0].tolist() == [2.5, 0.0, 0.0, 0.0] and got["f"].numpy()[1].tolist() == [True, False, False, False]
    assert got["i"].numpy()[0].tolist() == [-7, 0, 2 ** 63 - 1, 0] and got["i"].numpy()[1].tolist() == [True, False, True, False]
    assert got["s"].to_list() == [b"hello", b"", b"", b""]


def test_tfrecord_decode_errors(ktw):
    import sys
    sys.path.insert(0, helpers.ROOT)
    from oracle import example_oracle as E
    good = E.encode_rows([{"a": ("int64", 1)}, {"a": ("int64", 2)}, {"a": ("int64", 3)}])
    assert ktw.read_tfrecord(good, {"a": "int64"})["a"].numpy()[0].tolist() == [1, 2, 3]
    bad = bytearray(good)
    bad[14] ^= 0x40                                       # payload byte of record 0: data CRC mismatch
    with pytest.raises(ktw._native.NativeError, match="record 0.*CRC"):
        ktw.read_tfrecord(bytes(bad), {"a": "int64"})
    bad = bytearray(good)
    bad[len(good) // 3 + 2] ^= 0xFF                       # inside the second record's header: the chain breaks there
    with pytest.raises(ktw._native.NativeError, match="corrupt|does not start"):
        ktw.read_tfrecord(bytes(bad), {"a": "int64"})
    two = E.frame(bytes([0x0A, 0x0D, 0x0A, 0x0B, 0x0A, 0x01, 0x61, 0x12, 0x06, 0x1A, 0x04, 0x0A, 0x02, 0x01, 0x02]))  # a: Int64List [1, 2]
    with pytest.raises(ktw._native.NativeError, match="more than one value"):
        ktw.read_tfrecord(two, {"a": "int64"})
    with pytest.raises(ktw._native.NativeError, match="another kind"):
        ktw.read_tfrecord(good, {"a": "float32"})
    assert len(ktw.read_tfrecord(b"", {"a": "int64"})) == 0
