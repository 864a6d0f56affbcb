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
