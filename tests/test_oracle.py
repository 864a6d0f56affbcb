"""CPU tests: the oracle against every golden vector / known answer available for this path."""
import json
import os

import numpy as np
import pytest

import helpers
from helpers import O


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(helpers.GOLDEN, "farmhash_vectors.json")) as f:
        return json.load(f)


def test_fingerprint64_tf_known_answers(gold):
    """TensorFlow string_to_hash_bucket_fast expectations."""
    for s, fp in gold["tf_kat"]["full"].items():
        assert O.fingerprint64(s.encode()) == int(fp) == O.fingerprint64_c(s.encode())
    for s, m in gold["tf_kat"]["mod10"].items():
        assert O.fingerprint64(s.encode()) % 10 == m


def test_fingerprint64_matches_abseil_cityhash_up_to_32_bytes(gold):
    """Independent compiled implementation (abseil CityHash64) for every length 0..32."""
    lens = set()
    for v in gold["abseil_le32"]:
        s = bytes.fromhex(v["hex"])
        lens.add(len(s))
        assert O.fingerprint64(s) == int(v["fp"]), len(s)
        assert O.fingerprint64_c(s) == int(v["fp"]), len(s)
    assert lens == set(range(33))


def test_fingerprint64_long_inputs_python_vs_c(gold):
    for v in gold["survey_gt32"]:
        s = bytes.fromhex(v["hex"])
        assert O.fingerprint64(s) == int(v["fp"]) == O.fingerprint64_c(s)
    rng = np.random.default_rng(11)
    for n in list(range(33, 140)) + [191, 192, 193, 255, 256, 257, 1023, 4096]:
        s = bytes(rng.integers(0, 256, n, dtype=np.uint8))
        assert O.fingerprint64(s) == O.fingerprint64_c(s), n


def test_fingerprint64_bulk():
    items = [b"", b"a", b"Cash", b"Credit Card", b"x" * 70, None]
    col = O.StrCol.from_list(items)
    got = O.fingerprint64_bulk(col.offsets, col.pool)
    want = [O.fingerprint64(s or b"") for s in items]
    assert [int(x) for x in got] == want


def test_notebook_03_known_answers():
    """notebooks/03_TensorFlow_Transform.ipynb:102-105,227-233: x - mean = [-1, 0, 1];
    compute_and_apply_vocabulary(['hello', 'world', 'hello']) = [0, 1, 0]."""
    x = np.array([1, 2, 3], np.float32)
    n, mean, m2 = O.moments(x)
    assert np.array_equal(x - np.float32(mean), [-1, 0, 1])
    s = [b"hello", b"world", b"hello"]
    vocab = O.finalize_vocabulary(O.vocab_counts(s), top_k=None)
    assert vocab == [b"hello", b"world"]
    assert O.apply_vocabulary(s, vocab).tolist() == [0, 1, 0]


def test_taxi_fixture_matches_survey_appendix_b():
    """The 15 000-row fixture (oracle output) against the independently computed anchors."""
    cols, out, state = helpers.load_taxi_golden()
    assert len(cols["fare"]) == 15000
    assert int(out["tips_xf"].sum()) == 3382
    m = O.moments(O.fill_in_missing(cols["trip_miles"]))
    assert m[1] == pytest.approx(2.8732693313838293, rel=1e-12) and m[2] / m[0] == pytest.approx(233.37054180562674, rel=1e-12)
    m = O.moments(O.fill_in_missing(cols["fare"]))
    assert m[1] == pytest.approx(11.768835334837188, rel=1e-12) and m[2] / m[0] == pytest.approx(133.14593564467486, rel=1e-12)
    m = O.moments(O.fill_in_missing(cols["trip_seconds"]))
    assert m[1] == pytest.approx(777.4201333333333, rel=1e-12) and m[2] / m[0] == pytest.approx(955424.7281546488, rel=1e-12)
    assert state["vocab"]["payment_type"] == [b"Cash", b"Credit Card", b"No Charge", b"Unknown", b"Dispute", b"Pcard", b"Prcard"]
    comp = state["vocab"]["company"]
    assert len(comp) == 61 and comp[0] == b"Taxi Affiliation Services" and comp[10] == b"Chicago Elite Cab Corp."
    assert comp.index(b"5074 - 54002 Ahzmi Inc") + 1 == comp.index(b"0118 - 42111 Godfrey S.Awir")
    assert comp[-3:] == [b"2809 - 95474 C & D Cab Co Inc.", b"2192 - Zeymane Corp", b"2092 - 61288 Sbeih company"]
    assert int((out["company_xf"] == 64).sum()) == 5140 and int(out["company_xf"][3]) == 64
    np.testing.assert_allclose(state["boundaries"]["pickup_latitude"],
                               [41.87740707, 41.88099289, 41.88498688, 41.89204025, 41.89250946, 41.89960098, 41.90585709,
                                41.93465805, 41.96581268], rtol=0, atol=2e-6)
    assert all(len(b) == 9 for b in state["boundaries"].values())
    row0 = {k: out[k][0] for k in out if not isinstance(out[k], list)}
    assert row0["trip_miles_xf"] == pytest.approx(0.636713, abs=1e-6)
    assert row0["fare_xf"] == pytest.approx(1.324319, abs=1e-6)
    assert row0["trip_seconds_xf"] == pytest.approx(0.616476, abs=1e-6)
    assert [int(row0[k + "_xf"]) for k in O.BUCKET_KEYS] == [0, 3, 0, 9]
    # a recomputation from the stored inputs reproduces the stored outputs exactly
    out2, _ = O.analyze_and_transform(cols)
    for k, v in out.items():
        assert (out2[k] == v) if isinstance(v, list) else np.array_equal(out2[k], v), k


def test_label_uses_fp32_threshold():
    cols, out, _ = helpers.load_taxi_golden()
    fare = O.fill_in_missing(cols["fare"]).astype(np.float64)
    tips = O.fill_in_missing(cols["tips"]).astype(np.float64)
    assert int((tips > fare * 0.2).sum()) == 3445  # the fp64 rule differs on 63 rows
    assert O.tips_label(np.array([np.nan], np.float32), np.array([1.0], np.float32)).tolist() == [0]


def test_vocabulary_rules():
    c = O.vocab_counts([b"b", b"a", b"", b"", b"", b"x\ny", b"x\ny", b"c", b"a", b"b", b"d\r"])
    assert O.finalize_vocabulary(c) == [b"b", b"a", b"c"]          # ties: descending bytes; '' and newlines dropped
    assert O.finalize_vocabulary(c, top_k=1) == [b"b"]
    assert O.finalize_vocabulary(O.vocab_counts([b"", b"\n"])) == [O.EMPTY_VOCAB_SENTINEL]
    ids = O.apply_vocabulary([b"a", b"zzz", b""], [b"b", b"a"], num_oov_buckets=10)
    assert ids.tolist() == [1, 2 + O.fingerprint64(b"zzz") % 10, 2 + 3]
    assert O.apply_vocabulary([b"q"], [b"b"], num_oov_buckets=0).tolist() == [-1]


def test_quantile_rules():
    x = np.array([5, 1, 4, 2, 3, 3, 3, 9, 7, 8], np.float32)
    b = O.quantile_boundaries(x, 5)   # ranks ceil(i*10/5) = 2,4,6,8 -> sorted[1,3,5,7] = 2,3,4,7
    assert b.tolist() == [2.0, 3.0, 4.0, 7.0]
    assert O.bucketize(np.array([1.9, 2.0, 3.0, 6.9, 7.0, np.nan], np.float32), b).tolist() == [0, 1, 2, 3, 4, 4]
    # duplicates collapse: ranks 3,5,8 of [1,3,3,3,3,3,3,3,9,9] -> 3,3,3 -> one boundary
    assert O.quantile_boundaries(np.array([1, 3, 3, 3, 3, 3, 3, 3, 9, 9], np.float32), 4).tolist() == [3.0]
    assert O.quantile_boundaries(np.zeros(0, np.float32), 10).size == 0
    assert O.quantile_boundaries(np.array([-0.0, 0.0, 0.0], np.float32), 2).view(np.uint32).tolist() == [0]
    with pytest.raises(TypeError):
        O.quantile_boundaries(x, 2.5)
    with pytest.raises(ValueError):
        O.quantile_boundaries(x, 0)
    assert O.boundary_rank_ok(np.arange(1000, dtype=np.float32), O.quantile_boundaries(np.arange(1000, dtype=np.float32), 10), 10)
    k = O.order_key(np.array([-np.inf, -1.5, -0.0, 0.0, 1e-45, 2.0, np.inf, np.nan], np.float32))
    assert np.all(np.diff(k.astype(np.int64)) >= 0) and k[2] == k[3]
    assert np.array_equal(O.key_to_f32(k[:7]), np.array([-np.inf, -1.5, 0.0, 0.0, 1e-45, 2.0, np.inf], np.float32))


def test_sharded_analyze_merges_to_whole():
    """Partial analyzer states of row shards merge to the single-shard state (what the GPU ranks do)."""
    from kubeflow_tfx_workshop_b200 import synth
    host = synth.taxi_table(30_000, backend="numpy")
    cols = helpers.oracle_from_table(host)
    whole = O.analyze(cols)
    cuts = [0, 7_000, 7_000, 19_001, 30_000]  # includes an empty shard
    parts = []
    for lo, hi in zip(cuts, cuts[1:]):
        shard = {k: c.slice(lo, hi) for k, c in cols.items()}
        parts.append(O.analyze_partial(shard))
    merged = O.finalize_state(O.merge_partials(parts))
    assert merged["vocab"] == whole["vocab"]
    for k in O.BUCKET_KEYS:
        assert np.array_equal(merged["boundaries"][k], whole["boundaries"][k])
    for k in O.DENSE_FLOAT_KEYS:
        assert merged["zscore"][k][0] == pytest.approx(whole["zscore"][k][0], rel=1e-6)
        assert merged["zscore"][k][1] == pytest.approx(whole["zscore"][k][1], rel=1e-6)


def test_example_oracle_layout_and_crc():
    """oracle/example_oracle.py: CRC-32C check value and the wire layout the GPU encoder restates."""
    import sys
    sys.path.insert(0, helpers.ROOT)
    from oracle import example_oracle as E
    assert E.crc32c(b"123456789") == 0xE3069283  # published check value (Castagnoli polynomial)
    ex = E.make_example({"b": ("float32", 1.5), "a": ("int64", -3), "s": ("bytes", b"hi"), "m": ("float32", None)})
    data = ex.SerializeToString(deterministic=True)
    assert data.hex() == ("0a3a" "0a13" "0a0161" "120e" "1a0c0a0a" "fdffffffffffffffff01" "0a0d" "0a0162" "1208" "12060a04" "0000c03f"
                          "0a07" "0a016d" "1202" "1200" "0a0b" "0a0173" "1206" "0a040a02" "6869")
    rec = E.frame(data)
    assert len(rec) == len(data) + 16 and E.decode_tfrecord(rec + rec)[1].features.feature["a"].int64_list.value[0] == -3
