"""Property tests of the oracle (hypothesis; CPU only).

The oracle is a restatement of tft 0.13 semantics that no reference run can
pin on this box (parity unpinned).  What can be checked without tft are the properties
every correct implementation of those analyzers must have, whatever the unpinned details are:
permutation invariance, shard-merge == whole for arbitrary cut points (empty shards included),
monotone bucket ids inside [0, len(boundaries)], vocabulary ids inside the consumer's range
(reference airflow-dags/taxi_utils.py:168-185)."""
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from helpers import O

# The suite the driver runs is deterministic (the same examples every run); TXF_HYPOTHESIS_RANDOM=1 explores with
# fresh seeds and ten times the examples (a developer's run, not a gate).
_RANDOM = os.environ.get("TXF_HYPOTHESIS_RANDOM") == "1"
_settings = settings(max_examples=600 if _RANDOM else 60, deadline=None, derandomize=not _RANDOM, database=None)

floats32 = st.floats(width=32, allow_nan=False, allow_infinity=False, min_value=-1e6, max_value=1e6)
terms = st.binary(min_size=0, max_size=40)


@st.composite
def float_column(draw, min_size=1, max_size=300):
    vals = draw(st.lists(floats32, min_size=min_size, max_size=max_size))
    # heavy duplicates and an atom at 0.0, like the taxi lat/long columns after fill-missing
    dup = draw(st.lists(st.sampled_from([0.0, 41.88, -87.63]), max_size=60))
    arr = np.array(vals + dup, np.float32)
    return arr[draw(st.permutations(range(len(arr))))]


@st.composite
def cuts_of(draw, n, k=4):
    pts = sorted(draw(st.lists(st.integers(0, n), min_size=k, max_size=k)))
    return [0] + pts + [n]  # repeated points make empty shards


@given(float_column(), st.randoms(use_true_random=False))
@_settings
def test_moments_are_permutation_invariant(x, rnd):
    p = list(range(len(x)))
    rnd.shuffle(p)
    a, b = O.moments(x), O.moments(x[p])
    assert a[0] == b[0]
    assert a[1] == pytest.approx(b[1], rel=1e-12, abs=1e-9)
    assert a[2] == pytest.approx(b[2], rel=1e-9, abs=1e-6)


@given(float_column(min_size=2), st.data())
@_settings
def test_sharded_moments_merge_to_whole(x, data):
    cuts = data.draw(cuts_of(len(x)))
    acc = (0.0, 0.0, 0.0)
    for lo, hi in zip(cuts, cuts[1:]):
        acc = O.merge_moments(acc, O.moments(x[lo:hi]))
    whole = O.moments(x)
    assert acc[0] == whole[0]
    assert acc[1] == pytest.approx(whole[1], rel=1e-10, abs=1e-7)
    assert acc[2] == pytest.approx(whole[2], rel=1e-8, abs=1e-4)


@given(float_column(), st.integers(2, 50), st.randoms(use_true_random=False))
@_settings
def test_quantile_boundaries_permutation_invariant_distinct_and_data_values(x, nb, rnd):
    p = list(range(len(x)))
    rnd.shuffle(p)
    b = O.quantile_boundaries(x, nb)
    assert np.array_equal(b.view(np.uint32), O.quantile_boundaries(x[p], nb).view(np.uint32))
    assert np.all(np.diff(b) > 0)                         # distinct, ascending (Appendix A.4 rule 3)
    assert len(b) <= nb - 1
    canon = O.key_to_f32(O.order_key(x))
    assert set(b.view(np.uint32).tolist()) <= set(canon.view(np.uint32).tolist())   # every boundary is a data value
    if len(b) == nb - 1 and len(x) >= 100 * nb:           # full-length list, enough rows for the epsilon window to hold a rank
        assert O.boundary_rank_ok(x, b, nb)               # inside tft's stated rank error


@given(float_column(min_size=2), st.integers(2, 20), st.data())
@_settings
def test_sharded_quantile_keys_merge_to_whole(x, nb, data):
    cuts = data.draw(cuts_of(len(x)))
    keys = np.sort(np.concatenate([np.sort(O.order_key(x[lo:hi])) for lo, hi in zip(cuts, cuts[1:])]))
    sel = keys[O.quantile_ranks(len(x), nb) - 1]
    sel = sel[np.concatenate([[True], sel[1:] != sel[:-1]])]
    assert np.array_equal(O.key_to_f32(sel).view(np.uint32), O.quantile_boundaries(x, nb).view(np.uint32))


@given(float_column(), st.integers(2, 30))
@_settings
def test_bucketize_is_monotone_and_in_range(x, nb):
    b = O.quantile_boundaries(x, nb)
    xs = np.sort(x)
    ids = O.bucketize(xs, b)
    assert ids.dtype == np.int64
    assert np.all(np.diff(ids) >= 0)
    assert ids.min() >= 0 and ids.max() <= len(b)
    for i, edge in enumerate(b):                           # a value equal to a boundary goes to the upper bucket
        assert O.bucketize(np.array([edge], np.float32), b)[0] == i + 1
    assert O.bucketize(np.array([np.nan], np.float32), b)[0] == len(b)


@given(st.lists(terms, min_size=0, max_size=200), st.integers(1, 12), st.integers(1, 10), st.randoms(use_true_random=False))
@_settings
def test_vocabulary_is_permutation_invariant_and_ids_stay_in_the_consumer_range(items, top_k, oov, rnd):
    shuffled = list(items)
    rnd.shuffle(shuffled)
    v = O.finalize_vocabulary(O.vocab_counts(items), top_k=top_k)
    assert v == O.finalize_vocabulary(O.vocab_counts(shuffled), top_k=top_k)
    counts = O.vocab_counts(items)
    if v != [O.EMPTY_VOCAB_SENTINEL]:
        assert len(v) <= top_k and all(t != b"" and b"\n" not in t and b"\r" not in t for t in v)
        keys = [(counts[t], t) for t in v]
        assert keys == sorted(keys, reverse=True)          # (count, bytes) descending: Appendix A.2 rule 4
        rest = [t for t in counts if t not in v and O.vocab_keep(t)]
        assert all((counts[t], t) < keys[-1] for t in rest)  # nothing outside beats the last admitted term
    ids = O.apply_vocabulary(items, v, num_oov_buckets=oov)
    assert ids.dtype == np.int64 and (len(items) == 0 or (ids.min() >= 0 and ids.max() < len(v) + oov))
    for t, i in zip(items, ids.tolist()):
        assert (i < len(v)) == (t in v) and (i >= len(v) or v[i] == t)


@given(st.lists(terms, min_size=1, max_size=150), st.data())
@_settings
def test_sharded_vocabulary_counts_merge_to_whole(items, data):
    cuts = data.draw(cuts_of(len(items)))
    acc = O.vocab_counts([])
    for lo, hi in zip(cuts, cuts[1:]):
        acc = O.merge_vocab_counts(acc, O.vocab_counts(items[lo:hi]))
    assert dict(acc) == dict(O.vocab_counts(items))


@given(st.lists(st.tuples(floats32, floats32), min_size=1, max_size=100))
@_settings
def test_label_is_the_fp32_comparison(rows):
    fare = np.array([r[0] for r in rows], np.float32)
    tips = np.array([r[1] for r in rows], np.float32)
    got = O.tips_label(fare, tips, 0.2)
    want = (tips > (fare * np.float32(0.2)).astype(np.float32)).astype(np.int64)
    assert got.dtype == np.int64 and np.array_equal(got, want)


@given(st.binary(min_size=0, max_size=300))
@settings(max_examples=3000 if _RANDOM else 300, deadline=None, derandomize=not _RANDOM, database=None)
def test_fingerprint64_python_and_c_restatements_agree(s):
    assert O.fingerprint64(s) == O.fingerprint64_c(s)
