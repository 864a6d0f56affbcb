"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the oracle."""
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


@pytest.fixture(scope="module")
def fn(ktw):
    return helpers.load_example_preprocessing_fn(ktw).preprocessing_fn


def check_outputs(out, want, wstate):
    for key, w in want.items():
        col = out[key]
        if isinstance(w, list):
            assert col.to_list() == w, key
        elif w.dtype == np.float32:
            mean, var = wstate["zscore"][key[:-3]]
            got = col.values.cpu().numpy()
            assert helpers.zscore_close(got, w, mean, var), (key, np.abs(got - w).max())
        else:
            got = col.values.cpu().numpy()
            assert got.dtype == w.dtype and np.array_equal(got, w), (key, int((got != w).sum()))


def check_state(state, wstate, exact_moments=False):
    for k, (mean, var) in wstate["zscore"].items():
        m, v = state.mean_var_of(k)
        assert abs(m - float(mean)) <= 1e-6 * abs(float(mean)) + 1e-30, (k, m, mean)
        assert abs(v - float(var)) <= 1e-6 * abs(float(var)) + 1e-30, (k, v, var)
    for key, res in state.vocabs.items():
        assert res.terms() == wstate["vocab"][key[0]], key
    for nb, res in state.buckets.items():
        for k in res.keys:
            got = res.boundaries_of(k)
            want = wstate["boundaries"][k]
            assert got.dtype == np.float32 and np.array_equal(got.view(np.uint32), np.asarray(want, np.float32).view(np.uint32)), (k, got, want)


# ------------------------------------------------------------------------------------------------
# FarmHash on the device
# ------------------------------------------------------------------------------------------------
def test_fingerprint64_device_matches_oracle(ktw):
    import json, os
    nat = ktw._native
    with open(os.path.join(helpers.GOLDEN, "farmhash_vectors.json")) as f:
        gold = json.load(f)
    items = [(bytes.fromhex(v["hex"]), int(v["fp"])) for v in gold["abseil_le32"] + gold["survey_gt32"]]
    items += [(k.encode(), int(v)) for k, v in gold["tf_kat"]["full"].items()]
    rng = np.random.default_rng(7)
    extra = [bytes(rng.integers(0, 256, n, dtype=np.uint8)) for n in list(range(0, 200)) + [255, 256, 257, 1000, 4097]]
    items += [(s, O.fingerprint64_c(s)) for s in extra]
    col = ktw.StringColumn.from_list([s for s, _ in items]).to("cuda")
    out = torch.empty(len(items), dtype=torch.int64, device="cuda")
    nat.call("txf_fingerprint64_batch", col.offsets.data_ptr(), col.pool.data_ptr(), len(items), out.data_ptr(), None)
    got = out.cpu().numpy().view(np.uint64)
    want = np.array([fp for _, fp in items], np.uint64)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, [(len(items[i][0]), int(got[i]), int(want[i])) for i in bad[:5]]


# ------------------------------------------------------------------------------------------------
# config 1: the bundled 15 000 rows
# ------------------------------------------------------------------------------------------------
def test_taxi_golden_15000(ktw, fn):
    cols, want, wstate = helpers.load_taxi_golden()
    host = helpers.table_from_oracle(cols, ktw)
    plan = ktw.trace(fn, host.schema())
    eng = ktw.Engine(plan)
    out, state = eng.analyze_and_transform(host.to("cuda"))
    check_state(state, wstate)
    check_outputs(out, want, wstate)
    # anchors stated in SURVEY.md Appendix B
    assert int(out["tips_xf"].values.sum()) == 3382
    assert state.vocabs[("company", 1000, None)].size == 61 and state.vocabs[("payment_type", 1000, None)].size == 7
    assert int(out["company_xf"].values[3]) == 64


def test_taxi_golden_through_tft_beam_api(ktw, fn):
    """Same result through the reference-facing driver (notebooks/06_...:236-262 shape)."""
    cols, want, wstate = helpers.load_taxi_golden()
    host = helpers.table_from_oracle(cols, ktw)
    meta = ktw.tf_metadata.DatasetMetadata(ktw.tf_metadata.Schema(host.schema()))
    with ktw.tft_beam.Context(temp_dir="/tmp/unused"):
        (transformed, tmeta), transform_fn = (host, meta) | ktw.tft_beam.AnalyzeAndTransformDataset(fn)
    check_outputs(transformed, want, wstate)
    assert tmeta.schema.column_dtypes["company_xf"] == "int64"
    # TransformDataset on a second dataset re-applies the saved state (03a notebook :468)
    sub = {k: O.NumCol(c.values[:4096], c.valid[:4096]) if isinstance(c, O.NumCol) else c.slice(0, 4096) for k, c in cols.items()}
    first = helpers.table_from_oracle(sub, ktw)
    out2, _ = ((first, meta), transform_fn) | ktw.tft_beam.TransformDataset()
    check_outputs(out2, {k: (v[:4096]) for k, v in want.items()}, wstate)


def test_state_save_load_roundtrip(ktw, fn, tmp_path):
    cols, want, wstate = helpers.load_taxi_golden()
    host = helpers.table_from_oracle(cols, ktw)
    eng = ktw.Engine(ktw.trace(fn, host.schema()))
    dev = host.to("cuda")
    state = eng.analyze(dev)
    ktw.save_state(state, str(tmp_path / "tfn"))
    with open(tmp_path / "tfn" / "company.top1000.thrNone.vocab", "rb") as f:
        assert f.read().split(b"\n")[:-1] == wstate["vocab"]["company"]
    eng2, state2 = ktw.load_state(str(tmp_path / "tfn"))
    check_outputs(eng2.transform(dev, state2), want, wstate)


# ------------------------------------------------------------------------------------------------
# synthetic taxi rows, both quantile engines, ragged sizes
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,force_general", [(300_001, False), (300_001, True), (1_000_003, False), (4099, False), (7, False), (1, False)])
def test_synthetic_taxi(ktw, fn, n, force_general):
    from kubeflow_tfx_workshop_b200 import synth
    host = synth.taxi_table(n, backend="numpy")
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    eng.force_general_quantiles = force_general
    out, state = eng.analyze_and_transform(host.to("cuda"))
    want, wstate = O.analyze_and_transform(helpers.oracle_from_table(host))
    if n >= (1 << 18) and not force_general:
        assert eng.last_quantile_path[10] == "sample-select"
    else:
        assert eng.last_quantile_path[10] == "radix-select"
    check_state(state, wstate)
    check_outputs(out, want, wstate)


def test_ragged_sizes_around_kernel_tile_edges(ktw, fn):
    """Row counts around every tiling unit of the kernels (quads of 4, warps of 32, 96-row steps of
    the vocabulary count, 1024-quad transform tiles, 3072 = 32 warps x 96 rows): state and every
    output against the oracle, through both the coded and the string vocabulary path."""
    from kubeflow_tfx_workshop_b200 import synth
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    for n in (2, 3, 4, 5, 31, 32, 33, 95, 96, 97, 191, 192, 193, 1023, 1024, 1025, 3071, 3072, 3073, 4095, 4096, 4097, 8193, 12289):
        host = synth.taxi_table(n, seed=n, backend="numpy")
        dev = host.to("cuda")
        out, state = eng.analyze_and_transform(dev)
        want, wstate = O.analyze_and_transform(helpers.oracle_from_table(host))
        check_state(state, wstate)
        check_outputs(out, want, wstate)
        eng.use_codes = False
        check_outputs(eng.transform(dev, state), want, wstate)
        eng.use_codes = True


def test_coded_and_string_vocabulary_paths_agree(ktw, fn):
    """Transform of the analyzed table decodes dictionary codes; any other table re-reads the term
    bytes.  Both must give the oracle's ids."""
    from kubeflow_tfx_workshop_b200 import synth
    host = synth.taxi_table(123_457, backend="numpy")
    dev = host.to("cuda")
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    state = eng.analyze(dev)
    assert len(state.codes) == 2
    coded = eng.transform(dev, state)
    eng.use_codes = False
    plain = eng.transform(dev, state)
    want, _ = O.analyze_and_transform(helpers.oracle_from_table(host))
    for k in ("payment_type_xf", "company_xf"):
        assert np.array_equal(coded[k].values.cpu().numpy(), want[k]), k
        assert np.array_equal(plain[k].values.cpu().numpy(), want[k]), k


def test_codes_of_a_freed_table_are_never_applied_to_another_table(ktw, fn):
    """The dictionary-code path is tied to the analyzed table itself (object and buffers), not to
    its `id()`: analyze A, free it, transform a same-length B (CPython may hand B the same id),
    and transform A again after one of its string columns was replaced."""
    import gc
    from kubeflow_tfx_workshop_b200 import synth
    n = 50_000
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    host_a = synth.taxi_table(n, row0=0, backend="numpy")
    host_b = synth.taxi_table(n, row0=1_000_000, backend="numpy")
    a = host_a.to("cuda")
    state = eng.analyze(a)
    want_a, wstate = O.analyze_and_transform(helpers.oracle_from_table(host_a))
    want_b = O.transform(helpers.oracle_from_table(host_b), wstate)
    for _ in range(4):  # several allocations: one of them is likely to reuse A's address
        del a
        gc.collect()
        b = host_b.to("cuda")
        got = eng.transform(b, state)
        for k in ("payment_type_xf", "company_xf"):
            assert np.array_equal(got[k].values.cpu().numpy(), want_b[k]), k
        a = b
    # same table object, one string column replaced after analyze: its codes must not be used
    a = host_a.to("cuda")
    state = eng.analyze(a)
    a.columns["company"] = host_b.to("cuda")["company"]
    got = eng.transform(a, state)
    assert np.array_equal(got["company_xf"].values.cpu().numpy(), want_b["company_xf"])
    assert np.array_equal(got["payment_type_xf"].values.cpu().numpy(), want_a["payment_type_xf"])


@pytest.mark.parametrize("n", [1, 7, 1023, 1024, 1025, 2048, 4099, 148 * 4 * 2048 + 13, 3_000_001])
def test_streamed_state_free_ops_match_fused_kernel(ktw, fn, n):
    """txf_transform_stream (the small-footprint kernel that runs under the analyzers) against the
    oracle and against txf_transform, at row counts around its 1024 / 2048-cell tiles, on a table whose
    string offsets do not start at 0 and whose bitmaps are not 16-byte aligned (slices)."""
    from kubeflow_tfx_workshop_b200 import synth
    skip = 8 * 3  # slice start: offsets base != 0, validity pointers 3 bytes into their allocation
    host = synth.taxi_table(n + skip, backend="numpy")
    dev_full = host.to("cuda")
    cols = {}
    for k, c in dev_full.items():
        valid = None if c.valid is None else c.valid[skip // 8:]
        if c.kind == "numeric":
            v = c.values[skip:]
            if v.data_ptr() % 16:  # kernels take 16-byte aligned value arrays: re-home the slice
                v = v.clone()
            cols[k] = ktw.NumericColumn(v, valid)
        else:
            off = c.offsets[skip:]
            if off.data_ptr() % 16:
                off = off.clone()
            cols[k] = ktw.StringColumn(off, c.pool, valid, c.pool_bytes)
    dev = ktw.Table(cols)
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    got = eng.transform_state_free(dev)
    torch.cuda.synchronize()
    host_cols = helpers.oracle_from_table(host)
    sl = {k: (O.StrCol(c.offsets[skip:] - c.offsets[skip], c.pool[c.offsets[skip]:], c.valid[skip:]) if isinstance(c, O.StrCol)
              else O.NumCol(c.values[skip:], c.valid[skip:])) for k, c in host_cols.items()}
    want = {k: O.fill_in_missing(sl[k[:-3]]) for k in got if k != "tips_xf"}
    want["tips_xf"] = O.tips_label(O.fill_in_missing(sl["fare"]), O.fill_in_missing(sl["tips"]), 0.2)
    assert set(got) == {"trip_start_hour_xf", "trip_start_day_xf", "trip_start_month_xf", "pickup_census_tract_xf", "dropoff_census_tract_xf",
                        "pickup_community_area_xf", "dropoff_community_area_xf", "tips_xf"}
    for k, col in got.items():
        if col.kind == "string":
            assert col.to_list() == list(want[k]), k
        else:
            assert np.array_equal(col.values.cpu().numpy(), want[k]), k


def test_synthetic_generators_agree_on_gpu(ktw):
    """torch-on-GPU generator == numpy twin, bit for bit (the benchmark relies on it)."""
    from kubeflow_tfx_workshop_b200 import synth
    a = synth.taxi_table(50_001, backend="numpy")
    b = synth.taxi_table(50_001, backend="torch", device="cuda")
    for k in a.keys():
        ca, cb = a[k], b[k]
        if ca.kind == "string":
            assert torch.equal(ca.offsets, cb.offsets.cpu()) and torch.equal(ca.pool, cb.pool.cpu()), k
        else:
            va, vb = ca.values, cb.values.cpu()
            assert torch.equal(va.view(torch.int32) if va.dtype == torch.float32 else va,
                               vb.view(torch.int32) if vb.dtype == torch.float32 else vb), k
        assert (ca.valid is None) == (cb.valid is None) and (ca.valid is None or torch.equal(ca.valid, cb.valid.cpu())), k


def test_chunked_analyze_equals_whole(ktw, fn):
    """Analyzer state accumulates over chunks: two chunks == one table (exact for vocab/boundaries)."""
    from kubeflow_tfx_workshop_b200 import synth
    n1, n2 = 200_000, 150_003
    c1 = synth.taxi_table(n1, 0, backend="numpy")
    c2 = synth.taxi_table(n2, n1, backend="numpy")
    whole = synth.taxi_table(n1 + n2, backend="numpy")
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    state = eng.analyze([c1.to("cuda"), c2.to("cuda")]).detach()
    want, wstate = O.analyze_and_transform(helpers.oracle_from_table(whole))
    check_state(state, wstate)
    out2 = eng.transform(c2.to("cuda"), state)
    check_outputs(out2, {k: v[n1:] for k, v in want.items()}, wstate)


# ------------------------------------------------------------------------------------------------
# quantiles: config 5 style stress
# ------------------------------------------------------------------------------------------------
def _bucket_fn(ktw, keys, nb):
    tf, tft = ktw.tf, ktw.tft

    def f(inputs):
        def dense(x):
            return tf.squeeze(tf.sparse.to_dense(tf.SparseTensor(x.indices, x.values, [x.dense_shape[0], 1]), 0), axis=1)
        return {k + "_b": tft.bucketize(dense(inputs[k]), nb) for k in keys}
    return f


@pytest.mark.parametrize("nb", [2, 3, 10, 33, 65, 100, 1000])
def test_bucketize_stress(ktw, nb):
    rng = np.random.default_rng(nb)
    n = 400_003
    cent = rng.normal(41.8, 0.1, 300).astype(np.float32)
    data = {
        "centroid": cent[(rng.integers(0, 300, n) * rng.integers(0, 300, n)) // 300],
        "atom0": np.where(rng.random(n) < 0.4, 0.0, rng.normal(0, 1, n)).astype(np.float32),
        "fine": rng.normal(-87.6, 0.05, n).astype(np.float32),
        "uniform": rng.random(n).astype(np.float32),
        "const": np.full(n, 3.25, np.float32),
        "signed": (rng.standard_cauchy(n) * 10).astype(np.float32),
        "ints": rng.integers(-50, 50, n).astype(np.int64),
        "tiny": (rng.random(n) * 1e-40).astype(np.float32),  # denormals and +-0
    }
    data["tiny"][::7] = -0.0
    valid = {"fine": rng.random(n) > 0.03}
    cols = {k: O.NumCol(v, valid.get(k)) for k, v in data.items()}
    host = helpers.table_from_oracle(cols, ktw)
    keys = list(data)
    eng = ktw.Engine(ktw.trace(_bucket_fn(ktw, keys, nb), host.schema()))
    dev = host.to("cuda")
    out, state = eng.analyze_and_transform(dev)
    res = state.buckets[nb]
    if nb <= 33:
        assert eng.last_quantile_path[nb] == "sample-select"
    if nb >= 100:
        assert eng.last_quantile_path[nb] == "splitter-select"
    for k in keys:
        filled = O.fill_in_missing(cols[k])
        want_b = O.quantile_boundaries(filled, nb)
        got_b = res.boundaries_of(k)
        assert np.array_equal(got_b.view(np.uint32), want_b.view(np.uint32)), (k, nb, got_b[:5], want_b[:5])
        assert np.array_equal(out[k + "_b"].values.cpu().numpy(), O.bucketize(filled, want_b)), (k, nb)
    # the general engine and the two-pass splitter engine return the same boundaries
    for attr in ("force_general_quantiles", "force_split_quantiles"):
        setattr(eng, attr, True)
        state2 = eng.analyze(dev)
        setattr(eng, attr, False)
        assert eng.last_quantile_path[nb] == ("radix-select" if attr == "force_general_quantiles" else "splitter-select"), (attr, eng.last_quantile_path)
        for k in keys:
            assert np.array_equal(state2.buckets[nb].boundaries_of(k).view(np.uint32), O.quantile_boundaries(O.fill_in_missing(cols[k]), nb).view(np.uint32)), (k, attr)


def test_bucketize_nan_goes_last(ktw):
    x = np.array([0.5, np.nan, -1.0, 2.0, np.nan, 0.25, 1.0, 3.0] * 8, np.float32)
    cols = {"x": O.NumCol(x)}
    host = helpers.table_from_oracle(cols, ktw)
    eng = ktw.Engine(ktw.trace(_bucket_fn(ktw, ["x"], 4), host.schema()))
    out, state = eng.analyze_and_transform(host.to("cuda"))
    want_b = O.quantile_boundaries(x, 4)
    got_b = state.buckets[4].boundaries_of("x")
    assert np.array_equal(got_b, want_b, equal_nan=True)
    assert np.array_equal(out["x_b"].values.cpu().numpy(), O.bucketize(x, want_b))


# ------------------------------------------------------------------------------------------------
# vocabulary: config 4 style stress
# ------------------------------------------------------------------------------------------------
def _vocab_fn(ktw, key, top_k, oov, default_value=-1):
    tf, tft = ktw.tf, ktw.tft

    def f(inputs):
        x = inputs[key]
        d = tf.squeeze(tf.sparse.to_dense(tf.SparseTensor(x.indices, x.values, [x.dense_shape[0], 1]), ""), axis=1)
        return {key + "_id": tft.compute_and_apply_vocabulary(d, top_k=top_k, num_oov_buckets=oov, default_value=default_value)}
    return f


def _stress_terms(n_distinct, rng):
    """Distinct byte strings covering every FarmHash length class, with awkward bytes."""
    terms = set()
    lens = [1, 2, 3, 4, 7, 8, 15, 16, 17, 31, 32, 33, 63, 64, 65, 127, 128, 129, 300]
    i = 0
    while len(terms) < n_distinct:
        L = lens[i % len(lens)] if i < 4 * len(lens) else int(rng.integers(1, 48))
        t = bytes(rng.integers(1, 256, L, dtype=np.uint8))
        terms.add(t)
        i += 1
    terms = list(terms)
    terms[0] = b"has\nnewline"   # never admitted
    terms[1] = b"has\rreturn"    # never admitted
    terms[2] = b"ab"
    terms[3] = b"ab\x00"         # prefix ordering with an embedded NUL
    terms[4] = b"ab\x00\x00"
    return terms


@pytest.mark.parametrize("n_distinct,top_k,oov", [(50, 1000, 10), (3000, 1000, 10), (3000, 7, 3), (20000, 1000, 10), (20000, None, 0), (5000, 100, 70000)])
def test_vocabulary_stress(ktw, n_distinct, top_k, oov):
    rng = np.random.default_rng(n_distinct + (top_k or 0))
    terms = _stress_terms(n_distinct, rng)
    # constructed multiplicities: a long run of equal counts straddles the top_k cut-off
    counts = np.maximum(1, (n_distinct // (np.arange(n_distinct) + 1)) // 8).astype(np.int64)
    counts[:5] = [900, 800, 5, 5, 5]
    rows = np.repeat(np.arange(n_distinct), counts)
    rng.shuffle(rows)
    items = [terms[i] for i in rows]
    missing = rng.random(len(items)) < 0.1
    items = [None if m else t for t, m in zip(items, missing)]
    items[5] = b""  # a present-but-empty value behaves like a missing one
    col = O.StrCol.from_list(items)
    host = helpers.table_from_oracle({"s": col}, ktw)
    eng = ktw.Engine(ktw.trace(_vocab_fn(ktw, "s", top_k, oov), host.schema()), vocab_capacity=1 << 17, arena_bytes=1 << 23)
    out, state = eng.analyze_and_transform(host.to("cuda"))
    filled = O.fill_in_missing(col)
    want_vocab = O.finalize_vocabulary(O.vocab_counts(filled), top_k=top_k)
    got_vocab = state.vocabs[("s", top_k, None)].terms()
    assert got_vocab == want_vocab, next((i, a, b) for i, (a, b) in enumerate(zip(got_vocab, want_vocab)) if a != b)
    want_ids = O.apply_vocabulary(filled, want_vocab, oov)
    assert np.array_equal(out["s_id"].values.cpu().numpy(), want_ids)


def _vocab_roundtrip(ktw, items, top_k, oov, **engine_kw):
    """analyze+transform of one string column on the GPU (coded path and string path) -> (vocabulary, ids, ids of a
    re-application to a second table holding the same rows plus terms never analyzed, their oracle ids)."""
    col = O.StrCol.from_list(items)
    host = helpers.table_from_oracle({"s": col}, ktw)
    eng = ktw.Engine(ktw.trace(_vocab_fn(ktw, "s", top_k, oov), host.schema()), **engine_kw)
    dev = host.to("cuda")
    out, state = eng.analyze_and_transform(dev)
    filled = O.fill_in_missing(col)
    want_vocab = O.finalize_vocabulary(O.vocab_counts(filled), top_k=top_k)
    got_vocab = state.vocabs[("s", top_k, None)].terms()
    assert got_vocab == want_vocab, next((i, a, b) for i, (a, b) in enumerate(zip(got_vocab, want_vocab)) if a != b)
    want_ids = O.apply_vocabulary(filled, want_vocab, oov)
    assert np.array_equal(out["s_id"].values.cpu().numpy(), want_ids)                         # dictionary-code path
    eng.use_codes = False
    assert np.array_equal(eng.transform(dev, state)["s_id"].values.cpu().numpy(), want_ids)    # string path
    return eng, state, want_vocab


def test_terms_that_collided_under_the_round1_mixer_are_distinct(ktw):
    """Two 32-byte terms whose second window zeroes a factor of each fingerprint lane shared one
    identity under a bare-product mixer; they must be counted and numbered separately."""
    a, b = helpers.lane_zeroing_pair()
    items = [a] * 700 + [b] * 300 + [b"Cash"] * 50 + [None] * 20 + [a[:31]] * 5 + [b + b"x"] * 3
    rng = np.random.default_rng(3)
    rng.shuffle(items)
    _, state, vocab = _vocab_roundtrip(ktw, items, 1000, 10)
    assert vocab[:3] == [a, b, b"Cash"]


@pytest.mark.parametrize("bits", [1, 6, 20])
def test_fingerprint_collisions_never_merge_terms(ktw, bits):
    """Test hook txf_debug_set_identity_bits: with only `bits` hash bits kept, long terms collide all
    over every table (front table, count table, id lookup, decode table); every match is confirmed on
    the bytes, so vocabulary, counts (through the ranking), ids and OOV ids must equal the oracle's."""
    nat = ktw._native
    rng = np.random.default_rng(100 + bits)
    # long terms of every window class (16..32, 33..64, > 64), many of equal length; short terms; missing cells
    terms = []
    for L in (16, 17, 24, 24, 24, 31, 32, 32, 33, 40, 40, 64, 64, 65, 100, 130):
        terms += [bytes(rng.integers(32, 127, L, dtype=np.uint8)) for _ in range(120)]
    terms += [b"a" * L for L in (16, 20, 24, 32, 33, 48, 64, 65, 80)]          # equal windows, different lengths
    terms += [b"Cash", b"Credit Card", b"x" * 15]
    terms = list(dict.fromkeys(terms))
    counts = np.maximum(1, 3000 // (np.arange(len(terms)) + 1)).astype(np.int64)
    counts[400:460] = 7                                                          # a run of ties decided by the bytes
    rows = np.repeat(np.arange(len(terms)), counts)
    rng.shuffle(rows)
    items = [terms[i] for i in rows]
    for i in rng.integers(0, len(items), len(items) // 10):
        items[i] = None
    nat.call("txf_debug_set_identity_bits", bits)
    try:
        eng, state, vocab = _vocab_roundtrip(ktw, items, 1000, 10)
        # re-application to rows never analyzed (TransformDataset): unseen long terms that share fingerprints with
        # vocabulary terms must come out as OOV ids, never as vocabulary ids
        unseen = [bytes(rng.integers(32, 127, L, dtype=np.uint8)) for L in (16, 24, 32, 40, 64, 100) for _ in range(300)]
        items2 = unseen + items[:5000]
        col2 = O.StrCol.from_list(items2)
        got2 = eng.transform(helpers.table_from_oracle({"s": col2}, ktw).to("cuda"), state)["s_id"].values.cpu().numpy()
        assert np.array_equal(got2, O.apply_vocabulary(O.fill_in_missing(col2), vocab, 10))
        # a detached / reloaded state points its lookup table at its own copy of the terms
        det = state.detach()
        got3 = eng.transform(helpers.table_from_oracle({"s": col2}, ktw).to("cuda"), det)["s_id"].values.cpu().numpy()
        assert np.array_equal(got3, got2)
    finally:
        nat.call("txf_debug_set_identity_bits", 119)


def test_vocabulary_tables_grow_instead_of_raising(ktw):
    """Default Engine sizes (65 536 slots, 4 MB of term bytes) are a starting point: a column with more
    distinct terms / bytes grows the table and the arena and runs the pass again (tft has no such knob)."""
    rng = np.random.default_rng(11)
    n = 150_000
    items = [b"%06d-%s" % (i, bytes(rng.integers(97, 123, 30, dtype=np.uint8))) for i in range(n)]  # 37 bytes each: 5.5 MB
    items += items[:2000] * 3
    eng, state, vocab = _vocab_roundtrip(ktw, items, 1000, 10)
    assert eng.vocab_capacity >= 1 << 18 and eng.arena_bytes > 1 << 22 and len(eng.grow_events) >= 2
    assert len(vocab) == 1000


@pytest.mark.parametrize("top_k,thr", [(None, 3), (1000, 5), (10, 2), (None, 10 ** 9), (50, 1)])
def test_vocabulary_frequency_threshold(ktw, top_k, thr):
    """frequency_threshold keeps terms seen at least that many times (tft 0.13 keyword), before top_k."""
    tf, tft = ktw.tf, ktw.tft
    rng = np.random.default_rng(thr)
    terms = _stress_terms(4000, rng)
    counts = np.maximum(1, 2000 // (np.arange(4000) + 1)).astype(np.int64)
    rows = np.repeat(np.arange(4000), counts)
    rng.shuffle(rows)
    col = O.StrCol.from_list([terms[i] for i in rows])
    host = helpers.table_from_oracle({"s": col}, ktw)

    def fn(inputs):
        x = inputs["s"]
        d = tf.squeeze(tf.sparse.to_dense(tf.SparseTensor(x.indices, x.values, [x.dense_shape[0], 1]), ""), axis=1)
        return {"s_id": tft.compute_and_apply_vocabulary(d, top_k=top_k, frequency_threshold=thr, num_oov_buckets=4)}

    eng = ktw.Engine(ktw.trace(fn, host.schema()))
    out, state = eng.analyze_and_transform(host.to("cuda"))
    filled = O.fill_in_missing(col)
    want_vocab = O.finalize_vocabulary(O.vocab_counts(filled), top_k=top_k, frequency_threshold=thr)
    assert state.vocabs[("s", top_k, thr)].terms() == want_vocab
    want_ids = O.apply_vocabulary(filled, want_vocab, 4) if want_vocab != [O.EMPTY_VOCAB_SENTINEL] else None
    if want_ids is not None:
        assert np.array_equal(out["s_id"].values.cpu().numpy(), want_ids)


def test_vocabulary_empty_and_all_missing(ktw):
    col = O.StrCol.from_list([None, b"", None, b"\n", None])
    host = helpers.table_from_oracle({"s": col}, ktw)
    eng = ktw.Engine(ktw.trace(_vocab_fn(ktw, "s", 10, 4), host.schema()))
    out, state = eng.analyze_and_transform(host.to("cuda"))
    filled = O.fill_in_missing(col)
    want_vocab = O.finalize_vocabulary(O.vocab_counts(filled), top_k=10)
    assert want_vocab == [O.EMPTY_VOCAB_SENTINEL]
    assert state.vocabs[("s", 10, None)].terms() == want_vocab
    assert np.array_equal(out["s_id"].values.cpu().numpy(), O.apply_vocabulary(filled, want_vocab, 4))


# ------------------------------------------------------------------------------------------------
# edge cases of the element-wise ops
# ------------------------------------------------------------------------------------------------
def test_label_nan_fare_and_fp32_threshold(ktw, fn):
    """NaN fare -> 0; the threshold is the fp32 product fare * 0.2f (63 of the bundled rows differ
    from an fp64 threshold)."""
    from kubeflow_tfx_workshop_b200 import synth
    host = synth.taxi_table(64, backend="numpy")
    fare = host["fare"].values
    fare[:6] = torch.tensor([float("nan"), 5.0, 10.0, 0.0, float("inf"), 27.05])
    host["tips"].values[:6] = torch.tensor([1.0, 1.0, 2.0000002, 0.0, 1.0, 5.41])
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    out, _ = eng.analyze_and_transform(host.to("cuda"))
    want = O.tips_label(O.fill_in_missing(O.NumCol(fare.numpy())), O.fill_in_missing(O.NumCol(host["tips"].values.numpy())))
    assert np.array_equal(out["tips_xf"].values.cpu().numpy(), want)
    assert int(want[0]) == 0


def test_zscore_constant_column_and_missing(ktw):
    tf, tft = ktw.tf, ktw.tft

    def f(inputs):
        def dense(x):
            return tf.squeeze(tf.sparse.to_dense(tf.SparseTensor(x.indices, x.values, [x.dense_shape[0], 1]), 0), axis=1)
        return {k + "_z": tft.scale_to_z_score(dense(inputs[k])) for k in ("const", "holes", "big")}

    n = 10_007
    rng = np.random.default_rng(3)
    cols = {"const": O.NumCol(np.full(n, 7.5, np.float32)),
            "holes": O.NumCol(rng.normal(100, 3, n).astype(np.float32), rng.random(n) > 0.5),
            "big": O.NumCol(rng.integers(0, 1 << 23, n).astype(np.int64), rng.random(n) > 0.01)}
    host = helpers.table_from_oracle(cols, ktw)
    eng = ktw.Engine(ktw.trace(f, host.schema()))
    out, state = eng.analyze_and_transform(host.to("cuda"))
    for k, c in cols.items():
        filled = O.fill_in_missing(c)
        mean, var = O.finalize_moments(O.moments(filled))
        want = O.z_score(filled, mean, var)
        got = out[k + "_z"].values.cpu().numpy()
        assert helpers.zscore_close(got, want, mean, var), k
    assert state.mean_var_of("const") == (7.5, 0.0)
    assert np.array_equal(out["const_z"].values.cpu().numpy(), np.zeros(n, np.float32))


def test_empty_table(ktw, fn):
    from kubeflow_tfx_workshop_b200 import synth
    host = synth.taxi_table(0, backend="numpy")
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    out, state = eng.analyze_and_transform(host.to("cuda"))
    assert len(out) == 0
    assert state.vocabs[("company", 1000, None)].terms() == [O.EMPTY_VOCAB_SENTINEL]


def test_errors_are_python_exceptions(ktw):
    nat = ktw._native
    with pytest.raises(nat.NativeError, match="ncols"):
        nat.call("txf_moments_finalize", 1, 0, 1, None)
    with pytest.raises(nat.NativeError, match="power of two"):
        nat.call("txf_vocab_compact", 1, 1000, 1, 1, 1, 1, None)
    with pytest.raises(TypeError):
        ktw.tft.bucketize(ktw.graph.Const(1.0), 2.5)
    with pytest.raises(ValueError):
        ktw.tft.bucketize(ktw.graph.Const(1.0), 0)
