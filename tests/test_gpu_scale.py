"""GPU tests (-m gpu) at BASELINE.json's stress sizes (configs 4 and 5), checked through
size-independent properties with torch's own sort / unique / bucketize as the independent
implementation (the numpy oracle would take minutes at these sizes; it checks the cut-off set only)."""
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


def test_config4_ten_million_distinct_terms(ktw):
    """10 M distinct strings (decimal text of scrambled 64-bit ids, 1-20 bytes: both identity classes),
    constructed multiplicities with a run of equal counts straddling rank 1000 so that the byte
    tie-break decides the cut-off; top_k=1000, num_oov_buckets=10; ids bit-exact."""
    tf, tft = ktw.tf, ktw.tft
    dev = "cuda"
    D = 10_000_000
    g = torch.Generator(device=dev).manual_seed(4)
    ids = torch.randint(-2 ** 62, 2 ** 62, (D,), generator=g, device=dev, dtype=torch.int64)
    ids[: 1 << 20] >>= 40  # a million short terms (<= 8 digits) so the exact-identity class is populated too
    ids = torch.unique(ids)
    ids = ids[torch.randperm(ids.numel(), generator=g, device=dev)]
    D = ids.numel()
    counts = torch.ones(D, dtype=torch.int64, device=dev)
    counts[:900] = torch.arange(5000, 4100, -1, device=dev)      # clear winners
    counts[900:1300] = 37                                          # 400 terms tie across the cut at rank 1000
    counts[1300:200_000] = torch.randint(2, 30, (200_000 - 1300,), generator=g, device=dev)
    rows = torch.repeat_interleave(ids, counts)
    rows = rows[torch.randperm(rows.numel(), generator=g, device=dev)]
    n = rows.numel()
    valid_bits = None
    table = ktw.Table({"key": ktw.NumericColumn(rows, valid_bits)})

    def fn(inputs):
        return {"id": tft.compute_and_apply_vocabulary(tf.as_string(inputs["key"]), top_k=1000, num_oov_buckets=10)}

    eng = ktw.Engine(ktw.trace(fn, {"key": "int64"}, dense_keys={"key"}), vocab_capacity=1 << 25, arena_bytes=1 << 28)
    out, state = eng.analyze_and_transform(table)
    torch.cuda.synchronize()
    res = next(iter(state.vocabs.values()))
    got_vocab = res.terms()
    assert res.size == 1000 and len(got_vocab) == 1000
    # expected vocabulary: every term with count >= the cut-off count, ordered by the oracle's rule
    cand = torch.nonzero(counts >= 37).flatten()
    cand_counts = {str(int(i)).encode(): int(c) for i, c in zip(ids[cand].tolist(), counts[cand].tolist())}
    want_vocab = O.finalize_vocabulary(cand_counts, top_k=1000)
    assert got_vocab == want_vocab
    assert sum(1 for t in want_vocab if cand_counts[t] == 37) == 100  # the tie really straddles the cut
    # ids: in-vocabulary rows map to their rank; the rest hash into [1000, 1010) exactly like the oracle
    got = out["id"].values
    rank_of = {int(t): r for r, t in enumerate(want_vocab)}
    in_ids = torch.tensor(list(rank_of.keys()), dtype=torch.int64, device=dev)
    in_rank = torch.tensor(list(rank_of.values()), dtype=torch.int64, device=dev)
    order = torch.argsort(in_ids)
    pos = torch.searchsorted(in_ids[order], rows).clamp(max=in_ids.numel() - 1)
    hit = in_ids[order][pos] == rows
    assert torch.equal(got[hit], in_rank[order][pos][hit])
    oov = got[~hit]
    assert int(oov.min()) >= 1000 and int(oov.max()) < 1010
    sample = torch.nonzero(~hit).flatten()[:: max(int((~hit).sum()) // 20000, 1)]
    strings = [str(v).encode() for v in rows[sample].tolist()]
    want = np.array([1000 + O.fingerprint64_c(s) % 10 for s in strings], np.int64)
    assert np.array_equal(got[sample].cpu().numpy(), want)
    # conservation: every row got an id, every bucket of the OOV range is used
    assert got.numel() == n and torch.unique(oov).numel() == 10


@pytest.mark.parametrize("nb", [10, 100, 1000])
def test_config5_eight_float_columns(ktw, nb):
    """8 lat/long-style float32 columns x 24 M rows: 4 centroid-style (heavy duplicates, an atom at
    0.0 from missing cells) and 4 fine-grained.  Boundaries must be the exact order statistics at
    ranks ceil(i N / nb) (distinct values only) and ids must equal torch.bucketize(right=True)."""
    tf, tft = ktw.tf, ktw.tft
    dev = "cuda"
    n = 24_000_000
    g = torch.Generator(device=dev).manual_seed(nb)
    cols, data = {}, {}
    cent = (torch.randn(800, generator=g, device=dev) * 0.1 + 41.8).float()
    for c in range(8):
        if c < 4:
            idx = (torch.rand(n, generator=g, device=dev) ** 3 * 800).long().clamp(max=799)
            x = cent[idx] + c
            miss = torch.rand(n, generator=g, device=dev) < 0.032
            bits = torch.from_numpy(np.packbits((~miss).cpu().numpy(), bitorder="little")).to(dev)
            cols[f"c{c}"] = ktw.NumericColumn(x.contiguous(), bits)
            data[f"c{c}"] = torch.where(miss, torch.zeros_like(x), x)
        else:
            x = (torch.randn(n, generator=g, device=dev) * 0.05 - 87.6 + c).float()
            cols[f"c{c}"] = ktw.NumericColumn(x)
            data[f"c{c}"] = x
    table = ktw.Table(cols)

    def fn(inputs):
        def dense(x):
            return tf.squeeze(tf.sparse.to_dense(tf.SparseTensor(x.indices, x.values, [x.dense_shape[0], 1]), 0), axis=1)
        return {k + "_b": tft.bucketize(dense(inputs[k]), nb) for k in cols}

    eng = ktw.Engine(ktw.trace(fn, table.schema()))
    out, state = eng.analyze_and_transform(table)
    torch.cuda.synchronize()
    res = state.buckets[nb]
    ranks = torch.tensor([-(-i * n // nb) for i in range(1, nb)], dtype=torch.int64, device=dev)  # ceil(i n / nb), 1-based
    for k in cols:
        srt = torch.sort(data[k]).values
        want = torch.unique_consecutive(srt[ranks - 1])
        got = torch.from_numpy(res.boundaries_of(k)).to(dev)
        assert got.numel() == want.numel() and torch.equal(got.view(torch.int32), want.view(torch.int32)), (k, nb)
        ids = torch.bucketize(data[k], got, right=True)
        assert torch.equal(out[k + "_b"].values, ids), (k, nb)
        del srt, ids


def test_config2_hundred_million_rows(ktw):
    """BASELINE configs[1] at full size: 100 M synthetic taxi rows on one GPU.  Every analyzer result
    and every output cell is checked on the device against an independent torch computation (sort,
    bincount, bucketize, fp32 arithmetic); the vocabulary order comes from the oracle's rule applied
    to the exact counts."""
    from kubeflow_tfx_workshop_b200 import synth
    dev = "cuda"
    n = 100_000_000
    fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
    table = synth.taxi_table(n)
    eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA))
    out, state = eng.analyze_and_transform(table)
    torch.cuda.synchronize()

    def filled(key):
        c = table[key]
        if c.valid is None:
            return c.values
        bits = c.valid[: (n + 7) // 8]
        present = ((bits.unsqueeze(1) >> torch.arange(8, device=dev, dtype=torch.uint8)) & 1).reshape(-1)[:n].bool()
        return torch.where(present, c.values, torch.zeros_like(c.values))

    # ---- z-score: moments against fp64 torch reductions, outputs against fp32 torch arithmetic ----
    for key in ("trip_miles", "fare", "trip_seconds"):
        x = filled(key)
        xd = x.double()
        mean, var = float(xd.mean()), float(xd.var(unbiased=False))
        m, v = state.mean_var_of(key)
        assert abs(m - mean) <= 1e-6 * abs(mean) and abs(v - var) <= 1e-6 * var, (key, m, mean, v, var)
        mt = torch.tensor(m, dtype=torch.float32, device=dev)
        want = (x.float() - mt) / torch.sqrt(torch.tensor(v, dtype=torch.float32, device=dev))
        got = out[key + "_xf"].values
        atol = 4 * float(np.spacing(np.float32(abs(m)) + np.float32(1))) / float(np.sqrt(np.float32(v)))
        assert torch.allclose(got, want, rtol=1e-6, atol=atol), key
        assert abs(float(got.double().mean())) < 1e-3 and abs(float(got.double().var(unbiased=False)) - 1.0) < 1e-3, key
        del x, xd, want
    # ---- bucketize: exact order statistics and upper_bound ids --------------------------------------
    ranks = torch.tensor([-(-i * n // 10) for i in range(1, 10)], dtype=torch.int64, device=dev)
    res = state.buckets[10]
    for key in ("pickup_latitude", "pickup_longitude", "dropoff_latitude", "dropoff_longitude"):
        x = filled(key)
        srt = torch.sort(x).values
        want_b = torch.unique_consecutive(srt[ranks - 1])
        got_b = torch.from_numpy(res.boundaries_of(key)).to(dev)
        assert got_b.numel() == want_b.numel() and torch.equal(got_b.view(torch.int32), want_b.view(torch.int32)), key
        assert torch.equal(out[key + "_xf"].values, torch.bucketize(x, got_b, right=True)), key
        del x, srt
    # ---- vocabularies: regenerate the term index of every row, count exactly, order by the oracle's rule
    xp = synth._Th(dev)
    pidx = torch.empty(n, dtype=torch.int64, device=dev)
    cidx = torch.empty(n, dtype=torch.int64, device=dev)
    cmiss = torch.empty(n, dtype=torch.bool, device=dev)
    for lo in range(0, n, 1 << 23):
        hi = min(lo + (1 << 23), n)
        rows = xp.rows(lo, hi)
        r = synth._draw(xp, rows, synth.SEED, "payment_type") % 10000
        p = (r >= synth._PAYMENT_THRESHOLDS[0]) * 1
        for t in synth._PAYMENT_THRESHOLDS[1:]:
            p = p + (r >= t) * 1
        pidx[lo:hi] = p
        cidx[lo:hi] = synth._skewed_index(xp, rows, synth.SEED, "company", 0, synth.N_COMPANIES)
        cmiss[lo:hi] = synth._draw(xp, rows, synth.SEED, "company", 2) % 100 < 34
    empty_id = lambda size: size + O.fingerprint64_c(b"") % 10  # a missing cell is '' -> out of vocabulary
    for key, idx, miss, names in (("payment_type", pidx, None, synth.PAYMENT_TYPES), ("company", cidx, cmiss, synth.company_names())):
        live = idx if miss is None else idx[~miss]
        counts = torch.bincount(live, minlength=len(names)).cpu().tolist()
        want_vocab = O.finalize_vocabulary({t: c for t, c in zip(names, counts) if c}, top_k=1000)
        got_vocab = state.vocabs[(key, 1000, None)].terms()
        assert got_vocab == want_vocab, key
        rank = {t: i for i, t in enumerate(want_vocab)}
        lut = torch.tensor([rank.get(t, -7) for t in names], dtype=torch.int64, device=dev)  # names that never occur are never looked up
        want_ids = lut[idx]
        if miss is not None:
            want_ids = torch.where(miss, torch.full_like(want_ids, empty_id(len(want_vocab))), want_ids)
        assert torch.equal(out[key + "_xf"].values, want_ids), key
        del want_ids
    # ---- pass-through categoricals, the all-missing string feature, and the label ------------------
    for key in ("trip_start_hour", "trip_start_day", "trip_start_month", "pickup_community_area", "dropoff_community_area",
                "dropoff_census_tract"):
        assert torch.equal(out[key + "_xf"].values, filled(key)), key
    pct = out["pickup_census_tract_xf"]
    assert int(pct.offsets.abs().max()) == 0 and pct.offsets.numel() == n + 1
    fare, tips = filled("fare"), filled("tips")
    want = (tips > fare * torch.tensor(0.2, dtype=torch.float32, device=dev)).to(torch.int64)
    assert torch.equal(out["tips_xf"].values, want)
    assert 0.05 < float(want.float().mean()) < 0.5
