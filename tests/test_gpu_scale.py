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
