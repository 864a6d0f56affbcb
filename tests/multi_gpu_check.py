"""Multi-GPU parity check, launched by torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tests/multi_gpu_check.py [rows]

Every rank analyzes its contiguous row shard, the partial states are merged over NCCL, every rank
transforms its shard; rank 0 checks state and outputs against the oracle run on ALL rows, and
that every rank ended with the same state.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 700_001
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import helpers
    from helpers import O
    import kubeflow_tfx_workshop_b200 as ktw
    from kubeflow_tfx_workshop_b200 import synth

    fn = helpers.load_example_preprocessing_fn(ktw).preprocessing_fn
    comm = ktw.Comm()
    lo, hi = ktw.shard_range(n, rank, world)
    shard_host = synth.taxi_table(hi - lo, lo, backend="numpy")
    # (general quantile engine?, fixed-layout exchange?, tiny exchange sections that force the overflow fallbacks?)
    for force_general, fixed, tiny in ((False, True, False), (True, True, False), (False, False, False), (False, True, True)):
        eng = ktw.Engine(ktw.trace(fn, synth.TAXI_SCHEMA), device=f"cuda:{local}", comm=comm)
        eng.force_general_quantiles = force_general
        eng.fixed_exchange = fixed
        if tiny:
            eng.xchg_terms = 8
        # the caller sharded the rows itself: every second configuration passes the row counts instead of having them gathered
        rpr = [b - a for a, b in (ktw.shard_range(n, r, world) for r in range(world))] if fixed and not tiny else None
        out, state = eng.analyze_and_transform(shard_host.to(f"cuda:{local}"), rows_per_rank=rpr)
        torch.cuda.synchronize()
        # identical state on every rank
        sig = []
        for k in state.moment_keys:
            sig += list(state.mean_var_of(k))
        for nb, res in state.buckets.items():
            for k in res.keys:
                sig += res.boundaries_of(k).astype(np.float64).tolist()
        for key, res in state.vocabs.items():
            sig += [float(res.size), float(sum(len(t) for t in res.terms()))]
        t = torch.tensor(sig, dtype=torch.float64, device=f"cuda:{local}")
        g = comm.all_gather(t)
        assert bool((g == g[0]).all()), "analyzer state differs between ranks"
        # parity of this rank's shard against the oracle over ALL rows (recomputed per rank: small n)
        whole = synth.taxi_table(n, backend="numpy")
        want, wstate = O.analyze_and_transform(helpers.oracle_from_table(whole))
        for k, (mean, var) in wstate["zscore"].items():
            m, v = state.mean_var_of(k)
            assert abs(m - float(mean)) <= 1e-6 * abs(float(mean)) and abs(v - float(var)) <= 1e-6 * abs(float(var)), k
        for key, res in state.vocabs.items():
            assert res.terms() == wstate["vocab"][key[0]], key
        for nb, res in state.buckets.items():
            for k in res.keys:
                assert np.array_equal(res.boundaries_of(k).view(np.uint32), wstate["boundaries"][k].view(np.uint32)), k
        for key, w in want.items():
            col = out[key]
            if isinstance(w, list):
                assert col.to_list() == w[lo:hi], key
            elif w.dtype == np.float32:
                mean, var = wstate["zscore"][key[:-3]]
                assert helpers.zscore_close(col.values.cpu().numpy(), w[lo:hi], mean, var), key
            else:
                assert np.array_equal(col.values.cpu().numpy(), w[lo:hi]), key
        if rank == 0:
            print(f"multi-gpu parity ok: world={world} rows={n} fixed_exchange={fixed} tiny={tiny} "
                  f"quantile path={eng.last_quantile_path}", flush=True)
    # sibling analyzers across ranks: min/max merge, mean scalar, as_string vocabulary
    tf, tft = ktw.tf, ktw.tft

    def sib(inputs):
        fare = inputs["fare"]
        return {"unit": tft.scale_to_0_1(fare), "centered": fare - tft.mean(fare),
                "area": tft.compute_and_apply_vocabulary(tf.as_string(inputs["pickup_community_area"]), top_k=20, num_oov_buckets=3)}

    schema = {"fare": "float32", "pickup_community_area": "int64"}
    eng = ktw.Engine(ktw.trace(sib, schema, dense_keys=set(schema)), device=f"cuda:{local}", comm=comm)
    shard = shard_host.select(list(schema)).to(f"cuda:{local}")
    out, state = eng.analyze_and_transform(shard)
    torch.cuda.synchronize()
    cols = helpers.oracle_from_table(synth.taxi_table(n, backend="numpy").select(list(schema)))
    fare = O.fill_in_missing(cols["fare"])
    mn, mx = O.min_max(fare)
    assert np.array_equal(np.array(state.min_max_of("fare"), np.float32).view(np.uint32), np.array([mn, mx], np.float32).view(np.uint32))
    got = out["unit"].values.cpu().numpy()
    assert np.array_equal(got.view(np.uint32), O.scale_by_min_max(fare, mn, mx)[lo:hi].view(np.uint32))
    mean = np.float32(state.mean_var_of("fare")[0])
    assert abs(float(mean) - float(O.mean_f32(fare))) <= 1e-6 * abs(float(mean))
    assert np.array_equal(out["centered"].values.cpu().numpy().view(np.uint32), O.arith_f32("sub", fare, mean)[lo:hi].view(np.uint32))
    strings = [str(int(v)).encode() for v in O.fill_in_missing(cols["pickup_community_area"]).tolist()]
    vocab = O.finalize_vocabulary(O.vocab_counts(strings), top_k=20)
    assert np.array_equal(out["area"].values.cpu().numpy(), O.apply_vocabulary(strings, vocab, num_oov_buckets=3)[lo:hi])
    if rank == 0:
        print(f"multi-gpu sibling parity ok: world={world} rows={n}", flush=True)
    # configs[3] across ranks: millions of distinct terms per column -> the fixed-layout sections overflow, the sized
    # exchange runs, and (top_k vocabulary) every rank re-counts only the terms it owns.  Checked against a single-rank
    # engine that sees the rows of all ranks, and against the all-entries-everywhere merge.
    class NoComm:
        active, enabled, rank, world = False, False, 0, 1

    spec = __import__("importlib.util").util.spec_from_file_location("stress", os.path.join(ROOT, "examples", "stress_preprocessing.py"))
    with ktw.install_shims():
        stress = __import__("importlib.util").util.module_from_spec(spec)
        spec.loader.exec_module(stress)
    nv, D = 3_000_000, 2_000_000
    dev = f"cuda:{local}"
    mine = synth.vocab_stress_table(nv, row0=rank * nv, n_distinct=D, device=dev)
    plan = ktw.trace(stress.vocab_stress_fn, mine.schema())
    solo = ktw.Engine(plan, device=dev, comm=NoComm(), vocab_capacity=1 << 23, arena_bytes=1 << 27)
    everyone = [synth.vocab_stress_table(nv, row0=r * nv, n_distinct=D, device=dev) for r in range(world)]
    st1 = solo.analyze(everyone)
    want_ids = solo.transform(everyone[rank], st1)
    for partition in (True, False):
        eng = ktw.Engine(plan, device=dev, comm=comm, vocab_capacity=1 << 23, arena_bytes=1 << 27)
        eng.partition_vocab = partition
        out, state = eng.analyze_and_transform(mine, rows_per_rank=[nv] * world)
        torch.cuda.synchronize()
        assert set(eng.last_vocab_exchange.values()) == ({"owner-partitioned"} if partition else set()), eng.last_vocab_exchange
        for key, res in state.vocabs.items():
            assert res.terms() == st1.vocabs[key].terms(), (key, partition)
        for k in out.keys():
            assert torch.equal(out[k].values, want_ids[k].values), (k, partition)
    if rank == 0:
        print(f"multi-gpu vocabulary stress parity ok: world={world} rows/rank={nv} distinct={D}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
