"""CPU tests of the host side: C-ABI exports, shims/plan tracing, ingest, device-function logic
compiled for the host, and the multi-rank exchange over gloo."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

import helpers
from helpers import O, ROOT


@pytest.fixture(scope="module")
def ktw():
    import kubeflow_tfx_workshop_b200 as k
    return k


def test_library_exports_every_declared_symbol(ktw):
    """libtaxi_xf.so loads and exports exactly what include/taxi_xf.h declares."""
    header = open(os.path.join(ROOT, "include", "taxi_xf.h")).read()
    declared = sorted(set(re.findall(r"TXF_API [\w\s\*]*?(txf_\w+)\(", header)))
    out = subprocess.run(["nm", "-D", "--defined-only", ktw._native.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T txf_" in l)
    assert declared == exported == ktw._native.EXPORTED
    assert ktw._native.query("txf_abi_version") == 1
    assert ktw._native.query("txf_moments_workspace_bytes", 3) > 0
    assert ktw._native.query("txf_vocab_table_bytes", 1024) == 32 * 1024
    assert ktw._native.query("txf_msel_state_bytes", 4, 18) > ktw._native.query("txf_msel_hist_bytes", 4, 18) == 4 * 18 * 256 * 8


def test_no_cpu_fallback_without_library(tmp_path):
    """The product path fails loudly when the CUDA library is missing."""
    code = ("import sys; sys.path.insert(0, %r)\n"
            "import importlib.util, os\n"
            "p = os.path.join(%r, 'kubeflow-tfx-workshop_b200', '_native.py')\n"
            "src = open(p).read().replace('libtaxi_xf.so', 'libtaxi_xf_missing.so')\n"
            "ns = {'__file__': p, '__name__': 'native_probe'}\n"
            "try:\n    exec(compile(src, p, 'exec'), ns)\nexcept ImportError as e:\n    print('IMPORT-ERROR', e)\n" % (ROOT, ROOT))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True).stdout
    assert "IMPORT-ERROR" in out and "no CPU fallback" in out


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "kubeflow-tfx-workshop_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "tft_oracle" not in text and "liboracle" not in text and "from oracle" not in text, f


def test_engine_refuses_cpu_device(ktw):
    plan = ktw.trace(helpers.load_example_preprocessing_fn(ktw).preprocessing_fn, ktw.synth.TAXI_SCHEMA) if hasattr(ktw, "synth") else None
    from kubeflow_tfx_workshop_b200 import synth
    plan = ktw.trace(helpers.load_example_preprocessing_fn(ktw).preprocessing_fn, synth.TAXI_SCHEMA)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ktw.Engine(plan, device="cpu")


def test_trace_example_preprocessing_fn(ktw):
    from kubeflow_tfx_workshop_b200 import plan as P, synth
    mod = helpers.load_example_preprocessing_fn(ktw)
    plan = ktw.trace(mod.preprocessing_fn, synth.TAXI_SCHEMA)
    kinds = [type(s).__name__ for s in plan.specs]
    assert kinds == ["ZScoreSpec"] * 3 + ["VocabSpec"] * 2 + ["BucketSpec"] * 4 + ["FillSpec"] * 7 + ["RatioLabelSpec"]
    v = plan.of(P.VocabSpec)[0]
    assert (v.top_k, v.num_oov_buckets, v.default_value) == (1000, 10, -1)
    lab = plan.of(P.RatioLabelSpec)[0]
    assert (lab.fare_key, lab.tips_key) == ("fare", "tips") and lab.ratio == pytest.approx(0.2)
    assert plan.output_schema()["pickup_census_tract_xf"] == "bytes" and plan.output_schema()["trip_seconds_xf"] == "float32"
    assert "trip_start_timestamp" not in plan.input_keys()


@pytest.mark.skipif(not os.path.exists("/root/reference/airflow-dags/taxi_utils.py"), reason="reference checkout not present")
def test_reference_taxi_utils_runs_unmodified_through_the_shims(ktw):
    """airflow-dags/taxi_utils.py imports tensorflow / tft / tfma and its preprocessing_fn traces
    to the same plan as examples/taxi_preprocessing.py."""
    import importlib.util
    from kubeflow_tfx_workshop_b200 import synth
    with ktw.install_shims():
        spec = importlib.util.spec_from_file_location("ref_taxi_utils", "/root/reference/airflow-dags/taxi_utils.py")
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
    assert "tensorflow" not in sys.modules or sys.modules["tensorflow"].__name__ != "kubeflow_tfx_workshop_b200.tf_shim"
    schema = dict(synth.TAXI_SCHEMA, trip_start_timestamp="int64")
    a = ktw.trace(ref.preprocessing_fn, schema)
    b = ktw.trace(helpers.load_example_preprocessing_fn(ktw).preprocessing_fn, schema)
    assert a.specs == b.specs
    assert ref._transformed_name("fare") == "fare_xf"


def test_notebook_06_variant_with_sparse_to_dense(ktw):
    """notebooks/06_Airflow_Feature_Analysis.ipynb:176 uses tf.sparse_to_dense(indices, shape, values, default)."""
    tf, tft = ktw.tf, ktw.tft

    def fill(x):
        default_value = '' if x.dtype == tf.string else 0
        return tf.squeeze(tf.sparse_to_dense(x.indices, [x.dense_shape[0], 1], x.values, default_value), axis=1)

    plan = ktw.trace(lambda inputs: {"a": tft.scale_to_z_score(fill(inputs["x"])), "b": fill(inputs["s"])}, {"x": "float32", "s": "bytes"})
    assert [type(s).__name__ for s in plan.specs] == ["ZScoreSpec", "FillSpec"]


def test_unsupported_expressions_raise(ktw):
    tf, tft = ktw.tf, ktw.tft
    with pytest.raises(NotImplementedError):
        ktw.trace(lambda i: {"a": tf.multiply(tf.squeeze(tf.sparse.to_dense(i["x"], 0), axis=1), 2.0)}, {"x": "float32"})
    with pytest.raises(NotImplementedError):  # fill value other than '' / 0
        ktw.trace(lambda i: {"a": tf.squeeze(tf.sparse.to_dense(i["x"], 7), axis=1)}, {"x": "float32"})
    with pytest.raises(TypeError):
        tft.bucketize(ktw.graph.Const(1.0), "10")
    with pytest.raises(ValueError):
        tft.bucketize(ktw.graph.Const(1.0), 0)
    with pytest.raises(ValueError):
        tft.compute_and_apply_vocabulary(ktw.graph.Const("a"), top_k=0)
    with pytest.raises(TypeError):
        tft.scale_to_z_score(ktw.graph.Const("a"))


def test_table_ingest_rows_arrow_csv(ktw, tmp_path):
    rows = [{"f": [1.5], "i": [3], "s": [b"ab"]}, {"f": [], "i": None, "s": []}, {"f": 2.5, "i": [4], "s": "cd"}]
    t = ktw.Table.from_rows(rows, {"f": "float32", "i": "int64", "s": "bytes"})
    assert t["f"].values.tolist() == [1.5, 0.0, 2.5] and t["f"].numpy()[1].tolist() == [True, False, True]
    assert t["s"].to_list() == [b"ab", b"", b"cd"] and t["s"].numpy()[2].tolist() == [True, False, True]
    with pytest.raises(ValueError, match="at most one value"):
        ktw.Table.from_rows([{"f": [1.0, 2.0]}], {"f": "float32"})
    import pyarrow as pa
    at = pa.table({"f": pa.array([1.0, None, 3.0]), "i": pa.array([1, 2, None]), "s": pa.array(["x", None, "yz"]), "z": pa.nulls(3)})
    t2 = ktw.Table.from_arrow(at)
    assert t2["f"].dtype == "float32" and t2["i"].dtype == "int64" and t2["s"].to_list() == [b"x", b"", b"yz"] and t2["z"].dtype == "bytes"
    assert t2["s"].pool.numel() >= t2["s"].pool_bytes + 16
    p = tmp_path / "d.csv"
    p.write_text("a,b,c,d\n1,2.5,x,\n,3,y,\n7,,,\n")
    t3 = ktw.Table.from_csv(str(p))
    assert t3.schema() == {"a": "int64", "b": "float32", "c": "bytes", "d": "bytes"}
    assert t3["a"].values.tolist() == [1, 0, 7] and t3["c"].to_list() == [b"x", b"y", b""]


@pytest.mark.skipif(not os.path.exists("/root/reference/airflow-dags/data/taxi_data/data.csv"), reason="reference checkout not present")
def test_csv_ingest_of_bundled_data_matches_fixture(ktw):
    t = ktw.Table.from_csv("/root/reference/airflow-dags/data/taxi_data/data.csv")
    cols, _, _ = helpers.load_taxi_golden()
    for k, c in cols.items():
        got = t[k]
        if isinstance(c, O.StrCol):
            assert got.to_list() == c.to_list(), k
        else:
            v, valid = got.numpy()
            assert v.dtype == c.values.dtype and np.array_equal(valid, c.valid), k
            assert np.array_equal(np.where(valid, v, 0), np.where(c.valid, c.values, 0)), k


def test_device_hash_and_key_logic_compiled_for_host(tmp_path):
    """farmhash.cuh / common.cuh helpers are __host__ __device__: compile them with g++ through nvcc
    and check them against the oracle without a GPU."""
    src = tmp_path / "hostcheck.cu"
    src.write_text(r'''
#include <cstdio>
#include <cstring>
#include <vector>
#include "common.cuh"
#include "farmhash.cuh"
#include "vocab.cuh"
namespace txf { void set_error(const char*, ...) {} int check_cuda(cudaError_t, const char*) { return 0; } void count_launch(int) {} }
int main(int argc, char** argv) {
  FILE* f = fopen(argv[1], "rb");
  std::vector<unsigned char> buf(1 << 22, 0);
  size_t n = fread(buf.data() + 8, 1, buf.size() - 64, f);
  fclose(f);
  size_t pos = 8;
  while (pos < 8 + n) {
    unsigned len; memcpy(&len, &buf[pos], 4); pos += 4;
    txf::StrKey k = txf::make_key(&buf[pos], len);
    printf("%llu %llu %llu\n", (unsigned long long)txf::fingerprint64(&buf[pos], len), k.a, k.b);
    pos += len;
  }
  float probes[] = {-1.5f, -0.0f, 0.0f, 1e-45f, 2.0f};
  for (float p : probes) printf("K %u %08x\n", txf::f32_key(p), txf::f32_bits(txf::key_f32(txf::f32_key(p))));
  return 0;
}
''')
    exe = tmp_path / "hostcheck"
    csrc = os.path.join(ROOT, "kubeflow-tfx-workshop_b200", "csrc")
    subprocess.run(["nvcc", "-O1", "-std=c++17", "-I", csrc, "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    rng = np.random.default_rng(5)
    items = [bytes(rng.integers(0, 256, L, dtype=np.uint8)) for L in list(range(0, 130)) + [200, 511, 512, 513]]
    # identity corner cases: NUL padding, a leading 8-byte word of zeros, prefixes of each other
    items += [b"", b"\0", b"\0\0", b"\0" * 8, b"\0" * 9, b"\0" * 15, b"\0" * 16, b"\x01", b"ab", b"ab\0", b"ab\0\0",
              b"abcdefgh", b"abcdefgh\0", b"abcdefghi", b"Credit Card", b"Credit Cash", b"x" * 15, b"x" * 16, b"x" * 17]
    # two 32-byte terms whose second window zeroes a factor of each mixer lane (they collided under a bare-product mixer)
    items += helpers.lane_zeroing_pair()
    # same windows, different lengths: the length is part of a long term's key
    items += [b"a" * 20, b"a" * 24, b"a" * 33, b"a" * 64, b"a" * 65, b"a" * 127, b"a" * 128, b"a" * 129]
    blob = b"".join(len(s).to_bytes(4, "little") + s for s in items)
    (tmp_path / "in.bin").write_bytes(blob)
    out = subprocess.run([str(exe), str(tmp_path / "in.bin")], capture_output=True, text=True, check=True).stdout.split("\n")
    got = [int(x.split()[0]) for x in out[: len(items)]]
    assert got == [O.fingerprint64_c(s) for s in items]
    keys128 = [tuple(int(v) for v in x.split()[1:]) for x in out[: len(items)]]
    assert all(a != 0 and b != 0 for a, b in keys128)                      # 0 marks an unclaimed table slot
    assert len(set(keys128)) == len(set(items))                            # distinct terms, distinct identities
    for s_, (a, b) in zip(items, keys128):                                 # terms of <= 15 bytes are stored verbatim
        if len(s_) <= 15:
            raw = s_.ljust(16, b"\0")
            k0, k1 = int.from_bytes(raw[:8], "little"), int.from_bytes(raw[8:], "little") | ((len(s_) + 1) << 56)
            assert (a, b) == ((k0, k1) if k0 else (1, k1 | (0x40 << 56))), s_
    keys = [l.split() for l in out if l.startswith("K ")]
    probes = np.array([-1.5, -0.0, 0.0, 1e-45, 2.0], np.float32)
    assert [int(k[1]) for k in keys] == O.order_key(probes).tolist()
    assert [int(k[2], 16) for k in keys] == O.key_to_f32(O.order_key(probes)).view(np.uint32).tolist()


def test_fingerprint64_long_inputs_device_header_vs_c_oracle(tmp_path):
    """Fingerprint64 for inputs longer than 32 bytes has no reference-held vector.  What this box
    allows: the device header (csrc/farmhash.cuh, compiled for the host) against the separately written C oracle
    on 10^5 random strings of 33-4096 bytes.  Two restatements of the same published algorithm agreeing is
    consistency, not a pin."""
    src = tmp_path / "fp.cu"
    src.write_text(r'''
#include <cstdio>
#include <cstring>
#include <vector>
#include "common.cuh"
#include "farmhash.cuh"
namespace txf { void set_error(const char*, ...) {} int check_cuda(cudaError_t, const char*) { return 0; } void count_launch(int) {} }
int main(int argc, char** argv) {
  FILE* f = fopen(argv[1], "rb");
  fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
  std::vector<unsigned char> buf(sz + 64, 0);
  if (fread(buf.data(), 1, sz, f) != (size_t)sz) return 1;
  fclose(f);
  long n; memcpy(&n, buf.data(), 8);
  const long long* off = reinterpret_cast<const long long*>(buf.data() + 8);
  const unsigned char* pool = buf.data() + 8 + 8 * (n + 1);
  FILE* o = fopen(argv[2], "wb");
  for (long i = 0; i < n; ++i) {
    unsigned long long h = txf::fingerprint64(pool + off[i], (uint32_t)(off[i + 1] - off[i]));
    fwrite(&h, 8, 1, o);
  }
  fclose(o);
  return 0;
}
''')
    exe = tmp_path / "fp"
    csrc = os.path.join(ROOT, "kubeflow-tfx-workshop_b200", "csrc")
    subprocess.run(["nvcc", "-O2", "-std=c++17", "-I", csrc, "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    rng = np.random.default_rng(64)
    lens = np.concatenate([rng.integers(33, 201, 80_000), rng.integers(33, 4097, 20_000), np.arange(33, 200)]).astype(np.int64)
    lens = np.concatenate([lens, [64, 65, 127, 128, 129, 4096]])
    # pool starts 8-byte aligned inside the file buffer: [n][offsets (n+1)][bytes]; pad so that it does
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    pool = rng.integers(0, 256, int(offsets[-1]), dtype=np.uint8)
    (tmp_path / "in.bin").write_bytes(np.int64(len(lens)).tobytes() + offsets.tobytes() + pool.tobytes())
    subprocess.run([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], check=True)
    got = np.fromfile(tmp_path / "out.bin", dtype=np.uint64)
    want = O.fingerprint64_bulk(offsets.astype(np.int32), np.concatenate([pool, np.zeros(64, np.uint8)]))
    assert got.shape == want.shape and np.array_equal(got, want)
    for i in rng.integers(0, len(lens), 50):  # and the pure-Python restatement on a handful
        assert int(got[i]) == O.fingerprint64(pool[offsets[i]:offsets[i + 1]].tobytes())


def test_long_term_fingerprint_quality(tmp_path):
    """Avalanche and slice-collision statistics of the long-term fingerprint (vocab.cuh KeyMixer), on the
    host build of the header.  The fingerprint only pre-filters (matches are confirmed on the bytes), so
    this guards speed, not correctness: every input bit must flip every kept output bit with
    probability near 1/2, and 32-bit slices must collide about as often as random values do."""
    src = tmp_path / "mixer.cu"
    src.write_text(r'''
#include <cstdio>
#include <cstring>
#include <cstdint>
#include <vector>
#include <algorithm>
#include "common.cuh"
#include "farmhash.cuh"
#include "vocab.cuh"
namespace txf { void set_error(const char*, ...) {} int check_cuda(cudaError_t, const char*) { return 0; } void count_launch(int) {} }
static uint64_t rng_state = 88172645463325252ULL;
static uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }
int main() {
  const int lens[] = {16, 17, 24, 31, 32, 33, 40, 64, 65, 100};
  double lo = 1.0, hi = 0.0;
  for (int n : lens) {
    std::vector<uint32_t> flips(static_cast<size_t>(n) * 8 * 119, 0);
    const int trials = 3000;
    std::vector<unsigned char> buf(n + 64, 0);
    for (int t = 0; t < trials; ++t) {
      for (int i = 0; i < n; ++i) buf[i] = static_cast<unsigned char>(rnd());
      const txf::StrKey k0 = txf::make_key(buf.data(), n);
      for (int bit = 0; bit < n * 8; ++bit) {
        buf[bit >> 3] ^= 1u << (bit & 7);
        const txf::StrKey k1 = txf::make_key(buf.data(), n);
        buf[bit >> 3] ^= 1u << (bit & 7);
        const uint64_t da = k0.a ^ k1.a, db = (k0.b ^ k1.b) & ((1ULL << 55) - 1);
        uint32_t* f = &flips[static_cast<size_t>(bit) * 119];
        for (int o = 0; o < 64; ++o) f[o] += (da >> o) & 1;
        for (int o = 0; o < 55; ++o) f[64 + o] += (db >> o) & 1;
      }
    }
    for (uint32_t c : flips) { const double p = static_cast<double>(c) / trials; lo = std::min(lo, p); hi = std::max(hi, p); }
  }
  printf("AVALANCHE %.4f %.4f\n", lo, hi);
  // slice collisions over structured names (digits + words, like the synthetic company names)
  const int N = 1 << 20;
  std::vector<uint32_t> s0(N), s1(N), s2(N), s3(N);
  char name[96];
  for (int i = 0; i < N; ++i) {
    int n = snprintf(name, sizeof(name), "%04d - %05d Taxi Affiliation Services %d", i % 10000, (i * 7) % 100000, i / 10000);
    unsigned char b[160] = {0};
    memcpy(b, name, n);
    const txf::StrKey k = txf::make_key(b, n);
    s0[i] = static_cast<uint32_t>(k.a); s1[i] = static_cast<uint32_t>(k.a >> 32);
    s2[i] = static_cast<uint32_t>(k.b); s3[i] = static_cast<uint32_t>((k.b >> 20) ^ (k.a >> 13));
  }
  for (auto* v : {&s0, &s1, &s2, &s3}) {
    std::sort(v->begin(), v->end());
    long c = 0;
    for (int i = 1; i < N; ++i) c += (*v)[i] == (*v)[i - 1];
    printf("COLLISIONS %ld\n", c);
  }
  return 0;
}
''')
    exe = tmp_path / "mixer"
    csrc = os.path.join(ROOT, "kubeflow-tfx-workshop_b200", "csrc")
    subprocess.run(["nvcc", "-O2", "-std=c++17", "-I", csrc, "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split("\n")
    lo, hi = (float(x) for x in out[0].split()[1:])
    assert 0.42 <= lo and hi <= 0.58, (lo, hi)   # 3000 trials: sigma = 0.009, ~4e5 (input bit, output bit) cells
    expected = (1 << 20) ** 2 / 2 ** 33          # birthday estimate for 2^20 values in 2^32 bins = 128
    for line in out[1:5]:
        assert int(line.split()[1]) < 1.5 * expected, out


def test_shard_range(ktw):
    n = 1_000_003
    spans = [ktw.shard_range(n, r, 8) for r in range(8)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:])) and all(lo % 8 == 0 for lo, _ in spans)
    assert ktw.shard_range(5, 3, 4) == (5, 5)


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import kubeflow_tfx_workshop_b200 as ktw
        from kubeflow_tfx_workshop_b200 import synth
        comm = ktw.Comm()
        n = 20_001
        lo, hi = ktw.shard_range(n, rank, world)
        cols = helpers.oracle_from_table(synth.taxi_table(hi - lo, lo, backend="numpy"))
        part = O.analyze_partial(cols)
        # Welford triples: all-gather, merge in rank order (what txf_moments_merge does on the GPU)
        mine = torch.tensor([part["moments"][k] for k in O.DENSE_FLOAT_KEYS], dtype=torch.float64)
        allm = comm.all_gather(mine)
        merged = {}
        for j, k in enumerate(O.DENSE_FLOAT_KEYS):
            acc = (0.0, 0.0, 0.0)
            for r in range(world):
                acc = O.merge_moments(acc, tuple(allm[r, j].tolist()))
            merged[k] = acc
        # integer histograms of the quantile keys: all-reduce(sum)
        hist = torch.stack([torch.from_numpy(np.bincount(part["qkeys"][k] >> 16, minlength=1 << 16)) for k in O.BUCKET_KEYS])
        comm.all_reduce_sum_(hist)
        # variable-length exchange (candidate keys / vocabulary entries): rank-ordered concatenation
        keys0 = torch.from_numpy(part["qkeys"][O.BUCKET_KEYS[0]].astype(np.int64))
        allk, counts = comm.all_gather_var(torch.cat([keys0, torch.zeros(5, dtype=torch.int64)]), keys0.numel())
        terms = sorted(part["vocab"]["company"].items())
        pool = torch.from_numpy(np.frombuffer(b"".join(t for t, _ in terms) or b"\0", np.uint8).copy())
        lens = torch.tensor([len(t) for t, _ in terms], dtype=torch.int64)
        cnts = torch.tensor([c for _, c in terms], dtype=torch.int64)
        g_pool, _ = comm.all_gather_var(pool, int(lens.sum()))
        g_lens, _ = comm.all_gather_var(lens, len(terms))
        g_cnts, _ = comm.all_gather_var(cnts, len(terms))
        buf = g_pool.numpy().tobytes()
        ends = np.cumsum(g_lens.numpy())
        vocab = {}
        for e, L, c in zip(ends, g_lens.numpy(), g_cnts.numpy()):
            t = buf[e - L:e]
            vocab[t] = vocab.get(t, 0) + int(c)
        q.put((rank, merged, hist.sum(dim=1).tolist(), int(allk.numel()), counts, vocab))
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_over_gloo(ktw):
    """world_size 2, gloo: the collectives used between analyze and transform reproduce the
    single-process analyzer state."""
    import torch.multiprocessing as mp
    from kubeflow_tfx_workshop_b200 import synth
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted([q.get(timeout=180) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n = 20_001
    whole = O.analyze_partial(helpers.oracle_from_table(synth.taxi_table(n, backend="numpy")))
    for rank, merged, hist_tot, nk, counts, vocab in results:
        for k in O.DENSE_FLOAT_KEYS:
            assert merged[k][0] == n
            assert merged[k][1] == pytest.approx(whole["moments"][k][1], rel=1e-12)
            assert merged[k][2] == pytest.approx(whole["moments"][k][2], rel=1e-10)
        assert hist_tot == [n] * 4 and nk == n and sum(counts) == n
        assert vocab == dict(whole["vocab"]["company"])
    assert results[0][1] == results[1][1]  # identical on every rank (rank-ordered merge)


# ------------------------------------------------------------------------------------------------
# tracing of the sibling shapes (no GPU needed: tracing computes nothing)
# ------------------------------------------------------------------------------------------------
def test_oracle_reproduces_notebook03_known_answer():
    """notebooks/03_TensorFlow_Transform.ipynb:227-231 -- the one result the reference states."""
    r = O.notebook03_example([1, 2, 3], [1, 2, 3], [b"hello", b"world", b"hello"])
    assert r["x_centered"].tolist() == [-1.0, 0.0, 1.0]
    assert r["y_normalized"].tolist() == [0.0, 0.5, 1.0]
    assert r["s_integerized"].tolist() == [0, 1, 0]
    assert r["x_centered_times_y_normalized"].tolist() == [0.0, 0.0, 1.0]
    assert np.signbit(r["x_centered_times_y_normalized"]).tolist() == [True, False, False]
    assert O.scale_by_min_max(np.array([4, 4], np.float32), 4, 4).tolist() == [0.5, 0.5]


def test_trace_notebook03_and_kfp_variants(ktw):
    tf, tft = ktw.tf, ktw.tft

    def nb03(inputs):
        x_centered = inputs["x"] - tft.mean(inputs["x"])
        y_normalized = tft.scale_to_0_1(inputs["y"])
        return {"x_centered": x_centered, "y_normalized": y_normalized,
                "s_integerized": tft.compute_and_apply_vocabulary(inputs["s"]),
                "x_centered_times_y_normalized": x_centered * y_normalized}

    sch = ktw.tf_metadata.dataset_schema.from_feature_spec({
        "x": tf.FixedLenFeature([], tf.float32), "y": tf.FixedLenFeature([], tf.float32), "s": tf.FixedLenFeature([], tf.string)})
    assert sch.dense_keys == {"x", "y", "s"}
    plan = ktw.trace(nb03, sch)
    kinds = [type(s).__name__ for s in plan.specs]
    assert kinds == ["ArithSpec", "ScaleSpec", "VocabSpec", "ArithSpec"]
    assert plan.specs[0].a == ["col", "x"] and plan.specs[0].b == ["scalar", "mean", "x"]
    assert plan.specs[3].a == ["out", "x_centered"] and plan.specs[3].b == ["out", "y_normalized"]
    assert plan.output_schema() == {"x_centered": "float32", "y_normalized": "float32", "s_integerized": "int64",
                                    "x_centered_times_y_normalized": "float32"}

    # an operand that is not itself an output becomes a hidden temporary, emitted before its user
    plan2 = ktw.trace(lambda i: {"o": (i["x"] - tft.mean(i["x"])) * tft.scale_to_0_1(i["y"])}, sch)
    assert [s.out_key for s in plan2.specs] == ["__tmp0", "__tmp1", "o"] and list(plan2.output_schema()) == ["o"]

    # tft.vocabulary is an analyzer without an output column
    plan3 = ktw.trace(lambda i: (tft.vocabulary(i["s"], vocab_filename="s_vocab"), {"s": i["s"]})[1], sch)
    assert [type(s).__name__ for s in plan3.specs] == ["VocabOnlySpec", "FillSpec"] and list(plan3.output_schema()) == ["s"]

    with pytest.raises(NotImplementedError, match="at least one operand must be a column"):
        ktw.trace(lambda i: {"o": tft.mean(i["x"]) - 1.0}, sch)
    with pytest.raises(ValueError, match="output_min"):
        ktw.trace(lambda i: {"o": tft.scale_by_min_max(i["x"], 1.0, 1.0)}, sch)

    ref = "/root/reference/kubeflow-pipelines/taxi/preprocessing.py"
    if os.path.exists(ref):  # the reference's KFP variant, unmodified
        import importlib.util
        import json
        with ktw.install_shims():
            spec = importlib.util.spec_from_file_location("kfp_preprocessing", ref)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
        with open(os.path.join(os.path.dirname(ref), "schema.json")) as f:
            cols = {c["name"]: ("float32" if c["type"] == "NUMBER" else "bytes") for c in json.load(f)}
        plan4 = ktw.trace(mod.preprocess, cols, dense_keys=set(cols))
        kinds = [type(s).__name__ for s in plan4.specs]
        assert kinds == ["ZScoreSpec"] * 3 + ["VocabSpec"] * 6 + ["BucketSpec"] * 4 + ["CastSpec"] * 3 + ["RatioLabelSpec"]
        assert plan4.specs[3].vocab_filename == "vocab_pickup_census_tract" and not plan4.specs[3].as_string
        cols["pickup_community_area"] = "int64"
        plan5 = ktw.trace(mod.preprocess, cols, dense_keys=set(cols))
        assert [s.in_key for s in plan5.specs if getattr(s, "as_string", False)] == ["pickup_community_area"]


def test_exact_decimal_to_double_compiled_for_host(tmp_path):
    """csrc/decimal.cuh is __host__ __device__: the big-integer conversion the CSV ingest uses for
    long literals must give Python's float() bit for bit (it is correctly rounded)."""
    src = tmp_path / "dec.cu"
    src.write_text(r'''
#include <cstdio>
#include <cstring>
#include "decimal.cuh"
int main() {
  unsigned long long m; int e;
  while (scanf("%llu %d", &m, &e) == 2) {
    double d = txf::decimal_to_double_exact(m, e);
    unsigned long long u; memcpy(&u, &d, 8);
    printf("%016llx\n", u);
  }
}
''')
    exe = tmp_path / "dec"
    subprocess.run(["nvcc", "-O1", "-std=c++17", "-I", os.path.join(ROOT, "kubeflow-tfx-workshop_b200", "csrc"), str(src), "-o", str(exe)],
                   check=True, capture_output=True)
    rng = np.random.default_rng(11)
    cases = [(1, 0), (1, -77), (1, 77), (2 ** 64 - 1, -77), (2 ** 64 - 1, 0), (9007199254740993, 0), (9007199254740993, 5),
             (5, -1), (4503599627370497, -1), (17976931348623157, 60)]
    for _ in range(60_000):
        digits = int(rng.integers(1, 20))
        m = int(rng.integers(1, 10)) * 10 ** (digits - 1) + (int(rng.integers(0, 10 ** (digits - 1))) if digits > 1 else 0)
        cases.append((min(m, 2 ** 64 - 1), int(rng.integers(-77, 78))))
    import struct
    for _ in range(20_000):  # 17-digit reprs of random doubles: must round-trip
        x = float(np.ldexp(float(rng.integers(1, 2 ** 53)), int(rng.integers(-200, 150))))
        mant, exp = ("%.16e" % x).split("e")
        m, e = int(mant.replace(".", "")), int(exp) - 16
        if -77 <= e <= 77:
            cases.append((m, e))
    out = subprocess.run([str(exe)], input="".join(f"{m} {e}\n" for m, e in cases), text=True, capture_output=True, check=True).stdout.split()
    assert len(out) == len(cases)
    for (m, e), got in zip(cases, out):
        want = struct.pack(">d", float(f"{m}e{e}")).hex()
        assert got == want, (m, e, got, want)


def test_stream_kernel_pipeline_protocol_model():
    """tools/model/stream_pipeline_model.py: random schedules (with one consumer warp starved for long stretches)
    of stream_ops_kernel's full/empty protocol -- every warp sees every tile once, no deadlock; the same model
    run with a single arrival on the empty barrier does lose tiles, which is what the per-warp arrival is for."""
    import importlib.util
    import random
    spec = importlib.util.spec_from_file_location("stream_pipeline_model", os.path.join(ROOT, "tools", "model", "stream_pipeline_model.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    rng = random.Random(7)
    results = [m.run(rng.randint(0, 40), rng.choice([0.0, 0.3, 1.0]), False, True, rng) for _ in range(400)]
    assert set(results) == {"ok"}
    old = [m.run(rng.randint(20, 40), 0.0, True, True, rng) for _ in range(200)]
    assert any(r != "ok" for r in old)


def test_count_table_insert_protocol_model():
    """tools/model/count_table_model.py: random interleavings of the global count table's lock-free insert with many
    terms sharing a (truncated) fingerprint and a home slot -- one slot per distinct term, exact counts; without the
    byte compare the same schedules merge terms."""
    import importlib.util
    import random
    spec = importlib.util.spec_from_file_location("count_table_model", os.path.join(ROOT, "tools", "model", "count_table_model.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    rng = random.Random(3)
    for verify, expect_ok in ((True, True), (False, False)):
        results = [m.run(rng.randint(2, 8), 64, m.make_terms(rng, rng.randint(4, 14), 2, 4), rng, verify) for _ in range(200)]
        assert all(r == "ok" for r in results) == expect_ok, results[:5]


def test_unsupported_tft_keywords_raise_at_trace_time(ktw):
    """Keyword names follow tft 0.13; what this path does not implement raises where tft would be
    called -- at trace time -- never silently ignored."""
    tft, g = ktw.tft, ktw.graph
    s, x = g.Const("a"), g.Const(1.0)
    for call in (lambda: tft.scale_to_z_score(x, elementwise=True),
                 lambda: tft.compute_and_apply_vocabulary(s, weights=x),
                 lambda: tft.compute_and_apply_vocabulary(s, labels=x),
                 lambda: tft.compute_and_apply_vocabulary(x),              # numeric input
                 lambda: tft.vocabulary(s, store_frequency=True),
                 lambda: tft.mean(x, reduce_instance_dims=False),
                 lambda: tft.scale_to_0_1(x, elementwise=True),
                 lambda: tft.bucketize(x, 10, weights=x)):
        with pytest.raises(NotImplementedError):
            call()
    for call in (lambda: tft.compute_and_apply_vocabulary(s, num_oov_buckets=-1),
                 lambda: tft.compute_and_apply_vocabulary(s, top_k=True),
                 lambda: tft.compute_and_apply_vocabulary(s, frequency_threshold=-2),
                 lambda: tft.scale_by_min_max(x, 1.0, 1.0)):
        with pytest.raises(ValueError):
            call()
    with pytest.raises(TypeError):
        tft.bucketize(x, True)
    assert tft.string_to_int is tft.compute_and_apply_vocabulary and tft.uniques is tft.vocabulary


def test_committed_bench_lines_are_self_consistent():
    """Every bench.py record under profiles/ recomputes from its own primary fields (tools/check_bench_line.py):
    value from rows and ms_per_step, roofline fractions from achieved / peak, the end-to-end fraction of the
    measured host-link ceiling from the bytes copied."""
    import glob
    import importlib.util
    import json
    spec = importlib.util.spec_from_file_location("check_bench_line", os.path.join(ROOT, "tools", "check_bench_line.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    paths = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_bench*.json")))
    assert len(paths) >= 8
    for p in paths:
        line = json.loads(open(p).read().strip().splitlines()[-1])
        assert m.check(line) == [], p
        assert line.get("parity_check", {"ok": True})["ok"] is True, p
