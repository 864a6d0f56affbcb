"""GPU parity tests (-m gpu) for the "next" rows of SURVEY.md section 8(f): the sibling mappers of
the workshop notebooks (f4), the KFP ``preprocess()`` variant (f3) and state re-application (f1).
Every check goes through the C ABI and compares with the oracle."""
import importlib.util
import json
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


def bits(a):
    return np.asarray(a, np.float32).view(np.uint32)


def notebook03_fn(tft):
    """The body of notebooks/03_TensorFlow_Transform.ipynb:141-157 (an equivalent, not a copy)."""
    def preprocessing_fn(inputs):
        x, y, s = inputs["x"], inputs["y"], inputs["s"]
        x_centered = x - tft.mean(x)
        y_normalized = tft.scale_to_0_1(y)
        return {"x_centered": x_centered, "y_normalized": y_normalized,
                "s_integerized": tft.compute_and_apply_vocabulary(s),
                "x_centered_times_y_normalized": x_centered * y_normalized}
    return preprocessing_fn


def test_notebook03_known_answer(ktw):
    """The only known-answer the reference itself states (notebooks/03_...:227-231)."""
    tf, tft, tft_beam = ktw.tf, ktw.tft, ktw.tft_beam
    raw_data = [{"x": 1, "y": 1, "s": "hello"}, {"x": 2, "y": 2, "s": "world"}, {"x": 3, "y": 3, "s": "hello"}]
    meta = ktw.tf_metadata.dataset_metadata.DatasetMetadata(ktw.tf_metadata.dataset_schema.from_feature_spec({
        "y": tf.FixedLenFeature([], tf.float32), "x": tf.FixedLenFeature([], tf.float32), "s": tf.FixedLenFeature([], tf.string)}))
    with tft_beam.Context(temp_dir="/tmp/unused"):
        (transformed, _), _ = (raw_data, meta) | tft_beam.AnalyzeAndTransformDataset(notebook03_fn(tft))
    rows = tft_beam.rows(transformed)
    assert [r["x_centered"] for r in rows] == [-1.0, 0.0, 1.0]
    assert [r["y_normalized"] for r in rows] == [0.0, 0.5, 1.0]
    assert [r["s_integerized"] for r in rows] == [0, 1, 0]
    prod = np.array([r["x_centered_times_y_normalized"] for r in rows], np.float32)
    assert prod.tolist() == [0.0, 0.0, 1.0] and np.signbit(prod).tolist() == [True, False, False]  # [-0., 0., 1.]
    want = O.notebook03_example([1, 2, 3], [1, 2, 3], [b"hello", b"world", b"hello"])
    for k, w in want.items():
        got = transformed[k].values.cpu().numpy()
        assert np.array_equal(got.view(np.uint32) if w.dtype == np.float32 else got, w.view(np.uint32) if w.dtype == np.float32 else w), k


@pytest.mark.parametrize("n", [1, 3, 4, 1001, 250_003])
def test_scale_and_arithmetic_match_oracle(ktw, n):
    """min/max analyzer, scale_by_min_max, mean/var/min/max scalars and every arithmetic op, on
    sparse float32 / int64 features with missing cells and on a dense feature; ragged sizes."""
    tf, tft = ktw.tf, ktw.tft
    rng = np.random.default_rng(n)
    a = O.NumCol((rng.standard_normal(n) * 37 + 5).astype(np.float32), rng.random(n) > 0.1)
    b = O.NumCol(rng.integers(-1000, 5000, n).astype(np.int64), rng.random(n) > 0.2)
    c = O.NumCol((rng.random(n) * 3 + 0.5).astype(np.float32), None)
    d = O.NumCol(np.full(n, 7.25, np.float32), None)  # constant feature: min == max
    cols = {"a": a, "b": b, "c": c, "d": d}

    def fill(x):
        return tf.squeeze(tf.sparse.to_dense(tf.SparseTensor(x.indices, x.values, [x.dense_shape[0], 1])), axis=1)

    def fn(inputs):
        fa, fb, fc, fd = fill(inputs["a"]), fill(inputs["b"]), inputs["c"], inputs["d"]
        centered = fa - tft.mean(fa)
        unit = tft.scale_to_0_1(fa)
        return {
            "centered": centered,
            "unit": unit,
            "b_pm1": tft.scale_by_min_max(fb, -1.0, 1.0),
            "d_flat": tft.scale_to_0_1(fd),
            "ratio": centered / fc,
            "manual": (fa - tft.min(fa)) * 0.5,
            "zc": tft.scale_to_z_score(fa) + unit,
            "shift": 2.0 - tf.to_float(fb),
            "var_scaled": fc * tft.var(fa),
            "sum": fa + fc,
        }

    plan = ktw.trace(fn, {"a": "float32", "b": "int64", "c": "float32", "d": "float32"}, dense_keys={"c", "d"})
    eng = ktw.Engine(plan)
    table = helpers.table_from_oracle(cols, ktw).to("cuda")
    state = eng.analyze(table)
    out = eng.transform(table, state)
    torch.cuda.synchronize()
    fa, fb, fc, fd = (O.fill_in_missing(cols[k]) for k in "abcd")
    # analyzer state: min/max bit-exact, mean/var within the north_star tolerance
    assert bits(state.min_max_of("a")).tolist() == bits(O.min_max(fa)).tolist()
    assert bits(state.min_max_of("b")).tolist() == bits(O.min_max(fb)).tolist()
    assert bits(state.min_max_of("d")).tolist() == bits([7.25, 7.25]).tolist()
    mean, var = state.mean_var_of("a")
    wmean, wvar = O.finalize_moments(O.moments(fa))
    assert abs(mean - float(wmean)) <= 1e-6 * abs(float(wmean)) + 1e-30 and abs(var - float(wvar)) <= 1e-6 * abs(float(wvar)) + 1e-30
    # mappers: bit-exact given the device's own analyzer scalars
    mean, var = np.float32(mean), np.float32(var)
    mn, mx = O.min_max(fa)
    centered = O.arith_f32("sub", fa, mean)
    unit = O.scale_by_min_max(fa, mn, mx)
    want = {
        "centered": centered,
        "unit": unit,
        "b_pm1": O.scale_by_min_max(fb, *O.min_max(fb), -1.0, 1.0),
        "d_flat": np.full(n, 0.5, np.float32),
        "ratio": O.arith_f32("div", centered, fc),
        "manual": O.arith_f32("mul", O.arith_f32("sub", fa, mn), 0.5),
        "shift": O.arith_f32("sub", 2.0, fb.astype(np.float32)),
        "var_scaled": O.arith_f32("mul", fc, var),
        "sum": O.arith_f32("add", fa, fc),
    }
    assert list(out.keys()) == ["centered", "unit", "b_pm1", "d_flat", "ratio", "manual", "zc", "shift", "var_scaled", "sum"]
    for k, w in want.items():
        got = out[k].values.cpu().numpy()
        assert got.dtype == np.float32 and np.array_equal(bits(got), bits(w)), (k, n, int((bits(got) != bits(w)).sum()))
    z = out["zc"].values.cpu().numpy()
    wz = O.arith_f32("add", O.z_score(fa, mean, var), unit)
    assert np.allclose(z, wz, rtol=1e-6, atol=1e-6)


def kfp_equivalent(tf, transform):
    """Same structure as kubeflow-pipelines/taxi/preprocessing.py:65-101 (used when /root/reference is absent)."""
    dense_float = ["trip_miles", "fare", "trip_seconds"]
    vocab = ["pickup_census_tract", "dropoff_census_tract", "payment_type", "company", "pickup_community_area", "dropoff_community_area"]
    bucket = ["pickup_latitude", "pickup_longitude", "dropoff_latitude", "dropoff_longitude"]
    categorical = ["trip_start_hour", "trip_start_day", "trip_start_month"]

    def preprocess(inputs):
        outputs = {}
        for key in dense_float:
            outputs[key] = transform.scale_to_z_score(inputs[key])
        for key in vocab:
            t = inputs[key] if inputs[key].dtype == tf.string else tf.as_string(inputs[key])
            outputs[key] = transform.string_to_int(t, vocab_filename="vocab_" + key, top_k=1000, num_oov_buckets=10)
        for key in bucket:
            outputs[key] = transform.bucketize(inputs[key], 10)
        for key in categorical:
            outputs[key] = tf.to_int64(inputs[key])
        fare, tip = inputs["fare"], inputs["tips"]
        outputs["tips"] = tf.logical_and(tf.logical_not(tf.is_nan(fare)), tf.greater(tip, tf.multiply(fare, tf.constant(0.2))))
        return outputs
    return preprocess


def load_kfp_preprocess(ktw):
    ref = "/root/reference/kubeflow-pipelines/taxi/preprocessing.py"
    if os.path.exists(ref):  # the reference's own file, unmodified, through the stand-ins
        with ktw.install_shims():
            spec = importlib.util.spec_from_file_location("kfp_preprocessing", ref)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
        return mod.preprocess
    return kfp_equivalent(ktw.tf, ktw.tft)


def test_kfp_preprocess_variant(ktw, tmp_path):
    """f3: dense inputs (NUMBER -> float32, CATEGORY -> string; two category features kept as int64
    to drive the tf.as_string branch), string_to_int with vocab_filename, to_int64, boolean label."""
    from kubeflow_tfx_workshop_b200 import synth
    n = 60_011
    src = synth.taxi_table(n, seed=7, backend="numpy")
    oc = helpers.oracle_from_table(src)
    cols = {}
    number = ["fare", "trip_start_month", "trip_start_hour", "trip_start_day", "pickup_latitude", "pickup_longitude",
              "dropoff_latitude", "dropoff_longitude", "trip_miles", "trip_seconds", "tips"]
    for k in number:  # the KFP schema types every NUMBER as float32, missing -> 0.0
        cols[k] = O.NumCol(O.fill_in_missing(oc[k]).astype(np.float32), None)
    for k in ["payment_type", "company", "pickup_census_tract"]:
        cols[k] = O.StrCol.from_list(O.fill_in_missing(oc[k]))
    cols["dropoff_census_tract"] = O.StrCol.from_list([str(v).encode() if ok else b"" for v, ok in
                                                       zip(oc["dropoff_census_tract"].values.tolist(), oc["dropoff_census_tract"].valid)])
    rng = np.random.default_rng(5)
    neg = rng.integers(-3, 4, n) * 10 ** rng.integers(0, 18, n)
    cols["pickup_community_area"] = O.NumCol(O.fill_in_missing(oc["pickup_community_area"]).astype(np.int64), None)
    cols["dropoff_community_area"] = O.NumCol(np.where(rng.random(n) < 0.3, neg, oc["dropoff_community_area"].values).astype(np.int64),
                                              oc["dropoff_community_area"].valid)
    schema = {k: ("bytes" if isinstance(c, O.StrCol) else "float32" if c.values.dtype == np.float32 else "int64") for k, c in cols.items()}
    plan = ktw.trace(load_kfp_preprocess(ktw), schema, dense_keys=set(schema))
    assert sum(getattr(s, "as_string", False) for s in plan.specs) == 2
    eng = ktw.Engine(plan)
    table = helpers.table_from_oracle(cols, ktw).to("cuda")
    out, state = eng.analyze_and_transform(table)
    torch.cuda.synchronize()
    for k in ["trip_miles", "fare", "trip_seconds"]:
        x = O.fill_in_missing(cols[k])
        mean, var = O.finalize_moments(O.moments(x))
        assert helpers.zscore_close(out[k].values.cpu().numpy(), O.z_score(x, mean, var), mean, var), k
    for k in ["pickup_census_tract", "dropoff_census_tract", "payment_type", "company", "pickup_community_area", "dropoff_community_area"]:
        if isinstance(cols[k], O.StrCol):
            strings = O.fill_in_missing(cols[k])
        else:
            strings = [str(int(v)).encode() for v in O.fill_in_missing(cols[k]).tolist()]
        vocab = O.finalize_vocabulary(O.vocab_counts(strings), top_k=1000)
        want = O.apply_vocabulary(strings, vocab, num_oov_buckets=10)
        got = out[k].values.cpu().numpy()
        assert np.array_equal(got, want), (k, int((got != want).sum()))
    for k in ["pickup_latitude", "pickup_longitude", "dropoff_latitude", "dropoff_longitude"]:
        x = O.fill_in_missing(cols[k])
        assert np.array_equal(out[k].values.cpu().numpy(), O.bucketize(x, O.quantile_boundaries(x, 10))), k
    for k in ["trip_start_hour", "trip_start_day", "trip_start_month"]:
        got = out[k].values.cpu().numpy()
        assert got.dtype == np.int64 and np.array_equal(got, np.trunc(cols[k].values).astype(np.int64)), k
    want = O.tips_label(cols["fare"].values, cols["tips"].values, 0.2)
    assert np.array_equal(out["tips"].values.cpu().numpy(), want)
    # f1: the saved state re-applies to the same answers, and the named vocabulary files exist
    ktw.save_state(state, str(tmp_path))
    tfo = ktw.TFTransformOutput(str(tmp_path))
    again = tfo.transform_raw_features(table)
    torch.cuda.synchronize()
    for k in out.keys():
        assert torch.equal(again[k].values, out[k].values), k
    path = tfo.vocabulary_file_by_name("vocab_pickup_community_area")
    strings = [str(int(v)).encode() for v in cols["pickup_community_area"].values.tolist()]
    with open(path, "rb") as f:
        assert f.read().split(b"\n")[:-1] == O.finalize_vocabulary(O.vocab_counts(strings), top_k=1000)
    assert tfo.vocabulary_size_by_name("vocab_payment_type") == 7
    assert tfo.transformed_feature_spec()["tips"] == "int64"


def test_census_style_preprocessing(ktw, tmp_path):
    """The shape of notebooks/03a_TensorFlow_Transform_Advanced.ipynb:314-347: pass-through copies,
    scale_to_0_1 on dense and on a filled optional feature, tft.vocabulary(vocab_filename=) with no
    output column, and a static index table for the label."""
    tf, tft = ktw.tf, ktw.tft
    n = 20_001
    rng = np.random.default_rng(3)
    work = [b"Private", b"Self-emp", b"Federal-gov", b"?", b"Local-gov"]
    labels = [b">50K", b"<=50K", b"<=50K."]
    cols = {
        "age": O.NumCol(rng.integers(17, 91, n).astype(np.float32), None),
        "hours": O.NumCol(rng.integers(1, 100, n).astype(np.float32), None),
        "education-num": O.NumCol(rng.integers(1, 17, n).astype(np.float32), rng.random(n) > 0.25),
        "workclass": O.StrCol.from_list([work[i] for i in rng.choice(len(work), n, p=[.6, .2, .1, .06, .04])]),
        "label": O.StrCol.from_list([labels[i] for i in rng.integers(0, 3, n)]),
    }

    def fn(inputs):
        outputs = inputs.copy()
        for key in ("age", "hours"):
            outputs[key] = tft.scale_to_0_1(outputs[key])
        opt = outputs["education-num"]
        dense = tf.sparse_to_dense(opt.indices, [opt.dense_shape[0], 1], opt.values, default_value=0.)
        outputs["education-num"] = tft.scale_to_0_1(tf.squeeze(dense, axis=1))
        tft.vocabulary(inputs["workclass"], vocab_filename="workclass")
        table = tf.contrib.lookup.index_table_from_tensor([">50K", "<=50K"])
        outputs["label"] = table.lookup(outputs["label"])
        return outputs

    schema = {"age": "float32", "hours": "float32", "education-num": "float32", "workclass": "bytes", "label": "bytes"}
    plan = ktw.trace(fn, schema, dense_keys={"age", "hours", "workclass", "label"})
    eng = ktw.Engine(plan)
    table = helpers.table_from_oracle(cols, ktw).to("cuda")
    out, state = eng.analyze_and_transform(table)
    torch.cuda.synchronize()
    for k in ("age", "hours", "education-num"):
        x = O.fill_in_missing(cols[k])
        assert np.array_equal(bits(out[k].values.cpu().numpy()), bits(O.scale_by_min_max(x, *O.min_max(x)))), k
    assert out["workclass"].to_list() == cols["workclass"].to_list()
    want = O.apply_vocabulary(cols["label"].to_list(), [b">50K", b"<=50K"], num_oov_buckets=0, default_value=-1)
    assert np.array_equal(out["label"].values.cpu().numpy(), want) and (want == -1).any()
    ktw.save_state(state, str(tmp_path))
    tfo = ktw.TFTransformOutput(str(tmp_path))
    with open(tfo.vocabulary_file_by_name("workclass"), "rb") as f:
        assert f.read().split(b"\n")[:-1] == O.finalize_vocabulary(O.vocab_counts(cols["workclass"].to_list()))
    again = tfo.transform_raw_features(table)
    for k in ("age", "hours", "education-num", "label"):
        assert torch.equal(again[k].values, out[k].values), k


def test_sibling_abi_rejects_bad_arguments(ktw):
    nat = ktw._native
    x = torch.zeros(8, dtype=torch.float32, device="cuda")
    a = nat.TxfOperand(col=x.data_ptr(), dtype=nat.TXF_F32)
    both = nat.TxfOperand(col=x.data_ptr(), scalar=x.data_ptr(), dtype=nat.TXF_F32)
    with pytest.raises(nat.NativeError, match="exactly one"):
        nat.call("txf_map_binary", nat.ARITH_ADD, a, both, x.data_ptr(), 8, None)
    with pytest.raises(nat.NativeError, match="unknown opcode"):
        nat.call("txf_map_binary", 9, a, a, x.data_ptr(), 8, None)
    with pytest.raises(nat.NativeError, match="output_min"):
        nat.call("txf_map_scale_minmax", a, x.data_ptr(), 1.0, 1.0, x.data_ptr(), 8, None)
    with pytest.raises(nat.NativeError, match="16B aligned"):
        nat.call("txf_map_cast", a, nat.TXF_I64, x.data_ptr() + 4, 4, None)
    assert nat.query("txf_minmax_workspace_bytes", 0) == 0
