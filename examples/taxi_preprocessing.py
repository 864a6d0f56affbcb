"""A preprocessing_fn for the Chicago-taxi features, written against the tf / tft API exactly as a
TFX user would (same feature groups, same tft calls and arguments as the reference pipeline's
``airflow-dags/taxi_utils.py:86-145``).  The GPU box has no /root/reference, so the GPU tests and
the benchmark trace this module; ``tests/test_host.py::test_reference_taxi_utils_runs_unmodified_through_the_shims`` checks, where the reference file is
present, that both trace to the same plan.

Import it under ``kubeflow_tfx_workshop_b200.install_shims()`` (or with real TF/tft installed).
"""
import tensorflow as tf
import tensorflow_transform as tft

Z_SCORED = ("trip_miles", "fare", "trip_seconds")
VOCABULARY = ("payment_type", "company")
BUCKETIZED = ("pickup_latitude", "pickup_longitude", "dropoff_latitude", "dropoff_longitude")
PASSED_THROUGH = ("trip_start_hour", "trip_start_day", "trip_start_month", "pickup_census_tract",
                  "dropoff_census_tract", "pickup_community_area", "dropoff_community_area")
TOP_K, OOV_BUCKETS, BUCKETS = 1000, 10, 10
BIG_TIP_RATIO = 0.2


def out_name(feature):
    return feature + "_xf"


def densify(sparse):
    """[batch, <=1] sparse feature -> [batch] dense, '' or 0 where the value is absent."""
    blank = "" if sparse.dtype == tf.string else 0
    one_wide = tf.SparseTensor(sparse.indices, sparse.values, [sparse.dense_shape[0], 1])
    return tf.squeeze(tf.sparse.to_dense(one_wide, blank), axis=1)


def preprocessing_fn(inputs):
    xf = {}
    for name in Z_SCORED:
        xf[out_name(name)] = tft.scale_to_z_score(densify(inputs[name]))
    for name in VOCABULARY:
        xf[out_name(name)] = tft.compute_and_apply_vocabulary(densify(inputs[name]), top_k=TOP_K, num_oov_buckets=OOV_BUCKETS)
    for name in BUCKETIZED:
        xf[out_name(name)] = tft.bucketize(densify(inputs[name]), BUCKETS)
    for name in PASSED_THROUGH:
        xf[out_name(name)] = densify(inputs[name])
    fare, tips = densify(inputs["fare"]), densify(inputs["tips"])
    big_tip = tf.cast(tf.greater(tips, tf.multiply(fare, tf.constant(BIG_TIP_RATIO))), tf.int64)
    xf[out_name("tips")] = tf.where(tf.is_nan(fare), tf.cast(tf.zeros_like(fare), tf.int64), big_tip)
    return xf
