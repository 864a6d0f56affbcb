"""preprocessing_fns of the two stress workloads of BASELINE.json (configs[3]: vocabulary, configs[4]:
bucketize), written against the tf / tft API like ``taxi_preprocessing.py``: the same tft calls the
reference makes at ``airflow-dags/taxi_utils.py:121-130``, on other columns and sizes.

Import under ``kubeflow_tfx_workshop_b200.install_shims()`` (or with real TF/tft installed).
"""
import tensorflow as tf
import tensorflow_transform as tft

TOP_K, OOV_BUCKETS = 1000, 10


def _dense_string(sparse):
    one_wide = tf.SparseTensor(sparse.indices, sparse.values, [sparse.dense_shape[0], 1])
    return tf.squeeze(tf.sparse.to_dense(one_wide, ""), axis=1)


def vocab_stress_fn(inputs):
    """Two string columns with millions of distinct terms -> vocabulary ids (top_k=1000, 10 OOV buckets)."""
    return {name + "_id": tft.compute_and_apply_vocabulary(_dense_string(inputs[name]), top_k=TOP_K, num_oov_buckets=OOV_BUCKETS)
            for name in ("code", "name")}


def bucket_stress_fn(num_buckets):
    """Eight dense float columns -> quantile bucket ids."""
    def fn(inputs):
        return {name + "_b": tft.bucketize(inputs[name], num_buckets) for name in sorted(inputs)}
    return fn
