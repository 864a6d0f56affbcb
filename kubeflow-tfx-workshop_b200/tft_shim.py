"""Stand-in for ``tensorflow_transform`` (tft 0.13 API surface used by the workshop).

Reference call sites: ``airflow-dags/taxi_utils.py:118`` (``scale_to_z_score``), ``:123-126``
(``compute_and_apply_vocabulary(top_k=, num_oov_buckets=)``), ``:129-130`` (``bucketize``).
Keyword names and defaults follow tft 0.13; argument errors are raised
here, at trace time, like tft does.
"""
from __future__ import annotations

from . import graph as g


def _dense_numeric(x, who):
    x = g.as_node(x)
    if x.dtype == g.string:
        raise TypeError(f"{who}: expected a numeric tensor, got {x.dtype}")
    return x


def scale_to_z_score(x, elementwise=False, name=None, output_dtype=None):
    """(x - mean) / sqrt(var) with a full-pass mean/variance analyzer; float32 out."""
    if elementwise:
        raise NotImplementedError("scale_to_z_score(elementwise=True) is not supported")
    return g.ZScore(_dense_numeric(x, "scale_to_z_score"))


def compute_and_apply_vocabulary(x, default_value=-1, top_k=None, frequency_threshold=None, num_oov_buckets=0,
                                 vocab_filename=None, weights=None, labels=None, name=None):
    """Vocabulary ids by descending frequency; OOV terms hash into num_oov_buckets extra ids."""
    x = g.as_node(x)
    if x.dtype != g.string:
        raise NotImplementedError("compute_and_apply_vocabulary: only string inputs are supported (see INTEGRATION.md)")
    if weights is not None or labels is not None:
        raise NotImplementedError("compute_and_apply_vocabulary: weights/labels are not supported")
    if top_k is not None and (not isinstance(top_k, int) or isinstance(top_k, bool) or top_k <= 0):
        raise ValueError(f"top_k must be a positive int or None, got {top_k!r}")
    if frequency_threshold is not None and (not isinstance(frequency_threshold, int) or frequency_threshold < 0):
        raise ValueError(f"frequency_threshold must be a non-negative int or None, got {frequency_threshold!r}")
    if not isinstance(num_oov_buckets, int) or isinstance(num_oov_buckets, bool) or num_oov_buckets < 0:
        raise ValueError(f"num_oov_buckets must be a non-negative int, got {num_oov_buckets!r}")
    return g.Vocab(x, top_k, num_oov_buckets, default_value, frequency_threshold, vocab_filename)


def vocabulary(x, top_k=None, frequency_threshold=None, vocab_filename=None, store_frequency=False, weights=None,
               labels=None, name=None):
    """The analyzer half of compute_and_apply_vocabulary; returns the (deferred) vocabulary file."""
    x = g.as_node(x)
    if x.dtype != g.string:
        raise NotImplementedError("vocabulary: only string inputs are supported")
    if store_frequency or weights is not None or labels is not None:
        raise NotImplementedError("vocabulary: store_frequency/weights/labels are not supported")
    if top_k is not None and (not isinstance(top_k, int) or isinstance(top_k, bool) or top_k <= 0):
        raise ValueError(f"top_k must be a positive int or None, got {top_k!r}")
    node = g.VocabFile(x, top_k, frequency_threshold, vocab_filename)
    g.register_side_analyzer(node)
    return node


uniques = vocabulary  # the pre-0.13 name


def _reduce(kind):
    def fn(x, reduce_instance_dims=True, name=None, output_dtype=None):
        if not reduce_instance_dims:
            raise NotImplementedError(f"tft.{kind}(reduce_instance_dims=False) is not supported")
        return g.Reduce(kind, _dense_numeric(x, f"tft.{kind}"))

    fn.__name__ = kind
    fn.__doc__ = f"Full-pass {kind} of all (filled) values of x as a float32 scalar."
    return fn


mean, min, max, var = _reduce("mean"), _reduce("min"), _reduce("max"), _reduce("var")  # noqa: A001


def scale_by_min_max(x, output_min=0.0, output_max=1.0, elementwise=False, name=None):
    """(x - min) / (max - min) mapped onto [output_min, output_max]; 0.5 of the range when min == max."""
    if elementwise:
        raise NotImplementedError("scale_by_min_max(elementwise=True) is not supported")
    if not output_min < output_max:
        raise ValueError("output_min must be less than output_max")
    return g.ScaleMinMax(_dense_numeric(x, "scale_by_min_max"), output_min, output_max)


def scale_to_0_1(x, elementwise=False, name=None):
    """notebooks/03_TensorFlow_Transform.ipynb:148, 03a_TensorFlow_Transform_Advanced.ipynb:323,333."""
    return scale_by_min_max(x, 0.0, 1.0, elementwise, name)


# tft 0.13 keeps the old name as an alias (kubeflow-pipelines/taxi/preprocessing.py:75)
string_to_int = compute_and_apply_vocabulary


def bucketize(x, num_buckets, epsilon=None, weights=None, name=None):
    """Bucket ids from num_buckets-1 quantile boundaries; int64 out."""
    if not isinstance(num_buckets, int) or isinstance(num_buckets, bool):
        raise TypeError(f"num_buckets must be an int, got {type(num_buckets).__name__}")
    if num_buckets < 1:
        raise ValueError(f"Invalid num_buckets {num_buckets}")
    if weights is not None:
        raise NotImplementedError("bucketize: weights are not supported")
    return g.Bucketize(_dense_numeric(x, "bucketize"), num_buckets, epsilon)
