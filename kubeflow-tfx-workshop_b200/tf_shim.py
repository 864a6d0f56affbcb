"""Stand-in for the handful of ``tensorflow`` symbols that ``preprocessing_fn`` touches.

Reference call sites: ``airflow-dags/taxi_utils.py:98-103`` (``tf.string``, ``tf.squeeze``,
``tf.sparse.to_dense``, ``tf.SparseTensor``), ``:138-143`` (``tf.where``, ``tf.is_nan``,
``tf.cast``, ``tf.zeros_like``, ``tf.int64``, ``tf.greater``, ``tf.multiply``, ``tf.constant``) and
``notebooks/06_Airflow_Feature_Analysis.ipynb:176`` (``tf.sparse_to_dense``).  Every function
returns a ``graph`` node; nothing is computed here.
"""
from __future__ import annotations

import types

from . import graph as g

string, int64, int32, float32, float64 = g.string, g.int64, g.int32, g.float32, g.float64
bool = g.bool_  # noqa: A001  (mirrors tf.bool)


def _owner_of_shape(dense_shape):
    """[x.dense_shape[0], 1] -> x (the only shape _fill_in_missing builds, taxi_utils.py:100)."""
    if (isinstance(dense_shape, (list, tuple)) and len(dense_shape) == 2 and isinstance(dense_shape[0], g._ShapeDim)
            and dense_shape[0].axis == 0 and dense_shape[1] == 1):
        return dense_shape[0].owner
    raise NotImplementedError("only the dense shape [x.dense_shape[0], 1] is supported")


def SparseTensor(indices, values, dense_shape):  # noqa: N802
    if not (isinstance(indices, g._Part) and isinstance(values, g._Part) and indices.owner is values.owner
            and indices.what == "indices" and values.what == "values"):
        raise NotImplementedError("tf.SparseTensor: indices and values must come from the same input feature")
    if _owner_of_shape(dense_shape) is not indices.owner:
        raise NotImplementedError("tf.SparseTensor: dense_shape must come from the same input feature")
    return g.SparseView(indices.owner)


def _to_dense(sp_input, default_value=None, validate_indices=True, name=None):
    if isinstance(sp_input, g.SparseInput):
        sp_input = g.SparseView(sp_input)
    if not isinstance(sp_input, g.SparseView):
        raise TypeError("tf.sparse.to_dense expects a SparseTensor built from an input feature")
    if default_value is None:
        default_value = "" if sp_input.dtype == g.string else 0
    return g.Dense(sp_input.owner, default_value, rank=2)


sparse = types.SimpleNamespace(to_dense=_to_dense)
sparse_tensor_to_dense = _to_dense


def sparse_to_dense(sparse_indices, output_shape, sparse_values, default_value=0, validate_indices=True, name=None):
    return _to_dense(SparseTensor(sparse_indices, sparse_values, output_shape), default_value)


def squeeze(input, axis=None, name=None):  # noqa: A002
    if isinstance(input, g.Dense) and input.rank == 2 and axis in (1, [1], (1,), -1, None):
        return g.Dense(input.owner, "" if input.dtype == g.string else 0, rank=1)
    raise NotImplementedError("tf.squeeze is only supported on the [batch, 1] result of sparse.to_dense")


def constant(value, dtype=None, shape=None, name="Const"):
    if isinstance(value, float) or dtype == g.float32:
        return g.Const(float(value), g.float32)
    return g.Const(value, dtype)


def is_nan(x, name=None):
    return g.IsNan(g.as_node(x))


def zeros_like(tensor, dtype=None, name=None, optimize=True):
    return g.ZerosLike(g.as_node(tensor), dtype)


def cast(x, dtype, name=None):
    return g.Cast(g.as_node(x), dtype)


def greater(x, y, name=None):
    return g.Greater(g.as_node(x), g.as_node(y))


def multiply(x, y, name=None):
    return g.Multiply(g.as_node(x), g.as_node(y))


def where(condition, x=None, y=None, name=None):
    if x is None or y is None:
        raise NotImplementedError("tf.where needs all three arguments here")
    return g.Where(g.as_node(condition), g.as_node(x), g.as_node(y))


def add(x, y, name=None):
    return g.as_node(x) + y


def subtract(x, y, name=None):
    return g.as_node(x) - y


def divide(x, y, name=None):
    return g.as_node(x) / y


def to_float(x, name="ToFloat"):
    return g.Cast(g.as_node(x), g.float32)


def to_int64(x, name="ToInt64"):
    return g.Cast(g.as_node(x), g.int64)


def as_string(input, precision=-1, scientific=False, shortest=False, width=-1, fill="", name=None):  # noqa: A002
    x = g.as_node(input)
    if x.dtype not in (g.int64, g.int32) or (precision, scientific, shortest, width, fill) != (-1, False, False, -1, ""):
        raise NotImplementedError("tf.as_string: only integer tensors with default formatting are supported")
    return g.AsString(x)


def logical_not(x, name=None):
    return g.LogicalNot(g.as_node(x))


def logical_and(x, y, name=None):
    return g.LogicalAnd(g.as_node(x), g.as_node(y))


math = types.SimpleNamespace(is_nan=is_nan, greater=greater, multiply=multiply, add=add, subtract=subtract, divide=divide,
                             logical_and=logical_and, logical_not=logical_not)
strings = types.SimpleNamespace(as_string=as_string)


class _StaticTable:
    """tf.contrib.lookup.index_table_from_tensor(mapping) (03a_TensorFlow_Transform_Advanced.ipynb:344)."""

    def __init__(self, mapping, num_oov_buckets, default_value):
        self.mapping, self.num_oov_buckets, self.default_value = list(mapping), num_oov_buckets, default_value

    def lookup(self, keys, name=None):
        return g.StaticVocab(g.as_node(keys), self.mapping, self.num_oov_buckets, self.default_value)


def _index_table_from_tensor(mapping, num_oov_buckets=0, default_value=-1, hasher_spec=None, dtype=string, name=None):
    return _StaticTable(mapping, num_oov_buckets, default_value)


contrib = types.SimpleNamespace(lookup=types.SimpleNamespace(index_table_from_tensor=_index_table_from_tensor))
lookup = types.SimpleNamespace(index_table_from_tensor=_index_table_from_tensor)
logging = types.SimpleNamespace(ERROR=40, WARN=30, INFO=20, set_verbosity=lambda level: None)


class FixedLenFeature:
    def __init__(self, shape, dtype, default_value=None):
        self.shape, self.dtype, self.default_value = shape, dtype, default_value


class VarLenFeature:
    def __init__(self, dtype):
        self.dtype = dtype


io = types.SimpleNamespace(FixedLenFeature=FixedLenFeature, VarLenFeature=VarLenFeature)
