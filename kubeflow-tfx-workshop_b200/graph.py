"""Symbolic expression DAG recorded while ``preprocessing_fn`` runs ONCE over column handles.

tf.Transform calls the user's ``preprocessing_fn(inputs)`` once with symbolic tensors
(``airflow-dags/taxi_utils.py:106-145``) and splits the traced graph into analyzers and a per-row
transform.  The shims in ``tf_shim`` / ``tft_shim`` build these nodes instead of TF ops;
``plan.compile_outputs`` then maps each output expression onto the CUDA analyzers and onto one op
of the fused transform program.
"""
from __future__ import annotations

from typing import Optional, Tuple


class DType:
    def __init__(self, name: str):
        self.name = name

    def __eq__(self, other):
        return isinstance(other, DType) and other.name == self.name

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return hash(self.name)

    def __repr__(self):
        return f"tf.{self.name}"


string = DType("string")
int64 = DType("int64")
int32 = DType("int32")
float32 = DType("float32")
float64 = DType("float64")
bool_ = DType("bool")

_FROM_COLUMN = {"bytes": string, "int64": int64, "float32": float32}


class Node:
    dtype: DType
    rank: int = 1

    def children(self) -> Tuple["Node", ...]:
        return ()

    # tensor arithmetic (`x - tft.mean(x)`, `a * b`: notebooks/03_TensorFlow_Transform.ipynb:147,150)
    def __add__(self, o):
        return Arith("add", self, as_node(o))

    def __radd__(self, o):
        return Arith("add", as_node(o), self)

    def __sub__(self, o):
        return Arith("sub", self, as_node(o))

    def __rsub__(self, o):
        return Arith("sub", as_node(o), self)

    def __mul__(self, o):
        return Arith("mul", self, as_node(o))

    def __rmul__(self, o):
        return Arith("mul", as_node(o), self)

    def __truediv__(self, o):
        return Arith("div", self, as_node(o))

    def __rtruediv__(self, o):
        return Arith("div", as_node(o), self)

    __div__, __rdiv__ = __truediv__, __rtruediv__
    __hash__ = object.__hash__

    def describe(self) -> str:
        kids = ", ".join(c.describe() for c in self.children())
        return f"{type(self).__name__}({kids})"


class SparseInput(Node):
    """inputs[key]: a rank-2 SparseTensor with at most one value per row."""

    rank = 2

    def __init__(self, key: str, column_dtype: str):
        self.key = key
        self.dtype = _FROM_COLUMN[column_dtype]
        self.indices = _Part(self, "indices")
        self.values = _Part(self, "values")
        self.dense_shape = _Shape(self)

    def describe(self):
        return f"inputs[{self.key!r}]"


class _Part:
    def __init__(self, owner: SparseInput, what: str):
        self.owner, self.what = owner, what


class _Shape:
    def __init__(self, owner: SparseInput):
        self.owner = owner

    def __getitem__(self, i):
        return _ShapeDim(self.owner, i)


class _ShapeDim:
    def __init__(self, owner: SparseInput, axis: int):
        self.owner, self.axis = owner, axis


class SparseView(Node):
    """tf.SparseTensor(x.indices, x.values, [x.dense_shape[0], 1]) (taxi_utils.py:100)."""

    rank = 2

    def __init__(self, owner: SparseInput):
        self.owner = owner
        self.dtype = owner.dtype


class Dense(Node):
    """_fill_in_missing: sparse [batch, <=1] -> dense, default '' / 0 (taxi_utils.py:86-103)."""

    def __init__(self, owner: SparseInput, default, rank: int):
        want = "" if owner.dtype == string else 0
        if isinstance(default, bytes):
            default = default.decode()
        if default != want:
            raise NotImplementedError(
                f"fill value {default!r} for {owner.describe()}: only '' (strings) / 0 (numbers) are supported")
        self.owner = owner
        self.dtype = owner.dtype
        self.rank = rank

    def children(self):
        return (self.owner,)


class Const(Node):
    rank = 0

    def __init__(self, value, dtype: Optional[DType] = None):
        self.value = value
        self.dtype = dtype or (float32 if isinstance(value, float) else int32 if isinstance(value, int) else string)

    def describe(self):
        return f"Const({self.value!r})"


class Unary(Node):
    def __init__(self, x: Node, dtype: Optional[DType] = None):
        self.x = x
        self.dtype = dtype or x.dtype
        self.rank = x.rank

    def children(self):
        return (self.x,)


class Binary(Node):
    def __init__(self, a: Node, b: Node, dtype: Optional[DType] = None):
        self.a, self.b = a, b
        self.dtype = dtype or a.dtype
        self.rank = max(a.rank, b.rank)

    def children(self):
        return (self.a, self.b)


class IsNan(Unary):
    def __init__(self, x):
        super().__init__(x, bool_)


class ZerosLike(Unary):
    pass


class Cast(Unary):
    pass


class Greater(Binary):
    def __init__(self, a, b):
        super().__init__(a, b, bool_)


class Multiply(Binary):
    pass


class Where(Node):
    def __init__(self, cond: Node, a: Node, b: Node):
        self.cond, self.a, self.b = cond, a, b
        self.dtype = a.dtype
        self.rank = a.rank

    def children(self):
        return (self.cond, self.a, self.b)


# ---- tft mappers whose state comes from a full-pass analyzer -----------------------------------
class ZScore(Unary):
    """tft.scale_to_z_score(x) (taxi_utils.py:118)."""

    def __init__(self, x):
        super().__init__(x, float32)


class Vocab(Unary):
    """tft.compute_and_apply_vocabulary(x, top_k, num_oov_buckets, ...) (taxi_utils.py:123-126)."""

    def __init__(self, x, top_k, num_oov_buckets, default_value, frequency_threshold, vocab_filename):
        super().__init__(x, int64)
        self.top_k = top_k
        self.num_oov_buckets = num_oov_buckets
        self.default_value = default_value
        self.frequency_threshold = frequency_threshold
        self.vocab_filename = vocab_filename


class Bucketize(Unary):
    """tft.bucketize(x, num_buckets, epsilon=None) (taxi_utils.py:129-130)."""

    def __init__(self, x, num_buckets, epsilon):
        super().__init__(x, int64)
        self.num_buckets = num_buckets
        self.epsilon = epsilon


# ---- sibling analyzers / mappers of the workshop notebooks ---------------------
class Arith(Binary):
    """One float32 operation between two tensors (either may be a scalar)."""

    def __init__(self, op: str, a: Node, b: Node):
        super().__init__(a, b, float32)
        self.op = op

    def describe(self):
        return f"{self.op}({self.a.describe()}, {self.b.describe()})"


class Reduce(Unary):
    """tft.mean / tft.min / tft.max / tft.var: a full-pass analyzer whose result is a scalar."""

    rank = 0

    def __init__(self, kind: str, x: Node):
        super().__init__(x, float32)
        self.kind = kind
        self.rank = 0

    def describe(self):
        return f"tft.{self.kind}({self.x.describe()})"


class ScaleMinMax(Unary):
    """tft.scale_by_min_max(x, output_min, output_max) / tft.scale_to_0_1(x)
    (notebooks/03_TensorFlow_Transform.ipynb:148)."""

    def __init__(self, x, output_min: float, output_max: float):
        super().__init__(x, float32)
        self.output_min, self.output_max = float(output_min), float(output_max)


class VocabFile(Unary):
    """tft.vocabulary(x, ..., vocab_filename=): analyzer only; the value is a deferred file name
    (notebooks/03a_TensorFlow_Transform_Advanced.ipynb:341)."""

    rank = 0

    def __init__(self, x, top_k, frequency_threshold, vocab_filename):
        super().__init__(x, string)
        self.top_k, self.frequency_threshold, self.vocab_filename = top_k, frequency_threshold, vocab_filename
        self.rank = 0


class StaticVocab(Unary):
    """tf.contrib.lookup.index_table_from_tensor(mapping).lookup(x) (03a_...:344-345)."""

    def __init__(self, x, terms, num_oov_buckets, default_value):
        super().__init__(x, int64)
        self.terms, self.num_oov_buckets, self.default_value = list(terms), num_oov_buckets, default_value


# ---- KFP variant (kubeflow-pipelines/taxi/preprocessing.py:65-101) ---------------
class AsString(Unary):
    """tf.as_string(x) on an integer feature (:82): shortest decimal text."""

    def __init__(self, x):
        super().__init__(x, string)


class LogicalNot(Unary):
    def __init__(self, x):
        super().__init__(x, bool_)


class LogicalAnd(Binary):
    def __init__(self, a, b):
        super().__init__(a, b, bool_)


# Analyzers that are requested without being part of an output expression (tft.vocabulary) register
# themselves with the trace in progress.
_TRACES = []


def begin_trace():
    _TRACES.append([])


def end_trace():
    return _TRACES.pop()


def register_side_analyzer(node):
    if _TRACES:
        _TRACES[-1].append(node)


def as_node(x) -> Node:
    if isinstance(x, Node):
        return x
    if isinstance(x, (int, float, str, bytes)):
        return Const(x)
    raise TypeError(f"cannot use {type(x).__name__} as a symbolic tensor")
