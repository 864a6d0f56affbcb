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


def as_node(x) -> Node:
    if isinstance(x, Node):
        return x
    if isinstance(x, (int, float, str, bytes)):
        return Const(x)
    raise TypeError(f"cannot use {type(x).__name__} as a symbolic tensor")
