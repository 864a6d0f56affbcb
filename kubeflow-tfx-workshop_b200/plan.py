"""Trace ``preprocessing_fn`` once and compile its outputs into analyzers + a transform program.

The reference's ``preprocessing_fn`` (``airflow-dags/taxi_utils.py:106-145``) is called once with
symbolic inputs; each output expression must match one of the shapes below (they are exactly the
shapes that function builds -- anything else raises NotImplementedError naming the expression):

    tft.scale_to_z_score(fill(x))                          -> ZScoreSpec      (:116-119)
    tft.compute_and_apply_vocabulary(fill(x), ...)         -> VocabSpec       (:121-126)
    tft.bucketize(fill(x), n)                              -> BucketSpec      (:128-130)
    fill(x)                                                -> FillSpec        (:132-133)
    where(is_nan(fill(a)), cast(zeros_like(fill(a)), int64),
          cast(greater(fill(b), multiply(fill(a), const)), int64)) -> RatioLabelSpec (:135-143)

Sibling shapes used by the workshop notebooks and the KFP variant:

    tft.scale_to_0_1(x) / tft.scale_by_min_max(x, lo, hi)  -> ScaleSpec       (notebook 03:148, 03a:323,333)
    float32 arithmetic over features, other float32 results,
      tft.mean/min/max/var(x) scalars and constants        -> ArithSpec       (notebook 03:147,150)
    tft.vocabulary(x, vocab_filename=)                     -> VocabOnlySpec   (notebook 03a:341)
    index_table_from_tensor(terms).lookup(x)               -> StaticVocabSpec (notebook 03a:344-345)
    tf.to_int64(x) / tf.to_float(x)                        -> CastSpec        (kubeflow-pipelines/taxi/preprocessing.py:92)
    string_to_int(tf.as_string(x), ...)                    -> VocabSpec(as_string=True)   (:79-85)
    logical_and(logical_not(is_nan(a)), greater(b, multiply(a, const))) -> RatioLabelSpec (:94-99)

Features declared tf.FixedLenFeature are dense: ``inputs[key]`` is already the filled column.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional

from . import graph as g


@dataclass
class ZScoreSpec:
    out_key: str
    in_key: str


@dataclass
class VocabSpec:
    out_key: str
    in_key: str
    top_k: Optional[int]
    num_oov_buckets: int
    default_value: int
    frequency_threshold: Optional[int]
    vocab_filename: Optional[str]
    as_string: bool = False  # the input is an int64 feature formatted with tf.as_string first


@dataclass
class VocabOnlySpec:
    in_key: str
    top_k: Optional[int]
    frequency_threshold: Optional[int]
    vocab_filename: Optional[str]
    out_key: Optional[str] = None  # analyzer only: produces no column


@dataclass
class StaticVocabSpec:
    out_key: str
    in_key: str
    terms: List[str]  # latin-1 text of the byte strings (JSON-safe)
    num_oov_buckets: int
    default_value: int


@dataclass
class ScaleSpec:
    out_key: str
    in_key: str
    output_min: float
    output_max: float


@dataclass
class ArithSpec:
    """out = a (op) b in float32.  Operands: ["col", feature] | ["out", out_key of an earlier spec] |
    ["scalar", "mean"|"var"|"min"|"max", feature] | ["const", value]."""
    out_key: str
    op: str
    a: list
    b: list


@dataclass
class CastSpec:
    out_key: str
    in_key: str
    to: str  # 'int64' | 'float32'


TEMP_PREFIX = "__tmp"  # ArithSpec results that are only operands of other specs


@dataclass
class BucketSpec:
    out_key: str
    in_key: str
    num_buckets: int
    epsilon: Optional[float]


@dataclass
class FillSpec:
    out_key: str
    in_key: str


@dataclass
class RatioLabelSpec:
    out_key: str
    fare_key: str
    tips_key: str
    ratio: float


def spec_inputs(s) -> List[str]:
    """Input features one spec reads."""
    if isinstance(s, RatioLabelSpec):
        return [s.fare_key, s.tips_key]
    if isinstance(s, ArithSpec):
        return [o[-1] for o in (s.a, s.b) if o[0] in ("col", "scalar")]
    return [s.in_key]


@dataclass
class Plan:
    schema: Dict[str, str]                      # input feature -> 'float32' | 'int64' | 'bytes'
    specs: List[object] = field(default_factory=list)

    def of(self, cls):
        return [s for s in self.specs if isinstance(s, cls)]

    def input_keys(self) -> List[str]:
        keys = []
        for s in self.specs:
            for k in spec_inputs(s):
                if k not in keys:
                    keys.append(k)
        return keys

    def scalar_keys(self, kinds) -> List[str]:
        """Features whose analyzer scalars of the given kinds are used by ArithSpec operands."""
        keys = []
        for s in self.of(ArithSpec):
            for o in (s.a, s.b):
                if o[0] == "scalar" and o[1] in kinds and o[2] not in keys:
                    keys.append(o[2])
        return keys

    def output_schema(self) -> Dict[str, str]:
        out = {}
        for s in self.specs:
            if s.out_key is None or s.out_key.startswith(TEMP_PREFIX):
                continue
            if isinstance(s, (ZScoreSpec, ScaleSpec, ArithSpec)):
                out[s.out_key] = "float32"
            elif isinstance(s, FillSpec):
                out[s.out_key] = self.schema[s.in_key]
            elif isinstance(s, CastSpec):
                out[s.out_key] = s.to
            else:
                out[s.out_key] = "int64"
        return out


def _dense_key(node) -> Optional[str]:
    if isinstance(node, g.Dense) and node.rank == 1:
        return node.owner.key
    return None


def _need_dense(node, what) -> str:
    key = _dense_key(node)
    if key is None:
        raise NotImplementedError(f"{what}: the argument must be a filled input feature, got {node.describe()}")
    return key


def _match_ratio_label(node) -> Optional[tuple]:
    """where(is_nan(F), cast(zeros_like(F), int64), cast(greater(T, multiply(F, c)), int64))"""
    if not isinstance(node, g.Where):
        return None
    cond, a, b = node.cond, node.a, node.b
    if not (isinstance(cond, g.IsNan) and _dense_key(cond.x)):
        return None
    fare = _dense_key(cond.x)
    if not (isinstance(a, g.Cast) and a.dtype == g.int64 and isinstance(a.x, g.ZerosLike) and _dense_key(a.x.x) == fare):
        return None
    if not (isinstance(b, g.Cast) and b.dtype == g.int64 and isinstance(b.x, g.Greater)):
        return None
    gt = b.x
    tips = _dense_key(gt.a)
    mul = gt.b
    if tips is None or not isinstance(mul, g.Multiply):
        return None
    if _dense_key(mul.a) == fare and isinstance(mul.b, g.Const):
        ratio = mul.b.value
    elif _dense_key(mul.b) == fare and isinstance(mul.a, g.Const):
        ratio = mul.a.value
    else:
        return None
    return fare, tips, float(ratio)


def _match_bool_label(node) -> Optional[tuple]:
    """logical_and(logical_not(is_nan(F)), greater(T, multiply(F, c)))
    (kubeflow-pipelines/taxi/preprocessing.py:94-99): the same 0/1 as the ratio label."""
    if not isinstance(node, g.LogicalAnd):
        return None
    for notnan, gt in ((node.a, node.b), (node.b, node.a)):
        if not (isinstance(notnan, g.LogicalNot) and isinstance(notnan.x, g.IsNan) and isinstance(gt, g.Greater)):
            continue
        fare = _dense_key(notnan.x.x)
        tips = _dense_key(gt.a)
        mul = gt.b
        if fare is None or tips is None or not isinstance(mul, g.Multiply):
            continue
        if _dense_key(mul.a) == fare and isinstance(mul.b, g.Const):
            return fare, tips, float(mul.b.value)
        if _dense_key(mul.b) == fare and isinstance(mul.a, g.Const):
            return fare, tips, float(mul.a.value)
    return None


_F32_MAPPERS = (g.ZScore, g.ScaleMinMax, g.Arith)


class _Compiler:
    def __init__(self, outputs: Dict[str, g.Node], schema: Dict[str, str]):
        self.plan = Plan(schema=dict(schema))
        self.schema = schema
        self.named = {}
        for k, node in outputs.items():
            self.named.setdefault(id(node), k)
        self.done: Dict[int, str] = {}   # id(node) -> out_key of the spec that computes it
        self.ntemp = 0

    def _numeric_col(self, node, what) -> str:
        key = _need_dense(node, what)
        if self.schema[key] == "bytes":
            raise NotImplementedError(f"{what}: {key!r} is a string feature")
        return key

    def operand(self, node) -> list:
        if isinstance(node, g.Const):
            return ["const", float(node.value)]
        if isinstance(node, g.Reduce):
            return ["scalar", node.kind, self._numeric_col(node.x, f"tft.{node.kind}")]
        if isinstance(node, g.Cast) and node.dtype == g.float32 and _dense_key(node.x):
            node = node.x  # tf.to_float(feature): operands are converted to float32 anyway
        if _dense_key(node) is not None:
            return ["col", self._numeric_col(node, "arithmetic")]
        if isinstance(node, _F32_MAPPERS):
            return ["out", self.emit(node, None)]
        raise NotImplementedError(f"arithmetic operand {node.describe()} is outside the supported set")

    def emit(self, node, out_key: Optional[str] = None) -> str:
        """Append the spec computing `node` (after those of its operands) and return its out_key.
        out_key=None: the node is an operand -- reuse its spec, or name it after the output it is
        returned as, or give it a temporary name."""
        have = self.done.get(id(node))
        if out_key is None:
            if have is not None:
                return have
            out_key = self.named.get(id(node))
            if out_key is None:
                out_key = f"{TEMP_PREFIX}{self.ntemp}"
                self.ntemp += 1
        elif have == out_key:
            return out_key
        specs = self.plan.specs
        if isinstance(node, g.ZScore):
            specs.append(ZScoreSpec(out_key, _need_dense(node.x, "scale_to_z_score")))
        elif have is not None:  # the same tensor returned under a second name: x + 0
            specs.append(ArithSpec(out_key, "add", ["out", have], ["const", 0.0]))
            return out_key
        elif isinstance(node, g.ScaleMinMax):
            specs.append(ScaleSpec(out_key, self._numeric_col(node.x, "scale_by_min_max"), node.output_min, node.output_max))
        elif isinstance(node, g.Arith):
            a, b = self.operand(node.a), self.operand(node.b)
            if a[0] in ("const", "scalar") and b[0] in ("const", "scalar"):
                raise NotImplementedError(f"{node.describe()}: at least one operand must be a column")
            specs.append(ArithSpec(out_key, node.op, a, b))
        else:
            raise AssertionError(node)
        self.done.setdefault(id(node), out_key)
        return out_key


def compile_outputs(outputs: Dict[str, g.Node], schema: Dict[str, str], side=()) -> Plan:
    comp = _Compiler(outputs, schema)
    plan = comp.plan
    for node in side:
        if isinstance(node, g.VocabFile):
            plan.specs.append(VocabOnlySpec(_need_dense(node.x, "vocabulary"), node.top_k, node.frequency_threshold,
                                            node.vocab_filename))
    for out_key, node in outputs.items():
        if isinstance(node, _F32_MAPPERS):
            comp.emit(node, out_key)
        elif isinstance(node, g.Vocab) and isinstance(node.x, g.AsString):
            key = _need_dense(node.x.x, "as_string")
            if schema[key] != "int64":
                raise NotImplementedError("tf.as_string: only int64 features are supported")
            plan.specs.append(VocabSpec(out_key, key, node.top_k, node.num_oov_buckets, node.default_value,
                                        node.frequency_threshold, node.vocab_filename, as_string=True))
        elif isinstance(node, g.StaticVocab):
            terms = [t.decode("latin-1") if isinstance(t, bytes) else str(t).encode("utf-8").decode("latin-1") for t in node.terms]
            plan.specs.append(StaticVocabSpec(out_key, _need_dense(node.x, "table.lookup"), terms, node.num_oov_buckets,
                                              node.default_value))
        elif isinstance(node, g.Cast) and _dense_key(node.x) is not None and node.dtype in (g.int64, g.float32):
            key = _dense_key(node.x)
            to = "int64" if node.dtype == g.int64 else "float32"
            if schema[key] == "bytes":
                raise NotImplementedError(f"output {out_key!r}: cannot cast a string feature")
            plan.specs.append(FillSpec(out_key, key) if schema[key] == to else CastSpec(out_key, key, to))
        elif isinstance(node, g.Vocab):
            plan.specs.append(VocabSpec(out_key, _need_dense(node.x, "compute_and_apply_vocabulary"), node.top_k,
                                        node.num_oov_buckets, node.default_value, node.frequency_threshold,
                                        node.vocab_filename))
        elif isinstance(node, g.Bucketize):
            plan.specs.append(BucketSpec(out_key, _need_dense(node.x, "bucketize"), node.num_buckets, node.epsilon))
        elif _dense_key(node) is not None:
            plan.specs.append(FillSpec(out_key, _dense_key(node)))
        else:
            m = _match_ratio_label(node) or _match_bool_label(node)
            if m is None:
                raise NotImplementedError(
                    f"output {out_key!r}: expression {node.describe()} is outside the supported set "
                    "(z-score, vocabulary, bucketize, fill-missing pass-through, ratio label)")
            fare, tips, ratio = m
            if schema[fare] != "float32" or schema[tips] != "float32":
                raise NotImplementedError(f"output {out_key!r}: the ratio label needs float32 features")
            plan.specs.append(RatioLabelSpec(out_key, fare, tips, ratio))
    if len(plan.specs) > 32:
        raise NotImplementedError("at most 32 output features per transform program (TXF_MAX_OPS)")
    return plan


def trace(preprocessing_fn: Callable, schema: Dict[str, str], dense_keys=()) -> Plan:
    """Call preprocessing_fn ONCE with symbolic inputs (as tf.Transform does) and compile.
    Features in `dense_keys` (tf.FixedLenFeature) arrive as dense tensors, the rest as rank-2
    SparseTensors with at most one value per row (tf.VarLenFeature)."""
    dense_keys = set(getattr(schema, "dense_keys", ())) | set(dense_keys)
    schema = dict(getattr(schema, "column_dtypes", schema))
    inputs = {}
    for k, dt in schema.items():
        sp = g.SparseInput(k, dt)
        inputs[k] = g.Dense(sp, "" if dt == "bytes" else 0, rank=1) if k in dense_keys else sp
    g.begin_trace()
    try:
        outputs = preprocessing_fn(inputs)
    finally:
        side = g.end_trace()
    if not isinstance(outputs, dict):
        raise TypeError("preprocessing_fn must return a dict of output features")
    return compile_outputs(outputs, schema, side)
