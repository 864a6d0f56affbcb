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


@dataclass
class Plan:
    schema: Dict[str, str]                      # input feature -> 'float32' | 'int64' | 'bytes'
    specs: List[object] = field(default_factory=list)

    def of(self, cls):
        return [s for s in self.specs if isinstance(s, cls)]

    def input_keys(self) -> List[str]:
        keys = []
        for s in self.specs:
            for k in ([s.fare_key, s.tips_key] if isinstance(s, RatioLabelSpec) else [s.in_key]):
                if k not in keys:
                    keys.append(k)
        return keys

    def output_schema(self) -> Dict[str, str]:
        out = {}
        for s in self.specs:
            if isinstance(s, ZScoreSpec):
                out[s.out_key] = "float32"
            elif isinstance(s, FillSpec):
                out[s.out_key] = self.schema[s.in_key]
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


def compile_outputs(outputs: Dict[str, g.Node], schema: Dict[str, str]) -> Plan:
    plan = Plan(schema=dict(schema))
    for out_key, node in outputs.items():
        if isinstance(node, g.ZScore):
            plan.specs.append(ZScoreSpec(out_key, _need_dense(node.x, "scale_to_z_score")))
        elif isinstance(node, g.Vocab):
            plan.specs.append(VocabSpec(out_key, _need_dense(node.x, "compute_and_apply_vocabulary"), node.top_k,
                                        node.num_oov_buckets, node.default_value, node.frequency_threshold,
                                        node.vocab_filename))
        elif isinstance(node, g.Bucketize):
            plan.specs.append(BucketSpec(out_key, _need_dense(node.x, "bucketize"), node.num_buckets, node.epsilon))
        elif _dense_key(node) is not None:
            plan.specs.append(FillSpec(out_key, _dense_key(node)))
        else:
            m = _match_ratio_label(node)
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


def trace(preprocessing_fn: Callable, schema: Dict[str, str]) -> Plan:
    """Call preprocessing_fn ONCE with symbolic sparse inputs (as tf.Transform does) and compile."""
    inputs = {k: g.SparseInput(k, dt) for k, dt in schema.items()}
    outputs = preprocessing_fn(inputs)
    if not isinstance(outputs, dict):
        raise TypeError("preprocessing_fn must return a dict of output features")
    return compile_outputs(outputs, schema)
