"""Stand-ins for ``tensorflow_transform.tf_metadata`` (dataset_metadata, dataset_schema,
schema_utils) as used at ``airflow-dags/taxi_utils.py:29,75-76`` and
``notebooks/06_Airflow_Feature_Analysis.ipynb:104-105``: just enough to carry a feature spec."""
from __future__ import annotations

import types

from . import graph as g
from . import tf_shim

_TO_COLUMN = {"string": "bytes", "int64": "int64", "float32": "float32"}


class Schema:
    """Feature name -> column dtype ('float32' | 'int64' | 'bytes')."""

    def __init__(self, column_dtypes, dense_keys=()):
        self.column_dtypes = dict(column_dtypes)
        self.dense_keys = set(dense_keys)  # tf.FixedLenFeature features: dense tensors in preprocessing_fn

    def as_feature_spec(self):
        inv = {"bytes": g.string, "int64": g.int64, "float32": g.float32}
        return {k: (tf_shim.FixedLenFeature([], inv[v]) if k in self.dense_keys else tf_shim.VarLenFeature(inv[v]))
                for k, v in self.column_dtypes.items()}


def from_feature_spec(feature_spec):
    cols, dense = {}, set()
    for name, feat in feature_spec.items():
        dt = feat.dtype if isinstance(feat, (tf_shim.FixedLenFeature, tf_shim.VarLenFeature)) else feat
        cols[name] = _TO_COLUMN[dt.name] if isinstance(dt, g.DType) else str(dt)
        if isinstance(feat, tf_shim.FixedLenFeature):
            if list(feat.shape) not in ([], [1]):
                raise NotImplementedError(f"feature {name!r}: only scalar FixedLenFeature shapes are supported")
            dense.add(name)
    return Schema(cols, dense)


class DatasetMetadata:
    def __init__(self, schema):
        self.schema = schema if isinstance(schema, Schema) else Schema(schema)


def schema_as_feature_spec(schema):
    sch = schema if isinstance(schema, Schema) else Schema(schema)
    return types.SimpleNamespace(feature_spec=sch.as_feature_spec(), domains={})


dataset_schema = types.SimpleNamespace(from_feature_spec=from_feature_spec, Schema=Schema)
dataset_metadata = types.SimpleNamespace(DatasetMetadata=DatasetMetadata)
schema_utils = types.SimpleNamespace(schema_as_feature_spec=schema_as_feature_spec)
