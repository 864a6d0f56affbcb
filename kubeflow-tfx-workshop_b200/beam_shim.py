"""Stand-in for ``tensorflow_transform.beam`` (``tft_beam``): the driver API the notebooks use.

Reference call sites: ``notebooks/06_Airflow_Feature_Analysis.ipynb:236-262``,
``notebooks/03_TensorFlow_Transform.ipynb:197-200`` (``Context`` + ``AnalyzeAndTransformDataset``),
``notebooks/03a_TensorFlow_Transform_Advanced.ipynb:442,468,482`` (``TransformDataset``,
``WriteTransformFn``).  The PTransforms apply with Beam's ``value | transform`` syntax:

    with tft_beam.Context(temp_dir=tmp):
        (transformed, metadata), transform_fn = (
            (raw_data, raw_metadata) | tft_beam.AnalyzeAndTransformDataset(preprocessing_fn))

``raw_data`` is a list of dict rows (as in the reference) or a ``Table``; ``transformed`` is a
GPU-resident ``Table`` (``rows()`` below converts small ones back to dict rows).
"""
from __future__ import annotations

import threading
from typing import Callable, List

from .columns import StringColumn, Table
from .engine import Engine, TransformState
from .plan import Plan, trace
from .tf_metadata import DatasetMetadata, Schema

_ctx = threading.local()


class Context:
    """tft_beam.Context(temp_dir=...): only scopes options; nothing is written to temp_dir."""

    def __init__(self, temp_dir=None, desired_batch_size=None, passthrough_keys=None, use_deep_copy_optimization=None,
                 device="cuda", comm=None, **engine_kwargs):
        self.temp_dir = temp_dir
        self.device = device
        self.comm = comm
        self.engine_kwargs = engine_kwargs

    def __enter__(self):
        self._prev = getattr(_ctx, "current", None)
        _ctx.current = self
        return self

    def __exit__(self, *exc):
        _ctx.current = self._prev
        return False


def _context() -> Context:
    return getattr(_ctx, "current", None) or Context()


def _as_table(dataset, metadata) -> Table:
    schema = metadata.schema.column_dtypes if isinstance(metadata, DatasetMetadata) else dict(metadata)
    if isinstance(dataset, Table):
        return dataset
    return Table.from_rows(list(dataset), schema)


def _schema_of(dataset, metadata):
    if isinstance(metadata, DatasetMetadata):
        return metadata.schema.column_dtypes
    if isinstance(metadata, Schema):
        return metadata.column_dtypes
    if metadata is None and isinstance(dataset, Table):
        return dataset.schema()
    return dict(metadata)


def _dense_of(metadata):
    sch = metadata.schema if isinstance(metadata, DatasetMetadata) else metadata
    return set(getattr(sch, "dense_keys", ()))


class TransformFn:
    """What AnalyzeDataset returns: the traced plan plus the analyzer state (tft's transform_fn)."""

    def __init__(self, engine: Engine, state: TransformState):
        self.engine = engine
        self.state = state

    @property
    def plan(self) -> Plan:
        return self.engine.plan

    def __or__(self, other):
        return other.__ror__(self)


class _PTransform:
    def __ror__(self, left):
        return self.expand(left)


class AnalyzeDataset(_PTransform):
    def __init__(self, preprocessing_fn: Callable):
        self.preprocessing_fn = preprocessing_fn

    def _engine(self, dataset, metadata) -> Engine:
        ctx = _context()
        plan = trace(self.preprocessing_fn, _schema_of(dataset, metadata), _dense_of(metadata))
        return Engine(plan, device=ctx.device, comm=ctx.comm, **ctx.engine_kwargs)

    def expand(self, dataset_and_metadata) -> TransformFn:
        dataset, metadata = dataset_and_metadata
        engine = self._engine(dataset, metadata)
        table = _as_table(dataset, DatasetMetadata(_schema_of(dataset, metadata))).to(engine.device)
        return TransformFn(engine, engine.analyze(table).detach())


class TransformDataset(_PTransform):
    def expand(self, args):
        (dataset, metadata), transform_fn = args
        eng = transform_fn.engine
        table = _as_table(dataset, DatasetMetadata(_schema_of(dataset, metadata))).to(eng.device)
        out = eng.transform(table, transform_fn.state)
        return out, DatasetMetadata(eng.plan.output_schema())


class AnalyzeAndTransformDataset(AnalyzeDataset):
    def expand(self, dataset_and_metadata):
        dataset, metadata = dataset_and_metadata
        engine = self._engine(dataset, metadata)
        table = _as_table(dataset, DatasetMetadata(_schema_of(dataset, metadata))).to(engine.device)
        state = engine.analyze(table).detach()
        out = engine.transform(table, state)
        return (out, DatasetMetadata(engine.plan.output_schema())), TransformFn(engine, state)


class WriteTransformFn(_PTransform):
    def __init__(self, path: str):
        self.path = path

    def expand(self, transform_fn: TransformFn):
        from .state import save_state
        save_state(transform_fn.state, self.path)
        return self.path


def rows(table: Table) -> List[dict]:
    """Small tables back to the reference's row-major form (list of dicts of scalars)."""
    cols = {}
    for k, c in table.items():
        if isinstance(c, StringColumn):
            cols[k] = c.to_list()
        else:
            cols[k] = c.values.cpu().numpy().tolist()
    return [{k: v[i] for k, v in cols.items()} for i in range(len(table))]
