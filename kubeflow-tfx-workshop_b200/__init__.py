"""B200-native analyze + transform pass of the Chicago-taxi ``preprocessing_fn``
(``pablomendes/kubeflow-tfx-workshop``, ``airflow-dags/taxi_utils.py:86-145``).

Public surface:

* ``Table`` / ``NumericColumn`` / ``StringColumn`` -- columnar ingest (host or GPU buffers);
  ``read_csv`` parses CSV text on the GPU into such a table;
* ``tf``, ``tft``, ``tft_beam``, ``tf_metadata`` -- stand-ins for the symbols the reference's
  ``preprocessing_fn`` and notebooks touch; ``install_shims()`` binds them into ``sys.modules`` so
  that file runs unmodified;
* ``trace`` / ``Engine`` / ``TransformState`` -- the explicit API: trace once, ``analyze`` (full
  pass, multi-GPU merge), ``transform`` (one fused kernel);
* ``save_state`` / ``load_state`` -- persist and re-apply the analyzer state.

All arithmetic runs in ``libtaxi_xf.so`` (hand-written sm_100a CUDA behind the C ABI in
``include/taxi_xf.h``).  There is no CPU fallback: without the built library, import fails.
"""
from __future__ import annotations

import contextlib
import os
import sys
import types

# analyze_and_transform keeps four CUDA streams busy at once; with the default of 8 host->device work queues two of
# them can land on one queue and serialise falsely.  Only effective if the CUDA context does not exist yet.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from . import _native  # noqa: F401  (raises ImportError when libtaxi_xf.so is missing)
from . import beam_shim as tft_beam
from . import tf_metadata, tf_shim, tft_shim
from .columns import NumericColumn, StringColumn, Table
from .dist import Comm, shard_range
from .engine import Engine, TransformState
from .egress import encode_tfrecord, write_tfrecord
from .ingest import read_csv, read_tfrecord
from .plan import Plan, trace
from .state import TFTransformOutput, load_state, save_state

tf = tf_shim
tft = tft_shim

__all__ = ["Table", "NumericColumn", "StringColumn", "Engine", "TransformState", "Plan", "trace", "Comm", "shard_range",
           "tf", "tft", "tft_beam", "tf_metadata", "save_state", "load_state", "TFTransformOutput", "read_csv", "read_tfrecord", "encode_tfrecord", "write_tfrecord", "install_shims"]

_SHIM_MODULES = {
    "tensorflow": tf_shim,
    "tensorflow_transform": tft_shim,
    "tensorflow_transform.beam": tft_beam,
    "tensorflow_transform.beam.impl": tft_beam,
    "tensorflow_transform.tf_metadata": tf_metadata,
    "tensorflow_transform.tf_metadata.dataset_metadata": tf_metadata.dataset_metadata,
    "tensorflow_transform.tf_metadata.dataset_schema": tf_metadata.dataset_schema,
    "tensorflow_transform.tf_metadata.schema_utils": tf_metadata.schema_utils,
}


@contextlib.contextmanager
def install_shims():
    """Bind the stand-ins into ``sys.modules`` so the reference's ``taxi_utils.py`` (which does
    ``import tensorflow as tf`` / ``import tensorflow_transform as tft`` /
    ``import tensorflow_model_analysis as tfma``, lines 26-29) imports and runs unmodified."""
    tft_shim.beam = tft_beam
    tft_shim.tf_metadata = tf_metadata
    tft_shim.TFTransformOutput = TFTransformOutput
    mods = dict(_SHIM_MODULES)
    mods["tensorflow_model_analysis"] = types.ModuleType("tensorflow_model_analysis")  # imported, unused on this path
    saved = {k: sys.modules.get(k) for k in mods}
    sys.modules.update(mods)
    try:
        yield
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
