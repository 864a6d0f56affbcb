"""Persist / reload the analyzer state (what tft writes as ``transform_fn`` and
what ``tft.TFTransformOutput`` reads back at ``airflow-dags/taxi_utils.py:211-212,244-245,307``).

Layout of the directory:  ``state.json`` (plan, means/variances, boundaries) and one
``<feature>.vocab`` file per vocabulary -- one term per line in id order, the format of tft's own
vocabulary assets (which is why terms containing newlines never enter a vocabulary).
"""
from __future__ import annotations

import dataclasses
import json
import os
from typing import Dict

import torch

from . import plan as P
from .engine import BucketResult, Engine, TransformState, _vocab_key

_SPEC_TYPES = {c.__name__: c for c in (P.ZScoreSpec, P.VocabSpec, P.BucketSpec, P.FillSpec, P.RatioLabelSpec, P.VocabOnlySpec,
                                       P.StaticVocabSpec, P.ScaleSpec, P.ArithSpec, P.CastSpec)}


def _vocab_file(key) -> str:
    in_key, top_k, thr = key
    return f"{in_key}.top{top_k}.thr{thr}.vocab"


def save_state(state: TransformState, path: str) -> None:
    os.makedirs(path, exist_ok=True)
    doc = {
        "format": 1,
        "num_rows": state.num_rows,
        "schema": state.plan.schema,
        "specs": [dict(type=type(s).__name__, **dataclasses.asdict(s)) for s in state.plan.specs],
        "moments": {},
        "buckets": {},
        "minmax": {},
        "vocabs": [],
    }
    if state.mean_var is not None:
        mv = state.mean_var.cpu().numpy()
        m = state.moments.cpu().numpy()
        for i, k in enumerate(state.moment_keys):
            doc["moments"][k] = {"mean_f32_hex": float(mv[i, 0]).hex(), "var_f32_hex": float(mv[i, 1]).hex(),
                                 "count": float(m[i, 0]), "mean": float(m[i, 1]), "m2": float(m[i, 2])}
    if state.minmax is not None:
        mm = state.minmax.cpu().numpy()
        for i, k in enumerate(state.minmax_keys):
            doc["minmax"][k] = {"min_f32_hex": float(mm[i, 0]).hex(), "max_f32_hex": float(mm[i, 1]).hex()}
    named = {}  # vocab key -> file name the user asked for (vocab_filename=)
    for sp in state.plan.of(P.VocabSpec) + state.plan.of(P.VocabOnlySpec):
        if sp.vocab_filename:
            named.setdefault(_vocab_key(sp), sp.vocab_filename)
    for nb, res in state.buckets.items():
        doc["buckets"][str(nb)] = {k: [float(x).hex() for x in res.boundaries_of(k)] for k in res.keys}
    for key, res in state.vocabs.items():
        fname = _vocab_file(key)
        with open(os.path.join(path, fname), "wb") as f:
            for t in res.terms():
                f.write(t + b"\n")
        doc["vocabs"].append({"key": list(key), "file": fname, "size": res.size, "vocab_filename": named.get(key)})
        if key in named:  # tft stores the asset under the name the user gave (vocabulary_file_by_name)
            with open(os.path.join(path, named[key]), "wb") as f:
                for t in res.terms():
                    f.write(t + b"\n")
    with open(os.path.join(path, "state.json"), "w") as f:
        json.dump(doc, f, indent=1)


def load_state(path: str, device="cuda", comm=None):
    """-> (Engine, TransformState) ready for Engine.transform (tft's transform_raw_features)."""
    with open(os.path.join(path, "state.json")) as f:
        doc = json.load(f)
    specs = []
    for d in doc["specs"]:
        d = dict(d)
        specs.append(_SPEC_TYPES[d.pop("type")](**d))
    plan = P.Plan(schema=doc["schema"], specs=specs)
    eng = Engine(plan, device=device, comm=comm)
    dev = eng.device
    keys = list(doc["moments"].keys())
    mean_var = moments = None
    if keys:
        mean_var = torch.tensor([[float.fromhex(doc["moments"][k]["mean_f32_hex"]), float.fromhex(doc["moments"][k]["var_f32_hex"])]
                                 for k in keys], dtype=torch.float32, device=dev)
        moments = torch.tensor([[doc["moments"][k]["count"], doc["moments"][k]["mean"], doc["moments"][k]["m2"]] for k in keys],
                               dtype=torch.float64, device=dev)
    state = TransformState(plan=plan, moment_keys=keys, mean_var=mean_var, moments=moments, num_rows=doc["num_rows"])
    mk = list(doc.get("minmax", {}).keys())
    if mk:
        state.minmax_keys = mk
        state.minmax = torch.tensor([[float.fromhex(doc["minmax"][k]["min_f32_hex"]), float.fromhex(doc["minmax"][k]["max_f32_hex"])]
                                     for k in mk], dtype=torch.float32, device=dev)
    for nb_s, cols in doc["buckets"].items():
        nb = int(nb_s)
        ks = list(cols.keys())
        T = max(nb - 1, 1)
        b = torch.zeros((len(ks), T), dtype=torch.float32)
        nbound = torch.zeros(len(ks), dtype=torch.int32)
        for c, k in enumerate(ks):
            vals = [float.fromhex(h) for h in cols[k]]
            b[c, : len(vals)] = torch.tensor(vals, dtype=torch.float32)
            nbound[c] = len(vals)
        state.buckets[nb] = BucketResult(ks, nb, b.to(dev), nbound.to(dev), path="loaded")
    for v in doc["vocabs"]:
        with open(os.path.join(path, v["file"]), "rb") as f:
            data = f.read()
        terms = data.split(b"\n")[:-1] if data else []
        key = tuple(v["key"])
        state.vocabs[key] = eng._vocab_from_terms(key, terms)
    return eng, state


class TFTransformOutput:
    """``tft.TFTransformOutput(dir)`` (``airflow-dags/taxi_utils.py:307``,
    ``notebooks/03a_TensorFlow_Transform_Advanced.ipynb``): reads what WriteTransformFn wrote."""

    def __init__(self, transform_output_dir: str, device="cuda", comm=None):
        self.transform_output_dir = transform_output_dir
        self._device, self._comm = device, comm
        self._loaded = None
        with open(os.path.join(transform_output_dir, "state.json")) as f:
            self._doc = json.load(f)

    def _load(self):
        if self._loaded is None:
            self._loaded = load_state(self.transform_output_dir, self._device, self._comm)
        return self._loaded

    def transform_raw_features(self, raw_features):
        """Apply the saved transform to a GPU-resident Table of raw features (:211-212, :244-245)."""
        eng, state = self._load()
        return eng.transform(raw_features, state)

    def vocabulary_file_by_name(self, vocab_filename: str) -> str:
        for v in self._doc["vocabs"]:
            if v.get("vocab_filename") == vocab_filename:
                return os.path.join(self.transform_output_dir, vocab_filename)
        raise KeyError(f"no vocabulary was saved under the name {vocab_filename!r}")

    def vocabulary_size_by_name(self, vocab_filename: str) -> int:
        for v in self._doc["vocabs"]:
            if v.get("vocab_filename") == vocab_filename:
                return int(v["size"])
        raise KeyError(f"no vocabulary was saved under the name {vocab_filename!r}")

    def transformed_feature_spec(self) -> Dict[str, str]:
        """Output feature -> 'float32' | 'int64' | 'bytes'."""
        return self._load()[0].plan.output_schema()
