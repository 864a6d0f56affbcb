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

_SPEC_TYPES = {c.__name__: c for c in (P.ZScoreSpec, P.VocabSpec, P.BucketSpec, P.FillSpec, P.RatioLabelSpec)}


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
        "vocabs": [],
    }
    if state.mean_var is not None:
        mv = state.mean_var.cpu().numpy()
        m = state.moments.cpu().numpy()
        for i, k in enumerate(state.moment_keys):
            doc["moments"][k] = {"mean_f32_hex": float(mv[i, 0]).hex(), "var_f32_hex": float(mv[i, 1]).hex(),
                                 "count": float(m[i, 0]), "mean": float(m[i, 1]), "m2": float(m[i, 2])}
    for nb, res in state.buckets.items():
        doc["buckets"][str(nb)] = {k: [float(x).hex() for x in res.boundaries_of(k)] for k in res.keys}
    for key, res in state.vocabs.items():
        fname = _vocab_file(key)
        with open(os.path.join(path, fname), "wb") as f:
            for t in res.terms():
                f.write(t + b"\n")
        doc["vocabs"].append({"key": list(key), "file": fname, "size": res.size})
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
