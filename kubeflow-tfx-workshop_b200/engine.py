"""Host driver of the analyze + transform pass (the work tft_beam.AnalyzeAndTransformDataset does
for ``preprocessing_fn``, ``airflow-dags/taxi_utils.py:106-145``).

``Engine.analyze`` runs every analyzer of the plan in ONE pass over the rows held by this rank
(moments, vocabulary counts, quantile bracket counts), merges the partial states across ranks
with ``dist.Comm`` and finalises them on the GPU.  ``Engine.transform`` launches the single fused
transform kernel.  All arithmetic happens in libtaxi_xf.so; PyTorch only owns the buffers and the
stream, and there is no CPU fallback (``_native`` raises at import when the library is missing).
"""
from __future__ import annotations

import math
import os
import time
import weakref
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch

from . import _native as nat
from .columns import NumericColumn, StringColumn, Table
from .dist import Comm
from .plan import (ArithSpec, BucketSpec, CastSpec, FillSpec, Plan, RatioLabelSpec, ScaleSpec, StaticVocabSpec, TEMP_PREFIX,
                   VocabOnlySpec, VocabSpec, ZScoreSpec)

_NO_CAP = 1 << 62  # "no candidate-buffer limit" for txf_qresolve when overflow is tracked elsewhere
EMPTY_VOCAB_SENTINEL = b"49d0cd50-04bb-48c0-bc6f-5b575dce351a"  # tft's placeholder for an empty vocabulary
_DT = {"float32": nat.TXF_F32, "int64": nat.TXF_I64}


def _pow2_at_least(x: int) -> int:
    p = 1
    while p < x:
        p <<= 1
    return p


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class VocabResult:
    """One finished vocabulary.  The term bytes (id order) are exported from the count table only
    when somebody asks for them (terms(), detach(), save_state): the transform needs the lookup
    table alone, and the export is a handful of launch-bound kernels."""

    def __init__(self, spec_key: Tuple, size: int, lookup: torch.Tensor, lookup_cap: int, lookup_pool: torch.Tensor,
                 terms_offsets: Optional[torch.Tensor] = None, terms_pool: Optional[torch.Tensor] = None, export=None):
        self.spec_key = spec_key
        self.size = size                      # len(vocabulary) -- the real list length, not top_k
        self.lookup = lookup                  # uint8 bytes of the 32-byte-slot lookup table
        self.lookup_cap = lookup_cap
        self.lookup_pool = lookup_pool        # the bytes the table's slots point into (count-table arena or the terms' pool)
        self._offsets, self._pool, self._export = terms_offsets, terms_pool, export

    def _materialize(self):
        if self._offsets is None:
            self._offsets, self._pool = self._export()
            self._export = None

    @property
    def terms_offsets(self) -> torch.Tensor:  # int32 [size+1] (device), id order
        self._materialize()
        return self._offsets

    @property
    def terms_pool(self) -> torch.Tensor:     # uint8 (device)
        self._materialize()
        return self._pool

    def terms(self) -> List[bytes]:
        o = self.terms_offsets.cpu().numpy()
        buf = self.terms_pool[: int(o[-1])].cpu().numpy().tobytes()
        return [buf[o[i]:o[i + 1]] for i in range(self.size)]


@dataclass
class BucketResult:
    keys: List[str]
    num_buckets: int
    boundaries: torch.Tensor       # float32 [ncols, T] (device), first nbound[c] entries are valid
    nbound: torch.Tensor           # int32 [ncols] (device)
    path: str = "sample-select"

    def boundaries_of(self, key: str):
        c = self.keys.index(key)
        return self.boundaries[c, : int(self.nbound[c].item())].cpu().numpy()


@dataclass
class TransformState:
    """Analyzer results, device resident (what tft stores as constants of the transform graph)."""

    plan: Plan
    moment_keys: List[str]
    mean_var: Optional[torch.Tensor]                     # float32 [n, 2] (device)
    moments: Optional[torch.Tensor]                      # float64 [n, 3] (count, mean, M2) (device)
    vocabs: Dict[Tuple, VocabResult] = field(default_factory=dict)
    buckets: Dict[int, BucketResult] = field(default_factory=dict)
    num_rows: int = 0
    # dictionary codes of the analyzed chunks: (vocab key, chunk index) -> CodedChunk.  They alias Engine
    # workspaces and only serve Engine.transform on the very tables (same object, same buffers) that were analyzed.
    codes: Dict[Tuple, "CodedChunk"] = field(default_factory=dict)
    minmax_keys: List[str] = field(default_factory=list)
    minmax: Optional[torch.Tensor] = None                # float32 [n, 2] (min, max) (device)

    def mean_var_of(self, key: str):
        return tuple(self.mean_var[self.moment_keys.index(key)].cpu().tolist())

    def min_max_of(self, key: str):
        return tuple(self.minmax[self.minmax_keys.index(key)].cpu().tolist())

    def scalar(self, kind: str, key: str) -> torch.Tensor:
        """One-element device view of an analyzer scalar (tft.mean / var / min / max of a feature)."""
        if kind in ("mean", "var"):
            return self.mean_var[self.moment_keys.index(key), 0 if kind == "mean" else 1:][:1]
        return self.minmax[self.minmax_keys.index(key), 0 if kind == "min" else 1:][:1]

    def detach(self) -> "TransformState":
        """Copy the state out of the Engine's reusable workspaces.  A state returned by
        Engine.analyze aliases them and is overwritten by the next analyze() on that Engine."""
        c = lambda t: None if t is None else t.clone()
        return TransformState(
            plan=self.plan, moment_keys=list(self.moment_keys), mean_var=c(self.mean_var), moments=c(self.moments),
            vocabs={k: vocab_from_device_terms(v.spec_key, v.size, v.terms_offsets.clone(), v.terms_pool.clone())
                    for k, v in self.vocabs.items()},
            buckets={k: BucketResult(list(b.keys), b.num_buckets, b.boundaries.clone(), b.nbound.clone(), b.path)
                     for k, b in self.buckets.items()},
            num_rows=self.num_rows, minmax_keys=list(self.minmax_keys), minmax=c(self.minmax))


@dataclass
class CodedChunk:
    """Dictionary codes of one analyzed chunk.  The chunk is remembered by a weak reference plus the
    addresses and sizes of the source column's buffers: `id()` of a freed table can be reused by
    another table of the same length, and a column can be replaced on the table after analyze."""

    table: "weakref.ref"
    offsets_ptr: int
    pool_ptr: int
    valid_ptr: Optional[int]
    n: int
    codes: torch.Tensor       # int32 [n]
    decode: torch.Tensor      # int64 [local codes + capacity]

    def matches(self, table: Table, col: StringColumn) -> bool:
        return (self.table() is table and len(table) == self.n and col.offsets.data_ptr() == self.offsets_ptr
                and col.pool.data_ptr() == self.pool_ptr and _ptr(col.valid) == self.valid_ptr and self.codes.numel() == self.n)


def vocab_from_device_terms(spec_key: Tuple, k: int, offsets: torch.Tensor, pool: torch.Tensor, lookup: Optional[torch.Tensor] = None,
                            lookup_cap: Optional[int] = None) -> VocabResult:
    """Self-contained VocabResult over an explicit vocabulary (device int32 offsets [k+1] + byte pool with slack):
    the lookup table's slots point into `pool`."""
    if lookup is None:
        lookup_cap = _pow2_at_least(max(2 * k, 2))
        lookup = torch.empty(nat.query("txf_vocab_lookup_bytes", lookup_cap), dtype=torch.uint8, device=offsets.device)
    with torch.cuda.device(offsets.device):
        nat.call("txf_vocab_lookup_from_terms", offsets.data_ptr(), pool.data_ptr(), k, lookup.data_ptr(), lookup_cap,
                 torch.cuda.current_stream().cuda_stream)
    return VocabResult(spec_key, k, lookup, lookup_cap, pool, offsets, pool)


AS_STRING_SUFFIX = "@as_string"  # vocabulary over tf.as_string(int64 feature)


def _vocab_key(s) -> Tuple:
    """(source, top_k, frequency_threshold) of a VocabSpec / VocabOnlySpec: specs with the same key share one analyzer."""
    src = s.in_key + AS_STRING_SUFFIX if getattr(s, "as_string", False) else s.in_key
    return (src, s.top_k, s.frequency_threshold)


class _VocabGrow(Exception):
    """Raised (by every rank alike) when a vocabulary pass ran out of table slots or arena bytes."""

    def __init__(self, key: str, what: str, factor: int):
        super().__init__(f"{key}: {what}")
        self.key, self.what, self.factor = key, what, factor


class Engine:
    def __init__(self, plan: Plan, device: Union[str, torch.device] = "cuda", comm: Optional[Comm] = None,
                 vocab_capacity: int = 1 << 16, arena_bytes: int = 1 << 22, sample_size: int = 1 << 20,
                 bracket_sigmas: float = 5.0, seed: int = 20190501):
        self.plan = plan
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("Engine needs a CUDA device: the analyze/transform kernels have no CPU path")
        self.comm = comm or Comm()
        self.vocab_capacity = _pow2_at_least(vocab_capacity)   # starting sizes: analyze() grows them when a pass overflows
        self.arena_bytes = int(arena_bytes)
        self.max_vocab_capacity, self.max_arena_bytes = 1 << 30, 1 << 34
        self.grow_events: List[Tuple] = []
        self._count_warps = 0        # 24 while analyze_and_transform streams the state-independent ops next to the count pass
        self.overlap_transform = True  # analyze_and_transform: run the ops that need no analyzer result under the analyzers
        self._xstream = None
        self._step_end = None
        self.drain_between_steps_world = 4  # see analyze_and_transform
        self._side_priority = -1          # experiment knobs (tools/exp/ab_wait.py)
        self._host_waits_on_event = True
        self.sample_size = int(sample_size)
        self.bracket_sigmas = float(bracket_sigmas)
        self.seed = int(seed)
        self._ws: Dict[Tuple, torch.Tensor] = {}
        self.force_general_quantiles = False
        self.bracket_select = True           # candidates: one bracket-histogram pass (falls back to radix digits when a bracket is wide)
        self.last_quantile_select: Dict[int, str] = {}
        self.split_quantiles = True          # many targets: two passes over equi-depth splitters instead of the 4-pass radix select
        self.force_split_quantiles = False   # (tests) take the two-pass engine even where one pass would do
        self.use_codes = True   # analyze writes dictionary codes; transform of the same table decodes them
        self.overlap_streams = True
        self._side = None
        # multi-rank merge through ONE all-gather of a fixed-layout buffer (no sizes via the host);
        # anything that does not fit (more terms / longer terms / more candidates) falls back to the sized exchange
        self.fixed_exchange = True
        self.xchg_terms, self.xchg_term_bytes = 4096, 64
        self.partition_vocab = True   # sized exchange: owners re-count their share of the terms (top_k vocabularies)
        self.last_vocab_exchange: Dict[str, str] = {}
        self.timeline: Dict[str, float] = {}   # TXF_PROFILE=1: wall ms per analyze() segment (with device syncs)
        self._profile = os.environ.get("TXF_PROFILE") == "1"
        self._trace_on = os.environ.get("TXF_TRACE") == "1"
        self.trace_events: List[Tuple[str, torch.cuda.Event]] = []
        self.last_quantile_path: Dict[int, str] = {}

        self.moment_keys: List[str] = []
        for k in [s.in_key for s in plan.of(ZScoreSpec)] + plan.scalar_keys(("mean", "var")):
            if k not in self.moment_keys:
                self.moment_keys.append(k)
        self.minmax_keys: List[str] = []
        for k in [s.in_key for s in plan.of(ScaleSpec)] + plan.scalar_keys(("min", "max")):
            if k not in self.minmax_keys:
                self.minmax_keys.append(k)
        self.vocab_specs: Dict[Tuple, VocabSpec] = {}
        for s in plan.of(VocabSpec) + plan.of(VocabOnlySpec):
            self.vocab_specs.setdefault(_vocab_key(s), s)
        self._static_vocabs: Dict[str, VocabResult] = {}   # StaticVocabSpec.out_key -> lookup built from its terms
        self._formatted: Dict[Tuple, Tuple[Table, StringColumn]] = {}  # (id(table), feature) -> tf.as_string column
        self._consts: Dict[float, torch.Tensor] = {}
        self._host_consts: Dict[Tuple, torch.Tensor] = {}
        self.bucket_groups: Dict[int, List[str]] = {}
        for s in plan.of(BucketSpec):
            keys = self.bucket_groups.setdefault(s.num_buckets, [])
            if s.in_key not in keys:
                keys.append(s.in_key)
        for nb, keys in self.bucket_groups.items():
            if len(keys) > nat.TXF_MAX_COLS:
                raise NotImplementedError(f"at most {nat.TXF_MAX_COLS} bucketized features per num_buckets value")
        if len(self.moment_keys) > nat.TXF_MAX_COLS:
            raise NotImplementedError(f"at most {nat.TXF_MAX_COLS} z-scored features")
        if len(self.minmax_keys) > nat.TXF_MAX_COLS:
            raise NotImplementedError(f"at most {nat.TXF_MAX_COLS} min/max-scaled features")

    # ---- buffers ------------------------------------------------------------------------------
    def _buf(self, name: str, shape, dtype, zero: bool = False) -> torch.Tensor:
        """Workspace cached by (name, shape, dtype): repeated analyze/transform calls reuse it."""
        shape = tuple(int(x) for x in (shape if isinstance(shape, (tuple, list)) else (shape,)))
        key = (name, shape, dtype)
        t = self._ws.get(key)
        if t is None:
            t = torch.zeros(shape, dtype=dtype, device=self.device) if zero else torch.empty(shape, dtype=dtype, device=self.device)
            self._ws[key] = t
        elif zero:
            t.zero_()
        return t

    def _host_const(self, data, dtype) -> torch.Tensor:
        """Small device tensor with host-computed contents, cached by value.  torch.tensor(...,
        device=cuda) synchronises the current stream; inside the analyze pass that would stall the
        host behind every kernel queued so far, so these constants are built once and reused."""
        key = (dtype, repr(data))
        t = self._host_consts.get(key)
        if t is None:
            if len(self._host_consts) > 256:
                self._host_consts.clear()
            t = torch.tensor(data, dtype=dtype, device=self.device)
            self._host_consts[key] = t
        return t

    def _mark(self, name: str, stream=None):
        """TXF_TRACE=1: record a CUDA event on `stream` (default: current) -- a timeline without host syncs."""
        if self._trace_on:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(stream if stream is not None else torch.cuda.current_stream())
            self.trace_events.append((name, ev))

    def trace_report(self) -> Dict[str, float]:
        """ms from the first mark to every later one (call after a device synchronize)."""
        if not self.trace_events:
            return {}
        t0 = next((ev for name, ev in self.trace_events if name == "analyze_start"), self.trace_events[0][1])
        out = {name: round(t0.elapsed_time(ev), 3) for name, ev in self.trace_events if ev is not t0}
        self.trace_events = []
        return out

    def _tick(self, name: str):
        if self._profile:
            torch.cuda.synchronize(self.device)
            now = time.perf_counter()
            self.timeline[name] = self.timeline.get(name, 0.0) + (now - self._t0) * 1e3
            self._t0 = now

    @staticmethod
    def _stream() -> int:
        return torch.cuda.current_stream().cuda_stream

    def _check_table(self, t: Table):
        if t.device.type != "cuda":
            raise RuntimeError("kernels take GPU-resident tables: call table.to('cuda') first")
        for k in self.plan.input_keys():
            if k not in t:
                raise KeyError(f"input feature {k!r} is missing from the table")
            want = self.plan.schema[k]
            if t[k].dtype != want:
                raise TypeError(f"feature {k!r}: table has {t[k].dtype}, the plan was traced with {want}")

    @staticmethod
    def _num_args(cols: Sequence[NumericColumn]):
        return (nat.ptr_array([c.values.data_ptr() for c in cols]), nat.ptr_array([_ptr(c.valid) for c in cols]),
                nat.i32_array([_DT[c.dtype] for c in cols]))

    # ---- analyze ------------------------------------------------------------------------------
    def analyze(self, data: Union[Table, Sequence[Table]], rows_per_rank: Optional[Sequence[int]] = None,
                _side_work=None, _between=None) -> TransformState:
        """One analyze pass.  The vocabulary count table and its byte arena start small; a pass that
        fills either grows it and runs again (tft has no such knob), on every rank alike.
        `rows_per_rank`: the number of rows every rank holds, when the caller knows it (it sharded the rows,
        e.g. with `shard_range`); otherwise the counts are gathered, which costs a host round trip up front."""
        chunks = [data] if isinstance(data, Table) else list(data)
        for c in chunks:
            self._check_table(c)
        if rows_per_rank is not None:
            rows_per_rank = [int(v) for v in rows_per_rank]
            if len(rows_per_rank) != self.comm.world or rows_per_rank[self.comm.rank] != sum(len(c) for c in chunks):
                raise ValueError("rows_per_rank must list every rank's row count, this rank's included")
        while True:
            try:
                return self._analyze_once(chunks, rows_per_rank, _side_work, _between)
            except _VocabGrow as g:
                if g.what == "table":
                    if self.vocab_capacity >= self.max_vocab_capacity:
                        raise nat.NativeError(f"vocabulary of {g.key!r} needs more than {self.max_vocab_capacity} table slots") from None
                    self.vocab_capacity = min(self.vocab_capacity * g.factor, self.max_vocab_capacity)
                else:
                    if self.arena_bytes >= self.max_arena_bytes:
                        raise nat.NativeError(f"vocabulary of {g.key!r} needs more than {self.max_arena_bytes} arena bytes") from None
                    self.arena_bytes = min(self.arena_bytes * g.factor, self.max_arena_bytes)
                self.grow_events.append((g.key, g.what, self.vocab_capacity if g.what == "table" else self.arena_bytes))
                for k in [k for k in self._ws if k[0].startswith("v_")]:
                    del self._ws[k]  # table / arena / code buffers are sized by the old capacity

    def _analyze_once(self, chunks: List[Table], rows_per_rank: Optional[List[int]] = None, side_work=None,
                      between=None) -> TransformState:
        """between(state): called once moments and bucket boundaries are final and before the vocabularies are ranked;
        analyze_and_transform queues the ops that need only those there, so that the launch-bound ranking chain
        (compact, select, sort, lookup, decode tables -- a dozen tiny kernels and a few host reads) runs next to them.
        side_work(events): analyze_and_transform's launch of the state-free stream kernel.  Single rank: first thing.
        Multi-rank: only after the all-gather of the quantile samples (`events`) -- while that kernel sits on every SM a
        collective kernel does not become resident (its shared-memory split differs and an SM only changes split when idle)."""
        n_local = sum(len(c) for c in chunks)
        if side_work is not None and not self.comm.active:
            side_work([])
            side_work = None
        if self._profile:
            torch.cuda.synchronize(self.device)
            self._t0 = time.perf_counter()
        main = torch.cuda.current_stream()
        self._mark("analyze_start")
        self._formatted.clear()
        # the three analyzer families are independent: queue them on three streams so that the
        # HBM-bound moments pass and the issue-bound counting passes share the SMs
        pending_v = pending_q = None
        count_vocab = lambda: {vk: self._vocab_count(vk, spec, chunks) for vk, spec in self.vocab_specs.items()}
        if self.overlap_streams:
            if self._side is None:
                # high priority: their kernels (analyzers, then the short launch-bound finishing chains) are dispatched ahead of
                # the hundreds of thousands of transform tiles pending on the main stream, instead of after the last of them
                self._side = tuple(torch.cuda.Stream(device=self.device, priority=self._side_priority) for _ in range(2))
            sq, sv = self._side
            start = torch.cuda.Event()
            start.record(main)
            sv.wait_event(start)
            sq.wait_event(start)
            if not self.comm.active or rows_per_rank is None:
                # the vocabulary counts need nothing from the other ranks: they are queued before the
                # row counts are gathered, so the GPU is already busy while the host waits for those
                with torch.cuda.stream(sv):
                    pending_v = count_vocab()
        if rows_per_rank is not None:
            n_locals = rows_per_rank
        elif self.comm.active:
            nl = torch.tensor([n_local], dtype=torch.int64, device=self.device)
            n_locals = [int(v) for v in self.comm.all_gather(nl).flatten().tolist()]
        else:
            n_locals = [n_local]
        n_total = sum(n_locals)
        self._tick("row_counts")
        state = TransformState(plan=self.plan, moment_keys=list(self.moment_keys), mean_var=None, moments=None, num_rows=n_total)
        # multi-rank: the Welford triples and min/max pairs travel inside the one fixed-layout exchange after the local pass
        defer = self.comm.active and self.fixed_exchange and not self._profile
        if self.overlap_streams:
            with torch.cuda.stream(sq):
                pending_q = {nb: self._quantiles_sample(nb, keys, chunks, n_locals) for nb, keys in self.bucket_groups.items()}
            if side_work is not None:
                side_work([pq["gathered"] for pq in pending_q.values() if pq.get("gathered") is not None])
                side_work = None
            if pending_v is None:
                # Row counts came from the caller, so nothing made the host wait above.  The count pass is queued right BEHIND
                # the all-gather of the quantile samples (and before the rest of the quantile pipeline, which takes the host a
                # millisecond to enqueue): a collective kernel cannot become resident once the persistent count CTAs (plus the
                # stream kernel) hold every SM's registers, and the quantile pass would wait for the whole count pass.
                for pq in pending_q.values():
                    if pq.get("gathered") is not None:
                        sv.wait_event(pq["gathered"])
                with torch.cuda.stream(sv):
                    pending_v = count_vocab()
                    self._mark("vocab_counted")
            with torch.cuda.stream(sq):
                pending_q = {nb: self._quantiles_count(nb, pq, chunks) for nb, pq in pending_q.items()}
                self._mark("quantiles_counted")
            if self.comm.active and self.comm.world >= self.drain_between_steps_world:
                # >= 4 ranks: like the other persistent kernels the moments pass waits for the sample all-gather.  If its
                # wave of CTAs is on the SMs when the all-gather completes (a rank that waits for slower peers has queued
                # everything by then), the kernels released there are not placed at once but as SMs drain, in PRIORITY
                # order -- see the note at the stream kernel's stream.  Released with them it is placed last (normal
                # priority, no room next to the count pass), as in a step launched kernel by kernel (hypothesis, not measured).
                for pq in pending_q.values():
                    if pq.get("gathered") is not None:
                        main.wait_event(pq["gathered"])
            self._analyze_moments(chunks, state, defer)
            self._analyze_minmax(chunks, state, defer)
            self._mark("moments_done")
            # (one rank: the moments are final here, but z-score tiles queued now would crowd the quantile count pass, which
            #  is sized as one resident wave -- measured: no gain; they are queued with the bucket ops below instead)
            main.wait_stream(sv)
            main.wait_stream(sq)
            self._mark("local_pass_done")
        else:
            if side_work is not None:
                side_work([])
            self._analyze_moments(chunks, state, defer)
            self._analyze_minmax(chunks, state, defer)
            pending_q = {nb: self._quantiles_launch(nb, keys, chunks, n_locals) for nb, keys in self.bucket_groups.items()}
            pending_v = count_vocab()
        def queue_bucket_ops():
            """analyze_and_transform: the boundaries of the one- and two-pass engines are selected on the device without any
            host value, so the ops that need them are queued BEFORE the host waits (no launch gap after the sync).  The
            status words are read afterwards; a column that has to fall back to the radix select gets its ops queued again."""
            if between is None or self._profile:
                return
            for nb, pend in pending_q.items():
                if "bounds" in pend:
                    state.buckets[nb] = BucketResult(list(pend["keys"]), nb, pend["bounds"], pend["nbound"], path="pending")
            if len(state.buckets) == len(pending_q):
                between(state, "boundaries")

        # The host waits for the analyzer state (an event), NOT for the ops queued behind it: the finishing work below reads a
        # few control words and launches a dozen tiny kernels, on a side stream, while the main stream already transforms.
        fin = self._side[1] if (between is not None and self.overlap_streams and not self._profile) else main
        state_ready = torch.cuda.Event()

        def wait_for_state():
            state_ready.record(main)
            queue_bucket_ops()
            if fin is not main:
                fin.wait_event(state_ready)
            if self._host_waits_on_event:
                state_ready.synchronize()
            else:
                main.synchronize()

        if self.comm.active and self.fixed_exchange and not self._profile:
            # the fixed-layout exchange needs no value from the host: queue it behind the local pass
            # so that its launches are issued while the GPU is still busy, then wait once
            self._exchange_fixed(pending_q, pending_v, state, (lambda st: between(st, "moments")) if between is not None else None)
            wait_for_state()
        else:
            if not self.comm.active:
                wait_for_state()
            else:
                main.synchronize()
            # one host round trip: everything queued above has to finish before sizes/flags are read
            self._tick("local_pass")
            if self.comm.active:
                if self.fixed_exchange:
                    self._exchange_fixed(pending_q, pending_v, None)
                else:
                    self._exchange(pending_q, pending_v)
                self._tick("exchange")
                if fin is not main:
                    fin.wait_stream(main)
        redo_buckets = False
        with torch.cuda.stream(fin):
            for nb, pend in pending_q.items():
                state.buckets[nb] = self._quantiles_finish(nb, pend, chunks, n_total)
                redo_buckets |= state.buckets[nb].path == "radix-select" or bool(pend.get("recomputed"))
        if fin is not main:
            if redo_buckets or any("bounds" not in p for p in pending_q.values()):
                main.wait_stream(fin)  # boundaries were (re)computed on the side stream
            if between is not None:
                # first launch (general engine) or again (the optimistic launch used boundaries that were not final)
                between(state, "boundaries-again" if redo_buckets else "boundaries")
        self._tick("quantiles_finish")
        self._mark("early_ops_queued")
        with torch.cuda.stream(fin):
            for vk, pend in pending_v.items():
                state.vocabs[vk] = self._vocab_finish(vk, self.vocab_specs[vk], pend)
                for spec in self.plan.of(VocabSpec):  # decode tables depend on num_oov_buckets / default_value
                    if _vocab_key(spec) == vk:
                        self._decode_tables(vk, spec, pend, state.vocabs[vk], state)
                        break
            self._mark("vocab_ranked")
        if fin is not main:
            main.wait_stream(fin)
        self._mark("analyze_done")
        self._tick("vocab_finish")
        return state

    # -- moments
    def _analyze_moments(self, chunks, state: TransformState, defer_merge: bool = False):
        nm = len(self.moment_keys)
        if nm == 0:
            return
        s = self._stream()
        mstate = self._buf("mom_state", (nm, 3), torch.float64, zero=True)
        ws = self._buf("mom_ws", nat.query("txf_moments_workspace_bytes", nm), torch.uint8)
        for ch in chunks:
            if len(ch) == 0:
                continue
            v, m, d = self._num_args([ch[k] for k in self.moment_keys])
            nat.call("txf_moments_update", v, m, d, nm, len(ch), mstate.data_ptr(), ws.data_ptr(), s)
        state.moments = mstate
        if defer_merge:
            return
        gathered = self.comm.all_gather(mstate) if self.comm.active else None  # [W, nm, 3], rank order
        self._finish_moments(state, gathered)

    def _finish_moments(self, state: TransformState, gathered: Optional[torch.Tensor]):
        """Chan merge of the ranks' triples in rank order (bit-identical on every rank), then (mean, variance) as fp32."""
        nm = len(self.moment_keys)
        s = self._stream()
        mstate = state.moments
        if gathered is not None:
            nat.call("txf_moments_merge", gathered.data_ptr(), self.comm.world, nm, mstate.data_ptr(), s)
        mv = self._buf("mean_var", (nm, 2), torch.float32)
        nat.call("txf_moments_finalize", mstate.data_ptr(), nm, mv.data_ptr(), s)
        state.mean_var = mv

    # -- min/max (tft.scale_to_0_1 / scale_by_min_max, tft.min / tft.max)
    def _analyze_minmax(self, chunks, state: TransformState, defer_merge: bool = False):
        nm = len(self.minmax_keys)
        if nm == 0:
            return
        s = self._stream()
        mm = self._buf("minmax_state", (nm, 2), torch.float32)
        ws = self._buf("minmax_ws", nat.query("txf_minmax_workspace_bytes", nm), torch.uint8)
        nat.call("txf_minmax_begin", mm.data_ptr(), nm, s)
        for ch in chunks:
            if len(ch) == 0:
                continue
            v, m, d = self._num_args([ch[k] for k in self.minmax_keys])
            nat.call("txf_minmax_update", v, m, d, nm, len(ch), mm.data_ptr(), ws.data_ptr(), s)
        state.minmax_keys, state.minmax = list(self.minmax_keys), mm
        if self.comm.active and not defer_merge:
            self._finish_minmax(state, self.comm.all_gather(mm))  # [W, nm, 2]

    def _finish_minmax(self, state: TransformState, gathered: torch.Tensor):
        nm = len(self.minmax_keys)
        s = self._stream()
        nat.call("txf_minmax_begin", state.minmax.data_ptr(), nm, s)
        nat.call("txf_minmax_merge", gathered.data_ptr(), self.comm.world, nm, state.minmax.data_ptr(), s)

    # -- tf.as_string(int64 feature): the decimal text of every cell as a string column
    def _vocab_source(self, table: Table, spec) -> StringColumn:
        if not getattr(spec, "as_string", False):
            return table[spec.in_key]
        hit = self._formatted.get((id(table), spec.in_key))
        if hit is not None and hit[0] is table:
            return hit[1]
        col: NumericColumn = table[spec.in_key]
        n, s = len(table), self._stream()
        offsets = torch.zeros(n + 1, dtype=torch.int32, device=self.device)
        total = 0
        if n:
            lens = torch.empty(n, dtype=torch.int32, device=self.device)
            nat.call("txf_i64_format_lengths", col.values.data_ptr(), _ptr(col.valid), n, lens.data_ptr(), s)
            ends = torch.cumsum(lens, 0, dtype=torch.int64)
            total = int(ends[-1].item())
            if total >= 1 << 31:
                raise nat.NativeError(f"tf.as_string({spec.in_key!r}): more than 2 GiB of text in one chunk; split the table")
            offsets[1:] = ends.to(torch.int32)
        pool = torch.zeros(total + nat.TXF_POOL_SLACK, dtype=torch.uint8, device=self.device)
        if n:
            nat.call("txf_i64_format_write", col.values.data_ptr(), _ptr(col.valid), n, offsets.data_ptr(), pool.data_ptr(), s)
        out = StringColumn(offsets, pool, None, total)
        if len(self._formatted) >= 16:
            self._formatted.clear()
        self._formatted[(id(table), spec.in_key)] = (table, out)
        return out

    # -- exact multi-rank selection driver (general engine)
    def _msel(self, name: str, ncols: int, R: int, ranks_dev: torch.Tensor, passes_fn, reduce: bool) -> torch.Tensor:
        """ranks_dev: int64 [ncols, R] (device).  passes_fn(state_ptr, p) queues the histogram
        launches of pass p.  Returns uint32-as-int32 keys [ncols, R]."""
        s = self._stream()
        st = self._buf(name + "_msel", nat.query("txf_msel_state_bytes", ncols, R), torch.uint8)
        nat.call("txf_msel_begin", st.data_ptr(), ncols, R, ranks_dev.data_ptr(), s)
        hist = st[: nat.query("txf_msel_hist_bytes", ncols, R)].view(torch.int64)
        for p in range(4):
            passes_fn(st.data_ptr(), p)
            if reduce:
                self.comm.all_reduce_sum_(hist)
            nat.call("txf_msel_resolve", st.data_ptr(), ncols, R, p, s)
        keys = self._buf(name + "_keys", (ncols, R), torch.int32)
        nat.call("txf_msel_result", st.data_ptr(), ncols, R, keys.data_ptr(), s)
        return keys

    def _target_ranks(self, n_total: int, nb: int) -> List[int]:
        return [min(max((i * n_total + nb - 1) // nb, 1), max(n_total, 1)) for i in range(1, nb)]

    # -- quantiles
    def _quantiles_launch(self, nb: int, keys: List[str], chunks, n_locals: List[int]):
        """Queue the sample-select pipeline for one num_buckets group; returns what _finish needs."""
        return self._quantiles_count(nb, self._quantiles_sample(nb, keys, chunks, n_locals), chunks)

    def _quantiles_sample(self, nb: int, keys: List[str], chunks, n_locals: List[int]):
        """First half of the pipeline: the stratified sample of this rank's rows and, multi-rank, its
        all-gather (pend["gathered"]: event after that collective)."""
        T = nb - 1
        n_total = sum(n_locals)
        n_local = n_locals[self.comm.rank]
        ncols = len(keys)
        pend = {"keys": keys, "T": T, "fast": False}
        if T < 1 or n_total == 0:
            return pend
        M_total = min(self.sample_size, n_total // 8)
        if self.force_general_quantiles or n_total < (1 << 18) or M_total < (1 << 12) or T > 1024:
            return pend
        # per-rank sample sizes, proportional to the rows each rank holds (host arithmetic, no sync)
        M_ranks = [(M_total * nl) // n_total for nl in n_locals]
        M_sum = sum(M_ranks)
        ranks = self._target_ranks(n_total, nb)
        edges_ranks = []
        for r in ranks:
            q = r / n_total
            k = min(max(math.ceil(q * M_sum), 1), M_sum)
            d = int(math.ceil(self.bracket_sigmas * math.sqrt(M_sum * q * (1.0 - q)))) + 2
            edges_ranks += [max(1, k - d), min(M_sum, k + d)]
        # one pass (brackets around every target) while the brackets are few and disjoint; otherwise two passes over
        # equi-depth splitters (with T targets the brackets cover about 10 T sigma / M of the rows: everything for T in the hundreds)
        pend["mode"] = "bracket"
        if T > nat.TXF_QMAX_TARGETS or M_total < 1024 * T or any(b < a for a, b in zip(edges_ranks, edges_ranks[1:])) or self.force_split_quantiles:
            if not self.split_quantiles:
                return pend
            pend["mode"] = "split"
        s = self._stream()
        # (1) sample, chunk by chunk
        M_local = M_ranks[self.comm.rank]
        parts = []
        left = M_local
        for i, ch in enumerate(chunks):
            m_c = left if i == len(chunks) - 1 else (M_local * len(ch)) // max(n_local, 1)
            m_c = min(m_c, len(ch))
            left -= m_c
            if m_c <= 0:
                continue
            buf = self._buf(f"q{nb}_sample{i}", (ncols, m_c), torch.int32)
            v, m, d = self._num_args([ch[k] for k in keys])
            nat.call("txf_qsample", v, m, d, ncols, len(ch), m_c, self.seed + 7919 * i + 104729 * self.comm.rank, buf.data_ptr(), s)
            parts.append(buf)
        if parts:
            sample = parts[0] if len(parts) == 1 else torch.cat(parts, dim=1)
        else:
            sample = torch.empty((ncols, 0), dtype=torch.int32, device=self.device)
        M_have = sample.shape[1]
        if self.comm.active:
            Mmax = max(max(M_ranks), 1)
            pad = self._buf(f"q{nb}_sample_pad", (ncols, Mmax), torch.int32, zero=True)
            pad[:, :M_have] = sample
            g = self.comm.all_gather(pad)  # [W, ncols, Mmax]
            pend["gathered"] = torch.cuda.Event()
            pend["gathered"].record(torch.cuda.current_stream())
            self._mark("sample_gathered")
            # ranks whose chunks could not supply their share contribute what they have; the
            # order statistics below are taken over the slots every rank agreed on (host arithmetic)
            sample = torch.cat([g[r, :, : M_ranks[r]] for r in range(self.comm.world)], dim=1).contiguous()
            if M_have != M_local:
                raise RuntimeError("sample size mismatch between chunks and plan")
        pend["_ctx"] = (sample.contiguous(), edges_ranks, ranks, n_local)
        return pend

    def _quantiles_count(self, nb: int, pend, chunks):
        """Second half: bracket edges from the sample, then the one counting pass over the columns."""
        if "_ctx" not in pend:
            return pend
        sample, edges_ranks, ranks, n_local = pend.pop("_ctx")
        keys, T = pend["keys"], pend["T"]
        ncols = len(keys)
        s = self._stream()
        M_sum = sample.shape[1]
        if pend["mode"] == "split":
            return self._quantiles_split(nb, pend, chunks, sample, ranks, n_local)
        # (2) bracket edges = exact order statistics of the sample
        er = self._host_const([edges_ranks] * ncols, torch.int64)
        sptrs = nat.ptr_array([sample[c].data_ptr() for c in range(ncols)])

        def sample_pass(st, p):
            nat.call("txf_msel_hist", st, ncols, 2 * T, p, sptrs, None, None, 1, M_sum, None, s)

        edges = self._msel(f"q{nb}_edges", ncols, 2 * T, er, sample_pass, reduce=False)
        # (3) the one streaming pass over the column
        cells = self._buf(f"q{nb}_cells", (ncols, 4 * T + 1), torch.int64, zero=True)
        # keys strictly inside a bracket are kept: expect (bracket width / sample size) of the rows
        frac = sum(hi - lo + 2 for lo, hi in zip(edges_ranks[0::2], edges_ranks[1::2])) / max(M_sum, 1)
        cand_cap = max(1 << 16, min(int(1.5 * frac * n_local) + 4096, max(n_local, 1)))
        cand = self._buf(f"q{nb}_cand", (ncols, cand_cap), torch.int32)
        cand_count = self._buf(f"q{nb}_cand_count", (ncols,), torch.int64, zero=True)
        for ch in chunks:
            if len(ch) == 0:
                continue
            v, m, d = self._num_args([ch[k] for k in keys])
            nat.call("txf_qcount", v, m, d, ncols, len(ch), T, edges.data_ptr(), cells.data_ptr(), cand.data_ptr(),
                     cand_count.data_ptr(), cand_cap, s)
        pend.update(fast=True, edges=edges, cells=cells, cand=cand, cand_count=cand_count, cand_cap=cand_cap, ranks=ranks)
        if not self.comm.active:
            self._quantiles_resolve(nb, pend)
        return pend

    def _quantiles_split(self, nb: int, pend, chunks, sample: torch.Tensor, ranks: List[int], n_local: int):
        """Two passes for many targets (txf_qsplit_*): S equi-depth splitters from the sample, exact cell counts, then
        only the keys of the <= T target cells are collected for the final selection.  Multi-rank: the cell counts are
        all-reduced and the selection all-reduces its histograms; the collected keys stay where they are."""
        keys, T = pend["keys"], pend["T"]
        ncols = len(keys)
        s = self._stream()
        M_sum = sample.shape[1]
        S = 4096
        while 8 * S > M_sum:  # the splitter kernel sorts a subsample of 8 S keys per column in shared memory
            S >>= 1
        edges = self._buf(f"q{nb}_sp_edges", (ncols, S), torch.int32)
        nat.call("txf_qsplit_edges", sample.data_ptr(), ncols, M_sum, S, edges.data_ptr(), s)
        cells = self._buf(f"q{nb}_sp_cells", (ncols, 2 * S + 1), torch.int64, zero=True)
        args = [self._num_args([ch[k] for k in keys]) + (len(ch),) for ch in chunks if len(ch)]
        for v, m, d, n in args:
            nat.call("txf_qsplit_count", v, m, d, ncols, n, S, edges.data_ptr(), cells.data_ptr(), s)
        if self.comm.active:
            self.comm.all_reduce_sum_(cells)
        ranks_dev = self._host_const(ranks, torch.int64)
        direct = self._buf(f"q{nb}_direct", (ncols, T), torch.int32)
        cand_rank = self._buf(f"q{nb}_cand_rank", (ncols, T), torch.int64)
        flags = self._buf(f"q{nb}_sp_flags", (ncols, nat.query("txf_qsplit_flag_words", S)), torch.int32)
        cand_total = self._buf(f"q{nb}_sp_total", (ncols,), torch.int64)
        scratch = self._buf(f"q{nb}_sp_scratch", nat.query("txf_qsplit_scratch_bytes", ncols, S), torch.uint8)
        nat.call("txf_qsplit_resolve", cells.data_ptr(), edges.data_ptr(), ranks_dev.data_ptr(), ncols, S, T, direct.data_ptr(),
                 cand_rank.data_ptr(), flags.data_ptr(), cand_total.data_ptr(), scratch.data_ptr(), s)
        # a target cell holds about n / (S + 1) keys (heavy duplicates sit ON splitters and are answered directly)
        cand_cap = max(1 << 16, min(int(3.0 * min(T, S + 1) / (S + 1) * n_local) + 65536, max(n_local, 1)))
        cand = self._buf(f"q{nb}_cand", (ncols, cand_cap), torch.int32)
        cand_count = self._buf(f"q{nb}_cand_count", (ncols,), torch.int64, zero=True)
        for v, m, d, n in args:
            nat.call("txf_qsplit_collect", v, m, d, ncols, n, S, edges.data_ptr(), flags.data_ptr(), cand.data_ptr(),
                     cand_count.data_ptr(), cand_cap, s)
        status = (cand_count > cand_cap).to(torch.int32)  # overflow of the collected-key buffer: the general engine takes over
        if self.comm.active:
            self.comm.all_reduce_max_(status)
        cptrs = nat.ptr_array([cand[c].data_ptr() for c in range(ncols)])

        def cand_pass(st, p):
            nat.call("txf_msel_hist", st, ncols, T, p, cptrs, None, None, 1, cand.shape[1], cand_count.data_ptr(), s)

        sel = self._msel(f"q{nb}_sel", ncols, T, cand_rank, cand_pass, reduce=self.comm.active)
        bounds = self._buf(f"q{nb}_bounds", (ncols, T), torch.float32)
        nbound = self._buf(f"q{nb}_nbound", (ncols,), torch.int32)
        nat.call("txf_qfinal", direct.data_ptr(), cand_rank.data_ptr(), sel.data_ptr(), ncols, T, bounds.data_ptr(), nbound.data_ptr(), s)
        pend.update(split_done=True, status=status, bounds=bounds, nbound=nbound)
        return pend

    def _quantiles_resolve(self, nb: int, pend, global_count: Optional[torch.Tensor] = None):
        """cells + candidates -> boundaries; all on the stream.  Multi-rank: `cells` are already
        summed over ranks and `global_count` holds the total number of candidates per column; the
        candidates themselves stay where they are -- the selection all-reduces its histograms."""
        keys, T = pend["keys"], pend["T"]
        ncols = len(keys)
        s = self._stream()
        ranks_dev = self._host_const(pend["ranks"], torch.int64)
        direct = self._buf(f"q{nb}_direct", (ncols, T), torch.int32)
        cand_rank = self._buf(f"q{nb}_cand_rank", (ncols, T), torch.int64)
        status = self._buf(f"q{nb}_status", (ncols,), torch.int32)
        cand, cand_count, cand_cap = pend["cand"], pend["cand_count"], pend["cand_cap"]
        if global_count is None:
            nat.call("txf_qresolve", pend["cells"].data_ptr(), pend["edges"].data_ptr(), cand_count.data_ptr(), cand_cap,
                     ranks_dev.data_ptr(), ncols, T, direct.data_ptr(), cand_rank.data_ptr(), status.data_ptr(), s)
        else:  # overflow of any rank's buffer is encoded as a count above _NO_CAP
            nat.call("txf_qresolve", pend["cells"].data_ptr(), pend["edges"].data_ptr(), global_count.data_ptr(), _NO_CAP,
                     ranks_dev.data_ptr(), ncols, T, direct.data_ptr(), cand_rank.data_ptr(), status.data_ptr(), s)
        pend.update(status=status, direct=direct, cand_rank=cand_rank, reduce=global_count is not None)
        self._quantiles_select(nb, pend, use_bracket_hist=self.bracket_select)

    def _quantiles_select(self, nb: int, pend, use_bracket_hist: bool):
        """The exact order statistic inside every bracket, among the candidates: one histogram pass over the
        bracket's key range (txf_bsel_*; multi-rank: ONE all-reduce), or -- brackets wider than the histogram,
        flagged by status bit 4 -- the four-digit radix select (four all-reduces)."""
        keys, T = pend["keys"], pend["T"]
        ncols = len(keys)
        s = self._stream()
        cand, cand_count, cand_rank, status = pend["cand"], pend["cand_count"], pend["cand_rank"], pend["status"]
        if use_bracket_hist:
            B = 16384
            while B > 256 and 4 * ncols * T * B > (8 << 20):  # keep the histogram (and its all-reduce) under 8 MB
                B >>= 1
            hist = self._buf(f"q{nb}_bsel", nat.query("txf_bsel_hist_bytes", ncols, T, B) // 4, torch.int32)
            nat.call("txf_bsel_hist", cand.data_ptr(), cand.shape[1], cand_count.data_ptr(), pend["edges"].data_ptr(), ncols, T, B,
                     hist.data_ptr(), s)
            if pend["reduce"]:
                self.comm.all_reduce_sum_(hist)
            sel = self._buf(f"q{nb}_sel_keys", (ncols, T), torch.int32)
            nat.call("txf_bsel_resolve", hist.data_ptr(), pend["cells"].data_ptr(), pend["edges"].data_ptr(), cand_rank.data_ptr(), ncols, T, B,
                     sel.data_ptr(), status.data_ptr(), s)
        else:
            cptrs = nat.ptr_array([cand[c].data_ptr() for c in range(ncols)])

            def cand_pass(st, p):
                nat.call("txf_msel_hist", st, ncols, T, p, cptrs, None, None, 1, cand.shape[1], cand_count.data_ptr(), s)

            sel = self._msel(f"q{nb}_sel", ncols, T, cand_rank, cand_pass, reduce=pend["reduce"])
        bounds = self._buf(f"q{nb}_bounds", (ncols, T), torch.float32)
        nbound = self._buf(f"q{nb}_nbound", (ncols,), torch.int32)
        nat.call("txf_qfinal", pend["direct"].data_ptr(), cand_rank.data_ptr(), sel.data_ptr(), ncols, T, bounds.data_ptr(),
                 nbound.data_ptr(), s)
        pend.update(bounds=bounds, nbound=nbound)

    @staticmethod
    def _global_cand_count(counts_all: torch.Tensor, caps_all: torch.Tensor) -> torch.Tensor:
        """counts_all [W, ncols], caps_all [W, 1] (device) -> per-column total, or _NO_CAP + 1 where some
        rank overflowed its candidate buffer (txf_qresolve then flags the column)."""
        over = (counts_all > caps_all).any(dim=0)
        return torch.where(over, torch.full_like(counts_all[0], _NO_CAP + 1), counts_all.sum(dim=0)).contiguous()

    def _exchange_fixed(self, pending_q, pending_v, state: Optional[TransformState], after_moments=None):
        """The exchange step as one collective: every rank packs its vocabulary entries, candidate
        keys and cell counts into a fixed-layout int64 buffer on the device, the buffers are
        all-gathered, and each rank rebuilds the merged state from all rows -- no size ever goes
        through the host.  Overflow of a section is flagged on the device and handled later
        (vocabulary: sized exchange; quantiles: the general radix select)."""
        dev, W, s = self.device, self.comm.world, self._stream()
        K, S = self.xchg_terms, self.xchg_term_bytes
        fast = [(nb, p) for nb, p in pending_q.items() if p["fast"]]
        voc = list(pending_v.items())
        off, layout = 0, {}
        for vk, p in voc:
            layout[("v", vk)] = off
            off += 1 + 4 * K + K * S // 8
        for nb, p in fast:
            ncols, ncell = len(p["keys"]), 4 * p["T"] + 1
            layout[("q", nb)] = (off, off + ncols, off + ncols + 1)  # candidate counts, buffer capacity, cells
            off += ncols + 1 + ncols * ncell
        nmom = len(self.moment_keys) if state is not None and state.moments is not None else 0
        nmm = len(self.minmax_keys) if state is not None and state.minmax is not None else 0
        if nmom:
            layout["mom"] = off
            off += 3 * nmom           # fp64 triples as their bit patterns
        if nmm:
            layout["mm"] = off
            off += nmm                # (min, max) fp32 pairs as their bit patterns
        if off == 0:
            return
        blob = self._buf("xchg_blob", off, torch.int64)
        if nmom:
            blob[layout["mom"]: layout["mom"] + 3 * nmom].copy_(state.moments.view(torch.int64).flatten())
        if nmm:
            blob[layout["mm"]: layout["mm"] + nmm].copy_(state.minmax.view(torch.int64).flatten())
        for vk, p in voc:
            o = layout[("v", vk)]
            nat.call("txf_vocab_export_blob", p["table"].data_ptr(), p["arena"].data_ptr(), p["ent"].data_ptr(), p["ctrl"].data_ptr(),
                     K, S, blob[o:].data_ptr(), s)
        for nb, p in fast:
            ncols = len(p["keys"])
            oc, ok, ox = layout[("q", nb)]
            blob[oc:oc + ncols].copy_(p["cand_count"])
            blob[ok:ox].fill_(p["cand_cap"])
            blob[ox:ox + p["cells"].numel()].copy_(p["cells"].flatten())
        g = self.comm.all_gather(blob)  # [W, off] -- the one collective of the exchange
        self._mark("blob_gathered")
        if nmom:
            o = layout["mom"]
            self._finish_moments(state, g[:, o:o + 3 * nmom].contiguous().view(torch.float64))
        if nmm:
            o = layout["mm"]
            self._finish_minmax(state, g[:, o:o + nmm].contiguous().view(torch.float32))
        # The boundary selection is a chain of four small dependent all-reduces (latency, not bytes): it runs on the
        # quantile stream while this stream merges the vocabularies and -- analyze_and_transform -- already z-scores.
        main = torch.cuda.current_stream()
        qs = self._side[0] if (self.overlap_streams and self._side is not None and fast) else main
        if qs is not main:
            qs.wait_stream(main)
        with torch.cuda.stream(qs):
            for nb, p in fast:
                ncols = len(p["keys"])
                oc, ok, ox = layout[("q", nb)]
                p["cells"].copy_(g[:, ox:ox + p["cells"].numel()].sum(dim=0).view_as(p["cells"]))
                self._quantiles_resolve(nb, p, self._global_cand_count(g[:, oc:oc + ncols], g[:, ok:ox]))
            self._mark("boundaries_selected")
        if after_moments is not None and state is not None:
            after_moments(state)
        for vk, p in voc:
            cap2 = max(self.vocab_capacity, _pow2_at_least(2 * W * K))
            tag2 = p["tag"] + "_merged"
            table, arena, ctrl, ent, counts_buf = self._vocab_bufs(tag2, cap2)
            nat.call("txf_vocab_insert_blob", g.data_ptr(), off, layout[("v", vk)], W, K, S, table.data_ptr(), cap2, arena.data_ptr(),
                     self.arena_bytes, ctrl.data_ptr(), s)
            nat.call("txf_vocab_compact", table.data_ptr(), cap2, arena.data_ptr(), ent.data_ptr(), counts_buf.data_ptr(),
                     ctrl.data_ptr(), s)
            p["merged"] = dict(tag=tag2, cap=cap2, table=table, arena=arena, ctrl=ctrl, ent=ent, counts=counts_buf)
        self._mark("vocab_merged")
        if qs is not main:
            main.wait_stream(qs)
        self._mark("exchange_done")

    def _exchange(self, pending_q, pending_v):
        """Sized variant of the exchange (used when a section of the fixed layout overflows, or with
        fixed_exchange = False): quantile cell counts are all-reduced and candidate counts
        all-gathered (the candidate keys never move: the final selection all-reduces its
        histograms); vocabulary entries travel as one packed all-gather per vocabulary after their
        sizes went through the host."""
        dev, W = self.device, self.comm.world
        s = self._stream()
        fast = [(nb, p) for nb, p in pending_q.items() if p["fast"]]
        voc = list(pending_v.items())
        # vocabulary entries of this rank, serialised on the device (sizes known after the sync above)
        for vk, p in voc:
            kept = int(p["ctrl"][3].item())
            p["kept_local"] = kept
            p["export"] = self._export_entries(p, p["ent"], kept)
        for nb, p in fast:  # quantiles: sum the cells, total the candidate counts; candidates stay local
            ncols = len(p["keys"])
            self.comm.all_reduce_sum_(p["cells"])
            cc = self.comm.all_gather(torch.cat([p["cand_count"], torch.tensor([p["cand_cap"]], dtype=torch.int64, device=dev)]))
            self._quantiles_resolve(nb, p, self._global_cand_count(cc[:, :ncols], cc[:, ncols:]))
        sizes = []
        for vk, p in voc:
            sizes.append(torch.tensor([p["kept_local"]], dtype=torch.int64, device=dev))
            sizes.append(p["export"][3][-1:].clone())  # bytes of the exported terms
            sizes.append(p["ctrl"][2:3].clone())       # this rank's overflow flags: every rank must take the same decision
        if not sizes:
            return
        self._tick("x_export")
        all_sizes = self.comm.all_gather(torch.cat(sizes)).cpu()  # [W, K] -- one more host sync
        self._tick("x_sizes")
        k = 0
        for vk, p in voc:
            kepts = all_sizes[:, k].tolist()
            nbytes = all_sizes[:, k + 1].tolist()
            p["peer_flags"] = 0
            for fl in all_sizes[:, k + 2].tolist():
                p["peer_flags"] |= int(fl) & 0x3F
            k += 3
            fp, tg, cnt, offsets, pool = p["export"]
            kmax, bmax = max(max(kepts), 1), (max(max(nbytes), 1) + 7) // 8 * 8  # rows of the gathered matrix stay 8-byte aligned
            lens = torch.zeros(kmax, dtype=torch.int64, device=dev)
            kl = p["kept_local"]
            if kl:
                lens[:kl] = offsets[1:kl + 1] - offsets[:kl]
            pad8 = lambda t: torch.cat([t[:kl], torch.zeros(kmax - kl, dtype=torch.int64, device=dev)]).view(torch.uint8)
            packed = torch.cat([pad8(fp), pad8(tg), pad8(cnt), lens.view(torch.uint8),
                                pool[: nbytes[self.comm.rank]], torch.zeros(bmax - nbytes[self.comm.rank], dtype=torch.uint8, device=dev)])
            g = self.comm.all_gather(packed)  # [W, 32*kmax + bmax]
            col = lambda r, j: g[r, 8 * kmax * j: 8 * kmax * (j + 1)].view(torch.int64)[: kepts[r]]
            g_fp = torch.cat([col(r, 0) for r in range(W)])
            g_tg = torch.cat([col(r, 1) for r in range(W)])
            g_cnt = torch.cat([col(r, 2) for r in range(W)])
            g_len = torch.cat([col(r, 3) for r in range(W)])
            g_pool = torch.cat([g[r, 32 * kmax: 32 * kmax + nbytes[r]] for r in range(W)] +
                               [torch.zeros(nat.TXF_POOL_SLACK, dtype=torch.uint8, device=dev)])
            m = int(g_fp.numel())
            g_off = torch.zeros(m + 1, dtype=torch.int64, device=dev)
            if m:
                torch.cumsum(g_len, 0, out=g_off[1:])
            spec = self.vocab_specs[vk]
            if self.partition_vocab and spec.top_k is not None and not spec.frequency_threshold and W > 1:
                if self._exchange_owned(vk, spec, p, (g_fp, g_tg, g_cnt, g_off, g_pool, m)):
                    continue
            cap2 = max(self.vocab_capacity, _pow2_at_least(2 * max(m, 1)))
            tag2 = p["tag"] + "_merged"
            table, arena, ctrl, ent, counts_buf = self._vocab_bufs(tag2, cap2)
            nat.call("txf_vocab_insert", g_fp.data_ptr(), g_tg.data_ptr(), g_cnt.data_ptr(), g_off.data_ptr(), g_pool.data_ptr(), m,
                     1, 0, table.data_ptr(), cap2, arena.data_ptr(), self.arena_bytes, ctrl.data_ptr(), s)
            nat.call("txf_vocab_compact", table.data_ptr(), cap2, arena.data_ptr(), ent.data_ptr(), counts_buf.data_ptr(),
                     ctrl.data_ptr(), s)
            p["merged"] = dict(tag=tag2, cap=cap2, table=table, arena=arena, ctrl=ctrl, ent=ent, counts=counts_buf)

    def _exchange_owned(self, vk, spec: VocabSpec, p, gathered) -> bool:
        """Owner-partitioned merge of a large vocabulary: of the entries of all ranks each
        rank re-counts only the terms it OWNS (identity hash mod world), which gives exact global counts at
        1/world of the insert work; the owners' top_k lists (a superset of the global top_k) are then
        all-gathered in the fixed layout and merged everywhere.  Returns False (nothing changed) when an
        owner's list does not fit the fixed layout (a term longer than the byte slot): the caller then
        re-counts everything everywhere."""
        dev, W, s = self.device, self.comm.world, self._stream()
        g_fp, g_tg, g_cnt, g_off, g_pool, m = gathered
        capo = max(self.vocab_capacity, _pow2_at_least(2 * (m // W + 1024)))
        tago = p["tag"] + "_owned"
        table, arena, ctrl, ent, counts_buf = self._vocab_bufs(tago, capo)
        nat.call("txf_vocab_insert", g_fp.data_ptr(), g_tg.data_ptr(), g_cnt.data_ptr(), g_off.data_ptr(), g_pool.data_ptr(), m,
                 W, self.comm.rank, table.data_ptr(), capo, arena.data_ptr(), self.arena_bytes, ctrl.data_ptr(), s)
        nat.call("txf_vocab_compact", table.data_ptr(), capo, arena.data_ptr(), ent.data_ptr(), counts_buf.data_ptr(), ctrl.data_ptr(), s)
        owned = dict(tag=tago, cap=capo, table=table, arena=arena, ctrl=ctrl, ent=ent, counts=counts_buf)
        ctrl_host = ctrl.cpu().tolist()
        sel, k = self._select_top(owned, spec, int(ctrl_host[3]))
        # the owner's list in the fixed layout: K = top_k slots; the header carries this rank's flags and k
        K, S = int(spec.top_k), self.xchg_term_bytes
        words = 1 + 4 * K + K * S // 8
        sec = self._buf(f"{tago}_sec", words, torch.int64)
        hdr = ctrl.clone()
        hdr[3] = k
        nat.call("txf_vocab_export_blob", table.data_ptr(), arena.data_ptr(), sel.data_ptr(), hdr.data_ptr(), K, S, sec.data_ptr(), s)
        g = self.comm.all_gather(sec)  # [W, words]
        capf = max(1 << 12, _pow2_at_least(2 * W * K))
        tagf = p["tag"] + "_merged"
        tablef, arenaf, ctrlf, entf, countsf = self._vocab_bufs(tagf, capf)
        nat.call("txf_vocab_insert_blob", g.data_ptr(), words, 0, W, K, S, tablef.data_ptr(), capf, arenaf.data_ptr(), self.arena_bytes,
                 ctrlf.data_ptr(), s)
        nat.call("txf_vocab_compact", tablef.data_ptr(), capf, arenaf.data_ptr(), entf.data_ptr(), countsf.data_ptr(), ctrlf.data_ptr(), s)
        if int(ctrlf[2].item()) & 4:  # a term longer than the byte slot (the same verdict on every rank)
            return False
        p["merged"] = dict(tag=tagf, cap=capf, table=tablef, arena=arenaf, ctrl=ctrlf, ent=entf, counts=countsf)
        self.last_vocab_exchange[vk[0]] = "owner-partitioned"
        return True

    def _quantiles_general(self, nb: int, keys: List[str], chunks, n_total: int) -> BucketResult:
        """Exact radix select over the whole column (any number of boundaries)."""
        T = nb - 1
        ncols = len(keys)
        s = self._stream()
        ranks = self._target_ranks(n_total, nb)
        rd = self._host_const([ranks] * ncols, torch.int64)
        args = [self._num_args([ch[k] for k in keys]) + (len(ch),) for ch in chunks if len(ch)]

        def data_pass(st, p):
            for v, m, d, n in args:
                nat.call("txf_msel_hist", st, ncols, T, p, v, m, d, 0, n, None, s)

        sel = self._msel(f"q{nb}_gen", ncols, T, rd, data_pass, reduce=self.comm.active)
        bounds = self._buf(f"q{nb}_bounds", (ncols, T), torch.float32)
        nbound = self._buf(f"q{nb}_nbound", (ncols,), torch.int32)
        zeros = self._buf(f"q{nb}_direct", (ncols, T), torch.int32, zero=True)
        nat.call("txf_qfinal", zeros.data_ptr(), rd.data_ptr(), sel.data_ptr(), ncols, T, bounds.data_ptr(), nbound.data_ptr(), s)
        return BucketResult(list(keys), nb, bounds, nbound, path="radix-select")

    def _quantiles_finish(self, nb: int, pend, chunks, n_total: int) -> BucketResult:
        keys, T = pend["keys"], pend["T"]
        ncols = len(keys)
        if T < 1 or n_total == 0:
            self.last_quantile_path[nb] = "empty"
            return BucketResult(list(keys), nb, torch.zeros((ncols, max(T, 1)), dtype=torch.float32, device=self.device),
                                torch.zeros((ncols,), dtype=torch.int32, device=self.device), path="empty")
        if pend["fast"] or pend.get("split_done"):
            status = pend["status"].cpu()
            if pend["fast"] and int(status.max()) == 4 and int((status & ~4).abs().sum()) == 0:
                # a bracket spans more keys than the one-pass histogram holds (the same verdict on every rank):
                # select among the candidates by radix digits instead
                pend["status"] &= ~4
                pend["recomputed"] = True
                self._quantiles_select(nb, pend, use_bracket_hist=False)
                self.last_quantile_select[nb] = "radix digits"
                status = pend["status"].cpu()
            elif pend["fast"]:
                self.last_quantile_select[nb] = "bracket histogram" if self.bracket_select else "radix digits"
            if int(status.abs().sum()) == 0:
                path = "sample-select" if pend["fast"] else "splitter-select"
                self.last_quantile_path[nb] = path
                return BucketResult(list(keys), nb, pend["bounds"], pend["nbound"], path=path)
        self.last_quantile_path[nb] = "radix-select"
        return self._quantiles_general(nb, keys, chunks, n_total)

    # -- vocabulary
    def _vocab_bufs(self, tag: str, capacity: int):
        table = self._buf(f"{tag}_table", nat.query("txf_vocab_table_bytes", capacity), torch.uint8, zero=True)
        arena = self._buf(f"{tag}_arena", self.arena_bytes + nat.TXF_POOL_SLACK, torch.uint8)  # slack: 8-byte window reads past a term
        ctrl = self._buf(f"{tag}_ctrl", 8, torch.int64, zero=True)
        ent = self._buf(f"{tag}_ent", capacity, torch.int32)
        counts = self._buf(f"{tag}_counts", capacity, torch.int32)
        return table, arena, ctrl, ent, counts

    def _vocab_count(self, vk: Tuple, spec: VocabSpec, chunks):
        s = self._stream()
        tag = f"v_{vk[0]}_{spec.top_k}"
        cap = self.vocab_capacity
        table, arena, ctrl, ent, counts = self._vocab_bufs(tag, cap)
        coded = []
        nlocal = nat.query("txf_vocab_local_codes")
        for i, ch in enumerate(chunks):
            if len(ch) == 0:
                continue
            col: StringColumn = self._vocab_source(ch, spec)
            codes = self._buf(f"{tag}_codes{i}", len(ch), torch.int32)
            l2g = self._buf(f"{tag}_l2g{i}", nlocal, torch.int32)
            nat.call("txf_vocab_count", col.offsets.data_ptr(), col.pool.data_ptr(), _ptr(col.valid), len(ch), table.data_ptr(),
                     cap, arena.data_ptr(), self.arena_bytes, ctrl.data_ptr(), codes.data_ptr(), l2g.data_ptr(), self._count_warps, s)
            coded.append((i, ch, col, codes, l2g))
        nat.call("txf_vocab_compact", table.data_ptr(), cap, arena.data_ptr(), ent.data_ptr(), counts.data_ptr(), ctrl.data_ptr(), s)
        return dict(tag=tag, cap=cap, table=table, arena=arena, ctrl=ctrl, ent=ent, counts=counts, coded=coded)

    def _check_vocab_flags(self, flags: int, slots_used: Optional[int], spec: VocabSpec):
        """flags: OR of ctrl[2] of every table involved (the same value on every rank); slots_used: of this
        rank's table when it alone decides (single rank).  Grow-and-retry, or an error nothing can fix."""
        key = spec.in_key
        if flags & 2:
            raise _VocabGrow(key, "table", 8)
        if flags & 1:
            raise _VocabGrow(key, "arena", 8)
        if slots_used is not None and slots_used * 4 > self.vocab_capacity * 3:
            raise _VocabGrow(key, "table", 2)  # over 75 % full: probe chains get long
        if flags & 16:
            raise nat.NativeError(f"vocabulary of {key!r}: a term occurs more than 2^32-1 times; the top-k cut-off is selected on 32-bit counts")
        if flags & 32:
            raise nat.NativeError(f"vocabulary of {key!r}: a claimed table slot was never published (internal error)")

    def _export_entries(self, pend, ent: torch.Tensor, m: int):
        """(fp, tag, count, offsets int64 [m+1], pool) of the m listed entries, on the device."""
        s = self._stream()
        tag = pend["tag"]
        fp = self._buf(f"{tag}_x_fp", max(m, 1), torch.int64)
        tg = self._buf(f"{tag}_x_tag", max(m, 1), torch.int64)
        cnt = self._buf(f"{tag}_x_cnt", max(m, 1), torch.int64)
        ln = self._buf(f"{tag}_x_len", max(m, 1), torch.int32)
        nat.call("txf_vocab_export_meta", pend["table"].data_ptr(), ent.data_ptr(), m, fp.data_ptr(), tg.data_ptr(), cnt.data_ptr(),
                 ln.data_ptr(), s)
        offsets = torch.zeros(m + 1, dtype=torch.int64, device=self.device)
        if m:
            torch.cumsum(ln[:m].to(torch.int64), 0, out=offsets[1:])
        pool = self._buf(f"{tag}_x_pool", self.arena_bytes + nat.TXF_POOL_SLACK, torch.uint8)
        nat.call("txf_vocab_export_bytes", pend["table"].data_ptr(), pend["arena"].data_ptr(), ent.data_ptr(), m, offsets.data_ptr(),
                 pool.data_ptr(), s)
        return fp, tg, cnt, offsets, pool

    def _decode_tables(self, vk: Tuple, spec: VocabSpec, local, res: VocabResult, state: TransformState):
        """Per analyzed chunk: code -> final id table, from this rank's OWN count table and the final lookup."""
        s = self._stream()
        nlocal = nat.query("txf_vocab_local_codes")
        for i, ch, col, codes, l2g in local["coded"]:
            final = self._buf(f"{local['tag']}_final{i}", nlocal + local["cap"], torch.int64)
            nat.call("txf_vocab_final_codes", local["table"].data_ptr(), local["cap"], local["arena"].data_ptr(), l2g.data_ptr(),
                     res.lookup.data_ptr(), res.lookup_cap, res.lookup_pool.data_ptr(), res.size, spec.num_oov_buckets,
                     spec.default_value, final.data_ptr(), s)
            state.codes[(vk, i)] = CodedChunk(weakref.ref(ch), col.offsets.data_ptr(), col.pool.data_ptr(), _ptr(col.valid), len(ch),
                                              codes, final)

    def _select_top(self, pend, spec: VocabSpec, kept: int):
        """The first min(top_k, kept) entries of a compacted count table by (count, bytes) descending
        (after the frequency threshold): -> (entry list on the device, sorted; k)."""
        s = self._stream()
        tag, table, arena, ctrl, ent, counts = (pend[k] for k in ("tag", "table", "arena", "ctrl", "ent", "counts"))
        k = kept if spec.top_k is None else min(spec.top_k, kept)
        if spec.frequency_threshold and kept:
            # terms seen fewer than frequency_threshold times never enter the vocabulary; the list is
            # ordered by descending count, so they can only shorten it
            fthr = torch.tensor([min(int(spec.frequency_threshold), 0x7FFFFFFF)], dtype=torch.int32, device=self.device)
            scratch_sel = self._buf(f"{tag}_fsel", pend["cap"], torch.int32)
            nat.call("txf_vocab_filter_ge", ent.data_ptr(), counts.data_ptr(), kept, fthr.data_ptr(), scratch_sel.data_ptr(),
                     ctrl.data_ptr(), s)
            k = min(k, int(ctrl[5].item()))
        sel, m = ent, (kept if k else 0)
        if kept > k and k > 0:
            # count cut-off: the k-th largest count (exact), then only entries >= it need ordering
            rank = torch.tensor([[kept - k + 1]], dtype=torch.int64, device=self.device)
            cptr = nat.ptr_array([counts.data_ptr()])

            def count_pass(st, p):
                nat.call("txf_msel_hist", st, 1, 1, p, cptr, None, None, 1, kept, None, s)

            thr = self._msel(f"{tag}_thr", 1, 1, rank, count_pass, reduce=False)
            sel = self._buf(f"{tag}_sel", pend["cap"], torch.int32)
            nat.call("txf_vocab_filter_ge", ent.data_ptr(), counts.data_ptr(), kept, thr.data_ptr(), sel.data_ptr(), ctrl.data_ptr(), s)
            m = int(ctrl[5].item())
        scratch = self._buf(f"{tag}_sort", max(nat.query("txf_vocab_sort_scratch_bytes", m), 16), torch.uint8)
        nat.call("txf_vocab_sort", table.data_ptr(), arena.data_ptr(), sel.data_ptr(), m, scratch.data_ptr(), s)
        return sel, k

    def _vocab_finish(self, vk: Tuple, spec: VocabSpec, pend) -> VocabResult:
        s = self._stream()
        ctrl_host = pend["ctrl"].cpu().tolist()
        kept = int(ctrl_host[3])
        if not self.comm.active:
            self._check_vocab_flags(int(ctrl_host[2]), int(ctrl_host[1]), spec)
        else:
            # every decision below is taken on values that are identical on all ranks: the merged table's
            # flags, which carry the OR of every rank's local flags (bit 6, detail in bits 8..13)
            m = int(pend["merged"]["ctrl"][2].item())
            self._check_vocab_flags((((m >> 8) & 0x3F) | pend.get("peer_flags", 0)) & ~4, None, spec)  # some rank's local pass
            if m & 4:  # a fixed-layout section overflowed on some rank: sized exchange
                self._exchange({}, {vk: pend})
                m = int(pend["merged"]["ctrl"][2].item())  # owner tables of the partitioned merge report their flags the same way
                self._check_vocab_flags((((m >> 8) & 0x3F) | pend.get("peer_flags", 0)) & ~4, None, spec)
            ctrl_m = pend["merged"]["ctrl"].cpu().tolist()
            self._check_vocab_flags(int(ctrl_m[2]) & 0x3F & ~4, None, spec)                            # the merged table
            pend = pend["merged"]  # every rank's admitted entries, re-counted into one table
            kept = int(ctrl_m[3])
        tag, table, arena = pend["tag"], pend["table"], pend["arena"]
        sel, k = self._select_top(pend, spec, kept)
        lookup_cap = _pow2_at_least(max(2 * (spec.top_k if spec.top_k is not None else k), 2 * k, 2))
        lookup = self._buf(f"{tag}_lookup", nat.query("txf_vocab_lookup_bytes", lookup_cap), torch.uint8)
        if k == 0:
            return self._vocab_from_terms(vk, [EMPTY_VOCAB_SENTINEL], lookup, lookup_cap)
        nat.call("txf_vocab_build_lookup", table.data_ptr(), sel.data_ptr(), k, lookup.data_ptr(), lookup_cap, s)

        def export():  # valid until the next analyze() on this Engine overwrites the count table (like the rest of the state)
            _, _, _, offsets, pool = self._export_entries(pend, sel, k)
            return offsets.to(torch.int32), pool

        return VocabResult(vk, k, lookup, lookup_cap, arena, export=export)

    def _vocab_from_terms(self, vk: Tuple, terms: List[bytes], lookup: Optional[torch.Tensor] = None,
                          lookup_cap: Optional[int] = None) -> VocabResult:
        col = StringColumn.from_list(terms).to(self.device)
        return vocab_from_device_terms(vk, len(terms), col.offsets, col.pool, lookup, lookup_cap)

    # ---- transform ----------------------------------------------------------------------------
    def _static_vocab(self, spec: StaticVocabSpec) -> VocabResult:
        res = self._static_vocabs.get(spec.out_key)
        if res is None:
            res = self._vocab_from_terms(("static", spec.out_key), [t.encode("latin-1") for t in spec.terms])
            self._static_vocabs[spec.out_key] = res
        return res

    def _const(self, value: float) -> torch.Tensor:
        t = self._consts.get(value)
        if t is None:
            t = torch.tensor([value], dtype=torch.float32, device=self.device)
            self._consts[value] = t
        return t

    def _operand(self, o: list, table: Table, state: TransformState, made: Dict[str, torch.Tensor]) -> nat.TxfOperand:
        """txf_operand of one ArithSpec operand (see plan.ArithSpec)."""
        r = nat.TxfOperand()
        if o[0] == "col":
            col = table[o[1]]
            r.col, r.valid, r.dtype = col.values.data_ptr(), _ptr(col.valid), _DT[col.dtype]
        elif o[0] == "out":
            r.col, r.dtype = made[o[1]].data_ptr(), nat.TXF_F32
        elif o[0] == "scalar":
            r.scalar = state.scalar(o[1], o[2]).data_ptr()
        else:
            r.scalar = self._const(float(o[1])).data_ptr()
        return r

    def _run_fused(self, table: Table, state: Optional[TransformState], out: Optional[Table], fused, stream_kernel: bool):
        """Build the op program of `fused` specs and launch it (txf_transform, or txf_transform_stream for
        state-free ops); returns {out_key: column}."""
        n = len(table)
        dev = self.device
        ops = (nat.TxfOp * max(len(fused), 1))()
        cols: Dict[str, object] = {}
        copies = []
        def reuse(prev, dtype):
            """Buffer of an earlier output column with room for n rows (a ring of output sets may be larger), or a new one."""
            if prev is not None and prev.values.dtype == dtype and prev.values.numel() >= n:
                return prev.values[:n]
            return torch.empty(n, dtype=dtype, device=dev)

        for i, spec in enumerate(fused):
            op = ops[i]
            prev = out[spec.out_key] if out is not None and spec.out_key in out else None
            if isinstance(spec, RatioLabelSpec):
                fare, tips = table[spec.fare_key], table[spec.tips_key]
                o = reuse(prev, torch.int64)
                op.kind, op.in_dtype = nat.OP_RATIO_LABEL, nat.TXF_F32
                op.in0, op.valid0, op.in1, op.valid1 = fare.values.data_ptr(), _ptr(fare.valid), tips.values.data_ptr(), _ptr(tips.valid)
                op.f0 = spec.ratio
                op.out = o.data_ptr()
                cols[spec.out_key] = NumericColumn(o)
                continue
            col = table[spec.in_key]
            if isinstance(spec, ZScoreSpec):
                o = reuse(prev, torch.float32)
                op.kind, op.in_dtype = nat.OP_ZSCORE, _DT[col.dtype]
                op.in0, op.valid0 = col.values.data_ptr(), _ptr(col.valid)
                op.p0 = state.mean_var[state.moment_keys.index(spec.in_key)].data_ptr()
                cols[spec.out_key] = NumericColumn(o)
            elif isinstance(spec, BucketSpec):
                res = state.buckets[spec.num_buckets]
                c = res.keys.index(spec.in_key)
                o = reuse(prev, torch.int64)
                op.kind, op.in_dtype = nat.OP_BUCKETIZE, _DT[col.dtype]
                op.in0, op.valid0 = col.values.data_ptr(), _ptr(col.valid)
                op.p0, op.p1, op.a0 = res.boundaries[c].data_ptr(), res.nbound[c].data_ptr(), res.boundaries.shape[1]
                cols[spec.out_key] = NumericColumn(o)
            elif isinstance(spec, (VocabSpec, StaticVocabSpec)):
                o = reuse(prev, torch.int64)
                if isinstance(spec, StaticVocabSpec):
                    res = self._static_vocab(spec)
                else:
                    res = state.vocabs[_vocab_key(spec)]
                    col = self._vocab_source(table, spec)
                    first = next(sp for sp in self.plan.of(VocabSpec) if _vocab_key(sp) == _vocab_key(spec))
                    coded = None
                    if self.use_codes and (first.num_oov_buckets, first.default_value) == (spec.num_oov_buckets, spec.default_value):
                        coded = next((c for (vk, _), c in state.codes.items() if vk == _vocab_key(spec) and c.matches(table, col)), None)
                    if coded is not None:
                        op.kind = nat.OP_VOCAB_CODED  # this very table was analyzed: decode its dictionary codes
                        op.in0, op.p0 = coded.codes.data_ptr(), coded.decode.data_ptr()
                        op.out = o.data_ptr()
                        cols[spec.out_key] = NumericColumn(o)
                        continue
                op.kind = nat.OP_VOCAB
                op.in0, op.valid0, op.pool = col.offsets.data_ptr(), _ptr(col.valid), col.pool.data_ptr()
                op.p0, op.p1 = res.lookup.data_ptr(), res.lookup_pool.data_ptr()
                op.a0, op.a1, op.a2, op.a3 = res.lookup_cap, res.size, spec.num_oov_buckets, spec.default_value
                cols[spec.out_key] = NumericColumn(o)
            elif isinstance(spec, FillSpec):
                if col.kind == "string":
                    if prev is not None and prev.offsets.numel() >= n + 1 and prev.pool.numel() >= col.pool_bytes + nat.TXF_POOL_SLACK:
                        o_off, o_pool = prev.offsets[: n + 1], prev.pool
                    else:
                        o_off = torch.empty(n + 1, dtype=torch.int32, device=dev) if n else torch.zeros(1, dtype=torch.int32, device=dev)
                        o_pool = torch.empty(col.pool_bytes + nat.TXF_POOL_SLACK, dtype=torch.uint8, device=dev)
                    op.kind = nat.OP_FILL_STR
                    op.in0 = col.offsets.data_ptr()
                    op.out = o_off.data_ptr()
                    copies.append((o_pool, col))
                    cols[spec.out_key] = StringColumn(o_off, o_pool, None, col.pool_bytes)
                    continue
                o = reuse(prev, col.values.dtype)
                op.kind = nat.OP_FILL_F32 if col.dtype == "float32" else nat.OP_FILL_I64
                op.in_dtype = _DT[col.dtype]
                op.in0, op.valid0 = col.values.data_ptr(), _ptr(col.valid)
                cols[spec.out_key] = NumericColumn(o)
            else:  # pragma: no cover
                raise AssertionError(spec)
            op.out = o.data_ptr()
        if n and fused:
            if stream_kernel:
                ws = self._buf("xstream_ws", nat.query("txf_transform_stream_workspace_bytes"), torch.uint8)
                nat.call("txf_transform_stream", ops, len(fused), n, ws.data_ptr(), self._stream())
            else:
                nat.call("txf_transform", ops, len(fused), n, self._stream())
            for o_pool, col in copies:  # bytes of pass-through string features (the column may be a slice: its span is read on the device)
                nat.call("txf_copy_string_bytes", col.offsets.data_ptr(), col.pool.data_ptr(), n, col.pool_bytes, o_pool.data_ptr(),
                         self._stream())
        return cols

    @staticmethod
    def _state_free(spec) -> bool:
        """Ops that need no analyzer result: fill-missing pass-throughs and the tips label."""
        return isinstance(spec, (FillSpec, RatioLabelSpec))

    def transform_state_free(self, table: Table, out: Optional[Table] = None) -> Dict[str, object]:
        """The ops of the program that need no analyzer result, through txf_transform_stream (the
        small-footprint streaming kernel): queue it on a second stream while `analyze` runs on the same
        rows, then pass the columns it returns to `transform(..., done=...)`."""
        self._check_table(table)
        return self._run_fused(table, None, out, [sp for sp in self.plan.specs if self._state_free(sp)], stream_kernel=True)

    def transform(self, table: Table, state: TransformState, out: Optional[Table] = None,
                  done: Optional[Dict[str, object]] = None) -> Table:
        """One fused kernel for the ops of the taxi program over one GPU-resident chunk, then one
        small kernel per sibling mapper (scale, arithmetic, cast) in dependency order; returns the
        output table (reusing `out`'s buffers when given).  `done`: output columns already produced by
        `transform_state_free` for this table; their ops are skipped."""
        self._check_table(table)
        n = len(table)
        dev = self.device
        done = done or {}
        fused = [sp for sp in self.plan.specs if isinstance(sp, (ZScoreSpec, BucketSpec, VocabSpec, StaticVocabSpec, FillSpec,
                                                                   RatioLabelSpec)) and sp.out_key not in done]
        cols = dict(done)
        cols.update(self._run_fused(table, state, out, fused, stream_kernel=False))
        # sibling mappers: plan order already has every operand before its user
        made: Dict[str, torch.Tensor] = {k: c.values for k, c in cols.items() if isinstance(c, NumericColumn)}
        for spec in self.plan.specs:
            if not isinstance(spec, (ScaleSpec, ArithSpec, CastSpec)):
                continue
            prev = out[spec.out_key] if out is not None and spec.out_key in out else None
            dt = torch.int64 if isinstance(spec, CastSpec) and spec.to == "int64" else torch.float32
            o = prev.values[:n] if prev is not None and prev.values.dtype == dt and prev.values.numel() >= n else torch.empty(n, dtype=dt, device=dev)
            if n:
                if isinstance(spec, ScaleSpec):
                    x = self._operand(["col", spec.in_key], table, state, made)
                    nat.call("txf_map_scale_minmax", x, state.minmax[state.minmax_keys.index(spec.in_key)].data_ptr(),
                             spec.output_min, spec.output_max, o.data_ptr(), n, self._stream())
                elif isinstance(spec, ArithSpec):
                    code = {"add": nat.ARITH_ADD, "sub": nat.ARITH_SUB, "mul": nat.ARITH_MUL, "div": nat.ARITH_DIV}[spec.op]
                    nat.call("txf_map_binary", code, self._operand(spec.a, table, state, made),
                             self._operand(spec.b, table, state, made), o.data_ptr(), n, self._stream())
                else:
                    nat.call("txf_map_cast", self._operand(["col", spec.in_key], table, state, made), _DT[spec.to], o.data_ptr(), n,
                             self._stream())
            made[spec.out_key] = o
            if not spec.out_key.startswith(TEMP_PREFIX):
                cols[spec.out_key] = NumericColumn(o)
        order = [sp.out_key for sp in self.plan.specs if sp.out_key in cols and not sp.out_key.startswith(TEMP_PREFIX)]
        return Table({k: cols[k] for k in order})

    def analyze_and_transform(self, data: Union[Table, Sequence[Table]], out=None, rows_per_rank: Optional[Sequence[int]] = None):
        """The single pass tft_beam.AnalyzeAndTransformDataset makes (notebooks/06_...:236-262): analyze the
        rows and transform the same rows.  The ops that need no analyzer result (pass-throughs, label: half
        of the transform's bytes) are queued first, on a second stream, as the small-footprint streaming
        kernel; the analyzers -- issue- and latency-bound, a quarter of the HBM bandwidth -- run next to it
        (the count pass in its 24-warp shape), and the ops that need their results follow.  `out`: output
        table(s) of an earlier call whose buffers are reused; `rows_per_rank`: see `analyze`."""
        chunks = [data] if isinstance(data, Table) else list(data)
        outs_prev = [out] if isinstance(out, Table) else (list(out) if out is not None else [None] * len(chunks))
        free = [sp for sp in self.plan.specs if self._state_free(sp)]
        overlap = self.overlap_transform and bool(free) and any(len(c) for c in chunks)
        done = [None] * len(chunks)
        if overlap:
            for c in chunks:
                self._check_table(c)
            main = torch.cuda.current_stream()
            if self._xstream is None:
                # >= 4 ranks: the same priority as the analyzers' streams, so that among the kernels the sample all-gather
                # releases the launch order decides, as in a step whose kernels are launched one by one on an idle GPU: a CTA
                # of the quantile count pass (256 threads x 32 registers, 35 KB) FITS the slot next to a count-pass CTA that
                # the stream kernel's CTA is meant to take, and at a higher priority it would be placed first whenever both
                # are pending (hypothesis, not measured; with fewer ranks the measured configuration is kept)
                many = self.comm.active and self.comm.world >= self.drain_between_steps_world
                self._xstream = torch.cuda.Stream(device=self.device, priority=self._side_priority if many else 0)
            xs = self._xstream
            if self.comm.active and self.comm.world >= self.drain_between_steps_world and self._step_end is not None:
                # 4 and 8 ranks, measured (profiles/r2_ab_wait_4gpu.txt, r2_exchange_timeline_8gpu_head.txt): a step that starts on
                # an idle GPU takes 6.4-6.7 ms, a step queued while the previous one still runs 7.1-7.6 ms (2 ranks: no such
                # difference, 6.37 ms queued).  Presumably what a late sample all-gather releases then arrives all at once and
                # the SMs are taken in priority order, not in the order the pass was laid out in.  So with >= 4 ranks the host
                # lets the previous step drain first; the launch latency this exposes is part of the 6.4-6.7 ms.
                self._step_end.synchronize()
            # A timing event on the caller's stream at the start of every step.  Measured, not explained: when the CUDA context
            # was created with the default 8 host->device work queues (i.e. before this package could set
            # CUDA_DEVICE_MAX_CONNECTIONS), back-to-back steps serialise part of the four streams' work (7.55 ms per step);
            # with this record they do not (6.57 ms), the same as with 32 queues (tools/exp/conn.py).
            self._step_mark = torch.cuda.Event(enable_timing=True)
            self._step_mark.record(main)
            xs.wait_stream(main)

            def side_work(events):  # called by analyze at the point where the stream kernel may start (see _analyze_once)
                for ev in events:
                    xs.wait_event(ev)
                with torch.cuda.stream(xs):
                    done[:] = [self.transform_state_free(c, o) for c, o in zip(chunks, outs_prev)]
                    self._mark("stream_kernel_done")
                for d in done:  # buffers allocated under the side stream are handed to the caller's stream
                    for col in d.values():
                        for t in ((col.values,) if isinstance(col, NumericColumn) else (col.offsets, col.pool)):
                            t.record_stream(main)

            self._count_warps = 24
        launched = set()

        def between(state, phase):
            """Ops whose analyzer results are final are queued while the rest of the state is still being finished:
            z-scores once the moments are merged (multi-rank: while the boundary selection's all-reduces run),
            bucket ids once the boundaries are final (while the vocabularies are being ranked)."""
            kinds = (ZScoreSpec,) if phase == "moments" else (ZScoreSpec, BucketSpec)
            if phase == "boundaries-again":
                launched.difference_update(sp.out_key for sp in self.plan.specs if isinstance(sp, BucketSpec))
            early = [sp for sp in self.plan.specs if isinstance(sp, kinds) and sp.out_key not in launched]
            launched.update(sp.out_key for sp in early)
            for i, (c, o) in enumerate(zip(chunks, outs_prev)):
                done[i] = dict(done[i] or {}, **self._run_fused(c, state, o, early, stream_kernel=False))

        try:
            state = self.analyze(chunks, rows_per_rank, side_work if overlap else None, between if overlap else None)
        finally:
            self._count_warps = 0
        outs = [self.transform(c, state, o, d) for c, o, d in zip(chunks, outs_prev, done)]
        if overlap:
            torch.cuda.current_stream().wait_stream(xs)
        self._mark("step_done")
        if overlap and self.comm.active:
            self._step_end = torch.cuda.Event()
            self._step_end.record(torch.cuda.current_stream())
        return (outs[0] if isinstance(data, Table) else outs), state
