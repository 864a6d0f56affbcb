// vocab.cu -- vocabulary analyzer of tft.compute_and_apply_vocabulary (reference call site
// airflow-dags/taxi_utils.py:121-126, top_k=_VOCAB_SIZE :51, num_oov_buckets=_OOV_SIZE :54).
//
// Counting: one persistent CTA per SM whose 32 warps each walk their own 96-row steps.  A step's
// offsets and term bytes are landed in the warp's shared-memory stage by the bulk-copy engine
// (double-buffered per warp, no block-wide barriers); every row's term gets a 128-bit identity
// (vocab.cuh: the bytes themselves up to 15 bytes, a double hash beyond) and is counted in a
// per-CTA shared-memory table (2048 slots) after one warp-level aggregation round; terms that do
// not fit go straight to the global open-addressing table (32-byte slots, one DRAM sector per
// probe), into which the front tables are flushed at the end.  Terms of up to 15 bytes are their
// own key.  Longer terms carry a fingerprint (119 hash bits + the exact length): the global table --
// which every count reaches, through a flush, a direct add or a merge between ranks -- confirms every
// fingerprint match on the BYTES against its arena, so two terms never share a global slot; a CTA's
// front table is a cache in front of it that trusts the full fingerprint for the rows of one CTA.
// The thread that claims a global slot copies the term's bytes into the arena before it publishes
// the slot, so the table is self-contained: it can be exchanged between ranks and outlives the
// chunk it was built from.
// Ranking reproduces tft: drop '' and terms containing \n / \r, order by (count, bytes)
// descending, keep top_k.
// Algorithmic bytes of the count pass: 4 (offset) + term bytes per row + 1 validity bit.
#include "vocab.cuh"
#include "str_pipe.cuh"

namespace txf {

// ONE persistent CTA per SM; its warps run independent pipelines (no block barriers).  Two shapes: 32 warps
// (the whole SM: 64 registers x 1024 threads, 228 KB of shared memory) when the pass runs alone, and 24 warps
// (about 185 KB, 72 registers) when the caller streams other work next to it (txf_transform_stream): the count
// pass is latency-bound per warp (1.94 ms with 32 warps, 2.33 ms with 24 on 100 M taxi rows), the room it gives
// up lets 12 GB of pass-through bytes move underneath it.
constexpr int kVcWarpsFull = 32, kVcWarpsShared = 24;
constexpr int kVcSlots = 2048;            // shared-memory front table of the CTA: 2048 x 24 B = 48 KB
constexpr int kVcProbes = 8;
constexpr int kVcStepRows = 96;           // rows per warp and step: three per lane
constexpr int kVcRowsPerLane = kVcStepRows / 32;
constexpr int kVcOffSlots = 100;          // offsets staged per step: 96 + 1, padded to a multiple of 16 bytes
constexpr int kVcStageCap = 2304;         // staged term bytes per step (24 B per row on average; longer steps read the pool directly)
constexpr int kVcStageBytes = 16 + kVcOffSlots * 4 + kVcStageCap;  // header | offsets | bytes
constexpr int kVcWarpBytes = 2 * kVcStageBytes + kVcStepRows * 2;  // two stages + the warp's long-term queue
constexpr size_t vc_smem_bytes(int warps) { return kVcSlots * (8 + 8 + 4 + 4) + static_cast<size_t>(warps) * kVcWarpBytes + warps * 2 * 8; }
constexpr uint32_t kPadEnt = 0xFFFFFFFFu;
constexpr int kVcGrid = kSMs;
constexpr uint32_t kVcLocalCodes = kVcGrid * kVcSlots;  // codes below this: (CTA, front-table slot); above: global slot

__device__ __forceinline__ unsigned long long ldcg64(const unsigned long long* p) {
  return __ldcg(p);  // L2 only: never a stale L1 line for words other SMs update atomically
}

constexpr uint64_t kMaxProbes = 1u << 14;
constexpr uint32_t kSpinLimit = 1u << 22;  // a claimed slot is published within a few hundred cycles; this only bounds a bug

// the 8 bytes at p (any alignment) read with L2-only loads: arena lines are written by other SMs
// while this kernel runs, and a line may sit in this SM's L1 from before its neighbour term was added
__device__ __forceinline__ uint64_t fetch64_cg(const uint8_t* p) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const unsigned long long* q = reinterpret_cast<const unsigned long long*>(a & ~static_cast<uintptr_t>(7));
  const uint32_t sh = static_cast<uint32_t>(a & 7) * 8;
  const uint64_t lo = __ldcg(q), hi = __ldcg(q + 1);
  return (lo >> sh) | ((hi << 1) << (63 - sh));
}
__device__ bool arena_term_equals(const uint8_t* arena_term, const uint8_t* str, uint32_t n) {
  uint32_t i = 0;
  for (; i + 8 <= n; i += 8)
    if (fetch64_cg(arena_term + i) != fetch64(str + i)) return false;
  if (i < n) {
    const uint64_t m = (1ULL << (8 * (n - i))) - 1;
    if ((fetch64_cg(arena_term + i) ^ fetch64(str + i)) & m) return false;
  }
  return true;
}

// Add `cnt` occurrences of the term with identity (a, b), whose bytes are str[0..len), to the
// global table; returns the slot index of the term.  Lock-free and exact.  A slot is claimed by CAS
// on word a; the claimer copies the term into the arena, stores strref, and only then publishes
// word b.  Anyone who finds its a waits for b (a few hundred cycles), moves on if b differs, and --
// for a fingerprint -- compares the bytes with the arena copy and moves on if they differ.
__device__ __noinline__ uint32_t table_add(CountSlot* table, uint64_t mask, unsigned long long ka, unsigned long long kb,
                              unsigned long long cnt, const uint8_t* str, uint32_t len, uint8_t* arena, int64_t arena_cap,
                              unsigned long long* ctrl) {
  const bool fp = key_is_fingerprint(kb);
  uint64_t idx = home_slot(ka, kb, mask);
  const uint64_t max_probe = mask < kMaxProbes ? mask : kMaxProbes;  // a chain this long means the table is as good as full
  for (uint64_t probe = 0; probe <= max_probe; ++probe, idx = (idx + 1) & mask) {
    CountSlot* s = table + idx;
    unsigned long long cur = ldcg64(&s->a);
    if (cur == 0) {
      cur = atomicCAS(&s->a, 0ULL, ka);
      if (cur == 0) {  // this thread owns the slot: keep the term's bytes, then publish it
        atomicAdd(&ctrl[1], 1ULL);
        const unsigned long long off = atomicAdd(&ctrl[0], static_cast<unsigned long long>(len));
        unsigned long long ref = 0;
        if (len > kMaxTermLen || off + len > static_cast<unsigned long long>(arena_cap)) {
          atomicOr(&ctrl[2], 1ULL);
        } else {
          for (uint32_t i = 0; i < len; ++i) arena[off + i] = str[i];
          ref = (off << 24) | len;
        }
        __stcg(&s->strref, ref);
        atomicAdd(&s->count, cnt);
        __threadfence();
        atomicExch(&s->b, kb);
        return static_cast<uint32_t>(idx);
      }
    }
    if (cur != ka) continue;
    unsigned long long t = ldcg64(&s->b);
    for (uint32_t spin = 0; t == 0; ++spin) {
      if (spin > kSpinLimit) { atomicOr(&ctrl[2], 32ULL); return 0xFFFFFFFFu; }
      __nanosleep(32);
      t = ldcg64(&s->b);
    }
    if (t != kb) continue;
    if (fp) {  // same fingerprint: the bytes decide (ref == 0: the arena overflowed, the pass is void anyway)
      const unsigned long long ref = ldcg64(&s->strref);
      if (ref != 0 && (static_cast<uint32_t>(ref & 0xFFFFFFu) != len || !arena_term_equals(arena + (ref >> 24), str, len))) continue;
    }
    atomicAdd(&s->count, cnt);
    return static_cast<uint32_t>(idx);
  }
  atomicOr(&ctrl[2], 2ULL);  // table full
  return 0xFFFFFFFFu;
}

// Front-table state of one CTA (shared memory) plus what a row needs to be counted.
struct VcFront {
  unsigned long long* s_a;
  unsigned long long* s_b;
  uint32_t* s_cnt;
  uint32_t* s_row;
  CountSlot* table;
  uint64_t mask;
  uint8_t* arena;
  int64_t arena_cap;
  unsigned long long* ctrl;
  uint32_t* codes;
  const uint8_t* pool;
};

// Slow path of the front-table search: full claim protocol (CAS on word a, then on word b) over
// up to kVcProbes slots.  Kept out of line so the common "already there" case stays straight-line.
__device__ __noinline__ int vc_find_slow(unsigned long long* s_a, unsigned long long* s_b, uint32_t* s_row,
                                         unsigned long long ka, unsigned long long kb, uint32_t idx, uint32_t row) {
  struct { unsigned long long *s_a, *s_b; uint32_t* s_row; } f{s_a, s_b, s_row};
  const StrKey key{ka, kb};
  for (int p = 0; p < kVcProbes; ++p, idx = (idx + 1) & (kVcSlots - 1)) {
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&f.s_a[idx]);
    if (cur == 0) {
      cur = atomicCAS(&f.s_a[idx], 0ULL, key.a);
      if (cur == 0) cur = key.a;
    }
    if (cur != key.a) continue;
    unsigned long long t = *reinterpret_cast<volatile unsigned long long*>(&f.s_b[idx]);
    if (t == 0) {
      t = atomicCAS(&f.s_b[idx], 0ULL, key.b);
      if (t == 0) { t = key.b; f.s_row[idx] = row; }
    }
    if (t == key.b) return static_cast<int>(idx);
  }
  return -1;
}

// Count one term per lane (has == false: the lane only takes part in the warp votes).  Must be
// called by all 32 lanes of a warp together.  use_front == false sends the term straight to the
// byte-verified global table (test hook: fingerprints truncated on purpose).
__device__ __forceinline__ int vc_count(const VcFront& f, bool has, StrKey key, uint32_t row, int32_t b, uint32_t len,
                                        bool use_front = true) {
  int slot = -1;
  if (has && use_front) {
    const uint32_t idx = static_cast<uint32_t>(home_slot(key.a, key.b, kVcSlots - 1));
    // fast path: the term already sits in its home slot (both words are written before any count)
    const unsigned long long a0 = *reinterpret_cast<volatile unsigned long long*>(&f.s_a[idx]);
    const unsigned long long b0 = *reinterpret_cast<volatile unsigned long long*>(&f.s_b[idx]);
    slot = (a0 == key.a && b0 == key.b) ? static_cast<int>(idx)
                                        : vc_find_slow(f.s_a, f.s_b, f.s_row, key.a, key.b, idx, row);
  }
  // one warp-level aggregation round (absorbs the dominant term), then one atomic per lane
  const unsigned hits = __ballot_sync(0xffffffffu, slot >= 0);
  if (hits) {
    const int leader = __ffs(hits) - 1;
    const int lslot = __shfl_sync(0xffffffffu, slot, leader);
    const unsigned grp = __ballot_sync(0xffffffffu, slot == lslot);
    if (static_cast<int>(lane_id()) == leader) atomicAdd(&f.s_cnt[lslot], static_cast<uint32_t>(__popc(grp)));
    else if (slot >= 0 && slot != lslot) atomicAdd(&f.s_cnt[slot], 1u);
  }
  uint32_t code = blockIdx.x * static_cast<uint32_t>(kVcSlots) + static_cast<uint32_t>(slot);
  if (has && slot < 0)  // front table full around this term: count it in the global table
    code = kVcLocalCodes + table_add(f.table, f.mask, key.a, key.b, 1ULL, f.pool + b, len, f.arena, f.arena_cap, f.ctrl);
  if (has && f.codes != nullptr) __stcs(f.codes + row, code);  // the row's term, dictionary-encoded
  return slot;
}

// ---- per-warp pipeline ------------------------------------------------------------------------
// Every warp of the CTA owns two stages in shared memory and walks its own sequence of 96-row
// steps: lane 0 asks the bulk-copy engine for the step's offsets and term bytes (cp.async.bulk,
// completion on the stage's mbarrier), the warp waits on that barrier only, counts, and re-arms the
// stage for the step after next.  No block-wide barrier exists between the table initialisation and
// the final flush, so no warp ever waits for another one.
// Stage layout: [a0: pool offset of byte 0 (4) | pad (4) | staged (4) | meta (4)] [offsets x 100] [bytes].
struct VcStage {
  uint8_t* base;  // generic pointer to the stage
  __device__ __forceinline__ uint32_t& a0() const { return *reinterpret_cast<uint32_t*>(base); }  // pools hold < 2^31 bytes
  __device__ __forceinline__ int& staged() const { return *reinterpret_cast<int*>(base + 8); }
  __device__ __forceinline__ int& meta() const { return *reinterpret_cast<int*>(base + 12); }
  __device__ __forceinline__ int32_t* off() const { return reinterpret_cast<int32_t*>(base + 16); }
  __device__ __forceinline__ uint8_t* bytes() const { return base + 16 + kVcOffSlots * 4; }
};

struct VcSource {
  const int32_t* offsets;
  const uint8_t* pool;
  const uint8_t* valid;
  int64_t n;
  bool pool_ok, off_ok;  // 16-byte aligned bases (bulk copies need them)
};

// lane 0 of the owning warp: start the copies of step rows [r0, r1) into the stage
__device__ __forceinline__ void vc_issue(const VcSource& src, const VcStage& st, uint64_t* bar, uint32_t r0, uint32_t r1, uint32_t b0,
                                         uint32_t e0) {  // b0, e0 = offsets[r0], offsets[r1] (loaded by the caller well ahead of this call)
  const uint32_t start = b0 & ~15u;
  const uint32_t need = ((e0 - start + 15u) & ~15u) + 32u;  // window loads touch < 24 bytes past a term
  const bool do_bytes = src.pool_ok && need <= static_cast<uint32_t>(kVcStageCap);
  const bool do_meta = src.off_ok && (r1 - r0) == kVcStepRows && static_cast<int64_t>(r0) + kVcOffSlots <= src.n + 1;
  st.a0() = start;
  st.staged() = do_bytes;
  st.meta() = do_meta;
  const uint32_t total = (do_bytes ? need : 0u) + (do_meta ? kVcOffSlots * 4u : 0u);
  if (total == 0) {
    mbar_arrive(bar);
    return;
  }
  fence_proxy_async();  // the stage was read through ordinary loads until now
  mbar_arrive_expect_tx(bar, total);
  if (do_bytes) bulk_copy_g2s(st.bytes(), src.pool + start, need, bar);
  if (do_meta) bulk_copy_g2s(st.off(), src.offsets + r0, kVcOffSlots * 4u, bar);
}

// One step of one warp.  Phase A counts the short terms (exact identity, a handful of
// instructions) and queues the long ones; phase B hashes the queued terms with the lanes packed --
// columns that mix empty/short and long terms would otherwise idle most lanes during the hashing.
template <bool STAGED, bool META>
__device__ __forceinline__ void vc_step(const VcFront& f, const VcSource& src, const VcStage& st, uint32_t r0, uint32_t r1,
                                        uint16_t* wqueue, const uint4* s_mask, int& empty_slot, bool front_long) {
  const int32_t* soff = st.off();
  const uint8_t* base = STAGED ? st.bytes() - st.a0() : src.pool;       // the term at pool offset o is at base + o
  const uint32_t sbase = STAGED ? smem_u32(st.bytes()) - st.a0() : 0u;
  uint32_t nq = 0;
#pragma unroll
  for (int j = 0; j < kVcRowsPerLane; ++j) {
    const uint32_t row = r0 + j * 32 + lane_id();  // rows fit 31 bits (the entry point checks n)
    const bool in_step = row < r1;
    int32_t b = 0;
    uint32_t len = 0;
    if (in_step) {
      b = META ? soff[row - r0] : __ldg(src.offsets + row);
      if (row_valid(src.valid, row)) len = static_cast<uint32_t>((META ? soff[row - r0 + 1] : __ldg(src.offsets + row + 1)) - b);
    }
    const bool is_long = in_step && len > kExactKeyMax;
    const unsigned lm = __ballot_sync(0xffffffffu, is_long);
    if (is_long) wqueue[nq + __popc(lm & ((1u << lane_id()) - 1u))] = static_cast<uint16_t>(row - r0);
    nq += __popc(lm);  // warp-uniform
    // Missing cells are the empty term.  Once the warp knows that term's slot, a row group whose
    // short terms are all empty (every group of a column like `company`) needs no table search.
    const bool is_short = in_step && !is_long;
    const unsigned other = __ballot_sync(0xffffffffu, is_short && len != 0);
    if (other == 0 && empty_slot >= 0) {
      const unsigned em = __ballot_sync(0xffffffffu, is_short);
      if (em) {
        if (static_cast<int>(lane_id()) == __ffs(em) - 1) atomicAdd(&f.s_cnt[empty_slot], static_cast<uint32_t>(__popc(em)));
        if (is_short && f.codes != nullptr) __stcs(f.codes + row, blockIdx.x * static_cast<uint32_t>(kVcSlots) + static_cast<uint32_t>(empty_slot));
      }
    } else {
      StrKey key{0, 0};
      if (is_short)
        key = STAGED ? make_key_exact_smem(sbase + static_cast<uint32_t>(b), len, s_mask) : make_key_exact(PtrFetch{base + b}, len);
      const int slot = vc_count(f, is_short, key, row, b, len);
      const unsigned found = __ballot_sync(0xffffffffu, is_short && len == 0 && slot >= 0);
      if (found) empty_slot = __shfl_sync(0xffffffffu, slot, __ffs(found) - 1);
    }
  }
  __syncwarp();
  for (uint32_t t0 = 0; t0 < nq; t0 += 32) {  // warp-uniform trip count
    const uint32_t t = t0 + lane_id();
    const bool has = t < nq;
    uint32_t row = r0;
    int32_t b = 0;
    uint32_t len = 0;
    StrKey key{0, 0};
    if (has) {
      const uint32_t r = wqueue[t];
      row = r0 + r;
      b = META ? soff[r] : __ldg(src.offsets + row);
      len = static_cast<uint32_t>((META ? soff[r + 1] : __ldg(src.offsets + row + 1)) - b);
      key = STAGED ? make_key_hashed_smem(sbase + static_cast<uint32_t>(b), len) : make_key_hashed(PtrFetch{base + b}, len);
    }
    vc_count(f, has, key, row, b, len, front_long);
  }
}

template <int kVcWarps>
__global__ void __launch_bounds__(kVcWarps * 32, 1) __maxnreg__(kVcWarps == kVcWarpsFull ? 64 : 72)
    vocab_count_kernel(const int32_t* __restrict__ offsets,
                                                                  const uint8_t* __restrict__ pool,
                                                                  const uint8_t* __restrict__ valid, int64_t n,
                                                                  CountSlot* table, uint64_t mask, uint8_t* arena,
                                                                  int64_t arena_cap, unsigned long long* ctrl,
                                                                  uint32_t* __restrict__ codes,
                                                                  uint32_t* __restrict__ local2global) {
  constexpr int kVcBlock = kVcWarps * 32;
  extern __shared__ __align__(16) uint8_t vc_smem[];
  __shared__ uint4 s_mask[16];
  fill_short_masks(s_mask);
  VcFront f;
  f.s_a = reinterpret_cast<unsigned long long*>(vc_smem);
  f.s_b = f.s_a + kVcSlots;
  f.s_cnt = reinterpret_cast<uint32_t*>(f.s_b + kVcSlots);
  f.s_row = f.s_cnt + kVcSlots;
  f.table = table; f.mask = mask; f.arena = arena; f.arena_cap = arena_cap; f.ctrl = ctrl; f.codes = codes; f.pool = pool;
  // test hook: with fingerprints truncated on purpose the front table (which trusts them) is bypassed for long terms
  const bool front_long = identity_mask(1) == kFpBitsB && identity_mask(0) == ~0ULL;
  for (int i = threadIdx.x; i < kVcSlots; i += kVcBlock) { f.s_a[i] = 0; f.s_b[i] = 0; f.s_cnt[i] = 0; f.s_row[i] = 0; }
  const int warp = threadIdx.x >> 5;
  uint8_t* wbase = reinterpret_cast<uint8_t*>(f.s_row + kVcSlots) + static_cast<size_t>(warp) * kVcWarpBytes;
  uint16_t* wqueue = reinterpret_cast<uint16_t*>(wbase + 2 * kVcStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(f.s_row + kVcSlots) + static_cast<size_t>(kVcWarps) * kVcWarpBytes) +
                   2 * warp;
  VcSource src{offsets, pool, valid, n, (reinterpret_cast<uintptr_t>(pool) & 15) == 0, (reinterpret_cast<uintptr_t>(offsets) & 15) == 0};
  if (lane_id() == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_fence_init();
  }
  __syncthreads();  // table zeroed, barriers initialised: the only block-wide barrier before the flush

  // rows fit 31 bits (the entry point checks n), so steps and rows are 32-bit inside the loop
  const uint32_t n32 = static_cast<uint32_t>(n);
  const uint32_t nsteps = (n32 + kVcStepRows - 1) / kVcStepRows;
  const uint32_t gw = blockIdx.x * kVcWarps + warp, nw = gridDim.x * kVcWarps;
  auto step_end = [&](uint32_t t) { return (t + 1) * kVcStepRows < n32 ? (t + 1) * kVcStepRows : n32; };
  auto off_at = [&](uint32_t row) { return static_cast<uint32_t>(__ldg(offsets + row)); };
  if (lane_id() == 0) {
    if (gw < nsteps) vc_issue(src, VcStage{wbase}, &bars[0], gw * kVcStepRows, step_end(gw), off_at(gw * kVcStepRows), off_at(step_end(gw)));
    if (gw + nw < nsteps)
      vc_issue(src, VcStage{wbase + kVcStageBytes}, &bars[1], (gw + nw) * kVcStepRows, step_end(gw + nw), off_at((gw + nw) * kVcStepRows),
               off_at(step_end(gw + nw)));
  }
  uint32_t k = 0;
  int empty_slot = -1;  // front-table slot of the empty term once this warp has met it (warp-uniform)
  for (uint32_t t = gw; t < nsteps; t += nw, ++k) {
    const int s = k & 1;
    const uint32_t r0 = t * kVcStepRows, r1 = step_end(t);
    // the refill of this stage (issued after the step) needs two offsets from global memory: lane 0 asks for them now,
    // a whole step ahead of their use, so that the warp does not sit out two dependent global-load latencies per step
    const uint32_t nxt = t + 2 * nw;
    uint32_t nb0 = 0, ne0 = 0;
    if (lane_id() == 0 && nxt < nsteps) { nb0 = off_at(nxt * kVcStepRows); ne0 = off_at(step_end(nxt)); }
    mbar_wait(&bars[s], (k >> 1) & 1u);
    const VcStage st{wbase + s * kVcStageBytes};
    // two instantiations only (the kernel is issue-bound and its code must stay in the instruction cache): the
    // staged one for full, aligned steps; everything else (ragged last step, oversized or misaligned spans) reads
    // offsets and bytes from global memory
    if (st.staged() && st.meta()) vc_step<true, true>(f, src, st, r0, r1, wqueue, s_mask, empty_slot, front_long);
    else vc_step<false, false>(f, src, st, r0, r1, wqueue, s_mask, empty_slot, front_long);
    __syncwarp();  // every lane is done with the stage: refill it with the step after next
    if (lane_id() == 0 && nxt < nsteps) vc_issue(src, st, &bars[s], nxt * kVcStepRows, step_end(nxt), nb0, ne0);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kVcSlots; i += kVcBlock) {  // flush the front table
    uint32_t cnt = f.s_cnt[i];
    if (!cnt) continue;
    int64_t row = f.s_row[i];
    const uint8_t* str = pool;
    uint32_t len = 0;
    if (row_valid(valid, row)) {
      int32_t b = offsets[row];
      str = pool + b;
      len = static_cast<uint32_t>(offsets[row + 1] - b);
    }
    uint32_t g = table_add(table, mask, f.s_a[i], f.s_b[i], cnt, str, len, arena, arena_cap, ctrl);
    if (local2global != nullptr) local2global[blockIdx.x * static_cast<uint32_t>(kVcSlots) + i] = g;
  }
}

__global__ void vocab_insert_kernel(const uint64_t* __restrict__ fp, const uint64_t* __restrict__ tag,
                                    const uint64_t* __restrict__ count, const int64_t* __restrict__ offsets,
                                    const uint8_t* __restrict__ pool, int64_t m, uint32_t owners, uint32_t owner,
                                    CountSlot* table, uint64_t mask, uint8_t* arena, int64_t arena_cap, unsigned long long* ctrl) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= m) return;
  // owner-partitioned merge: every rank sees all entries and re-counts the terms whose identity hashes to it
  if (owners > 1 && static_cast<uint32_t>(home_slot(fp[i], tag[i], 0xFFFFFFFFULL) >> 7) % owners != owner) return;
  table_add(table, mask, fp[i], tag[i], count[i], pool + offsets[i], static_cast<uint32_t>(offsets[i + 1] - offsets[i]), arena,
            arena_cap, ctrl);
}

// ---- fixed-layout exchange section ("blob") of one vocabulary -------------------------------
// words: [0] = number of admitted entries (bit 62 set: some term longer than the byte slot);
// then a[K], b[K], count[K], len[K] and K byte slots of S bytes.  Every rank fills one such section,
// the sections are all-gathered in one collective, and every rank re-counts all of them into one
// merged table -- no sizes have to travel through the host.
__global__ void vocab_export_blob_kernel(const CountSlot* __restrict__ table, const uint8_t* __restrict__ arena,
                                         const uint32_t* __restrict__ ent, const unsigned long long* __restrict__ ctrl,
                                         int64_t K, int64_t S, unsigned long long* __restrict__ sec) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  const unsigned long long kept = ctrl[3];
  // caller zeroes word 0; bit 62 may be OR-ed in concurrently; bits 56..61 = this rank's own overflow flags (ctrl[2])
  if (i == 0) atomicOr(&sec[0], kept | ((ctrl[2] & 0x3FULL) << 56));
  if (i >= K || static_cast<unsigned long long>(i) >= kept) return;
  const CountSlot s = table[ent[i]];
  const uint32_t len = static_cast<uint32_t>(s.strref & 0xFFFFFFu);
  sec[1 + i] = s.a;
  sec[1 + K + i] = s.b;
  sec[1 + 2 * K + i] = s.count;
  sec[1 + 3 * K + i] = len;
  if (len > S) { atomicOr(&sec[0], 1ULL << 62); return; }
  uint8_t* dst = reinterpret_cast<uint8_t*>(sec + 1 + 4 * K) + i * S;
  const uint8_t* src = arena + (s.strref >> 24);
  for (uint32_t k = 0; k < len; ++k) dst[k] = src[k];
}

__global__ void vocab_insert_blob_kernel(const unsigned long long* __restrict__ blob, int64_t stride_words, int64_t sec_off,
                                         int W, int64_t K, int64_t S, CountSlot* table, uint64_t mask, uint8_t* arena,
                                         int64_t arena_cap, unsigned long long* ctrl) {
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= static_cast<int64_t>(W) * K) return;
  const int r = static_cast<int>(idx / K);
  const int64_t i = idx % K;
  const unsigned long long* sec = blob + r * stride_words + sec_off;
  const unsigned long long head = sec[0];
  const unsigned long long kept = head & ((1ULL << 56) - 1);
  if (i == 0 && (kept > static_cast<unsigned long long>(K) || ((head >> 62) & 1))) atomicOr(&ctrl[2], 4ULL);  // section overflowed
  if (i == 0 && ((head >> 56) & 0x3FULL)) atomicOr(&ctrl[2], 64ULL | (((head >> 56) & 0x3FULL) << 8));  // a rank's local pass overflowed
  if (static_cast<unsigned long long>(i) >= kept) return;
  const uint32_t len = static_cast<uint32_t>(sec[1 + 3 * K + i]);
  if (len > S) return;
  table_add(table, mask, sec[1 + i], sec[1 + K + i], sec[1 + 2 * K + i],
            reinterpret_cast<const uint8_t*>(sec + 1 + 4 * K) + i * S, len, arena, arena_cap, ctrl);
}

// warp-aggregated append: returns this lane's position in the list counted by *counter
__device__ __forceinline__ unsigned long long warp_append(bool pred, unsigned long long* counter) {
  unsigned m = __ballot_sync(0xffffffffu, pred);
  unsigned long long base = 0;
  int leader = __ffs(m) - 1;
  if (m && static_cast<int>(lane_id()) == leader) base = atomicAdd(counter, static_cast<unsigned long long>(__popc(m)));
  base = __shfl_sync(0xffffffffu, base, leader < 0 ? 0 : leader);
  return base + __popc(m & ((1u << lane_id()) - 1u));
}

__global__ void vocab_compact_kernel(const CountSlot* __restrict__ table, int64_t capacity,
                                     const uint8_t* __restrict__ arena, uint32_t* __restrict__ ent,
                                     uint32_t* __restrict__ counts, unsigned long long* ctrl) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  bool occ = false, keep = false;
  unsigned long long cnt = 0;
  if (i < capacity) {
    CountSlot s = table[i];
    occ = s.a != 0 && s.b != 0;
    if (occ) {
      cnt = s.count;
      uint32_t len = static_cast<uint32_t>(s.strref & 0xFFFFFFu);
      const uint8_t* p = arena + (s.strref >> 24);
      keep = len > 0;  // tft never admits '' ...
      for (uint32_t k = 0; k < len && keep; ++k) keep = p[k] != '\n' && p[k] != '\r';  // ... nor multi-line terms
    }
  }
  unsigned long long kp = warp_append(occ && keep, &ctrl[3]);
  unsigned long long dp = warp_append(occ && !keep, &ctrl[6]);
  if (occ && keep) {
    ent[kp] = static_cast<uint32_t>(i);
    counts[kp] = cnt > 0xFFFFFFFFULL ? 0xFFFFFFFFu : static_cast<uint32_t>(cnt);
    if (cnt > 0xFFFFFFFFULL) atomicOr(&ctrl[2], 16ULL);  // the top-k cut-off is selected on 32-bit counts: refuse, never mis-rank
  } else if (occ) {
    ent[capacity - 1 - dp] = static_cast<uint32_t>(i);
  }
  unsigned occm = __ballot_sync(0xffffffffu, occ);
  if (lane_id() == 0 && occm) atomicAdd(&ctrl[4], static_cast<unsigned long long>(__popc(occm)));
}

__global__ void vocab_export_meta_kernel(const CountSlot* __restrict__ table, const uint32_t* __restrict__ ent, int64_t m,
                                         uint64_t* __restrict__ fp, uint64_t* __restrict__ tag,
                                         uint64_t* __restrict__ count, int32_t* __restrict__ len) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= m) return;
  CountSlot s = table[ent[i]];
  fp[i] = s.a; tag[i] = s.b; count[i] = s.count;
  len[i] = static_cast<int32_t>(s.strref & 0xFFFFFFu);
}

__global__ void vocab_export_bytes_kernel(const CountSlot* __restrict__ table, const uint8_t* __restrict__ arena,
                                          const uint32_t* __restrict__ ent, int64_t m,
                                          const int64_t* __restrict__ out_offsets, uint8_t* __restrict__ out_pool) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= m) return;
  CountSlot s = table[ent[i]];
  uint32_t len = static_cast<uint32_t>(s.strref & 0xFFFFFFu);
  const uint8_t* p = arena + (s.strref >> 24);
  uint8_t* o = out_pool + out_offsets[i];
  for (uint32_t k = 0; k < len; ++k) o[k] = p[k];
}

__global__ void vocab_filter_ge_kernel(const uint32_t* __restrict__ ent, const uint32_t* __restrict__ counts, int64_t m,
                                       const uint32_t* __restrict__ thr, uint32_t* __restrict__ out,
                                       unsigned long long* ctrl) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  bool keep = i < m && counts[i] >= *thr;
  unsigned long long pos = warp_append(keep, &ctrl[5]);
  if (keep) out[pos] = ent[i];
}

// (count, bytes) descending; kPadEnt sorts last
__device__ bool term_greater(const CountSlot* __restrict__ table, const uint8_t* __restrict__ arena, uint32_t a, uint32_t b) {
  if (a == b || a == kPadEnt) return false;
  if (b == kPadEnt) return true;
  CountSlot sa = table[a], sb = table[b];
  if (sa.count != sb.count) return sa.count > sb.count;
  uint32_t la = static_cast<uint32_t>(sa.strref & 0xFFFFFFu), lb = static_cast<uint32_t>(sb.strref & 0xFFFFFFu);
  const uint8_t* pa = arena + (sa.strref >> 24);
  const uint8_t* pb = arena + (sb.strref >> 24);
  uint32_t l = la < lb ? la : lb;
  for (uint32_t k = 0; k < l; ++k)
    if (pa[k] != pb[k]) return pa[k] > pb[k];
  return la > lb;
}

constexpr int kSortLocalMax = 4096;
constexpr int kSortBlock = 512;

__global__ void __launch_bounds__(kSortBlock) vocab_sort_local_kernel(const CountSlot* __restrict__ table,
                                                                      const uint8_t* __restrict__ arena,
                                                                      uint32_t* __restrict__ ent, int m, int mpad) {
  __shared__ uint32_t s[kSortLocalMax];
  for (int i = threadIdx.x; i < mpad; i += kSortBlock) s[i] = i < m ? ent[i] : kPadEnt;
  __syncthreads();
  for (int k = 2; k <= mpad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < mpad; i += kSortBlock) {
        int l = i ^ j;
        if (l > i) {
          uint32_t a = s[i], b = s[l];
          bool first_block = (i & k) == 0;  // descending runs in "first" blocks
          bool swap = first_block ? term_greater(table, arena, b, a) : term_greater(table, arena, a, b);
          if (swap) { s[i] = b; s[l] = a; }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < m; i += kSortBlock) ent[i] = s[i];
}

__global__ void vocab_sort_pad_kernel(const uint32_t* __restrict__ ent, int64_t m, int64_t mpad, uint32_t* __restrict__ buf) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i < mpad) buf[i] = i < m ? ent[i] : kPadEnt;
}
__global__ void vocab_sort_step_kernel(const CountSlot* __restrict__ table, const uint8_t* __restrict__ arena,
                                       uint32_t* __restrict__ buf, int64_t mpad, int64_t k, int64_t j) {
  int64_t t = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;  // one thread per pair
  if (t >= (mpad >> 1)) return;
  int64_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // index with bit j clear
  int64_t l = i | j;
  uint32_t a = buf[i], b = buf[l];
  bool first_block = (i & k) == 0;
  bool swap = first_block ? term_greater(table, arena, b, a) : term_greater(table, arena, a, b);
  if (swap) { buf[i] = b; buf[l] = a; }
}
__global__ void vocab_sort_unpad_kernel(const uint32_t* __restrict__ buf, int64_t m, uint32_t* __restrict__ ent) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i < m) ent[i] = buf[i];
}

// Every vocabulary entry gets a slot of its own (entries are distinct terms); strref points at the
// entry's bytes so that a fingerprint hit can be confirmed by whoever looks a term up.
__device__ void lookup_put(LookupSlot* lut, uint64_t mask, unsigned long long ka, unsigned long long kb, unsigned long long strref,
                           int32_t id) {
  uint64_t idx = home_slot(ka, kb, mask);
  for (uint64_t probe = 0; probe <= mask; ++probe, idx = (idx + 1) & mask) {
    unsigned long long cur = atomicCAS(&lut[idx].b, 0ULL, kb);  // b is never 0: it marks occupancy
    if (cur == 0) {
      lut[idx].a = ka;
      lut[idx].strref = strref;
      lut[idx].id = id;
      return;
    }
  }
}

__global__ void vocab_build_lookup_kernel(const CountSlot* __restrict__ table, const uint32_t* __restrict__ ent, int64_t k,
                                          LookupSlot* lut, uint64_t mask) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= k) return;
  CountSlot s = table[ent[i]];
  lookup_put(lut, mask, s.a, s.b, s.strref, static_cast<int32_t>(i));
}

__global__ void vocab_lookup_from_terms_kernel(const int32_t* __restrict__ offsets, const uint8_t* __restrict__ pool,
                                               int64_t k, LookupSlot* lut, uint64_t mask) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= k) return;
  const uint32_t len = static_cast<uint32_t>(offsets[i + 1] - offsets[i]);
  StrKey key = make_key(pool + offsets[i], len);
  lookup_put(lut, mask, key.a, key.b, (static_cast<unsigned long long>(offsets[i]) << 24) | (len & 0xFFFFFFu), static_cast<int32_t>(i));
}

// Dictionary decode table: for every code the count pass may have written (front-table slots of
// every CTA, then global slots) the final int64 output of compute_and_apply_vocabulary: the
// vocabulary id, or vocab_size + Fingerprint64(term) % num_oov_buckets, or default_value.
__global__ void vocab_final_codes_kernel(const CountSlot* __restrict__ table, int64_t capacity,
                                         const uint8_t* __restrict__ arena, const uint32_t* __restrict__ local2global,
                                         const LookupSlot* __restrict__ lut, uint64_t lmask, const uint8_t* __restrict__ lut_pool,
                                         long long vocab_size, long long num_oov, long long dflt, long long* __restrict__ final_code) {
  int64_t code = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (code >= static_cast<int64_t>(kVcLocalCodes) + capacity) return;
  uint32_t g = code < kVcLocalCodes ? local2global[code] : static_cast<uint32_t>(code - kVcLocalCodes);
  if (g == 0xFFFFFFFFu || g >= capacity) return;
  CountSlot s = table[g];
  if (s.a == 0 || s.b == 0) return;
  const uint8_t* str = arena + (s.strref >> 24);
  const uint32_t len = static_cast<uint32_t>(s.strref & 0xFFFFFFu);
  const bool fp = key_is_fingerprint(s.b);
  long long id = -1;
  uint64_t idx = home_slot(s.a, s.b, lmask);
  for (uint64_t probe = 0; probe <= lmask; ++probe, idx = (idx + 1) & lmask) {
    const LookupSlot l = lut[idx];
    if (l.b == 0) break;
    if (l.b != s.b || l.a != s.a) continue;
    if (fp && (static_cast<uint32_t>(l.strref & 0xFFFFFFu) != len || !term_bytes_equal(lut_pool + (l.strref >> 24), str, len))) continue;
    id = l.id;
    break;
  }
  if (id < 0) {
    if (num_oov > 0) {
      uint64_t h = fingerprint64(str, len);
      id = vocab_size + static_cast<long long>(h % static_cast<uint64_t>(num_oov));
    } else {
      id = dflt;
    }
  }
  final_code[code] = id;
}

int transform_set_identity_mask(unsigned long long ma, unsigned long long mb);  // transform.cu's copy of the test-hook masks

static bool pow2(int64_t x) { return x > 0 && (x & (x - 1)) == 0; }
static int64_t next_pow2(int64_t x) { int64_t p = 1; while (p < x) p <<= 1; return p; }
static unsigned blocks_for(int64_t n, int block) { return static_cast<unsigned>((n + block - 1) / block); }

}  // namespace txf

extern "C" {

size_t txf_vocab_table_bytes(int64_t capacity) { return capacity > 0 ? static_cast<size_t>(capacity) * sizeof(txf::CountSlot) : 0; }
size_t txf_vocab_lookup_bytes(int64_t lookup_cap) { return lookup_cap > 0 ? static_cast<size_t>(lookup_cap) * sizeof(txf::LookupSlot) : 0; }
size_t txf_vocab_sort_scratch_bytes(int64_t m) { return m > 0 ? static_cast<size_t>(txf::next_pow2(m)) * sizeof(uint32_t) : 0; }

int64_t txf_vocab_local_codes(void) { return txf::kVcLocalCodes; }

int txf_debug_set_identity_bits(int bits) {
  using namespace txf;
  TXF_REQUIRE(bits >= 1 && bits <= 119, "debug_set_identity_bits: bits=%d out of [1,119]", bits);
  const unsigned long long ma = bits >= 64 ? ~0ULL : (1ULL << bits) - 1;
  const unsigned long long mb = bits <= 64 ? 0ULL : (1ULL << (bits - 64)) - 1;
  TXF_CUDA(set_identity_mask_this_tu(ma, mb));
  return transform_set_identity_mask(ma, mb);
}

int txf_vocab_count(const int32_t* offsets_dev, const uint8_t* pool_dev, const uint8_t* valid_dev, int64_t n, void* table_dev,
                    int64_t capacity, uint8_t* arena_dev, int64_t arena_cap, uint64_t* ctrl_dev, uint32_t* codes_dev,
                    uint32_t* local2global_dev, int cta_warps, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "vocab_count: n out of range");
  TXF_REQUIRE(cta_warps == 0 || cta_warps == kVcWarpsFull || cta_warps == kVcWarpsShared, "vocab_count: cta_warps must be 0, %d or %d",
              kVcWarpsShared, kVcWarpsFull);
  TXF_REQUIRE(pow2(capacity) && capacity < (1LL << 32), "vocab_count: capacity must be a power of two below 2^32");
  TXF_REQUIRE(offsets_dev && pool_dev && table_dev && arena_dev && ctrl_dev && arena_cap > 0, "vocab_count: null pointer");
  TXF_REQUIRE(aligned(table_dev, 32), "vocab_count: table must be 32-byte aligned");
  TXF_REQUIRE((codes_dev == nullptr) == (local2global_dev == nullptr), "vocab_count: codes and local2global go together");
  if (local2global_dev != nullptr)
    TXF_CUDA(cudaMemsetAsync(local2global_dev, 0xFF, static_cast<size_t>(kVcLocalCodes) * sizeof(uint32_t),
                             static_cast<cudaStream_t>(stream)));
  if (n == 0) return TXF_OK;
  const int warps = cta_warps == 0 ? kVcWarpsFull : cta_warps;
  int64_t want = (n + static_cast<int64_t>(kVcStepRows) * warps - 1) / (static_cast<int64_t>(kVcStepRows) * warps);
  int grid = static_cast<int>(want < kVcGrid ? want : kVcGrid);
  auto launch = [&](auto kernel, int w) -> int {
    TXF_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(vc_smem_bytes(w))));
    TXF_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    kernel<<<grid, w * 32, vc_smem_bytes(w), static_cast<cudaStream_t>(stream)>>>(
        offsets_dev, pool_dev, valid_dev, n, static_cast<CountSlot*>(table_dev), static_cast<uint64_t>(capacity - 1), arena_dev,
        arena_cap, reinterpret_cast<unsigned long long*>(ctrl_dev), codes_dev, local2global_dev);
    return TXF_OK;
  };
  const int rc = warps == kVcWarpsFull ? launch(vocab_count_kernel<kVcWarpsFull>, kVcWarpsFull)
                                       : launch(vocab_count_kernel<kVcWarpsShared>, kVcWarpsShared);
  if (rc) return rc;
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_final_codes(const void* table_dev, int64_t capacity, const uint8_t* arena_dev, const uint32_t* local2global_dev,
                          const void* lookup_dev, int64_t lookup_cap, const uint8_t* lookup_pool_dev, int64_t vocab_size,
                          int64_t num_oov_buckets, int64_t default_value, int64_t* final_code_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(pow2(capacity) && capacity < (1LL << 31), "vocab_final_codes: capacity must be a power of two below 2^31");
  TXF_REQUIRE(pow2(lookup_cap), "vocab_final_codes: lookup_cap must be a power of two");
  TXF_REQUIRE(table_dev && arena_dev && local2global_dev && lookup_dev && lookup_pool_dev && final_code_dev,
              "vocab_final_codes: null pointer");
  int64_t total = static_cast<int64_t>(kVcLocalCodes) + capacity;
  vocab_final_codes_kernel<<<blocks_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const CountSlot*>(table_dev), capacity, arena_dev, local2global_dev, static_cast<const LookupSlot*>(lookup_dev),
      static_cast<uint64_t>(lookup_cap - 1), lookup_pool_dev, vocab_size, num_oov_buckets, default_value,
      reinterpret_cast<long long*>(final_code_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_export_blob(const void* table_dev, const uint8_t* arena_dev, const uint32_t* ent_dev, const uint64_t* ctrl_dev,
                          int64_t K, int64_t S, uint64_t* section_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(K >= 1 && S >= 8 && (S & 7) == 0, "vocab_export_blob: need K >= 1 and a byte slot that is a multiple of 8");
  TXF_REQUIRE(table_dev && arena_dev && ent_dev && ctrl_dev && section_dev, "vocab_export_blob: null pointer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(section_dev, 0, sizeof(uint64_t), s));
  vocab_export_blob_kernel<<<blocks_for(K, 256), 256, 0, s>>>(static_cast<const CountSlot*>(table_dev), arena_dev, ent_dev,
                                                              reinterpret_cast<const unsigned long long*>(ctrl_dev), K, S,
                                                              reinterpret_cast<unsigned long long*>(section_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_insert_blob(const uint64_t* blob_dev, int64_t stride_words, int64_t section_offset_words, int nranks, int64_t K,
                          int64_t S, void* table_dev, int64_t capacity, uint8_t* arena_dev, int64_t arena_cap,
                          uint64_t* ctrl_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(nranks >= 1 && K >= 1 && S >= 8 && (S & 7) == 0, "vocab_insert_blob: bad sizes");
  TXF_REQUIRE(pow2(capacity) && capacity < (1LL << 32), "vocab_insert_blob: capacity must be a power of two below 2^32");
  TXF_REQUIRE(blob_dev && table_dev && arena_dev && ctrl_dev, "vocab_insert_blob: null pointer");
  vocab_insert_blob_kernel<<<blocks_for(static_cast<int64_t>(nranks) * K, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const unsigned long long*>(blob_dev), stride_words, section_offset_words, nranks, K, S,
      static_cast<CountSlot*>(table_dev), static_cast<uint64_t>(capacity - 1), arena_dev, arena_cap,
      reinterpret_cast<unsigned long long*>(ctrl_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_insert(const uint64_t* fp_dev, const uint64_t* tag_dev, const uint64_t* count_dev, const int64_t* offsets_dev,
                     const uint8_t* pool_dev, int64_t m, int owners, int owner, void* table_dev, int64_t capacity, uint8_t* arena_dev,
                     int64_t arena_cap, uint64_t* ctrl_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(m >= 0, "vocab_insert: negative m");
  TXF_REQUIRE(owners >= 1 && owner >= 0 && owner < owners, "vocab_insert: owner %d out of [0,%d)", owner, owners);
  TXF_REQUIRE(pow2(capacity) && capacity < (1LL << 32), "vocab_insert: capacity must be a power of two below 2^32");
  if (m == 0) return TXF_OK;
  TXF_REQUIRE(fp_dev && tag_dev && count_dev && offsets_dev && pool_dev && table_dev && arena_dev && ctrl_dev, "vocab_insert: null pointer");
  vocab_insert_kernel<<<blocks_for(m, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      fp_dev, tag_dev, count_dev, offsets_dev, pool_dev, m, static_cast<uint32_t>(owners), static_cast<uint32_t>(owner),
      static_cast<CountSlot*>(table_dev), static_cast<uint64_t>(capacity - 1), arena_dev, arena_cap,
      reinterpret_cast<unsigned long long*>(ctrl_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_compact(const void* table_dev, int64_t capacity, const uint8_t* arena_dev, uint32_t* ent_dev, uint32_t* counts_dev,
                      uint64_t* ctrl_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(pow2(capacity) && capacity < (1LL << 32), "vocab_compact: capacity must be a power of two below 2^32");
  TXF_REQUIRE(table_dev && arena_dev && ent_dev && counts_dev && ctrl_dev, "vocab_compact: null pointer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(ctrl_dev + 3, 0, 5 * sizeof(uint64_t), s));
  vocab_compact_kernel<<<blocks_for(capacity, 256), 256, 0, s>>>(static_cast<const CountSlot*>(table_dev), capacity, arena_dev,
                                                                 ent_dev, counts_dev, reinterpret_cast<unsigned long long*>(ctrl_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_export_meta(const void* table_dev, const uint32_t* ent_dev, int64_t m, uint64_t* fp_dev, uint64_t* tag_dev,
                          uint64_t* count_dev, int32_t* len_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(m >= 0, "vocab_export_meta: negative m");
  if (m == 0) return TXF_OK;
  TXF_REQUIRE(table_dev && ent_dev && fp_dev && tag_dev && count_dev && len_dev, "vocab_export_meta: null pointer");
  vocab_export_meta_kernel<<<blocks_for(m, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const CountSlot*>(table_dev), ent_dev, m, fp_dev, tag_dev, count_dev, len_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_export_bytes(const void* table_dev, const uint8_t* arena_dev, const uint32_t* ent_dev, int64_t m,
                           const int64_t* out_offsets_dev, uint8_t* out_pool_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(m >= 0, "vocab_export_bytes: negative m");
  if (m == 0) return TXF_OK;
  TXF_REQUIRE(table_dev && arena_dev && ent_dev && out_offsets_dev && out_pool_dev, "vocab_export_bytes: null pointer");
  vocab_export_bytes_kernel<<<blocks_for(m, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const CountSlot*>(table_dev), arena_dev, ent_dev, m, out_offsets_dev, out_pool_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_filter_ge(const uint32_t* ent_dev, const uint32_t* counts_dev, int64_t m, const uint32_t* threshold_key_dev,
                        uint32_t* ent_out_dev, uint64_t* ctrl_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(m >= 0, "vocab_filter_ge: negative m");
  TXF_REQUIRE(ctrl_dev != nullptr, "vocab_filter_ge: null ctrl");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(ctrl_dev + 5, 0, sizeof(uint64_t), s));
  if (m == 0) return TXF_OK;
  TXF_REQUIRE(ent_dev && counts_dev && threshold_key_dev && ent_out_dev, "vocab_filter_ge: null pointer");
  vocab_filter_ge_kernel<<<blocks_for(m, 256), 256, 0, s>>>(ent_dev, counts_dev, m, threshold_key_dev, ent_out_dev,
                                                            reinterpret_cast<unsigned long long*>(ctrl_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_sort(const void* table_dev, const uint8_t* arena_dev, uint32_t* ent_dev, int64_t m, void* scratch_dev,
                   txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(m >= 0 && m < (1LL << 31), "vocab_sort: m out of range");
  if (m <= 1) return TXF_OK;
  TXF_REQUIRE(table_dev && arena_dev && ent_dev, "vocab_sort: null pointer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int64_t mpad = next_pow2(m);
  const CountSlot* table = static_cast<const CountSlot*>(table_dev);
  if (mpad <= kSortLocalMax) {
    vocab_sort_local_kernel<<<1, kSortBlock, 0, s>>>(table, arena_dev, ent_dev, static_cast<int>(m), static_cast<int>(mpad));
    TXF_LAUNCH_CHECK();
    count_launch(1);
    return TXF_OK;
  }
  TXF_REQUIRE(scratch_dev != nullptr, "vocab_sort: scratch required for m > %d", kSortLocalMax);
  uint32_t* buf = static_cast<uint32_t*>(scratch_dev);
  vocab_sort_pad_kernel<<<blocks_for(mpad, 256), 256, 0, s>>>(ent_dev, m, mpad, buf);
  int launches = 2;
  for (int64_t k = 2; k <= mpad; k <<= 1)
    for (int64_t j = k >> 1; j > 0; j >>= 1) {
      vocab_sort_step_kernel<<<blocks_for(mpad >> 1, 256), 256, 0, s>>>(table, arena_dev, buf, mpad, k, j);
      ++launches;
    }
  vocab_sort_unpad_kernel<<<blocks_for(m, 256), 256, 0, s>>>(buf, m, ent_dev);
  TXF_LAUNCH_CHECK();
  count_launch(launches);
  return TXF_OK;
}

int txf_vocab_build_lookup(const void* table_dev, const uint32_t* ent_sorted_dev, int64_t k, void* lookup_dev,
                           int64_t lookup_cap, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(k >= 0, "vocab_build_lookup: negative k");
  TXF_REQUIRE(pow2(lookup_cap) && lookup_cap >= 2 * k && lookup_cap >= 2, "vocab_build_lookup: lookup_cap must be a power of two >= 2k");
  TXF_REQUIRE(lookup_dev != nullptr && aligned(lookup_dev, 8), "vocab_build_lookup: lookup table null or misaligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(lookup_dev, 0, txf_vocab_lookup_bytes(lookup_cap), s));
  if (k == 0) return TXF_OK;
  TXF_REQUIRE(table_dev && ent_sorted_dev, "vocab_build_lookup: null pointer");
  vocab_build_lookup_kernel<<<blocks_for(k, 256), 256, 0, s>>>(static_cast<const CountSlot*>(table_dev), ent_sorted_dev, k,
                                                               static_cast<LookupSlot*>(lookup_dev),
                                                               static_cast<uint64_t>(lookup_cap - 1));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_lookup_from_terms(const int32_t* voc_offsets_dev, const uint8_t* voc_pool_dev, int64_t k, void* lookup_dev,
                                int64_t lookup_cap, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(k >= 0, "vocab_lookup_from_terms: negative k");
  TXF_REQUIRE(pow2(lookup_cap) && lookup_cap >= 2 * k && lookup_cap >= 2, "vocab_lookup_from_terms: lookup_cap must be a power of two >= 2k");
  TXF_REQUIRE(lookup_dev != nullptr && aligned(lookup_dev, 8), "vocab_lookup_from_terms: lookup table null or misaligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(lookup_dev, 0, txf_vocab_lookup_bytes(lookup_cap), s));
  if (k == 0) return TXF_OK;
  TXF_REQUIRE(voc_offsets_dev && voc_pool_dev, "vocab_lookup_from_terms: null pointer");
  vocab_lookup_from_terms_kernel<<<blocks_for(k, 256), 256, 0, s>>>(voc_offsets_dev, voc_pool_dev, k,
                                                                    static_cast<LookupSlot*>(lookup_dev),
                                                                    static_cast<uint64_t>(lookup_cap - 1));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
