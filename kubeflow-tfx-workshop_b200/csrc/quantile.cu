// quantile.cu -- quantile-boundary analyzer of tft.bucketize (reference call site
// airflow-dags/taxi_utils.py:128-130; fill rule :86-103).
//
// Two engines, both EXACT (they return true order statistics, so the result is independent of
// chunking, grid size and the number of GPUs; per-shard state merges by integer addition):
//
//  (1) txf_msel_*: multi-rank radix select, 4 passes x 8 bits over an order-preserving uint32
//      key.  Works for any number of ranks; every element costs a shared-memory (or, for many
//      ranks, global) atomic per pass.  Used on small inputs (the sample, the candidates, the
//      vocabulary counts) and as the general fallback.
//  (2) txf_qsample / txf_qcount / txf_qresolve / txf_qfinal: sample-select.  A stratified
//      sample fixes a narrow bracket [lo_i, hi_i] around every target rank; ONE streaming pass
//      over the column counts, per thread and without atomics, how many keys fall in each cell
//      of the bracket grid and appends the few keys strictly inside a bracket to a candidate
//      list; the exact answer is then an edge or a small selection among the candidates.
//      Algorithmic bytes of that pass: 4 per row per f32 column + 1 validity bit.
#include "common.cuh"

namespace txf {

// --------------------------------------------------------------------------------------------
// element access shared by both engines
// --------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cell_key_f32(float v, bool present) { return f32_key(present ? v : 0.f); }
__device__ __forceinline__ uint32_t cell_key_i64(long long v, bool present) {
  return f32_key(present ? i64_to_f32(v) : 0.f);
}

// Calls f(key) for every row of column (values, valid, dtype) assigned to this thread;
// quads are interleaved across the whole grid.x so loads coalesce into full lines.
template <typename F>
__device__ __forceinline__ void for_each_key(const void* values, const uint8_t* valid, int dtype, int64_t n, F f) {
  const int64_t nquads = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  int64_t q = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (dtype == TXF_F32) {
    const float* __restrict__ v = static_cast<const float*>(values);
    for (; q + stride < nquads; q += 2 * stride) {
      float4 a = ldg_f4(v + 4 * q), b = ldg_f4(v + 4 * (q + stride));
      uint32_t ma = quad_valid(valid, 4 * q), mb = quad_valid(valid, 4 * (q + stride));
      f(cell_key_f32(a.x, ma & 1)); f(cell_key_f32(a.y, ma & 2)); f(cell_key_f32(a.z, ma & 4)); f(cell_key_f32(a.w, ma & 8));
      f(cell_key_f32(b.x, mb & 1)); f(cell_key_f32(b.y, mb & 2)); f(cell_key_f32(b.z, mb & 4)); f(cell_key_f32(b.w, mb & 8));
    }
    for (; q < nquads; q += stride) {
      float4 a = ldg_f4(v + 4 * q);
      uint32_t ma = quad_valid(valid, 4 * q);
      f(cell_key_f32(a.x, ma & 1)); f(cell_key_f32(a.y, ma & 2)); f(cell_key_f32(a.z, ma & 4)); f(cell_key_f32(a.w, ma & 8));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
      int64_t i = (nquads << 2) + threadIdx.x;
      f(cell_key_f32(v[i], row_valid(valid, i)));
    }
  } else {
    const long long* __restrict__ v = static_cast<const long long*>(values);
    for (; q < nquads; q += stride) {
      longlong2 a = ldg_l2(v + 4 * q), b = ldg_l2(v + 4 * q + 2);
      uint32_t ma = quad_valid(valid, 4 * q);
      f(cell_key_i64(a.x, ma & 1)); f(cell_key_i64(a.y, ma & 2)); f(cell_key_i64(b.x, ma & 4)); f(cell_key_i64(b.y, ma & 8));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
      int64_t i = (nquads << 2) + threadIdx.x;
      f(cell_key_i64(v[i], row_valid(valid, i)));
    }
  }
}

template <typename F>
__device__ __forceinline__ void for_each_raw_key(const uint32_t* __restrict__ keys, int64_t n, F f) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n; i += stride) f(keys[i]);
}

// --------------------------------------------------------------------------------------------
// (1) multi-rank radix select
// --------------------------------------------------------------------------------------------
struct MselLayout {
  size_t hist, rem, prefix, uprefix, slot, ucount, total;
};
__host__ __device__ inline MselLayout msel_layout(int ncols, int R) {
  MselLayout L;
  size_t cr = static_cast<size_t>(ncols) * R;
  L.hist = 0;
  L.rem = L.hist + cr * 256 * sizeof(uint64_t);
  L.prefix = L.rem + cr * sizeof(uint64_t);
  L.uprefix = L.prefix + cr * sizeof(uint32_t);
  L.slot = L.uprefix + cr * sizeof(uint32_t);
  L.ucount = L.slot + cr * sizeof(int32_t);
  L.total = L.ucount + static_cast<size_t>(ncols) * sizeof(int32_t);
  L.total = (L.total + 255) & ~static_cast<size_t>(255);
  return L;
}

constexpr int kMselBlock = 256;
constexpr int kMselSmemSlots = 32;  // privatised histogram for up to 32 distinct prefixes (32 KB)

__global__ void msel_begin_kernel(uint8_t* state, int ncols, int R, const uint64_t* __restrict__ ranks) {
  MselLayout L = msel_layout(ncols, R);
  int c = blockIdx.x;
  uint64_t* rem = reinterpret_cast<uint64_t*>(state + L.rem) + static_cast<size_t>(c) * R;
  uint32_t* prefix = reinterpret_cast<uint32_t*>(state + L.prefix) + static_cast<size_t>(c) * R;
  uint32_t* uprefix = reinterpret_cast<uint32_t*>(state + L.uprefix) + static_cast<size_t>(c) * R;
  int32_t* slot = reinterpret_cast<int32_t*>(state + L.slot) + static_cast<size_t>(c) * R;
  int32_t* ucount = reinterpret_cast<int32_t*>(state + L.ucount);
  __shared__ int any;
  if (threadIdx.x == 0) any = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < R; i += blockDim.x) {
    uint64_t r = ranks[static_cast<size_t>(c) * R + i];
    rem[i] = r;
    prefix[i] = 0;
    uprefix[i] = 0;
    slot[i] = r ? 0 : -1;
    if (r) any = 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) ucount[c] = any;
}

template <int KIND>
__global__ void __launch_bounds__(kMselBlock) msel_hist_kernel(uint8_t* state, int ncols, int R, int pass,
                                                               NumColsView cols, int64_t n,
                                                               const uint64_t* __restrict__ n_dev) {
  extern __shared__ uint32_t smem[];
  MselLayout L = msel_layout(ncols, R);
  const int c = blockIdx.y;
  const int U = reinterpret_cast<const int32_t*>(state + L.ucount)[c];
  if (U == 0) return;
  const uint32_t* g_upref = reinterpret_cast<const uint32_t*>(state + L.uprefix) + static_cast<size_t>(c) * R;
  unsigned long long* hist = reinterpret_cast<unsigned long long*>(state + L.hist) + static_cast<size_t>(c) * R * 256;
  uint32_t* s_upref = smem;                    // [R]
  uint32_t* s_hist = smem + R;                 // [min(U,32)][256]
  const bool priv = U <= kMselSmemSlots;
  for (int i = threadIdx.x; i < U; i += kMselBlock) s_upref[i] = g_upref[i];
  if (priv)
    for (int i = threadIdx.x; i < U * 256; i += kMselBlock) s_hist[i] = 0;
  __syncthreads();
  const int pshift = 32 - 8 * pass;  // prefix = key >> pshift (pass 0: everything shares prefix 0)
  const int dshift = 24 - 8 * pass;
  if (KIND == 1 && n_dev != nullptr) {
    int64_t nd = static_cast<int64_t>(n_dev[c]);
    n = nd < n ? nd : n;
  }
  auto visit = [&](uint32_t key) {
    uint32_t pre = pass == 0 ? 0u : (key >> pshift);
    int lo = 0, len = U;  // lower_bound over the sorted distinct prefixes
    while (len > 0) {
      int half = len >> 1;
      bool lt = s_upref[lo + half] < pre;
      lo = lt ? lo + half + 1 : lo;
      len = lt ? len - half - 1 : half;
    }
    if (lo < U && s_upref[lo] == pre) {
      uint32_t bin = (static_cast<uint32_t>(lo) << 8) | ((key >> dshift) & 255u);
      if (priv) {
        atomicAdd(&s_hist[bin], 1u);
      } else {  // many live prefixes: aggregate equal bins inside the warp, one global atomic each
        unsigned peers = __match_any_sync(__activemask(), bin);
        if (static_cast<int>(lane_id()) == __ffs(peers) - 1) atomicAdd(&hist[bin], static_cast<unsigned long long>(__popc(peers)));
      }
    }
  };
  if (KIND == 0)
    for_each_key(cols.values[c], cols.valid[c], cols.dtype[c], n, visit);
  else
    for_each_raw_key(static_cast<const uint32_t*>(cols.values[c]), n, visit);
  if (priv) {
    __syncthreads();
    for (int i = threadIdx.x; i < U * 256; i += kMselBlock) {
      uint32_t v = s_hist[i];
      if (v) atomicAdd(&hist[i], static_cast<unsigned long long>(v));
    }
  }
}

__global__ void msel_resolve_kernel(uint8_t* state, int ncols, int R, int pass) {
  MselLayout L = msel_layout(ncols, R);
  const int c = blockIdx.x;
  unsigned long long* hist = reinterpret_cast<unsigned long long*>(state + L.hist) + static_cast<size_t>(c) * R * 256;
  uint64_t* rem = reinterpret_cast<uint64_t*>(state + L.rem) + static_cast<size_t>(c) * R;
  uint32_t* prefix = reinterpret_cast<uint32_t*>(state + L.prefix) + static_cast<size_t>(c) * R;
  uint32_t* uprefix = reinterpret_cast<uint32_t*>(state + L.uprefix) + static_cast<size_t>(c) * R;
  int32_t* slot = reinterpret_cast<int32_t*>(state + L.slot) + static_cast<size_t>(c) * R;
  int32_t* ucount = reinterpret_cast<int32_t*>(state + L.ucount);
  const int U = ucount[c];
  // one warp per rank: lane l owns bins 8l..8l+7; a warp scan finds the bin holding the rank
  const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int i = warp; i < R; i += nwarps) {
    const int s = slot[i];
    if (s < 0) continue;  // warp-uniform
    const unsigned long long* h = hist + static_cast<size_t>(s) * 256 + 8 * lane_id();
    unsigned long long v[8], mine = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) { v[b] = h[b]; mine += v[b]; }
    unsigned long long incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      unsigned long long o = __shfl_up_sync(0xffffffffu, incl, off);
      if (static_cast<int>(lane_id()) >= off) incl += o;
    }
    const unsigned long long r = rem[i];
    const unsigned long long excl = incl - mine;
    const unsigned ok = __ballot_sync(0xffffffffu, r <= incl);  // first lane whose prefix reaches the rank
    const int owner = ok ? __ffs(ok) - 1 : 31;                  // rank beyond the data: last bin, result meaningless
    if (static_cast<int>(lane_id()) == owner) {
      unsigned long long cum = excl;
      int d = 7;
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        if (r <= cum + v[b]) { d = b; break; }
        cum += v[b];
      }
      if (!ok) cum = 0;
      prefix[i] = (prefix[i] << 8) | static_cast<uint32_t>(8 * owner + d);
      rem[i] = r - cum;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {  // rebuild the distinct-prefix table; ranks ascend, so prefixes ascend
    int u = 0;
    bool first = true;
    uint32_t last = 0;
    for (int i = 0; i < R; ++i) {
      if (slot[i] < 0) continue;
      uint32_t p = prefix[i];
      if (first || p != last) { uprefix[u++] = p; last = p; first = false; }
      slot[i] = u - 1;
    }
    ucount[c] = u;
  }
  for (int i = threadIdx.x; i < U * 256; i += blockDim.x) hist[i] = 0;  // ready for the next pass
  (void)pass;
}

__global__ void msel_result_kernel(const uint8_t* state, int ncols, int R, uint32_t* __restrict__ out) {
  MselLayout L = msel_layout(ncols, R);
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ncols * R) return;
  const uint32_t* prefix = reinterpret_cast<const uint32_t*>(state + L.prefix);
  const int32_t* slot = reinterpret_cast<const int32_t*>(state + L.slot);
  out[i] = slot[i] >= 0 ? prefix[i] : 0u;
}

// --------------------------------------------------------------------------------------------
// (2) sample-select
// --------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}

__global__ void qsample_kernel(NumColsView cols, int64_t n, int64_t M, uint64_t seed, uint32_t* __restrict__ out) {
  const int c = blockIdx.y;
  int64_t j = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (j >= M) return;
  const int64_t stride = n / M;  // >= 1 (caller guarantees M <= n)
  int64_t row = j * stride + static_cast<int64_t>(mix64(seed ^ (static_cast<uint64_t>(j) * 0x2545F4914F6CDD1DULL)) %
                                                  static_cast<uint64_t>(stride));
  bool present = row_valid(cols.valid[c], row);
  uint32_t key = cols.dtype[c] == TXF_F32
                     ? cell_key_f32(static_cast<const float*>(cols.values[c])[row], present)
                     : cell_key_i64(static_cast<const long long*>(cols.values[c])[row], present);
  out[static_cast<size_t>(c) * M + j] = key;
}

constexpr int kQBlock = 256;
constexpr int kQLutBits = 12;            // 4096-entry bin table over [first edge, last edge]
constexpr int kQTileQuads = 2 * kQBlock; // a tile = 2 quads (8 keys) per thread
constexpr int kQCandBuf = 2048;          // CTA-local candidate buffer (keys); flushed when half full (8 KB: five CTAs per SM)

template <bool F32>
__device__ __forceinline__ float4 load_quad_as_f32(const void* values, int64_t q) {
  if (F32) return ldg_f4(static_cast<const float*>(values) + 4 * q);
  const long long* v = static_cast<const long long*>(values) + 4 * q;
  longlong2 a = ldg_l2(v), b = ldg_l2(v + 2);
  return make_float4(i64_to_f32(a.x), i64_to_f32(a.y), i64_to_f32(b.x), i64_to_f32(b.y));
}

// cells: 2p = strictly between edge p-1 and edge p (p = 0..E), 2p+1 = equal to edge p (p < E).
// Locating a key among the edges is a table lookup, not a search: the key range spanned by the
// edges is cut into 1024 equal bins; each bin stores how many edges lie below it and whether an
// edge falls inside it (only then are a few edges compared).  Keys strictly inside a bracket are
// collected in a CTA-local shared-memory buffer and written out in batches, so the streaming
// loop never waits on a global atomic; the next tile's loads are issued before the current tile
// is processed.
template <bool F32>
__global__ void __launch_bounds__(kQBlock) qcount_kernel(NumColsView cols, int64_t n, int T,
                                                         const uint32_t* __restrict__ edges_all,
                                                         unsigned long long* __restrict__ cells_all,
                                                         uint32_t* __restrict__ cand_all,
                                                         unsigned long long* __restrict__ cand_count, int64_t cand_cap) {
  extern __shared__ uint32_t smem[];
  __shared__ uint32_t s_ncand;
  __shared__ unsigned long long s_base;
  const int c = blockIdx.y;
  if ((cols.dtype[c] == TXF_F32) != F32) return;  // the other instantiation handles this column
  const int E = 2 * T, ncell = 2 * E + 1;
  uint32_t* s_edges = smem;                                                     // [E]
  uint32_t* s_cand = smem + ((E + 1) & ~1);                                     // [kQCandBuf]
  uint16_t* s_lut = reinterpret_cast<uint16_t*>(s_cand + kQCandBuf);            // [1 << kQLutBits]
  uint16_t* s_cnt = s_lut + (1 << kQLutBits);                                   // [ncell][kQBlock], private per thread
  for (int i = threadIdx.x; i < E; i += kQBlock) s_edges[i] = edges_all[static_cast<size_t>(c) * E + i];
  for (int i = threadIdx.x; i < ncell * kQBlock; i += kQBlock) s_cnt[i] = 0;
  if (threadIdx.x == 0) s_ncand = 0;
  __syncthreads();
  const uint32_t first = s_edges[0], last = s_edges[E - 1];
  const uint32_t span = last - first;
  const int shift = span >> kQLutBits ? 32 - kQLutBits - __clz(span) : 0;  // (span >> shift) < 1024
  for (int j = threadIdx.x; j < (1 << kQLutBits); j += kQBlock) {
    const uint64_t lo_key = static_cast<uint64_t>(first) + (static_cast<uint64_t>(j) << shift);
    const uint64_t hi_key = lo_key + ((1ULL << shift) - 1);
    int p = 0;  // edges below the bin
    while (p < E && s_edges[p] < lo_key) ++p;
    const bool inside = p < E && s_edges[p] <= hi_key;
    s_lut[j] = static_cast<uint16_t>(p | (inside ? 0x8000 : 0));
  }
  __syncthreads();
  uint32_t* cand = cand_all + static_cast<size_t>(c) * cand_cap;
  const void* values = cols.values[c];
  const uint8_t* __restrict__ valid = cols.valid[c];
  // private 16-bit counter of (cell, thread): lanes of a warp sit in 32 different banks, warps w
  // and w+4 share the words (low/high half), so the read-modify-write never bank-conflicts
  const int my_slot = (((threadIdx.x >> 5) & 3) * 32 + static_cast<int>(lane_id())) * 2 + (threadIdx.x >> 7);
  const uint32_t cnt_addr = smem_addr(s_cnt) + 2u * static_cast<uint32_t>(my_slot);
  // table entry of a key: number of edges below it (exact unless bit 15 says the bin holds edges)
  auto lookup = [&](uint32_t key) -> uint32_t {
    const uint32_t d = key - first;  // wraps below the first edge; the table index is clamped, the result overridden
    uint32_t bin = d >> shift;
    bin = bin < (1u << kQLutBits) ? bin : (1u << kQLutBits) - 1;
    uint32_t t = s_lut[bin];
    t = key < first ? 0u : t;
    return key > last ? static_cast<uint32_t>(E) : t;
  };
  auto count = [&](uint32_t key, uint32_t t) {
    int lo = t & 0x7FFF;
    bool eq = false;
    if (t & 0x8000) {
      uint32_t e = s_edges[lo];
      while (e < key) {
        if (++lo == E) break;
        e = s_edges[lo];
      }
      eq = lo < E && e == key;
    }
    const int cell = 2 * lo + (eq ? 1 : 0);
    {  // private counter += 1 through explicit shared-window addresses (<= 65535 keys per thread per launch: grid sizing)
      const uint32_t a = cnt_addr + static_cast<uint32_t>(cell) * (2u * kQBlock);
      sts_u16(a, lds_u16(a) + 1u);
    }
    if ((lo & 1) && !eq) {  // strictly inside bracket (lo-1)/2: keep the key
      const uint32_t pos = atomicAdd(&s_ncand, 1u);
      if (pos < kQCandBuf) {
        s_cand[pos] = key;
      } else {  // CTA buffer full between two flushes (rare): append to the global list directly
        const unsigned long long g = atomicAdd(&cand_count[c], 1ULL);
        if (g < static_cast<unsigned long long>(cand_cap)) cand[g] = key;
      }
    }
  };
  auto visit = [&](uint32_t key) { count(key, lookup(key)); };
  // the four table lookups of a quad are issued together (independent loads), then the counters
  auto visit_quad = [&](float4 x, uint32_t vbyte, uint32_t q) {  // vbyte: raw bitmap byte of the quad (prefetched)
    const uint32_t m = valid == nullptr ? 0xFu : (vbyte >> ((4 * q) & 4)) & 0xFu;
    const uint32_t k0 = f32_key(m & 1 ? x.x : 0.f), k1 = f32_key(m & 2 ? x.y : 0.f), k2 = f32_key(m & 4 ? x.z : 0.f),
                   k3 = f32_key(m & 8 ? x.w : 0.f);
    const uint32_t t0 = lookup(k0), t1 = lookup(k1), t2 = lookup(k2), t3 = lookup(k3);
    count(k0, t0); count(k1, t1); count(k2, t2); count(k3, t3);
  };
  auto flush = [&]() {  // all threads; s_ncand is stable on entry
    if (threadIdx.x == 0) s_base = atomicAdd(&cand_count[c], static_cast<unsigned long long>(s_ncand < kQCandBuf ? s_ncand : kQCandBuf));
    __syncthreads();
    const uint32_t m = s_ncand < kQCandBuf ? s_ncand : kQCandBuf;
    for (uint32_t i = threadIdx.x; i < m; i += kQBlock) {
      unsigned long long pos = s_base + i;
      if (pos < static_cast<unsigned long long>(cand_cap)) cand[pos] = s_cand[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) s_ncand = 0;
    __syncthreads();
  };

  // n < 2^31 (checked by the entry point): quads and tiles are 32-bit inside the loop
  const uint32_t nquads = static_cast<uint32_t>(n >> 2);
  const uint32_t ntiles = (nquads + kQTileQuads - 1) / kQTileQuads;
  uint32_t tile = blockIdx.x;
  float4 xa = make_float4(0, 0, 0, 0), xb = xa;
  uint32_t va = 0, vb = 0;  // raw validity bytes: decoded at use, so the prefetch never waits on them
  bool ha = false, hb = false;
  auto load_tile = [&](uint32_t t) {
    const uint32_t qa = t * kQTileQuads + threadIdx.x, qb = qa + kQBlock;
    ha = t < ntiles && qa < nquads;
    hb = t < ntiles && qb < nquads;
    if (ha) { xa = load_quad_as_f32<F32>(values, qa); if (valid) va = __ldg(valid + (qa >> 1)); }
    if (hb) { xb = load_quad_as_f32<F32>(values, qb); if (valid) vb = __ldg(valid + (qb >> 1)); }
  };
  load_tile(tile);
  for (uint32_t it = 1; tile < ntiles; tile += gridDim.x, ++it) {
    const float4 ca = xa, cb = xb;
    const uint32_t cva = va, cvb = vb;
    const bool cha = ha, chb = hb;
    const uint32_t qa = tile * kQTileQuads + threadIdx.x;
    load_tile(tile + gridDim.x);  // prefetch
    if (cha) visit_quad(ca, cva, qa);
    if (chb) visit_quad(cb, cvb, qa + kQBlock);
    // the CTA buffer is drained every 4 tiles when more than half full (overflow goes to global directly)
    if ((it & 3) == 0 && __syncthreads_or(s_ncand > kQCandBuf / 2)) flush();
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // rows past the last full quad
    const int64_t i = (static_cast<int64_t>(nquads) << 2) + threadIdx.x;
    const bool present = row_valid(valid, i);
    visit(F32 ? cell_key_f32(static_cast<const float*>(values)[i], present)
              : cell_key_i64(static_cast<const long long*>(values)[i], present));
  }
  __syncthreads();
  flush();
  // fold the private counters: warp w takes cells w, w+8, ...; lanes stride over the 256 threads
  const int warp = threadIdx.x >> 5;
  for (int cell = warp; cell < ncell; cell += kQBlock / 32) {
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < kQBlock / 32; ++j) s += s_cnt[cell * kQBlock + j * 32 + lane_id()];
#pragma unroll
    for (int off = 16; off; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
    if (lane_id() == 0 && s) atomicAdd(&cells_all[static_cast<size_t>(c) * ncell + cell], static_cast<unsigned long long>(s));
  }
}

__global__ void qresolve_kernel(const unsigned long long* __restrict__ cells_all, const uint32_t* __restrict__ edges_all,
                                const unsigned long long* __restrict__ cand_count, int64_t cand_cap,
                                const uint64_t* __restrict__ ranks, int ncols, int T, uint32_t* __restrict__ direct_key,
                                uint64_t* __restrict__ cand_rank, int32_t* __restrict__ status) {
  const int c = blockIdx.x;
  const int E = 2 * T, ncell = 2 * E + 1;
  const unsigned long long* cells = cells_all + static_cast<size_t>(c) * ncell;
  const uint32_t* edges = edges_all + static_cast<size_t>(c) * E;
  __shared__ int bad;
  if (threadIdx.x == 0) bad = (cand_count[c] > static_cast<unsigned long long>(cand_cap)) ? 2 : 0;
  __syncthreads();
  for (int i = threadIdx.x; i < T; i += blockDim.x) {
    const int jl = 2 * i, jh = 2 * i + 1;
    const uint32_t el = edges[jl], eh = edges[jh];
    // number of keys < edge j and == edge j, with duplicate edges collapsing onto the first copy
    auto first_of = [&](int j) { while (j > 0 && edges[j - 1] == edges[j]) --j; return j; };
    auto count_lt = [&](int j) {
      int f = first_of(j);
      unsigned long long s = cells[2 * f];
      for (int p = 0; p < f; ++p) s += cells[2 * p] + cells[2 * p + 1];
      return s;
    };
    auto count_eq = [&](int j) { return cells[2 * first_of(j) + 1]; };
    const unsigned long long r = ranks[i];
    const unsigned long long Ll = count_lt(jl), El = count_eq(jl);
    uint32_t dk = 0;
    uint64_t cr = 0;
    int miss = 0;
    if (r <= Ll) {
      miss = 1;
    } else if (r <= Ll + El) {
      dk = el;
    } else if (eh == el) {
      miss = 1;
    } else {
      const unsigned long long Lh = count_lt(jh), Eh = count_eq(jh);
      if (r <= Lh) {
        unsigned long long before = 0;  // candidates of earlier brackets precede ours in key order
        for (int k = 0; k < i; ++k) before += cells[2 * (2 * k + 1)];
        cr = before + (r - Ll - El);
      } else if (r <= Lh + Eh) {
        dk = eh;
      } else {
        miss = 1;
      }
    }
    direct_key[static_cast<size_t>(c) * T + i] = dk;
    cand_rank[static_cast<size_t>(c) * T + i] = cr;
    if (miss) bad = 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) status[c] = bad;
}

// --------------------------------------------------------------------------------------------
// Bracket select: the exact order statistic inside each bracket from ONE histogram of the candidates.
// A bracket [lo, hi] around a target is narrow in KEY space on real columns (a handful of distinct
// values between two order statistics of the sample that are 0.5 % of the rows apart), so the candidate
// keys of bracket p are histogrammed by (key - lo - 1) into B bins; if every bracket spans at most B keys
// the histogram IS the exact distribution and one prefix scan finds the target -- one integer sum to merge
// across ranks instead of the four dependent digit passes of the radix select.  A wider bracket raises the
// column's overflow word and the caller falls back to txf_msel_* on the candidates.
// hist layout (u32): [ncols][T][B] counts, then [ncols] overflow words.
// --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bsel_hist_kernel(const uint32_t* __restrict__ cand_all, int64_t cand_cap,
                                                        const unsigned long long* __restrict__ cand_count,
                                                        const uint32_t* __restrict__ edges_all, int ncols, int T, int B,
                                                        uint32_t* __restrict__ hist) {
  extern __shared__ uint32_t bs_edges[];  // [2T]
  // Columns with few distinct values (centroids) send all candidates of a bracket to a handful of bins: counted with
  // global atomics they would serialise on those addresses.  A small CTA-local table absorbs them (open addressing,
  // 4 probes); bins that do not fit go to the global histogram directly -- fine-grained columns spread out anyway.
  constexpr int kSlots = 1024;
  __shared__ uint32_t s_key[kSlots], s_cnt[kSlots];
  const int c = blockIdx.y;
  const int E = 2 * T;
  for (int i = threadIdx.x; i < E; i += 256) bs_edges[i] = edges_all[static_cast<size_t>(c) * E + i];
  for (int i = threadIdx.x; i < kSlots; i += 256) { s_key[i] = 0xFFFFFFFFu; s_cnt[i] = 0; }
  __syncthreads();
  const unsigned long long have = cand_count[c];
  const int64_t m = have < static_cast<unsigned long long>(cand_cap) ? static_cast<int64_t>(have) : cand_cap;
  const uint32_t* cand = cand_all + static_cast<size_t>(c) * cand_cap;
  uint32_t* h = hist + static_cast<size_t>(c) * T * B;
  uint32_t* over = hist + static_cast<size_t>(ncols) * T * B + c;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * 256;
  for (int64_t i = blockIdx.x * 256LL + threadIdx.x; i < m; i += stride) {
    const uint32_t key = cand[i];
    int lo = 0, len = E;  // number of edges strictly below the key
    while (len > 0) {
      const int half = len >> 1;
      const bool lt = bs_edges[lo + half] < key;
      lo = lt ? lo + half + 1 : lo;
      len = lt ? len - half - 1 : half;
    }
    const int p = (lo & 1) ? (lo - 1) >> 1 : 0;  // candidates lie strictly inside a bracket: lo is odd
    const uint32_t bin = key - bs_edges[2 * p] - 1u;
    if (!(lo & 1) || bin >= static_cast<uint32_t>(B)) {
      atomicOr(over, 1u);
      continue;
    }
    const uint32_t idx = static_cast<uint32_t>(p) * static_cast<uint32_t>(B) + bin;
    uint32_t slot = (idx * 2654435761u) >> 22;
    bool placed = false;
#pragma unroll
    for (int probe = 0; probe < 4 && !placed; ++probe, slot = (slot + 1) & (kSlots - 1)) {
      uint32_t cur = s_key[slot];
      if (cur == 0xFFFFFFFFu) cur = atomicCAS(&s_key[slot], 0xFFFFFFFFu, idx);
      if (cur == 0xFFFFFFFFu || cur == idx) { atomicAdd(&s_cnt[slot], 1u); placed = true; }
    }
    if (!placed) atomicAdd(&h[idx], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kSlots; i += 256)
    if (s_cnt[i]) atomicAdd(&h[s_key[i]], s_cnt[i]);
}

// one CTA per (target, column): rank of the target among its bracket's candidates -> bin -> key
__global__ void __launch_bounds__(256) bsel_resolve_kernel(const uint32_t* __restrict__ hist, const unsigned long long* __restrict__ cells_all,
                                                           const uint32_t* __restrict__ edges_all, const uint64_t* __restrict__ cand_rank,
                                                           int ncols, int T, int B, uint32_t* __restrict__ selected,
                                                           int32_t* __restrict__ status) {
  const int p = blockIdx.x, c = blockIdx.y;
  const int E = 2 * T, ncell = 2 * E + 1;
  const size_t k = static_cast<size_t>(c) * T + p;
  const uint64_t cr = cand_rank[k];
  if (cr == 0) {  // answered by an edge
    if (threadIdx.x == 0) selected[k] = 0;
    return;
  }
  const uint32_t over = hist[static_cast<size_t>(ncols) * T * B + c];
  if (over) {
    if (threadIdx.x == 0) { selected[k] = 0; atomicOr(&status[c], 4); }
    return;
  }
  const unsigned long long* cells = cells_all + static_cast<size_t>(c) * ncell;
  unsigned long long before = 0;  // candidates of earlier brackets precede this bracket's in key order
  for (int q = 0; q < p; ++q) before += cells[2 * (2 * q + 1)];
  const unsigned long long r = cr - before;  // 1-based rank inside the bracket
  const uint32_t* h = hist + k * B;
  __shared__ unsigned long long s_part[256];
  const int per = (B + 255) / 256, b0 = threadIdx.x * per;
  unsigned long long sum = 0;
  for (int i = b0; i < b0 + per && i < B; ++i) sum += h[i];
  s_part[threadIdx.x] = sum;
  __syncthreads();
  for (int off = 1; off < 256; off <<= 1) {
    const unsigned long long v = threadIdx.x >= off ? s_part[threadIdx.x - off] : 0;
    __syncthreads();
    s_part[threadIdx.x] += v;
    __syncthreads();
  }
  const unsigned long long excl = threadIdx.x ? s_part[threadIdx.x - 1] : 0;
  if (r > excl && r <= s_part[threadIdx.x]) {  // exactly one thread owns the target
    unsigned long long run = excl;
    for (int i = b0; i < b0 + per && i < B; ++i) {
      run += h[i];
      if (r <= run) { selected[k] = edges_all[static_cast<size_t>(c) * E + 2 * p] + 1u + static_cast<uint32_t>(i); break; }
    }
  }
  if (threadIdx.x == 255 && r > s_part[255]) atomicOr(&status[c], 8);  // fewer candidates than the rank: inconsistent counts
}

__global__ void qfinal_kernel(const uint32_t* __restrict__ direct_key, const uint64_t* __restrict__ cand_rank,
                              const uint32_t* __restrict__ selected, int ncols, int T, float* __restrict__ boundaries,
                              int32_t* __restrict__ nbound) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncols) return;
  int m = 0;
  uint32_t last = 0;
  for (int i = 0; i < T; ++i) {
    size_t k = static_cast<size_t>(c) * T + i;
    uint32_t key = cand_rank[k] ? selected[k] : direct_key[k];
    if (m == 0 || key != last) {  // tft returns DISTINCT boundaries (Appendix A.4 rule 3)
      boundaries[static_cast<size_t>(c) * T + m++] = key_f32(key);
      last = key;
    }
  }
  nbound[c] = m;
}

// --------------------------------------------------------------------------------------------
// (3) splitter-select: exact order statistics for MANY targets (num_buckets 100 .. 1000) in two
//     passes over the column.  One pass cannot do it: with T targets the brackets of the
//     sample-select engine cover T * 10 sigma / M of the rows, i.e. everything once T reaches a few
//     hundred.  Instead S (<= 4096) equi-depth splitters are taken from the sample; pass 1 counts,
//     exactly, how many keys fall strictly between consecutive splitters and on each splitter
//     (2S+1 cells, shared-memory atomics); the cell of every target rank is then known, and pass 2
//     collects only the keys of those <= T cells (about T/S of the rows) for the final selection
//     (txf_msel_* on the collected keys).  Cell counts are integers: shards and ranks merge by addition.
// --------------------------------------------------------------------------------------------
constexpr int kSpBlock = 512;
constexpr int kSpLutBits = 12;
constexpr int kSpMaxEdges = 4096;
constexpr int kSpCandBuf = 4096;  // a 512-thread iteration adds at most 4096 keys... of which only flagged cells' keys land here

struct SpLocator {  // shared-memory tables that map a key to its cell among E sorted edges (duplicates allowed)
  const uint32_t* s_edges;  // [E]
  const uint16_t* s_lut;    // [(1 << kSpLutBits) + 1]: number of edges below the start of each bin (last entry: E)
  uint32_t first, last;
  int shift, E;
  // cell 2p = strictly between edge p-1 and edge p, 2p+1 = equal to edge p (first copy of a duplicated edge)
  __device__ __forceinline__ int cell(uint32_t key) const {
    if (key < first) return 0;
    if (key > last) return 2 * E;
    uint32_t bin = (key - first) >> shift;
    int lo = s_lut[bin], hi = s_lut[bin + 1];
    while (lo < hi) {  // lower_bound inside the bin's edge range (usually zero or one edge)
      const int mid = (lo + hi) >> 1;
      const bool lt = s_edges[mid] < key;
      lo = lt ? mid + 1 : lo;
      hi = lt ? hi : mid;
    }
    return 2 * lo + ((lo < E && s_edges[lo] == key) ? 1 : 0);
  }
};

// builds the locator tables of column c in shared memory (all threads; ends with a barrier)
__device__ __forceinline__ SpLocator sp_build(const uint32_t* __restrict__ edges, int E, uint32_t* s_edges, uint16_t* s_lut) {
  for (int i = threadIdx.x; i < E; i += kSpBlock) s_edges[i] = edges[i];
  __syncthreads();
  SpLocator L;
  L.s_edges = s_edges; L.s_lut = s_lut; L.E = E;
  L.first = s_edges[0]; L.last = s_edges[E - 1];
  const uint32_t span = L.last - L.first;
  L.shift = span >> kSpLutBits ? 32 - kSpLutBits - __clz(span) : 0;
  for (int j = threadIdx.x; j <= (1 << kSpLutBits); j += kSpBlock) {
    const uint64_t lo_key = static_cast<uint64_t>(L.first) + (static_cast<uint64_t>(j) << L.shift);
    int lo = 0, hi = E;  // number of edges below the bin start
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      const bool lt = s_edges[mid] < lo_key;
      lo = lt ? mid + 1 : lo;
      hi = lt ? hi : mid;
    }
    s_lut[j] = static_cast<uint16_t>(lo);
  }
  __syncthreads();
  return L;
}

// Splitters: S distinct-as-far-as-possible equi-depth order statistics of the sample.  One CTA per column sorts a
// strided subsample of 8 S keys in shared memory (bitonic), takes every 8th, drops repeated values and pads the
// tail with the largest one (repeated splitters are harmless to the locator; they only have to be sorted).
constexpr int kSeBlock = 1024;
__global__ void __launch_bounds__(kSeBlock) qsplit_edges_kernel(const uint32_t* __restrict__ sample, int64_t M, int S,
                                                                uint32_t* __restrict__ edges_all) {
  extern __shared__ uint32_t se_s[];
  const int c = blockIdx.x;
  const int msub = 8 * S;
  const uint32_t* src = sample + static_cast<size_t>(c) * M;
  const int64_t stride = M / msub;
  for (int i = threadIdx.x; i < msub; i += kSeBlock) se_s[i] = src[i * stride];
  __syncthreads();
  for (int k = 2; k <= msub; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (msub >> 1); t += kSeBlock) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // index with bit j clear
        const int l = i | j;
        const uint32_t a = se_s[i], b = se_s[l];
        const bool up = (i & k) == 0;
        if (up ? a > b : a < b) { se_s[i] = b; se_s[l] = a; }
      }
      __syncthreads();
    }
  }
  uint32_t* edges = edges_all + static_cast<size_t>(c) * S;
  if (threadIdx.x == 0) {
    int m = 0;
    uint32_t last = 0;
    for (int j = 0; j < S; ++j) {
      const uint32_t v = se_s[8 * j + 4];
      if (m == 0 || v != last) { edges[m++] = v; last = v; }
    }
    for (; m < S; ++m) edges[m] = last;
  }
}

// MODE 0: count cells.  MODE 1: append the keys whose cell is flagged to the candidate list.
template <int MODE>
__global__ void __launch_bounds__(kSpBlock) qsplit_kernel(NumColsView cols, int64_t n, int E, const uint32_t* __restrict__ edges_all,
                                                          unsigned long long* __restrict__ cells_all,
                                                          const uint32_t* __restrict__ flags_all, uint32_t* __restrict__ cand_all,
                                                          unsigned long long* __restrict__ cand_count, int64_t cand_cap) {
  extern __shared__ uint32_t sp_smem[];
  __shared__ uint32_t s_ncand;
  __shared__ unsigned long long s_base;
  const int c = blockIdx.y;
  const int ncell = 2 * E + 1;
  uint32_t* s_edges = sp_smem;                                             // [E]
  uint16_t* s_lut = reinterpret_cast<uint16_t*>(s_edges + E);              // [(1 << bits) + 1], padded to 4 bytes
  uint32_t* s_aux = reinterpret_cast<uint32_t*>(s_lut + (1 << kSpLutBits) + 2);  // MODE 0: counters [ncell]; MODE 1: flag words + buffer
  const SpLocator L = sp_build(edges_all + static_cast<size_t>(c) * E, E, s_edges, s_lut);
  const int nflag = (ncell + 31) / 32;
  if (MODE == 0) {
    for (int i = threadIdx.x; i < ncell; i += kSpBlock) s_aux[i] = 0;
  } else {
    for (int i = threadIdx.x; i < nflag; i += kSpBlock) s_aux[i] = flags_all[static_cast<size_t>(c) * nflag + i];
    if (threadIdx.x == 0) s_ncand = 0;
  }
  __syncthreads();
  uint32_t* s_cand = s_aux + nflag;  // MODE 1
  uint32_t* cand = cand_all + static_cast<size_t>(c) * cand_cap;
  auto flush = [&]() {  // MODE 1, all threads
    __syncthreads();
    if (threadIdx.x == 0) s_base = atomicAdd(&cand_count[c], static_cast<unsigned long long>(s_ncand < kSpCandBuf ? s_ncand : kSpCandBuf));
    __syncthreads();
    const uint32_t m = s_ncand < kSpCandBuf ? s_ncand : kSpCandBuf;
    for (uint32_t i = threadIdx.x; i < m; i += kSpBlock) {
      const unsigned long long pos = s_base + i;
      if (pos < static_cast<unsigned long long>(cand_cap)) cand[pos] = s_cand[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) s_ncand = 0;
    __syncthreads();
  };
  auto visit = [&](uint32_t key) {
    const int cell = L.cell(key);
    if (MODE == 0) {
      atomicAdd(&s_aux[cell], 1u);
    } else if ((s_aux[cell >> 5] >> (cell & 31)) & 1u) {
      const uint32_t pos = atomicAdd(&s_ncand, 1u);
      if (pos < kSpCandBuf) {
        s_cand[pos] = key;
      } else {  // CTA buffer full between two flushes: append to the global list directly
        const unsigned long long g = atomicAdd(&cand_count[c], 1ULL);
        if (g < static_cast<unsigned long long>(cand_cap)) cand[g] = key;
      }
    }
  };
  // quads interleaved over the grid; the candidate buffer is drained every few iterations (block-uniform trip count)
  const void* values = cols.values[c];
  const uint8_t* __restrict__ valid = cols.valid[c];
  const bool f32 = cols.dtype[c] == TXF_F32;
  const int64_t nquads = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kSpBlock;
  const int64_t iters = (nquads + 2 * stride - 1) / (2 * stride);
  for (int64_t it = 0; it < iters; ++it) {  // two quads per thread and iteration: both loads are issued before either is used
    const int64_t qa = 2 * it * stride + blockIdx.x * static_cast<int64_t>(kSpBlock) + threadIdx.x, qb = qa + stride;
    float4 xa = make_float4(0, 0, 0, 0), xb = xa;
    uint32_t ma = 0, mb = 0;
    if (qa < nquads) { xa = f32 ? load_quad_as_f32<true>(values, qa) : load_quad_as_f32<false>(values, qa); ma = quad_valid(valid, 4 * qa); }
    if (qb < nquads) { xb = f32 ? load_quad_as_f32<true>(values, qb) : load_quad_as_f32<false>(values, qb); mb = quad_valid(valid, 4 * qb); }
    if (qa < nquads) {
      visit(f32_key(ma & 1 ? xa.x : 0.f)); visit(f32_key(ma & 2 ? xa.y : 0.f)); visit(f32_key(ma & 4 ? xa.z : 0.f)); visit(f32_key(ma & 8 ? xa.w : 0.f));
    }
    if (qb < nquads) {
      visit(f32_key(mb & 1 ? xb.x : 0.f)); visit(f32_key(mb & 2 ? xb.y : 0.f)); visit(f32_key(mb & 4 ? xb.z : 0.f)); visit(f32_key(mb & 8 ? xb.w : 0.f));
    }
    if (MODE == 1 && __syncthreads_or(s_ncand > kSpCandBuf / 2)) flush();
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // rows past the last full quad
    const int64_t i = (nquads << 2) + threadIdx.x;
    const bool present = row_valid(valid, i);
    visit(f32 ? cell_key_f32(static_cast<const float*>(values)[i], present) : cell_key_i64(static_cast<const long long*>(values)[i], present));
  }
  if (MODE == 1) {
    flush();
  } else {
    __syncthreads();
    for (int i = threadIdx.x; i < ncell; i += kSpBlock) {
      const uint32_t v = s_aux[i];
      if (v) atomicAdd(&cells_all[static_cast<size_t>(c) * ncell + i], static_cast<unsigned long long>(v));
    }
  }
}

// One CTA per column: prefix sums over the cells, then per target rank either the splitter it falls on (direct) or
// its cell (flagged for pass 2) and its rank among all collected keys of the column (cells are in key order, so the
// collected keys of earlier flagged cells precede those of later ones).
__global__ void __launch_bounds__(1024) qsplit_resolve_kernel(const unsigned long long* __restrict__ cells_all,
                                                              const uint32_t* __restrict__ edges_all, const uint64_t* __restrict__ ranks,
                                                              int E, int T, uint32_t* __restrict__ direct_key,
                                                              uint64_t* __restrict__ cand_rank, uint32_t* __restrict__ flags_all,
                                                              unsigned long long* __restrict__ cand_total, unsigned long long* scratch) {
  const int c = blockIdx.x;
  const int ncell = 2 * E + 1, nflag = (ncell + 31) / 32;
  const unsigned long long* cells = cells_all + static_cast<size_t>(c) * ncell;
  unsigned long long* incl = scratch + static_cast<size_t>(c) * 2 * ncell;   // inclusive prefix of the cell counts
  unsigned long long* fincl = incl + ncell;                                   // inclusive prefix of the FLAGGED cells' counts
  uint32_t* flags = flags_all + static_cast<size_t>(c) * nflag;
  __shared__ unsigned long long s_part[1024];
  __shared__ int s_tcell[1024];  // cell of each target (T <= 1024 per launch chunk is enforced by the entry point)
  const int per = (ncell + 1023) / 1024;
  auto block_scan = [&](auto value_of, unsigned long long* out) {
    const int b0 = threadIdx.x * per;
    unsigned long long sum = 0;
    for (int i = b0; i < b0 + per && i < ncell; ++i) sum += value_of(i);
    s_part[threadIdx.x] = sum;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
      unsigned long long v = threadIdx.x >= off ? s_part[threadIdx.x - off] : 0;
      __syncthreads();
      s_part[threadIdx.x] += v;
      __syncthreads();
    }
    unsigned long long run = threadIdx.x ? s_part[threadIdx.x - 1] : 0;
    for (int i = b0; i < b0 + per && i < ncell; ++i) { run += value_of(i); out[i] = run; }
    __syncthreads();
  };
  block_scan([&](int i) { return cells[i]; }, incl);
  for (int i = threadIdx.x; i < nflag; i += 1024) flags[i] = 0;
  __syncthreads();
  const uint32_t* edges = edges_all + static_cast<size_t>(c) * E;
  for (int i = threadIdx.x; i < T; i += 1024) {
    const unsigned long long r = ranks[i];  // 1-based
    int lo = 0, hi = ncell - 1;             // first cell whose inclusive prefix reaches r
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (incl[mid] < r) lo = mid + 1; else hi = mid;
    }
    s_tcell[i] = lo;
    if (lo & 1) {
      direct_key[static_cast<size_t>(c) * T + i] = edges[lo >> 1];
    } else {
      direct_key[static_cast<size_t>(c) * T + i] = 0;
      atomicOr(&flags[lo >> 5], 1u << (lo & 31));
    }
  }
  __syncthreads();
  block_scan([&](int i) { return ((flags[i >> 5] >> (i & 31)) & 1u) ? cells[i] : 0ULL; }, fincl);
  for (int i = threadIdx.x; i < T; i += 1024) {
    const int cell = s_tcell[i];
    unsigned long long cr = 0;
    if (!(cell & 1)) {
      const unsigned long long before_cells = cell ? incl[cell - 1] : 0ULL;      // keys in all earlier cells
      const unsigned long long before_cand = cell ? fincl[cell - 1] : 0ULL;      // collected keys in earlier flagged cells
      cr = before_cand + (ranks[i] - before_cells);
    }
    cand_rank[static_cast<size_t>(c) * T + i] = cr;
  }
  if (threadIdx.x == 0) cand_total[c] = fincl[ncell - 1];
}

static int fill_view(NumColsView& v, const void* const* values, const uint8_t* const* valid, const int32_t* dtype,
                     int ncols, const char* who, bool raw) {
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS, "%s: ncols=%d out of [1,%d]", who, ncols, TXF_MAX_COLS);
  TXF_REQUIRE(values != nullptr, "%s: null column table", who);
  for (int c = 0; c < ncols; ++c) {
    TXF_REQUIRE(values[c] != nullptr && aligned(values[c], raw ? 4 : 16), "%s: column %d values null or misaligned", who, c);
    v.values[c] = values[c];
    v.valid[c] = (valid && !raw) ? valid[c] : nullptr;
    v.dtype[c] = (dtype && !raw) ? dtype[c] : TXF_F32;
    TXF_REQUIRE(v.dtype[c] == TXF_F32 || v.dtype[c] == TXF_I64, "%s: column %d bad dtype", who, c);
  }
  return TXF_OK;
}

// grid.x for a (grid.x, ncols) streaming launch: one resident wave, at least 128 CTAs per column
static int stream_grid(int64_t n, int block, int ctas_per_sm, int ncols) {
  int64_t want = ((n >> 2) + block) / block;
  int64_t cap = static_cast<int64_t>(kSMs) * ctas_per_sm / ncols;
  if (cap < 128) cap = 128;
  return static_cast<int>(want < cap ? (want < 1 ? 1 : want) : cap);
}

}  // namespace txf

extern "C" {

size_t txf_msel_state_bytes(int ncols, int R) { return txf::msel_layout(ncols, R).total; }
size_t txf_msel_hist_bytes(int ncols, int R) { return txf::msel_layout(ncols, R).rem; }

int txf_msel_begin(void* state_dev, int ncols, int R, const uint64_t* ranks_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS && R >= 1 && R <= 4096, "msel_begin: bad sizes ncols=%d R=%d", ncols, R);
  TXF_REQUIRE(state_dev && ranks_dev, "msel_begin: null pointer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(state_dev, 0, msel_layout(ncols, R).rem, s));
  msel_begin_kernel<<<ncols, 128, 0, s>>>(static_cast<uint8_t*>(state_dev), ncols, R, ranks_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_msel_hist(void* state_dev, int ncols, int R, int pass, const void* const* data, const uint8_t* const* valid,
                  const int32_t* dtype, int kind, int64_t n, const uint64_t* n_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(state_dev != nullptr, "msel_hist: null state");
  TXF_REQUIRE(R >= 1 && R <= 4096 && pass >= 0 && pass < 4, "msel_hist: bad R/pass");
  TXF_REQUIRE(kind == 0 || kind == 1, "msel_hist: kind must be 0 (cells) or 1 (raw keys)");
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "msel_hist: n out of range");
  if (n == 0) return TXF_OK;
  NumColsView v{};
  int rc = fill_view(v, data, valid, dtype, ncols, "msel_hist", kind == 1);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  size_t smem = (static_cast<size_t>(R) + static_cast<size_t>(R < kMselSmemSlots ? R : kMselSmemSlots) * 256) * sizeof(uint32_t);
  dim3 grid(stream_grid(kind == 0 ? n : n * 4, kMselBlock, 4, ncols), ncols);
  if (kind == 0)
    msel_hist_kernel<0><<<grid, kMselBlock, smem, s>>>(static_cast<uint8_t*>(state_dev), ncols, R, pass, v, n, n_dev);
  else
    msel_hist_kernel<1><<<grid, kMselBlock, smem, s>>>(static_cast<uint8_t*>(state_dev), ncols, R, pass, v, n, n_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_msel_resolve(void* state_dev, int ncols, int R, int pass, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(state_dev != nullptr && ncols >= 1 && ncols <= TXF_MAX_COLS && R >= 1 && R <= 4096, "msel_resolve: bad args");
  msel_resolve_kernel<<<ncols, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<uint8_t*>(state_dev), ncols, R, pass);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_msel_result(const void* state_dev, int ncols, int R, uint32_t* keys_out_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(state_dev && keys_out_dev && ncols >= 1 && ncols <= TXF_MAX_COLS && R >= 1 && R <= 4096, "msel_result: bad args");
  int total = ncols * R;
  msel_result_kernel<<<(total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(state_dev), ncols, R, keys_out_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_qsample(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols, int64_t n,
                int64_t M, uint64_t seed, uint32_t* sample_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(n >= 1 && n < (1LL << 31) && M >= 1 && M <= n, "qsample: need 1 <= M <= n < 2^31");
  TXF_REQUIRE(sample_dev != nullptr, "qsample: null output");
  NumColsView v{};
  int rc = fill_view(v, values, valid, dtype, ncols, "qsample", false);
  if (rc) return rc;
  dim3 grid(static_cast<unsigned>((M + 255) / 256), ncols);
  qsample_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(v, n, M, seed, sample_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_qcount(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols, int64_t n, int T,
               const uint32_t* edges_dev, uint64_t* cells_dev, uint32_t* cand_dev, uint64_t* cand_count_dev,
               int64_t cand_cap, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(T >= 1 && T <= TXF_QMAX_TARGETS, "qcount: T=%d out of [1,%d]", T, TXF_QMAX_TARGETS);
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "qcount: n out of range");
  TXF_REQUIRE(edges_dev && cells_dev && cand_dev && cand_count_dev && cand_cap >= 1, "qcount: null pointer / empty candidate buffer");
  if (n == 0) return TXF_OK;
  NumColsView v{};
  int rc = fill_view(v, values, valid, dtype, ncols, "qcount", false);
  if (rc) return rc;
  const int E = 2 * T, ncell = 2 * E + 1;
  size_t smem = static_cast<size_t>((E + 1) & ~1) * 4 + kQCandBuf * 4 + (2u << kQLutBits) + static_cast<size_t>(ncell) * kQBlock * 2;
  bool any_f32 = false, any_i64 = false;
  for (int c = 0; c < ncols; ++c) (v.dtype[c] == TXF_F32 ? any_f32 : any_i64) = true;
  if (smem > 48 * 1024) {
    TXF_CUDA(cudaFuncSetAttribute(qcount_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    TXF_CUDA(cudaFuncSetAttribute(qcount_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  }
  // same shared-memory split as the other kernels of the pass (a change of split waits for an idle SM)
  TXF_CUDA(cudaFuncSetAttribute(qcount_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  TXF_CUDA(cudaFuncSetAttribute(qcount_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  int per_sm = static_cast<int>((200 * 1024) / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 6 ? 6 : per_sm);
  dim3 grid(stream_grid(n, kQBlock, per_sm, ncols), ncols);
  // private 16-bit counters: every thread must see fewer than 65536 keys
  TXF_REQUIRE(n / (static_cast<int64_t>(grid.x) * kQBlock) + 8 < 65536, "qcount: chunk too large for 16-bit counters");
  if (any_f32) {
    qcount_kernel<true><<<grid, kQBlock, smem, static_cast<cudaStream_t>(stream)>>>(
        v, n, T, edges_dev, reinterpret_cast<unsigned long long*>(cells_dev), cand_dev,
        reinterpret_cast<unsigned long long*>(cand_count_dev), cand_cap);
    count_launch(1);
  }
  if (any_i64) {
    qcount_kernel<false><<<grid, kQBlock, smem, static_cast<cudaStream_t>(stream)>>>(
        v, n, T, edges_dev, reinterpret_cast<unsigned long long*>(cells_dev), cand_dev,
        reinterpret_cast<unsigned long long*>(cand_count_dev), cand_cap);
    count_launch(1);
  }
  TXF_LAUNCH_CHECK();
  return TXF_OK;
}

size_t txf_qsplit_scratch_bytes(int ncols, int E) { return static_cast<size_t>(ncols) * 2 * (2 * static_cast<size_t>(E) + 1) * sizeof(uint64_t); }
size_t txf_qsplit_flag_words(int E) { return (2 * static_cast<size_t>(E) + 1 + 31) / 32; }

static int qsplit_launch(int mode, const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols, int64_t n, int E,
                         const uint32_t* edges_dev, uint64_t* cells_dev, const uint32_t* flags_dev, uint32_t* cand_dev,
                         uint64_t* cand_count_dev, int64_t cand_cap, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(E >= 1 && E <= kSpMaxEdges, "qsplit: E=%d out of [1,%d]", E, kSpMaxEdges);
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "qsplit: n out of range");
  TXF_REQUIRE(edges_dev != nullptr, "qsplit: null edges");
  if (n == 0) return TXF_OK;
  NumColsView v{};
  int rc = fill_view(v, values, valid, dtype, ncols, "qsplit", false);
  if (rc) return rc;
  const size_t ncell = 2 * static_cast<size_t>(E) + 1;
  const size_t aux = mode == 0 ? ncell * 4 : ((ncell + 31) / 32) * 4 + kSpCandBuf * 4;
  const size_t smem = static_cast<size_t>(E) * 4 + ((1 << kSpLutBits) + 2) * 2 + aux;
  int per_sm = static_cast<int>((200 * 1024) / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 6 ? 6 : per_sm);
  dim3 grid(stream_grid(n, kSpBlock, per_sm, ncols), ncols);
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (mode == 0) {
    TXF_REQUIRE(cells_dev != nullptr, "qsplit_count: null cells");
    TXF_CUDA(cudaFuncSetAttribute(qsplit_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    qsplit_kernel<0><<<grid, kSpBlock, smem, cs>>>(v, n, E, edges_dev, reinterpret_cast<unsigned long long*>(cells_dev), nullptr, nullptr,
                                                   nullptr, 0);
  } else {
    TXF_REQUIRE(flags_dev && cand_dev && cand_count_dev && cand_cap >= 1, "qsplit_collect: null pointer / empty candidate buffer");
    TXF_CUDA(cudaFuncSetAttribute(qsplit_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    qsplit_kernel<1><<<grid, kSpBlock, smem, cs>>>(v, n, E, edges_dev, nullptr, flags_dev, cand_dev,
                                                   reinterpret_cast<unsigned long long*>(cand_count_dev), cand_cap);
  }
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_qsplit_edges(const uint32_t* sample_dev, int ncols, int64_t M, int S, uint32_t* edges_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS, "qsplit_edges: ncols out of range");
  TXF_REQUIRE(S >= 16 && S <= kSpMaxEdges && (S & (S - 1)) == 0, "qsplit_edges: S must be a power of two in [16,%d]", kSpMaxEdges);
  TXF_REQUIRE(M >= 8LL * S, "qsplit_edges: the sample must hold at least 8 S keys per column");
  TXF_REQUIRE(sample_dev && edges_dev, "qsplit_edges: null pointer");
  const int smem = 8 * S * 4;
  TXF_CUDA(cudaFuncSetAttribute(qsplit_edges_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  qsplit_edges_kernel<<<ncols, kSeBlock, smem, static_cast<cudaStream_t>(stream)>>>(sample_dev, M, S, edges_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_qsplit_count(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols, int64_t n, int E,
                     const uint32_t* edges_dev, uint64_t* cells_dev, txf_stream_t stream) {
  return qsplit_launch(0, values, valid, dtype, ncols, n, E, edges_dev, cells_dev, nullptr, nullptr, nullptr, 0, stream);
}

int txf_qsplit_collect(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols, int64_t n, int E,
                       const uint32_t* edges_dev, const uint32_t* flags_dev, uint32_t* cand_dev, uint64_t* cand_count_dev,
                       int64_t cand_cap, txf_stream_t stream) {
  return qsplit_launch(1, values, valid, dtype, ncols, n, E, edges_dev, nullptr, flags_dev, cand_dev, cand_count_dev, cand_cap, stream);
}

int txf_qsplit_resolve(const uint64_t* cells_dev, const uint32_t* edges_dev, const uint64_t* target_ranks_dev, int ncols, int E, int T,
                       uint32_t* direct_key_dev, uint64_t* cand_rank_dev, uint32_t* flags_dev, uint64_t* cand_total_dev,
                       void* scratch_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS && E >= 1 && E <= kSpMaxEdges && T >= 1 && T <= 1024, "qsplit_resolve: bad sizes");
  TXF_REQUIRE(cells_dev && edges_dev && target_ranks_dev && direct_key_dev && cand_rank_dev && flags_dev && cand_total_dev && scratch_dev,
              "qsplit_resolve: null pointer");
  qsplit_resolve_kernel<<<ncols, 1024, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const unsigned long long*>(cells_dev), edges_dev, target_ranks_dev, E, T, direct_key_dev, cand_rank_dev, flags_dev,
      reinterpret_cast<unsigned long long*>(cand_total_dev), static_cast<unsigned long long*>(scratch_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_qresolve(const uint64_t* cells_dev, const uint32_t* edges_dev, const uint64_t* cand_count_dev, int64_t cand_cap,
                 const uint64_t* target_ranks_dev, int ncols, int T, uint32_t* direct_key_dev, uint64_t* cand_rank_dev,
                 int32_t* status_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS && T >= 1 && T <= TXF_QMAX_TARGETS, "qresolve: bad sizes");
  TXF_REQUIRE(cells_dev && edges_dev && cand_count_dev && target_ranks_dev && direct_key_dev && cand_rank_dev && status_dev,
              "qresolve: null pointer");
  qresolve_kernel<<<ncols, 64, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const unsigned long long*>(cells_dev), edges_dev,
      reinterpret_cast<const unsigned long long*>(cand_count_dev), cand_cap, target_ranks_dev, ncols, T, direct_key_dev,
      cand_rank_dev, status_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

size_t txf_bsel_hist_bytes(int ncols, int T, int B) {
  return (static_cast<size_t>(ncols) * T * B + ncols) * sizeof(uint32_t);
}

int txf_bsel_hist(const uint32_t* cand_dev, int64_t cand_cap, const uint64_t* cand_count_dev, const uint32_t* edges_dev, int ncols, int T,
                  int B, uint32_t* hist_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS && T >= 1 && T <= TXF_QMAX_TARGETS && B >= 16 && B <= (1 << 20), "bsel_hist: bad sizes");
  TXF_REQUIRE(cand_dev && cand_count_dev && edges_dev && hist_dev && cand_cap >= 1, "bsel_hist: null pointer");
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(hist_dev, 0, txf_bsel_hist_bytes(ncols, T, B), cs));
  int64_t blocks = (cand_cap + 255) / 256;
  const int64_t cap = 4LL * kSMs * 8 / ncols + 1;
  if (blocks > cap) blocks = cap;
  dim3 grid(static_cast<unsigned>(blocks), ncols);
  bsel_hist_kernel<<<grid, 256, 2 * T * sizeof(uint32_t), cs>>>(cand_dev, cand_cap, reinterpret_cast<const unsigned long long*>(cand_count_dev),
                                                                edges_dev, ncols, T, B, hist_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_bsel_resolve(const uint32_t* hist_dev, const uint64_t* cells_dev, const uint32_t* edges_dev, const uint64_t* cand_rank_dev,
                     int ncols, int T, int B, uint32_t* selected_key_dev, int32_t* status_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS && T >= 1 && T <= TXF_QMAX_TARGETS && B >= 16, "bsel_resolve: bad sizes");
  TXF_REQUIRE(hist_dev && cells_dev && edges_dev && cand_rank_dev && selected_key_dev && status_dev, "bsel_resolve: null pointer");
  dim3 grid(T, ncols);
  bsel_resolve_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(hist_dev, reinterpret_cast<const unsigned long long*>(cells_dev),
                                                                         edges_dev, cand_rank_dev, ncols, T, B, selected_key_dev, status_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_qfinal(const uint32_t* direct_key_dev, const uint64_t* cand_rank_dev, const uint32_t* selected_key_dev, int ncols,
               int T, float* boundaries_dev, int32_t* nbound_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS && T >= 1 && T <= 4096, "qfinal: bad sizes");
  TXF_REQUIRE(direct_key_dev && cand_rank_dev && selected_key_dev && boundaries_dev && nbound_dev, "qfinal: null pointer");
  qfinal_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(direct_key_dev, cand_rank_dev, selected_key_dev, ncols, T,
                                                               boundaries_dev, nbound_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
