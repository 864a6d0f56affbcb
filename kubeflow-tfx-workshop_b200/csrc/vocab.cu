// vocab.cu -- vocabulary analyzer of tft.compute_and_apply_vocabulary (reference call site
// airflow-dags/taxi_utils.py:121-126, top_k=_VOCAB_SIZE :51, num_oov_buckets=_OOV_SIZE :54).
//
// Counting: every row's term is identified by (Fingerprint64, tag).  A per-CTA shared-memory
// table absorbs the hot terms (taxi columns have 7 / ~1000 distinct values) after a warp-level
// pre-aggregation of equal terms; terms that do not fit go straight to the global
// open-addressing table (32-byte slots, one DRAM sector per probe).  Counts are exact.  The
// first thread to claim a slot copies the term's bytes into an arena, so the table is
// self-contained: it can be exchanged between ranks and outlives the chunk it was built from.
// Ranking reproduces tft: drop '' and terms containing \n / \r, order by (count, bytes)
// descending, keep top_k.
// Algorithmic bytes of the count pass: 4 (offset) + term bytes per row + 1 validity bit.
#include "vocab.cuh"

namespace txf {

constexpr int kVcBlock = 256;
constexpr int kVcSlots = 1024;   // shared-memory front table (24 KB)
constexpr int kVcProbes = 8;
constexpr int kVcAggRounds = 3;  // warp pre-aggregation rounds (covers skewed columns)
constexpr uint32_t kPadEnt = 0xFFFFFFFFu;

__device__ __forceinline__ unsigned long long ldcg64(const unsigned long long* p) {
  return __ldcg(p);  // L2 only: never a stale L1 line for words other SMs update atomically
}

// Add `cnt` occurrences of the term (kfp, ktag) whose bytes are str[0..len) to the global table.
__device__ void table_add(CountSlot* table, uint64_t mask, unsigned long long kfp, unsigned long long ktag,
                          unsigned long long cnt, const uint8_t* str, uint32_t len, uint8_t* arena, int64_t arena_cap,
                          unsigned long long* ctrl) {
  uint64_t idx = home_slot(kfp, mask);
  for (uint64_t probe = 0; probe <= mask; ++probe, idx = (idx + 1) & mask) {
    CountSlot* s = table + idx;
    unsigned long long cur = ldcg64(&s->fp);
    if (cur == 0) {
      cur = atomicCAS(&s->fp, 0ULL, kfp);
      if (cur == 0) { cur = kfp; atomicAdd(&ctrl[1], 1ULL); }
    }
    if (cur != kfp) continue;
    unsigned long long t = ldcg64(&s->tag);
    if (t == 0) {
      t = atomicCAS(&s->tag, 0ULL, ktag);
      if (t == 0) {  // this thread owns the slot: keep the term's bytes
        t = ktag;
        unsigned long long off = atomicAdd(&ctrl[0], static_cast<unsigned long long>(len));
        if (len > kMaxTermLen || off + len > static_cast<unsigned long long>(arena_cap)) {
          atomicOr(&ctrl[2], 1ULL);
          s->strref = 0;
        } else {
          for (uint32_t i = 0; i < len; ++i) arena[off + i] = str[i];
          s->strref = (off << 24) | len;
        }
      }
    }
    if (t == ktag) {
      atomicAdd(&s->count, cnt);
      return;
    }
  }
  atomicOr(&ctrl[2], 2ULL);  // table full
}

__global__ void __launch_bounds__(kVcBlock) vocab_count_kernel(const int32_t* __restrict__ offsets,
                                                               const uint8_t* __restrict__ pool,
                                                               const uint8_t* __restrict__ valid, int64_t n,
                                                               CountSlot* table, uint64_t mask, uint8_t* arena,
                                                               int64_t arena_cap, unsigned long long* ctrl) {
  __shared__ unsigned long long s_fp[kVcSlots];
  __shared__ unsigned long long s_tag[kVcSlots];
  __shared__ uint32_t s_cnt[kVcSlots];
  __shared__ uint32_t s_row[kVcSlots];
  for (int i = threadIdx.x; i < kVcSlots; i += kVcBlock) { s_fp[i] = 0; s_tag[i] = 0; s_cnt[i] = 0; s_row[i] = 0; }
  __syncthreads();

  const int64_t stride = static_cast<int64_t>(gridDim.x) * kVcBlock;
  for (int64_t base = blockIdx.x * static_cast<int64_t>(kVcBlock); base < n; base += stride) {
    const int64_t row = base + threadIdx.x;
    const bool has = row < n;
    unsigned long long kfp = 0, ktag = 0;
    const uint8_t* str = pool;
    uint32_t len = 0;
    if (has) {
      int32_t b = __ldg(offsets + row), e = __ldg(offsets + row + 1);
      if (row_valid(valid, row)) { str = pool + b; len = static_cast<uint32_t>(e - b); }
      uint64_t fp, tag;
      fingerprint128(str, len, fp, tag);
      table_identity(fp, tag, kfp, ktag);
    }
    // warp pre-aggregation: a few rounds of "everyone equal to the first pending lane"
    uint32_t mine = has ? 1u : 0u;
    unsigned pending = __ballot_sync(0xffffffffu, has);
#pragma unroll
    for (int r = 0; r < kVcAggRounds; ++r) {
      if (!pending) break;
      int leader = __ffs(pending) - 1;
      unsigned long long lfp = __shfl_sync(0xffffffffu, kfp, leader);
      unsigned long long ltag = __shfl_sync(0xffffffffu, ktag, leader);
      bool same = has && ((pending >> lane_id()) & 1u) && kfp == lfp && ktag == ltag;
      unsigned grp = __ballot_sync(0xffffffffu, same);
      if (static_cast<int>(lane_id()) == leader) mine = __popc(grp);
      else if (same) mine = 0;
      pending &= ~grp;
    }
    if (mine) {
      bool done = false;
      uint32_t idx = static_cast<uint32_t>(home_slot(kfp, kVcSlots - 1));
#pragma unroll 1
      for (int p = 0; p < kVcProbes && !done; ++p, idx = (idx + 1) & (kVcSlots - 1)) {
        unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&s_fp[idx]);
        if (cur == 0) {
          cur = atomicCAS(&s_fp[idx], 0ULL, kfp);
          if (cur == 0) cur = kfp;
        }
        if (cur != kfp) continue;
        unsigned long long t = *reinterpret_cast<volatile unsigned long long*>(&s_tag[idx]);
        if (t == 0) {
          t = atomicCAS(&s_tag[idx], 0ULL, ktag);
          if (t == 0) { t = ktag; s_row[idx] = static_cast<uint32_t>(row); }
        }
        if (t == ktag) { atomicAdd(&s_cnt[idx], mine); done = true; }
      }
      if (!done) table_add(table, mask, kfp, ktag, mine, str, len, arena, arena_cap, ctrl);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kVcSlots; i += kVcBlock) {  // flush the front table
    uint32_t cnt = s_cnt[i];
    if (!cnt) continue;
    int64_t row = s_row[i];
    const uint8_t* str = pool;
    uint32_t len = 0;
    if (row_valid(valid, row)) {
      int32_t b = offsets[row];
      str = pool + b;
      len = static_cast<uint32_t>(offsets[row + 1] - b);
    }
    table_add(table, mask, s_fp[i], s_tag[i], cnt, str, len, arena, arena_cap, ctrl);
  }
}

__global__ void vocab_insert_kernel(const uint64_t* __restrict__ fp, const uint64_t* __restrict__ tag,
                                    const uint64_t* __restrict__ count, const int64_t* __restrict__ offsets,
                                    const uint8_t* __restrict__ pool, int64_t m, CountSlot* table, uint64_t mask,
                                    uint8_t* arena, int64_t arena_cap, unsigned long long* ctrl) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= m) return;
  unsigned long long kfp, ktag;
  table_identity(fp[i], tag[i], kfp, ktag);
  table_add(table, mask, kfp, ktag, count[i], pool + offsets[i], static_cast<uint32_t>(offsets[i + 1] - offsets[i]), arena,
            arena_cap, ctrl);
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
    occ = s.fp != 0 && s.tag != 0;
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
  fp[i] = s.fp; tag[i] = s.tag; count[i] = s.count;
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

__device__ void lookup_put(LookupSlot* lut, uint64_t mask, unsigned long long kfp, uint32_t tag32, int32_t id) {
  uint64_t idx = home_slot(kfp, mask);
  for (uint64_t probe = 0; probe <= mask; ++probe, idx = (idx + 1) & mask) {
    unsigned long long cur = atomicCAS(&lut[idx].fp, 0ULL, kfp);
    if (cur == 0) {  // claimed an empty slot (equal fingerprints of different terms take separate slots)
      lut[idx].tag32 = tag32;
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
  lookup_put(lut, mask, s.fp, static_cast<uint32_t>(s.tag), static_cast<int32_t>(i));
}

__global__ void vocab_lookup_from_terms_kernel(const int32_t* __restrict__ offsets, const uint8_t* __restrict__ pool,
                                               int64_t k, LookupSlot* lut, uint64_t mask) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= k) return;
  uint64_t fp, tag;
  fingerprint128(pool + offsets[i], static_cast<uint32_t>(offsets[i + 1] - offsets[i]), fp, tag);
  unsigned long long kfp, ktag;
  table_identity(fp, tag, kfp, ktag);
  lookup_put(lut, mask, kfp, static_cast<uint32_t>(ktag), static_cast<int32_t>(i));
}

static bool pow2(int64_t x) { return x > 0 && (x & (x - 1)) == 0; }
static int64_t next_pow2(int64_t x) { int64_t p = 1; while (p < x) p <<= 1; return p; }
static unsigned blocks_for(int64_t n, int block) { return static_cast<unsigned>((n + block - 1) / block); }

}  // namespace txf

extern "C" {

size_t txf_vocab_table_bytes(int64_t capacity) { return capacity > 0 ? static_cast<size_t>(capacity) * sizeof(txf::CountSlot) : 0; }
size_t txf_vocab_lookup_bytes(int64_t lookup_cap) { return lookup_cap > 0 ? static_cast<size_t>(lookup_cap) * sizeof(txf::LookupSlot) : 0; }
size_t txf_vocab_sort_scratch_bytes(int64_t m) { return m > 0 ? static_cast<size_t>(txf::next_pow2(m)) * sizeof(uint32_t) : 0; }

int txf_vocab_count(const int32_t* offsets_dev, const uint8_t* pool_dev, const uint8_t* valid_dev, int64_t n, void* table_dev,
                    int64_t capacity, uint8_t* arena_dev, int64_t arena_cap, uint64_t* ctrl_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "vocab_count: n out of range");
  TXF_REQUIRE(pow2(capacity) && capacity < (1LL << 32), "vocab_count: capacity must be a power of two below 2^32");
  TXF_REQUIRE(offsets_dev && pool_dev && table_dev && arena_dev && ctrl_dev && arena_cap > 0, "vocab_count: null pointer");
  TXF_REQUIRE(aligned(table_dev, 32), "vocab_count: table must be 32-byte aligned");
  if (n == 0) return TXF_OK;
  int64_t want = (n + kVcBlock - 1) / kVcBlock;
  int grid = static_cast<int>(want < kSMs * 8 ? want : kSMs * 8);
  vocab_count_kernel<<<grid, kVcBlock, 0, static_cast<cudaStream_t>(stream)>>>(
      offsets_dev, pool_dev, valid_dev, n, static_cast<CountSlot*>(table_dev), static_cast<uint64_t>(capacity - 1), arena_dev,
      arena_cap, reinterpret_cast<unsigned long long*>(ctrl_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_vocab_insert(const uint64_t* fp_dev, const uint64_t* tag_dev, const uint64_t* count_dev, const int64_t* offsets_dev,
                     const uint8_t* pool_dev, int64_t m, void* table_dev, int64_t capacity, uint8_t* arena_dev,
                     int64_t arena_cap, uint64_t* ctrl_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(m >= 0, "vocab_insert: negative m");
  TXF_REQUIRE(pow2(capacity) && capacity < (1LL << 32), "vocab_insert: capacity must be a power of two below 2^32");
  if (m == 0) return TXF_OK;
  TXF_REQUIRE(fp_dev && tag_dev && count_dev && offsets_dev && pool_dev && table_dev && arena_dev && ctrl_dev, "vocab_insert: null pointer");
  vocab_insert_kernel<<<blocks_for(m, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      fp_dev, tag_dev, count_dev, offsets_dev, pool_dev, m, static_cast<CountSlot*>(table_dev),
      static_cast<uint64_t>(capacity - 1), arena_dev, arena_cap, reinterpret_cast<unsigned long long*>(ctrl_dev));
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
  TXF_REQUIRE(lookup_dev != nullptr && aligned(lookup_dev, 16), "vocab_build_lookup: lookup table null or misaligned");
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
  TXF_REQUIRE(lookup_dev != nullptr && aligned(lookup_dev, 16), "vocab_lookup_from_terms: lookup table null or misaligned");
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
