// transform.cu -- the fused per-row phase of preprocessing_fn (reference:
// airflow-dags/taxi_utils.py:115-145): fill-missing (:86-103), z-score (:116-119), vocabulary
// ids (:121-126), bucketize (:128-130), categorical pass-through (:132-133) and the tips label
// (:135-143) as ONE program of column ops.
//
// All ops except string lookups run as ONE launch whose blocks are (op, tile) pairs in op-major
// order (transform_tiles_kernel): at any moment the running blocks stream one or two input
// columns and one or two output columns, and the hardware block scheduler balances SMs and DRAM
// channels.  String lookups (a vocabulary applied to a table that was not analyzed) run in a
// second, persistent kernel with the term bytes staged by the bulk-copy engine.
//
// Memory-bound by construction: every input cell is read once with 16-byte loads, every output
// written once with 16-byte streaming stores.
// Algorithmic bytes per row (taxi schema): read 98.1 + Lp + Lc, write 120; the
// label re-reads `fare` (4 B/row) because ops run column by column.
#include "vocab.cuh"
#include "str_pipe.cuh"

namespace txf {

constexpr int kXfBlock = 1024;
constexpr int kXfStageBytes = 48 * 1024;                       // per buffer: staged term bytes of one vocabulary tile
constexpr int kXfPipeBytes = (StrTilePipe::smem_bytes(kXfStageBytes, 2 * kXfBlock) + 15) & ~15;
constexpr int kXfVocabTileRows = 2 * kXfBlock;
constexpr int kXfSmemBudget = 200 * 1024;

struct XfOp {                // device-side op (txf_op + staging offsets)
  int32_t kind, in_dtype;
  const void* in0;
  const uint8_t* valid0;
  const void* in1;
  const uint8_t* valid1;
  const uint8_t* pool;
  void* out;
  const void* p0;
  const void* p1;
  int64_t a0, a1, a2, a3;
  float f0;
  int32_t smem_off;          // byte offset of this op's staged state, -1 = read it from global
};
struct XfProgram {
  XfOp ops[TXF_MAX_OPS];
  int32_t nops;
  int32_t smem_bytes;
  int32_t stage_off;  // byte offset of the string staging buffer (only present with vocabulary ops)
  int32_t pad;
};

// ---- element math (shared by the vector path and the scalar tail) ---------------------------
// (x - mean) / den.  div.rn compiles to an inline fast path plus a CALL to a slow-path routine;
// that call site made ptxas spill every live register of the streaming loop.  Since den is the
// same for every row, its correctly rounded reciprocal is computed once per op, and the quotient
// is one multiply plus a Markstein FMA correction: the IEEE-rounded quotient whenever nothing
// over/underflows (non-finite quotients are returned as the plain product).
struct ZScale {
  float mean, den, rden;
};
__device__ __forceinline__ ZScale make_zscale(const float* mv) {
  ZScale z;
  z.mean = mv[0];
  z.den = __fsqrt_rn(mv[1]);
  z.rden = z.den != 0.f ? __frcp_rn(z.den) : 0.f;
  return z;
}
__device__ __forceinline__ float zscore1(float x, const ZScale& z) {
  float num = __fsub_rn(x, z.mean);
  if (z.den == 0.f) return num;  // var == 0 -> x - mean (tft rule)
  float q = __fmul_rn(num, z.rden);
  float r = __fmaf_rn(-z.den, q, num);
  float q2 = __fmaf_rn(r, z.rden, q);
  return fabsf(q) < __int_as_float(0x7f800000) ? q2 : q;
}
__device__ __forceinline__ long long bucket1(float x, const float* __restrict__ b, int nb) {
  int lo = 0, len = nb;  // upper_bound: count of leading boundaries with !(x < b); NaN -> nb
  while (len > 0) {
    int half = len >> 1;
    bool go = !(x < b[lo + half]);
    lo = go ? lo + half + 1 : lo;
    len = go ? len - half - 1 : half;
  }
  return lo;
}
__device__ __forceinline__ long long label1(float fare, float tips, float ratio) {
  float thr = __fmul_rn(fare, ratio);  // fp32 product, never contracted
  return (fare != fare) ? 0LL : (tips > thr ? 1LL : 0LL);
}
// x % d without the 64-bit division subroutine (a CALL inside the row loop).  d <= 65535: three
// 32-bit remainders; otherwise shift-subtract long division (only for huge num_oov_buckets).
__device__ __forceinline__ uint64_t mod_u64(uint64_t x, uint64_t d) {
  if (d <= 65535u) {
    uint32_t dd = static_cast<uint32_t>(d);
    uint32_t hi = static_cast<uint32_t>(x >> 32), lo = static_cast<uint32_t>(x);
    uint32_t c = (0xFFFFFFFFu % dd + 1u) % dd;  // 2^32 mod d
    return ((hi % dd) * c % dd + lo % dd) % dd;
  }
  uint64_t r = 0;
#pragma unroll 1
  for (int i = 63; i >= 0; --i) {
    uint64_t top = r >> 63;
    r = (r << 1) | ((x >> i) & 1ULL);
    if (top || r >= d) r -= d;
  }
  return r;
}
__device__ __forceinline__ long long vocab1(const uint8_t* str, uint32_t len, const LookupSlot* __restrict__ lut,
                                            uint64_t mask, const uint8_t* __restrict__ lut_pool, long long vocab_size, long long num_oov,
                                            long long dflt) {
  const StrKey key = make_key(str, len);
  const bool fp = key_is_fingerprint(key.b);
  uint64_t idx = home_slot(key.a, key.b, mask);
  for (uint64_t probe = 0; probe <= mask; ++probe, idx = (idx + 1) & mask) {
    const LookupSlot* s = lut + idx;
    const unsigned long long sb = s->b;
    if (sb == 0) break;
    if (sb != key.b || s->a != key.a) continue;
    if (fp) {  // a fingerprint hit is confirmed on the vocabulary's bytes
      const unsigned long long ref = s->strref;
      if (static_cast<uint32_t>(ref & 0xFFFFFFu) != len || !term_bytes_equal(lut_pool + (ref >> 24), str, len)) continue;
    }
    return s->id;
  }
  if (num_oov <= 0) return dflt;
  // out of vocabulary: TF's StringToHashBucketFast = Fingerprint64(term) mod num_oov_buckets
  return vocab_size + static_cast<long long>(mod_u64(fingerprint64(str, len), static_cast<uint64_t>(num_oov)));
}

__device__ __forceinline__ float cell_f32(const XfOp& op, int64_t row) {  // fill + cast of one cell
  bool v = row_valid(op.valid0, row);
  if (op.in_dtype == TXF_F32) return v ? static_cast<const float*>(op.in0)[row] : 0.f;
  return v ? i64_to_f32(static_cast<const long long*>(op.in0)[row]) : 0.f;
}

// one row of one op (tail rows only)
__device__ void op_row(const XfOp& op, const uint8_t* smem, int64_t row, int64_t n) {
  switch (op.kind) {
    case TXF_OP_ZSCORE: {
      static_cast<float*>(op.out)[row] = zscore1(cell_f32(op, row), make_zscale(static_cast<const float*>(op.p0)));
      break;
    }
    case TXF_OP_BUCKETIZE: {
      const float* b = op.smem_off >= 0 ? reinterpret_cast<const float*>(smem + op.smem_off) : static_cast<const float*>(op.p0);
      int nb = *static_cast<const int32_t*>(op.p1);
      static_cast<long long*>(op.out)[row] = bucket1(cell_f32(op, row), b, nb);
      break;
    }
    case TXF_OP_FILL_I64:
      static_cast<long long*>(op.out)[row] = row_valid(op.valid0, row) ? static_cast<const long long*>(op.in0)[row] : 0LL;
      break;
    case TXF_OP_FILL_F32:
      static_cast<float*>(op.out)[row] = row_valid(op.valid0, row) ? static_cast<const float*>(op.in0)[row] : 0.f;
      break;
    case TXF_OP_FILL_STR: {
      const int32_t* off = static_cast<const int32_t*>(op.in0);
      int32_t base = off[0];
      static_cast<int32_t*>(op.out)[row] = off[row] - base;
      if (row == n - 1) static_cast<int32_t*>(op.out)[n] = off[n] - base;
      break;
    }
    case TXF_OP_RATIO_LABEL: {
      float fare = row_valid(op.valid0, row) ? static_cast<const float*>(op.in0)[row] : 0.f;
      float tips = row_valid(op.valid1, row) ? static_cast<const float*>(op.in1)[row] : 0.f;
      static_cast<long long*>(op.out)[row] = label1(fare, tips, op.f0);
      break;
    }
    case TXF_OP_VOCAB_CODED:
      static_cast<long long*>(op.out)[row] = static_cast<const long long*>(op.p0)[static_cast<const uint32_t*>(op.in0)[row]];
      break;
    default: break;  // TXF_OP_VOCAB has no quad path: its row loop covers the tail
  }
}

__device__ __forceinline__ void store_i64x4(long long* out, int64_t r, long long a, long long b, long long c, long long d) {
  stcs_l2(out + r, make_longlong2(a, b));
  stcs_l2(out + r + 2, make_longlong2(c, d));
}

__device__ __forceinline__ float4 fill_f4(float4 r, uint32_t m) {
  return make_float4(m & 1 ? r.x : 0.f, m & 2 ? r.y : 0.f, m & 4 ? r.z : 0.f, m & 8 ? r.w : 0.f);
}

constexpr int kLinearBuckets = 32;  // up to 32 boundaries: count them for a whole quad, one broadcast LDS each
constexpr int kSmemBuckets = 1024;  // up to 1024 boundaries (num_buckets <= 1000 and then some) are staged in shared memory per block
struct NumParams {  // per-op constants of the numeric kinds
  ZScale z;
  const float* b;
  int nb;
  float ratio;
  int32_t base;
};
// ids of the 4 cells of a quad: id = number of leading boundaries with !(x < b) (NaN -> nb)
__device__ __forceinline__ void bucket_quad(float4 x, const NumParams& p, long long (&id)[4]) {
  if (p.nb > kLinearBuckets) {
    id[0] = bucket1(x.x, p.b, p.nb); id[1] = bucket1(x.y, p.b, p.nb); id[2] = bucket1(x.z, p.b, p.nb); id[3] = bucket1(x.w, p.b, p.nb);
    return;
  }
  int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
  for (int j = 0; j < p.nb; ++j) {
    const float bj = p.b[j];
    c0 += !(x.x < bj) ? 1 : 0; c1 += !(x.y < bj) ? 1 : 0; c2 += !(x.z < bj) ? 1 : 0; c3 += !(x.w < bj) ? 1 : 0;
  }
  id[0] = c0; id[1] = c1; id[2] = c2; id[3] = c3;
}
// Vocabulary op: row tiles of kXfVocabTileRows rows; the tile's term bytes are staged in shared
// memory (coalesced 16-byte loads), then one row per lane: coalesced offsets in, coalesced ids out.
__device__ __forceinline__ void run_vocab(const XfOp& op, const uint8_t* smem, uint8_t* stage, int64_t n) {
  const int32_t* __restrict__ off = static_cast<const int32_t*>(op.in0);
  const uint8_t* __restrict__ valid = op.valid0;
  const uint8_t* __restrict__ pool = op.pool;
  const LookupSlot* lut = op.smem_off >= 0 ? reinterpret_cast<const LookupSlot*>(smem + op.smem_off)
                                            : static_cast<const LookupSlot*>(op.p0);
  const uint64_t mask = static_cast<uint64_t>(op.a0 - 1);
  const uint8_t* __restrict__ lut_pool = static_cast<const uint8_t*>(op.p1);
  const long long vocab_size = op.a1, num_oov = op.a2, dflt = op.a3;
  long long* out = static_cast<long long*>(op.out);
  StrTilePipe pipe;
  pipe.init(stage, kXfStageBytes, kXfVocabTileRows, off, pool, valid, n);
  const int64_t ntiles = (n + kXfVocabTileRows - 1) / kXfVocabTileRows;
  auto tile_end = [&](int64_t t) { return (t + 1) * kXfVocabTileRows < n ? (t + 1) * kXfVocabTileRows : n; };
  if (threadIdx.x == 0) {
    if (blockIdx.x < ntiles) pipe.issue(blockIdx.x * static_cast<int64_t>(kXfVocabTileRows), tile_end(blockIdx.x), 0);
    if (blockIdx.x + gridDim.x < ntiles)
      pipe.issue((blockIdx.x + gridDim.x) * static_cast<int64_t>(kXfVocabTileRows), tile_end(blockIdx.x + gridDim.x), 1);
  }
  uint32_t k = 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++k) {
    const int64_t r0 = tile * kXfVocabTileRows;
    const int64_t r1 = tile_end(tile);
    const uint8_t* base = pipe.acquire(k & 1, k >> 1);
#pragma unroll
    for (int j = 0; j < kXfVocabTileRows / kXfBlock; ++j) {
      const int64_t row = r0 + j * kXfBlock + threadIdx.x;
      if (row < r1) {
        int32_t b = __ldg(off + row), e = __ldg(off + row + 1);
        bool v = row_valid(valid, row);
        __stcs(out + row, vocab1(base + b, v ? static_cast<uint32_t>(e - b) : 0u, lut, mask, lut_pool, vocab_size, num_oov, dflt));
      }
    }
    __syncthreads();  // buffer free: refill it with the tile after next
    const int64_t nxt = tile + 2 * static_cast<int64_t>(gridDim.x);
    if (threadIdx.x == 0 && nxt < ntiles) pipe.issue(nxt * kXfVocabTileRows, tile_end(nxt), k & 1);
  }
  __syncthreads();  // the pipe's shared memory is re-initialised by the next vocabulary op
}

// String-lookup ops (TXF_OP_VOCAB: a table that was not analyzed, e.g. TransformDataset on new data
// or a loaded state): one persistent CTA per SM; lookup tables are staged in shared memory once.
__global__ void __launch_bounds__(kXfBlock, 1) transform_strings_kernel(const __grid_constant__ XfProgram prog, int64_t n) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ XfOp s_ops[TXF_MAX_OPS];
  const int nops = prog.nops;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&prog.ops[0]);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&s_ops[0]);
    for (int i = threadIdx.x; i < nops * static_cast<int>(sizeof(XfOp) / 4); i += kXfBlock) dst[i] = src[i];
  }
  __syncthreads();
  for (int o = 0; o < nops; ++o) {
    const XfOp& op = s_ops[o];
    if (op.smem_off < 0) continue;
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(smem + op.smem_off);
    const unsigned long long* src = static_cast<const unsigned long long*>(op.p0);
    for (int64_t i = threadIdx.x; i < op.a0 * static_cast<int64_t>(sizeof(LookupSlot) / 8); i += kXfBlock) dst[i] = src[i];
  }
  __syncthreads();
  for (int o = 0; o < nops; ++o) run_vocab(s_ops[o], smem, smem + prog.stage_off, n);
}

// ---- numeric ops as one grid of (op, tile) blocks ----------------------------------------------
// Measured on this GPU (tools/probes/copy_probe.cu): a plain int64 copy runs at 6.6-6.7 TB/s as a
// one-shot grid of 256-thread blocks, 64 bytes in flight per thread, and at 5.8 TB/s as 148
// persistent 1024-thread blocks walking tiles round-robin -- the hardware block scheduler balances
// the SMs and the DRAM channels better than a static assignment.  So every op of the program that
// is not a string lookup runs as ONE launch whose blocks are (op, tile) pairs in op-major order:
// block b belongs to the op whose tile range contains b; at any moment the running blocks stream
// through one or two columns.  A tile is 256 threads x 64 bytes of input cells.
constexpr int kTileThreads = 256;
constexpr int kTileQuads4 = kTileThreads * 4;  // 4-byte cells: 4 quads of 16 bytes per thread
constexpr int kTileQuads8 = kTileThreads * 2;  // 8-byte cells: 2 quads of 32 bytes per thread

struct TileProgram {
  XfProgram prog;
  int64_t first_block[TXF_MAX_OPS + 1];  // op o owns blocks [first_block[o], first_block[o + 1])
};

__device__ __forceinline__ bool op_in8(const XfOp& op) {
  return op.kind == TXF_OP_FILL_I64 || ((op.kind == TXF_OP_ZSCORE || op.kind == TXF_OP_BUCKETIZE) && op.in_dtype == TXF_I64);
}

__global__ void __launch_bounds__(kTileThreads) transform_tiles_kernel(const __grid_constant__ TileProgram tp, int64_t n) {
  __shared__ float s_bounds[kSmemBuckets];
  const int nops = tp.prog.nops;
  int o = 0;
  while (o + 1 < nops && static_cast<int64_t>(blockIdx.x) >= tp.first_block[o + 1]) ++o;  // block-uniform, <= 32 steps
  const XfOp& op = tp.prog.ops[o];
  const int64_t tile = static_cast<int64_t>(blockIdx.x) - tp.first_block[o];
  const int kind = op.kind;
  const int64_t nquads = n >> 2;
  NumParams p{};
  if (kind == TXF_OP_ZSCORE) p.z = make_zscale(static_cast<const float*>(op.p0));
  if (kind == TXF_OP_RATIO_LABEL) p.ratio = op.f0;
  if (kind == TXF_OP_FILL_STR) p.base = __ldg(static_cast<const int32_t*>(op.in0));
  if (kind == TXF_OP_BUCKETIZE) {
    p.nb = *static_cast<const int32_t*>(op.p1);
    p.b = static_cast<const float*>(op.p0);
    if (p.nb <= kSmemBuckets) {  // one shared copy per block: the linear count (<= 32) and the binary search both read it there
      for (int i = threadIdx.x; i < p.nb; i += kTileThreads) s_bounds[i] = p.b[i];
      __syncthreads();
      p.b = s_bounds;
    }
  }
  void* out = op.out;
  if (!op_in8(op)) {
    const float* __restrict__ v = static_cast<const float*>(op.in0);
    const uint8_t* __restrict__ valid = (kind == TXF_OP_FILL_STR || kind == TXF_OP_VOCAB_CODED) ? nullptr : op.valid0;
    // Without a validity bitmap a thread's four quads are 256 quads apart (unit-stride warps, the
    // fastest shape).  With a bitmap a thread owns two PAIRS of adjacent quads instead: the 8 rows of
    // a pair share exactly one validity byte, so a warp reads its validity as 32 consecutive bytes,
    // one load per pair -- a byte load per quad cost 25-30 % on every 4-byte op, the 32-byte stride
    // of the pairs costs 8 %.
    const bool paired = valid != nullptr;  // block-uniform
    const int64_t q0 = tile * kTileQuads4 + threadIdx.x, p0 = tile * (kTileQuads4 / 2) + threadIdx.x;
    auto quad_of = [&](int u) -> int64_t { return paired ? 2 * (p0 + (u >> 1) * kTileThreads) + (u & 1) : q0 + u * kTileThreads; };
    float4 raw[4];
    uint32_t m[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t q = quad_of(u);
      raw[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      m[u] = 0;
      if (q < nquads) { raw[u] = ldg_f4(v + 4 * q); m[u] = 0xFu; }
    }
    if (paired) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int64_t q = quad_of(2 * j);
        if (q < nquads) {
          const uint32_t vb = __ldg(valid + (q >> 1));
          m[2 * j] = vb & 0xFu;
          m[2 * j + 1] &= vb >> 4;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t q = quad_of(u);
      if (q >= nquads) continue;
      const float4 x = fill_f4(raw[u], m[u]);
      switch (kind) {
        case TXF_OP_ZSCORE:
          stcs_f4(static_cast<float*>(out) + 4 * q, make_float4(zscore1(x.x, p.z), zscore1(x.y, p.z), zscore1(x.z, p.z), zscore1(x.w, p.z)));
          break;
        case TXF_OP_FILL_F32:
          stcs_f4(static_cast<float*>(out) + 4 * q, x);
          break;
        case TXF_OP_FILL_STR:
          stcs_i4(static_cast<int32_t*>(out) + 4 * q, make_int4(__float_as_int(raw[u].x) - p.base, __float_as_int(raw[u].y) - p.base,
                                                                 __float_as_int(raw[u].z) - p.base, __float_as_int(raw[u].w) - p.base));
          break;
        case TXF_OP_BUCKETIZE: {
          long long id[4];
          bucket_quad(x, p, id);
          store_i64x4(static_cast<long long*>(out), 4 * q, id[0], id[1], id[2], id[3]);
          break;
        }
        case TXF_OP_VOCAB_CODED: {  // dictionary decode: codes written by the count pass -> final ids
          const long long* __restrict__ tab = static_cast<const long long*>(op.p0);
          store_i64x4(static_cast<long long*>(out), 4 * q, __ldg(tab + __float_as_uint(raw[u].x)), __ldg(tab + __float_as_uint(raw[u].y)),
                      __ldg(tab + __float_as_uint(raw[u].z)), __ldg(tab + __float_as_uint(raw[u].w)));
          break;
        }
        default: {  // TXF_OP_RATIO_LABEL: second fp32 column (tips)
          const float4 t = fill_f4(ldg_f4(static_cast<const float*>(op.in1) + 4 * q), quad_valid(op.valid1, 4 * q));
          store_i64x4(static_cast<long long*>(out), 4 * q, label1(x.x, t.x, p.ratio), label1(x.y, t.y, p.ratio), label1(x.z, t.z, p.ratio),
                      label1(x.w, t.w, p.ratio));
          break;
        }
      }
    }
    // offsets columns have n + 1 entries: entry 4 * nquads is written here, the rest of the tail below
    if (kind == TXF_OP_FILL_STR && tile == 0 && threadIdx.x == 0) {
      const int32_t* off = static_cast<const int32_t*>(op.in0);
      static_cast<int32_t*>(out)[4 * nquads] = off[4 * nquads] - p.base;
    }
  } else {
    const long long* __restrict__ v = static_cast<const long long*>(op.in0);
    const uint8_t* __restrict__ valid = op.valid0;
    const int64_t q0 = tile * kTileQuads8 + threadIdx.x;
    longlong2 ra[2], rb[2];
    uint32_t m[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int64_t q = q0 + u * kTileThreads;
      ra[u] = make_longlong2(0, 0); rb[u] = make_longlong2(0, 0); m[u] = 0;
      if (q < nquads) { ra[u] = ldg_l2(v + 4 * q); rb[u] = ldg_l2(v + 4 * q + 2); m[u] = quad_valid(valid, 4 * q); }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int64_t q = q0 + u * kTileThreads;
      if (q >= nquads) continue;
      const long long a = m[u] & 1 ? ra[u].x : 0LL, b = m[u] & 2 ? ra[u].y : 0LL, c = m[u] & 4 ? rb[u].x : 0LL,
                      d = m[u] & 8 ? rb[u].y : 0LL;
      switch (kind) {
        case TXF_OP_FILL_I64:
          store_i64x4(static_cast<long long*>(out), 4 * q, a, b, c, d);
          break;
        case TXF_OP_ZSCORE:
          stcs_f4(static_cast<float*>(out) + 4 * q, make_float4(zscore1(i64_to_f32(a), p.z), zscore1(i64_to_f32(b), p.z),
                                                                zscore1(i64_to_f32(c), p.z), zscore1(i64_to_f32(d), p.z)));
          break;
        default: {  // TXF_OP_BUCKETIZE
          long long id[4];
          bucket_quad(make_float4(i64_to_f32(a), i64_to_f32(b), i64_to_f32(c), i64_to_f32(d)), p, id);
          store_i64x4(static_cast<long long*>(out), 4 * q, id[0], id[1], id[2], id[3]);
          break;
        }
      }
    }
  }
  // ---- rows past the last full quad (at most 3): the op's first block takes them ----------------
  if (tile == 0 && threadIdx.x < (n & 3)) op_row(op, nullptr, (nquads << 2) + threadIdx.x, n);
}

// Bytes of a pass-through string column: out[0 .. off[n] - off[0]) = pool[off[0] .. off[n]).  The
// offsets of a sliced column do not start at 0 and the host does not know them without a round trip,
// so the span is read on the device.  8 bytes per thread and step; the source may be unaligned.
__global__ void copy_string_bytes_kernel(const int32_t* __restrict__ off, const uint8_t* __restrict__ pool, int64_t n,
                                         uint8_t* __restrict__ out) {
  const int64_t b0 = off[0], total = static_cast<int64_t>(off[n]) - b0;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 8;
  for (int64_t i = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) * 8; i < total; i += stride) {
    const uint64_t w = fetch64(pool + b0 + i);
    if (i + 8 <= total) {
      *reinterpret_cast<uint64_t*>(out + i) = w;
    } else {
      for (int k = 0; i + k < total; ++k) out[i + k] = static_cast<uint8_t>(w >> (8 * k));
    }
  }
}

static int g_last_grid = 0, g_last_block = 0, g_last_smem = 0;

int transform_set_identity_mask(unsigned long long ma, unsigned long long mb) {
  TXF_CUDA(set_identity_mask_this_tu(ma, mb));
  return TXF_OK;
}

}  // namespace txf

extern "C" {

int txf_transform(const txf_op* ops, int nops, int64_t n, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ops != nullptr && nops >= 1 && nops <= TXF_MAX_OPS, "transform: nops=%d out of [1,%d]", nops, TXF_MAX_OPS);
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "transform: n out of range");
  if (n == 0) return TXF_OK;
  XfProgram prog{};
  prog.nops = nops;
  bool have_vocab = false;
  for (int o = 0; o < nops; ++o) have_vocab = have_vocab || ops[o].kind == TXF_OP_VOCAB;
  int smem = have_vocab ? kXfPipeBytes : 0;  // [string tile pipe][per-op staged state ...]
  prog.stage_off = 0;
  for (int o = 0; o < nops; ++o) {
    const txf_op& s = ops[o];
    XfOp& d = prog.ops[o];
    d.kind = s.kind; d.in_dtype = s.in_dtype; d.in0 = s.in0; d.valid0 = s.valid0; d.in1 = s.in1; d.valid1 = s.valid1;
    d.pool = s.pool; d.out = s.out; d.p0 = s.p0; d.p1 = s.p1; d.a0 = s.a0; d.a1 = s.a1; d.a2 = s.a2; d.a3 = s.a3; d.f0 = s.f0;
    d.smem_off = -1;
    TXF_REQUIRE(s.in0 != nullptr && s.out != nullptr, "transform: op %d has a null input/output", o);
    TXF_REQUIRE(aligned(s.in0, 16) && aligned(s.out, 16), "transform: op %d input/output not 16B aligned", o);
    switch (s.kind) {
      case TXF_OP_ZSCORE:
        TXF_REQUIRE(s.p0 != nullptr, "transform: op %d (z-score) needs p0 = {mean,var}", o);
        TXF_REQUIRE(s.in_dtype == TXF_F32 || s.in_dtype == TXF_I64, "transform: op %d bad dtype", o);
        break;
      case TXF_OP_BUCKETIZE: {
        TXF_REQUIRE(s.p0 != nullptr && s.p1 != nullptr && s.a0 >= 0, "transform: op %d (bucketize) needs boundaries, count and capacity a0", o);
        TXF_REQUIRE(s.in_dtype == TXF_F32 || s.in_dtype == TXF_I64, "transform: op %d bad dtype", o);
        break;
      }
      case TXF_OP_VOCAB: {
        TXF_REQUIRE(s.p0 != nullptr && s.p1 != nullptr && s.pool != nullptr,
                    "transform: op %d (vocabulary) needs the lookup table (p0), the vocabulary's byte pool (p1) and the column's pool", o);
        TXF_REQUIRE(s.a0 >= 2 && (s.a0 & (s.a0 - 1)) == 0, "transform: op %d lookup capacity must be a power of two", o);
        TXF_REQUIRE(aligned(s.p0, 8), "transform: op %d lookup table not 8B aligned", o);
        have_vocab = true;
        int64_t need = (s.a0 * static_cast<int64_t>(sizeof(LookupSlot)) + 15) & ~15LL;
        if (smem + need <= kXfSmemBudget) { d.smem_off = smem; smem += static_cast<int>(need); }
        break;
      }
      case TXF_OP_FILL_I64: case TXF_OP_FILL_F32: case TXF_OP_FILL_STR: break;
      case TXF_OP_VOCAB_CODED:
        TXF_REQUIRE(s.p0 != nullptr, "transform: op %d (coded vocabulary) needs the decode table p0", o);
        break;
      case TXF_OP_RATIO_LABEL:
        TXF_REQUIRE(s.in1 != nullptr && aligned(s.in1, 16), "transform: op %d (label) needs in1 (16B aligned)", o);
        TXF_REQUIRE(s.in_dtype == TXF_F32, "transform: op %d (label) takes fp32 columns", o);
        break;
      default:
        TXF_REQUIRE(false, "transform: op %d has unknown kind %d", o, s.kind);
    }
  }
  // ---- numeric ops: one grid of (op, tile) blocks ------------------------------------------
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  const int64_t nquads = n >> 2;
  TileProgram tp{};
  XfProgram vprog{};
  int64_t blocks = 0;
  for (int o = 0; o < nops; ++o) {
    const XfOp& d = prog.ops[o];
    if (d.kind == TXF_OP_VOCAB) {
      vprog.ops[vprog.nops++] = d;
      continue;
    }
    XfOp& t = tp.prog.ops[tp.prog.nops];
    t = d;
    t.smem_off = -1;  // boundaries are read from p0 (the kernel keeps a per-block copy)
    const bool in8 = d.kind == TXF_OP_FILL_I64 || ((d.kind == TXF_OP_ZSCORE || d.kind == TXF_OP_BUCKETIZE) && d.in_dtype == TXF_I64);
    const int64_t tq = in8 ? kTileQuads8 : kTileQuads4;
    int64_t tiles = (nquads + tq - 1) / tq;
    if (tiles < 1) tiles = 1;  // the op's first block also takes the rows past the last full quad
    tp.first_block[tp.prog.nops] = blocks;
    blocks += tiles;
    ++tp.prog.nops;
  }
  tp.first_block[tp.prog.nops] = blocks;
  if (tp.prog.nops) {
    TXF_REQUIRE(blocks < (1LL << 31), "transform: too many tiles");
    transform_tiles_kernel<<<static_cast<unsigned>(blocks), kTileThreads, 0, cs>>>(tp, n);
    TXF_LAUNCH_CHECK();
    count_launch(1);
    g_last_grid = static_cast<int>(blocks); g_last_block = kTileThreads; g_last_smem = 0;
  }
  // ---- string-lookup ops: the persistent kernel with the staged string tiles -------------------
  if (vprog.nops) {
    vprog.smem_bytes = smem;
    vprog.stage_off = 0;
    if (smem > 48 * 1024)
      TXF_CUDA(cudaFuncSetAttribute(transform_strings_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int64_t ntiles = (n + kXfVocabTileRows - 1) / kXfVocabTileRows;
    int grid = static_cast<int>(ntiles < kSMs ? ntiles : kSMs);
    transform_strings_kernel<<<grid, kXfBlock, smem, cs>>>(vprog, n);
    TXF_LAUNCH_CHECK();
    count_launch(1);
    if (!tp.prog.nops) { g_last_grid = grid; g_last_block = kXfBlock; g_last_smem = smem; }
  }
  return TXF_OK;
}

int txf_copy_string_bytes(const int32_t* offsets_dev, const uint8_t* pool_dev, int64_t n, int64_t max_bytes, uint8_t* out_pool_dev,
                          txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(n >= 0 && n < (1LL << 31) && max_bytes >= 0, "copy_string_bytes: n / max_bytes out of range");
  if (n == 0 || max_bytes == 0) return TXF_OK;
  TXF_REQUIRE(offsets_dev && pool_dev && out_pool_dev && aligned(out_pool_dev, 8), "copy_string_bytes: null or misaligned pointer");
  int64_t blocks = (max_bytes / 8 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 8 * kSMs) blocks = 8 * kSMs;
  copy_string_bytes_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(offsets_dev, pool_dev, n, out_pool_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_transform_launch_info(int* grid, int* block, int* smem_bytes) {
  if (grid) *grid = txf::g_last_grid;
  if (block) *block = txf::g_last_block;
  if (smem_bytes) *smem_bytes = txf::g_last_smem;
  return TXF_OK;
}

}  // extern "C"
