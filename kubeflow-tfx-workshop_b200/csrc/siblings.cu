// siblings.cu -- the sibling analyzers/mappers the workshop notebooks use next to the taxi path
// (SURVEY.md 8 f4): tft.scale_to_0_1 / scale_by_min_max with its min/max analyzer
// (notebooks/03_TensorFlow_Transform.ipynb:148, 03a_TensorFlow_Transform_Advanced.ipynb:323,333)
// and fp32 arithmetic between columns and analyzer scalars such as `x - tft.mean(x)` and
// `x_centered * y_normalized` (03_...:147,150).
//
// These run as their own small kernels instead of being new cases of the fused transform kernel:
// that kernel's register budget is tuned for the taxi program, and these ops are plain streaming
// maps (4 or 8 bytes in, 4 bytes out per row and operand) that reach the same HBM bound alone.
#include "common.cuh"

namespace txf {

constexpr int kSibBlock = 256;
constexpr int kSibGrid = kSMs * 8;

// ---- min/max analyzer -------------------------------------------------------------------------
// State per column: (min, max) as fp32 of the FILLED values (an absent cell counts as 0, like
// every other analyzer on this path).  NaN cells are skipped (fminf/fmaxf); the reference's
// behaviour for NaN input is unpinned.
__device__ __forceinline__ void mm_add(float& lo, float& hi, float x) {
  lo = fminf(lo, x);
  hi = fmaxf(hi, x);
}

__global__ void __launch_bounds__(kSibBlock) minmax_partial_kernel(NumColsView cols, int64_t n, float* __restrict__ partials) {
  const int c = blockIdx.y;
  const uint8_t* __restrict__ valid = cols.valid[c];
  float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
  const int64_t nquads = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kSibBlock;
  int64_t q = blockIdx.x * static_cast<int64_t>(kSibBlock) + threadIdx.x;
  if (cols.dtype[c] == TXF_F32) {
    const float* __restrict__ v = static_cast<const float*>(cols.values[c]);
    for (; q < nquads; q += stride) {
      const float4 a = ldg_f4(v + 4 * q);
      const uint32_t m = quad_valid(valid, 4 * q);
      mm_add(lo, hi, m & 1 ? a.x : 0.f); mm_add(lo, hi, m & 2 ? a.y : 0.f);
      mm_add(lo, hi, m & 4 ? a.z : 0.f); mm_add(lo, hi, m & 8 ? a.w : 0.f);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
      const int64_t i = (nquads << 2) + threadIdx.x;
      mm_add(lo, hi, row_valid(valid, i) ? v[i] : 0.f);
    }
  } else {
    const long long* __restrict__ v = static_cast<const long long*>(cols.values[c]);
    for (; q < nquads; q += stride) {
      const longlong2 a = ldg_l2(v + 4 * q), b = ldg_l2(v + 4 * q + 2);
      const uint32_t m = quad_valid(valid, 4 * q);
      mm_add(lo, hi, m & 1 ? i64_to_f32(a.x) : 0.f); mm_add(lo, hi, m & 2 ? i64_to_f32(a.y) : 0.f);
      mm_add(lo, hi, m & 4 ? i64_to_f32(b.x) : 0.f); mm_add(lo, hi, m & 8 ? i64_to_f32(b.y) : 0.f);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
      const int64_t i = (nquads << 2) + threadIdx.x;
      mm_add(lo, hi, row_valid(valid, i) ? i64_to_f32(v[i]) : 0.f);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  __shared__ float s_lo[kSibBlock / 32], s_hi[kSibBlock / 32];
  if (lane_id() == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kSibBlock / 32; ++w) { lo = fminf(lo, s_lo[w]); hi = fmaxf(hi, s_hi[w]); }
    float* p = partials + (static_cast<int64_t>(c) * gridDim.x + blockIdx.x) * 2;
    p[0] = lo;
    p[1] = hi;
  }
}

// state[c] = fold(state[c], partials of column c); one warp per column.  Also used to merge the
// states of several ranks (npart states per column, laid out [part][col][2]).
__global__ void minmax_fold_kernel(const float* __restrict__ partials, int npart, int64_t part_stride, int64_t col_stride,
                                   float* __restrict__ state) {
  const int c = blockIdx.x;
  float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
  for (int i = threadIdx.x; i < npart; i += 32) {
    const float* p = partials + i * part_stride + c * col_stride;
    lo = fminf(lo, p[0]);
    hi = fmaxf(hi, p[1]);
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if (threadIdx.x == 0) {
    state[2 * c] = fminf(state[2 * c], lo);
    state[2 * c + 1] = fmaxf(state[2 * c + 1], hi);
  }
}

__global__ void minmax_begin_kernel(float* state, int ncols) {
  const int c = threadIdx.x;
  if (c < ncols) {
    state[2 * c] = __int_as_float(0x7f800000);
    state[2 * c + 1] = __int_as_float(0xff800000);
  }
}

// ---- elementwise mappers -----------------------------------------------------------------------
struct Operand {
  const void* col;        // column values (fp32 or int64) or nullptr
  const uint8_t* valid;   // its validity bitmap or nullptr
  const float* scalar;    // device scalar used when col == nullptr
  int32_t dtype;
};
__device__ __forceinline__ float4 load_quad(const Operand& o, int64_t q, float s) {
  if (o.col == nullptr) return make_float4(s, s, s, s);
  const uint32_t m = quad_valid(o.valid, 4 * q);
  if (o.dtype == TXF_F32) {
    const float4 a = ldg_f4(static_cast<const float*>(o.col) + 4 * q);
    return make_float4(m & 1 ? a.x : 0.f, m & 2 ? a.y : 0.f, m & 4 ? a.z : 0.f, m & 8 ? a.w : 0.f);
  }
  const long long* v = static_cast<const long long*>(o.col);
  const longlong2 a = ldg_l2(v + 4 * q), b = ldg_l2(v + 4 * q + 2);
  return make_float4(m & 1 ? i64_to_f32(a.x) : 0.f, m & 2 ? i64_to_f32(a.y) : 0.f, m & 4 ? i64_to_f32(b.x) : 0.f,
                     m & 8 ? i64_to_f32(b.y) : 0.f);
}
__device__ __forceinline__ float load_cell(const Operand& o, int64_t i, float s) {
  if (o.col == nullptr) return s;
  if (!row_valid(o.valid, i)) return 0.f;
  return o.dtype == TXF_F32 ? static_cast<const float*>(o.col)[i] : i64_to_f32(static_cast<const long long*>(o.col)[i]);
}
// one IEEE fp32 operation, never contracted with a neighbour
template <int OP>
__device__ __forceinline__ float arith1(float a, float b) {
  if (OP == TXF_ARITH_ADD) return __fadd_rn(a, b);
  if (OP == TXF_ARITH_SUB) return __fsub_rn(a, b);
  if (OP == TXF_ARITH_MUL) return __fmul_rn(a, b);
  return __fdiv_rn(a, b);
}

template <int OP>
__global__ void __launch_bounds__(kSibBlock) map_binary_kernel(Operand a, Operand b, float* __restrict__ out, int64_t n) {
  const float sa = a.col == nullptr ? __ldg(a.scalar) : 0.f, sb = b.col == nullptr ? __ldg(b.scalar) : 0.f;
  const int64_t nquads = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kSibBlock;
  for (int64_t q = blockIdx.x * static_cast<int64_t>(kSibBlock) + threadIdx.x; q < nquads; q += stride) {
    const float4 x = load_quad(a, q, sa), y = load_quad(b, q, sb);
    stcs_f4(out + 4 * q, make_float4(arith1<OP>(x.x, y.x), arith1<OP>(x.y, y.y), arith1<OP>(x.z, y.z), arith1<OP>(x.w, y.w)));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const int64_t i = (nquads << 2) + threadIdx.x;
    out[i] = arith1<OP>(load_cell(a, i, sa), load_cell(b, i, sb));
  }
}

// tft.scale_by_min_max: min < max ? (x - min) / (max - min) : 0.5, then * (omax - omin) + omin
__device__ __forceinline__ float scale1(float x, float lo, float range, bool ok, float oscale, float omin) {
  const float r = ok ? __fdiv_rn(__fsub_rn(x, lo), range) : 0.5f;
  return __fadd_rn(__fmul_rn(r, oscale), omin);
}
__global__ void __launch_bounds__(kSibBlock) map_scale_kernel(Operand a, const float* __restrict__ minmax, float omin, float omax,
                                                              float* __restrict__ out, int64_t n) {
  const float lo = __ldg(minmax), hi = __ldg(minmax + 1);
  const bool ok = lo < hi;
  const float range = __fsub_rn(hi, lo), oscale = __fsub_rn(omax, omin);
  const int64_t nquads = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kSibBlock;
  for (int64_t q = blockIdx.x * static_cast<int64_t>(kSibBlock) + threadIdx.x; q < nquads; q += stride) {
    const float4 x = load_quad(a, q, 0.f);
    stcs_f4(out + 4 * q, make_float4(scale1(x.x, lo, range, ok, oscale, omin), scale1(x.y, lo, range, ok, oscale, omin),
                                     scale1(x.z, lo, range, ok, oscale, omin), scale1(x.w, lo, range, ok, oscale, omin)));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const int64_t i = (nquads << 2) + threadIdx.x;
    out[i] = scale1(load_cell(a, i, 0.f), lo, range, ok, oscale, omin);
  }
}

// tf.to_int64 (float32 -> int64, truncation toward zero like the C++ cast TF uses) and tf.to_float
__global__ void __launch_bounds__(kSibBlock) map_cast_kernel(Operand a, int to_dtype, void* __restrict__ out, int64_t n) {
  const int64_t nquads = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kSibBlock;
  for (int64_t q = blockIdx.x * static_cast<int64_t>(kSibBlock) + threadIdx.x; q < nquads; q += stride) {
    const float4 x = load_quad(a, q, 0.f);
    if (to_dtype == TXF_F32) {
      stcs_f4(static_cast<float*>(out) + 4 * q, x);
    } else {
      long long* o = static_cast<long long*>(out) + 4 * q;
      stcs_l2(o, make_longlong2(__float2ll_rz(x.x), __float2ll_rz(x.y)));
      stcs_l2(o + 2, make_longlong2(__float2ll_rz(x.z), __float2ll_rz(x.w)));
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const int64_t i = (nquads << 2) + threadIdx.x;
    const float x = load_cell(a, i, 0.f);
    if (to_dtype == TXF_F32) static_cast<float*>(out)[i] = x;
    else static_cast<long long*>(out)[i] = __float2ll_rz(x);
  }
}

// tf.as_string on an int64 feature (kubeflow-pipelines/taxi/preprocessing.py:82): shortest decimal
// text, '-' for negatives; an absent cell is the filled value 0, i.e. "0".
__device__ __forceinline__ int dec_len(long long v) {
  unsigned long long m = v < 0 ? 0ULL - static_cast<unsigned long long>(v) : static_cast<unsigned long long>(v);
  int len = v < 0 ? 2 : 1;
  while (m >= 10ULL) { m /= 10ULL; ++len; }
  return len;
}
__global__ void i64_format_len_kernel(const long long* __restrict__ v, const uint8_t* __restrict__ valid, int64_t n,
                                      int32_t* __restrict__ len) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i < n) len[i] = dec_len(row_valid(valid, i) ? v[i] : 0LL);
}
__global__ void i64_format_write_kernel(const long long* __restrict__ v, const uint8_t* __restrict__ valid, int64_t n,
                                        const int32_t* __restrict__ offsets, uint8_t* __restrict__ pool) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const long long x = row_valid(valid, i) ? v[i] : 0LL;
  unsigned long long m = x < 0 ? 0ULL - static_cast<unsigned long long>(x) : static_cast<unsigned long long>(x);
  uint8_t* b = pool + offsets[i];
  int e = offsets[i + 1] - offsets[i];
  if (x < 0) b[0] = '-';
  do {
    b[--e] = static_cast<uint8_t>('0' + m % 10ULL);
    m /= 10ULL;
  } while (m);
}

// one-shot grid for the streaming maps: one quad per thread (tools/probes/copy_probe.cu: one-shot
// 256-thread blocks move more bytes per second here than a capped, grid-striding launch)
static unsigned map_grid(int64_t n) {
  const int64_t need = ((n >> 2) + kSibBlock - 1) / kSibBlock;
  return static_cast<unsigned>(need < 1 ? 1 : need);
}

static int sib_grid(int64_t n) {
  const int64_t need = ((n >> 2) + kSibBlock - 1) / kSibBlock;
  return static_cast<int>(need < 1 ? 1 : (need < kSibGrid ? need : kSibGrid));
}

static int check_operand(const txf_operand* o, const char* who, const char* name, Operand* d) {
  TXF_REQUIRE(o != nullptr, "%s: operand %s is null", who, name);
  TXF_REQUIRE((o->col != nullptr) != (o->scalar != nullptr), "%s: operand %s needs exactly one of col / scalar", who, name);
  if (o->col != nullptr) {
    TXF_REQUIRE(o->dtype == TXF_F32 || o->dtype == TXF_I64, "%s: operand %s has bad dtype %d", who, name, o->dtype);
    TXF_REQUIRE(aligned(o->col, 16), "%s: operand %s column not 16B aligned", who, name);
  }
  d->col = o->col;
  d->valid = static_cast<const uint8_t*>(o->valid);
  d->scalar = o->scalar;
  d->dtype = o->dtype;
  return TXF_OK;
}

}  // namespace txf

extern "C" {

size_t txf_minmax_workspace_bytes(int ncols) {
  if (ncols < 1 || ncols > TXF_MAX_COLS) return 0;
  return static_cast<size_t>(ncols) * txf::kSibGrid * 2 * sizeof(float);
}

int txf_minmax_begin(float* state_dev, int ncols, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(state_dev != nullptr && ncols >= 1 && ncols <= TXF_MAX_COLS, "minmax_begin: bad arguments");
  minmax_begin_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(state_dev, ncols);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_minmax_update(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols, int64_t n,
                      float* state_dev, void* workspace_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(values && dtype && state_dev && workspace_dev, "minmax_update: null argument");
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS, "minmax_update: ncols=%d out of [1,%d]", ncols, TXF_MAX_COLS);
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "minmax_update: n out of range");
  if (n == 0) return TXF_OK;
  NumColsView v{};
  for (int c = 0; c < ncols; ++c) {
    TXF_REQUIRE(values[c] != nullptr && aligned(values[c], 16), "minmax_update: column %d null or not 16B aligned", c);
    TXF_REQUIRE(dtype[c] == TXF_F32 || dtype[c] == TXF_I64, "minmax_update: column %d has bad dtype", c);
    v.values[c] = values[c];
    v.valid[c] = valid ? valid[c] : nullptr;
    v.dtype[c] = dtype[c];
  }
  const int grid = sib_grid(n);
  float* partials = static_cast<float*>(workspace_dev);
  minmax_partial_kernel<<<dim3(grid, ncols), kSibBlock, 0, static_cast<cudaStream_t>(stream)>>>(v, n, partials);
  TXF_LAUNCH_CHECK();
  minmax_fold_kernel<<<ncols, 32, 0, static_cast<cudaStream_t>(stream)>>>(partials, grid, 2, static_cast<int64_t>(grid) * 2, state_dev);
  TXF_LAUNCH_CHECK();
  count_launch(2);
  return TXF_OK;
}

int txf_minmax_merge(const float* states_dev, int nstates, int ncols, float* state_out_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(states_dev && state_out_dev && nstates >= 1 && ncols >= 1 && ncols <= TXF_MAX_COLS, "minmax_merge: bad arguments");
  minmax_fold_kernel<<<ncols, 32, 0, static_cast<cudaStream_t>(stream)>>>(states_dev, nstates, static_cast<int64_t>(ncols) * 2, 2,
                                                                          state_out_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_map_binary(int opcode, const txf_operand* a, const txf_operand* b, float* out_dev, int64_t n, txf_stream_t stream) {
  using namespace txf;
  Operand da, db;
  int rc = check_operand(a, "map_binary", "a", &da);
  if (rc) return rc;
  rc = check_operand(b, "map_binary", "b", &db);
  if (rc) return rc;
  TXF_REQUIRE(out_dev != nullptr && aligned(out_dev, 16), "map_binary: output null or not 16B aligned");
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "map_binary: n out of range");
  if (n == 0) return TXF_OK;
  const unsigned grid = map_grid(n);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (opcode) {
    case TXF_ARITH_ADD: map_binary_kernel<TXF_ARITH_ADD><<<grid, kSibBlock, 0, s>>>(da, db, out_dev, n); break;
    case TXF_ARITH_SUB: map_binary_kernel<TXF_ARITH_SUB><<<grid, kSibBlock, 0, s>>>(da, db, out_dev, n); break;
    case TXF_ARITH_MUL: map_binary_kernel<TXF_ARITH_MUL><<<grid, kSibBlock, 0, s>>>(da, db, out_dev, n); break;
    case TXF_ARITH_DIV: map_binary_kernel<TXF_ARITH_DIV><<<grid, kSibBlock, 0, s>>>(da, db, out_dev, n); break;
    default: TXF_REQUIRE(false, "map_binary: unknown opcode %d", opcode);
  }
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_map_cast(const txf_operand* x, int to_dtype, void* out_dev, int64_t n, txf_stream_t stream) {
  using namespace txf;
  Operand dx;
  int rc = check_operand(x, "map_cast", "x", &dx);
  if (rc) return rc;
  TXF_REQUIRE(dx.col != nullptr, "map_cast: x must be a column");
  TXF_REQUIRE(to_dtype == TXF_F32 || to_dtype == TXF_I64, "map_cast: bad target dtype %d", to_dtype);
  TXF_REQUIRE(out_dev != nullptr && aligned(out_dev, 16), "map_cast: output null or not 16B aligned");
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "map_cast: n out of range");
  if (n == 0) return TXF_OK;
  map_cast_kernel<<<map_grid(n), kSibBlock, 0, static_cast<cudaStream_t>(stream)>>>(dx, to_dtype, out_dev, n);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_i64_format_lengths(const int64_t* values_dev, const uint8_t* valid_dev, int64_t n, int32_t* len_out_dev,
                           txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(values_dev && len_out_dev, "i64_format_lengths: null argument");
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "i64_format_lengths: n out of range");
  if (n == 0) return TXF_OK;
  i64_format_len_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const long long*>(values_dev), valid_dev, n, len_out_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_i64_format_write(const int64_t* values_dev, const uint8_t* valid_dev, int64_t n, const int32_t* offsets_dev,
                         uint8_t* pool_out_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(values_dev && offsets_dev && pool_out_dev, "i64_format_write: null argument");
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "i64_format_write: n out of range");
  if (n == 0) return TXF_OK;
  i64_format_write_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const long long*>(values_dev), valid_dev, n, offsets_dev, pool_out_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_map_scale_minmax(const txf_operand* x, const float* minmax_dev, float out_min, float out_max, float* out_dev, int64_t n,
                         txf_stream_t stream) {
  using namespace txf;
  Operand dx;
  int rc = check_operand(x, "map_scale_minmax", "x", &dx);
  if (rc) return rc;
  TXF_REQUIRE(dx.col != nullptr, "map_scale_minmax: x must be a column");
  TXF_REQUIRE(minmax_dev != nullptr, "map_scale_minmax: needs the (min, max) state");
  TXF_REQUIRE(out_min < out_max, "map_scale_minmax: output_min must be less than output_max");
  TXF_REQUIRE(out_dev != nullptr && aligned(out_dev, 16), "map_scale_minmax: output null or not 16B aligned");
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "map_scale_minmax: n out of range");
  if (n == 0) return TXF_OK;
  map_scale_kernel<<<map_grid(n), kSibBlock, 0, static_cast<cudaStream_t>(stream)>>>(dx, minmax_dev, out_min, out_max, out_dev, n);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
