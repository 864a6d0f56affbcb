// moments.cu -- mean/variance analyzer of tft.scale_to_z_score (reference call site
// airflow-dags/taxi_utils.py:116-119; fill rule :86-103).
//
// HBM-bound streaming reduction: each thread reads 16-byte quads, accumulates shifted sums in
// fp64, converts to a Welford triple (n, mean, M2) and Chan-merges lanes by shuffles, warps
// through shared memory, blocks through a workspace that a single-CTA kernel folds -- in fixed
// index order, so the result is bit-reproducible for a given (n, grid).
// Algorithmic bytes: 4 (f32) or 8 (i64) per row per column + 1 validity bit.
#include <mutex>

#include "common.cuh"

namespace txf {

constexpr int kMomBlock = 256;
constexpr int kMomBlocksPerCol = kSMs * 8;  // workspace bound: one resident wave (8 CTAs of 256 per SM)

struct Triple { double n, mean, m2; };

__device__ __forceinline__ Triple chan(Triple a, Triple b) {
  if (b.n == 0.0) return a;
  if (a.n == 0.0) return b;
  Triple r;
  r.n = a.n + b.n;
  double d = b.mean - a.mean;
  r.mean = a.mean + d * (b.n / r.n);
  r.m2 = a.m2 + b.m2 + d * d * (a.n * b.n / r.n);
  return r;
}

struct Acc {
  double k, s1, s2;
  uint32_t cnt;
  __device__ __forceinline__ void add(float x) {
    double xd = static_cast<double>(x);
    if (cnt == 0) k = xd;
    double d = xd - k;
    s1 += d;
    s2 = fma(d, d, s2);
    ++cnt;
  }
  __device__ __forceinline__ Triple triple() const {
    Triple t;
    t.n = static_cast<double>(cnt);
    if (cnt == 0) { t.mean = 0.0; t.m2 = 0.0; return t; }
    t.mean = k + s1 / t.n;
    t.m2 = fmax(s2 - s1 * s1 / t.n, 0.0);
    return t;
  }
};

__global__ void __launch_bounds__(kMomBlock) moments_partial_kernel(NumColsView cols, int64_t n,
                                                                    double* __restrict__ partials) {
  const int c = blockIdx.y;
  const uint8_t* __restrict__ valid = cols.valid[c];
  Acc acc{0.0, 0.0, 0.0, 0u};
  const int64_t nquads = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kMomBlock;
  int64_t q = blockIdx.x * static_cast<int64_t>(kMomBlock) + threadIdx.x;
  if (cols.dtype[c] == TXF_F32) {
    const float* __restrict__ v = static_cast<const float*>(cols.values[c]);
    for (; q + stride < nquads; q += 2 * stride) {  // two quads in flight
      float4 a = ldg_f4(v + 4 * q), b = ldg_f4(v + 4 * (q + stride));
      uint32_t ma = quad_valid(valid, 4 * q), mb = quad_valid(valid, 4 * (q + stride));
      acc.add(ma & 1 ? a.x : 0.f); acc.add(ma & 2 ? a.y : 0.f); acc.add(ma & 4 ? a.z : 0.f); acc.add(ma & 8 ? a.w : 0.f);
      acc.add(mb & 1 ? b.x : 0.f); acc.add(mb & 2 ? b.y : 0.f); acc.add(mb & 4 ? b.z : 0.f); acc.add(mb & 8 ? b.w : 0.f);
    }
    for (; q < nquads; q += stride) {
      float4 a = ldg_f4(v + 4 * q);
      uint32_t ma = quad_valid(valid, 4 * q);
      acc.add(ma & 1 ? a.x : 0.f); acc.add(ma & 2 ? a.y : 0.f); acc.add(ma & 4 ? a.z : 0.f); acc.add(ma & 8 ? a.w : 0.f);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // tail rows
      int64_t i = (nquads << 2) + threadIdx.x;
      acc.add(row_valid(valid, i) ? v[i] : 0.f);
    }
  } else {
    const long long* __restrict__ v = static_cast<const long long*>(cols.values[c]);
    for (; q < nquads; q += stride) {
      longlong2 a = ldg_l2(v + 4 * q), b = ldg_l2(v + 4 * q + 2);
      uint32_t ma = quad_valid(valid, 4 * q);
      acc.add(ma & 1 ? i64_to_f32(a.x) : 0.f); acc.add(ma & 2 ? i64_to_f32(a.y) : 0.f);
      acc.add(ma & 4 ? i64_to_f32(b.x) : 0.f); acc.add(ma & 8 ? i64_to_f32(b.y) : 0.f);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
      int64_t i = (nquads << 2) + threadIdx.x;
      acc.add(row_valid(valid, i) ? i64_to_f32(v[i]) : 0.f);
    }
  }
  Triple t = acc.triple();
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {  // butterfly: every lane ends with the same merge tree
    Triple o;
    o.n = __shfl_xor_sync(0xffffffffu, t.n, off);
    o.mean = __shfl_xor_sync(0xffffffffu, t.mean, off);
    o.m2 = __shfl_xor_sync(0xffffffffu, t.m2, off);
    // merge (lower lane, upper lane) in that order on both sides so the result is identical
    t = (lane_id() & off) ? chan(o, t) : chan(t, o);
  }
  __shared__ Triple warp_t[kMomBlock / 32];
  if (lane_id() == 0) warp_t[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    Triple r = warp_t[0];
#pragma unroll
    for (int w = 1; w < kMomBlock / 32; ++w) r = chan(r, warp_t[w]);
    double* p = partials + (static_cast<int64_t>(c) * gridDim.x + blockIdx.x) * 3;
    p[0] = r.n; p[1] = r.mean; p[2] = r.m2;
  }
}

// one warp per column folds the block partials: lane l takes partials l, l+32, ... in order, then
// the lanes merge in a fixed butterfly -- a fixed tree, so the result is reproducible -- and the
// running state is merged last
__global__ void moments_fold_kernel(const double* __restrict__ partials, int nblocks, int ncols,
                                    double* __restrict__ state) {
  const int c = blockIdx.x;
  if (c >= ncols) return;
  Triple t{0.0, 0.0, 0.0};
  const double* p = partials + static_cast<int64_t>(c) * nblocks * 3;
  for (int b = lane_id(); b < nblocks; b += 32) t = chan(t, Triple{p[3 * b], p[3 * b + 1], p[3 * b + 2]});
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    Triple o;
    o.n = __shfl_xor_sync(0xffffffffu, t.n, off);
    o.mean = __shfl_xor_sync(0xffffffffu, t.mean, off);
    o.m2 = __shfl_xor_sync(0xffffffffu, t.m2, off);
    t = (lane_id() & off) ? chan(o, t) : chan(t, o);
  }
  if (lane_id() == 0) {
    Triple s{state[3 * c], state[3 * c + 1], state[3 * c + 2]};
    s = chan(s, t);
    state[3 * c] = s.n; state[3 * c + 1] = s.mean; state[3 * c + 2] = s.m2;
  }
}

__global__ void moments_merge_kernel(const double* __restrict__ states, int nstates, int ncols,
                                     double* __restrict__ out) {
  int c = threadIdx.x;
  if (c >= ncols) return;
  Triple r{0.0, 0.0, 0.0};
  for (int s = 0; s < nstates; ++s) {
    const double* p = states + (static_cast<int64_t>(s) * ncols + c) * 3;
    r = chan(r, Triple{p[0], p[1], p[2]});
  }
  out[3 * c] = r.n; out[3 * c + 1] = r.mean; out[3 * c + 2] = r.m2;
}

__global__ void moments_finalize_kernel(const double* __restrict__ state, int ncols, float* __restrict__ mv) {
  int c = threadIdx.x;
  if (c >= ncols) return;
  double n = state[3 * c];
  mv[2 * c] = static_cast<float>(state[3 * c + 1]);
  mv[2 * c + 1] = static_cast<float>(n > 0.0 ? state[3 * c + 2] / n : 0.0);
}

// ---- scratch for txf_moments_update_noalloc -------------------------------------------------
// Block partials of up to TXF_NOALLOC_MAX_COLS columns live in this file-scope device array
// instead of a caller-provided workspace (~111 KB, held by every GPU that loads the library until
// the process exits).  In host code the symbol is only a stand-in: its device address has to be
// asked for, once per device and with that device current.
static __device__ __align__(256) double g_noalloc_partials[TXF_NOALLOC_MAX_COLS * kMomBlocksPerCol * 3];

constexpr int kMaxDevices = 64;
static std::mutex g_noalloc_mutex;
static double* g_noalloc_ptr[kMaxDevices] = {};  // device address of g_noalloc_partials, per device

// Device address of the scratch array on the CURRENT device (looked up on first use).
static int noalloc_scratch(double** out) {
  int dev = -1;
  TXF_CUDA(cudaGetDevice(&dev));
  TXF_REQUIRE(dev >= 0 && dev < kMaxDevices, "moments_update_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    TXF_CUDA(cudaGetSymbolAddress(&p, g_noalloc_partials));
    g_noalloc_ptr[dev] = static_cast<double*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return TXF_OK;
}

static int moments_grid(int64_t n, int ncols) {
  int64_t quads = (n >> 2) + 1;
  int64_t want = (quads + kMomBlock - 1) / kMomBlock;
  int64_t cap = kMomBlocksPerCol / ncols;  // all columns together fill the machine once
  if (cap < 1) cap = 1;
  return static_cast<int>(want < cap ? want : cap);
}

}  // namespace txf

extern "C" {

size_t txf_moments_workspace_bytes(int ncols) {
  return static_cast<size_t>(ncols > 0 ? ncols : 0) * txf::kMomBlocksPerCol * 3 * sizeof(double);
}

int txf_moments_update(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols,
                       int64_t n, double* state_dev, void* workspace_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS, "moments_update: ncols=%d out of [1,%d]", ncols, TXF_MAX_COLS);
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "moments_update: n out of range");
  TXF_REQUIRE(state_dev && workspace_dev && values && dtype, "moments_update: null pointer");
  if (n == 0) return TXF_OK;
  NumColsView v{};
  for (int c = 0; c < ncols; ++c) {
    TXF_REQUIRE(values[c] && aligned(values[c], 16), "moments_update: column %d values null or not 16B aligned", c);
    TXF_REQUIRE(dtype[c] == TXF_F32 || dtype[c] == TXF_I64, "moments_update: column %d bad dtype", c);
    v.values[c] = values[c];
    v.valid[c] = valid ? valid[c] : nullptr;
    v.dtype[c] = dtype[c];
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int gx = moments_grid(n, ncols);
  moments_partial_kernel<<<dim3(gx, ncols), kMomBlock, 0, s>>>(v, n, static_cast<double*>(workspace_dev));
  TXF_LAUNCH_CHECK();
  moments_fold_kernel<<<ncols, 32, 0, s>>>(static_cast<const double*>(workspace_dev), gx, ncols, state_dev);
  TXF_LAUNCH_CHECK();
  count_launch(2);
  return TXF_OK;
}

int txf_moments_update_noalloc(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols,
                               int64_t n, double* state_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_NOALLOC_MAX_COLS,
              "moments_update_noalloc: ncols=%d exceeds the built-in scratch (at most %d columns per call; "
              "use txf_moments_update with a caller workspace for more)", ncols, TXF_NOALLOC_MAX_COLS);
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "moments_update_noalloc: n out of range");
  TXF_REQUIRE(state_dev && values && dtype, "moments_update_noalloc: null pointer");
  if (n == 0) return TXF_OK;
  NumColsView v{};
  for (int c = 0; c < ncols; ++c) {
    TXF_REQUIRE(values[c] && aligned(values[c], 16), "moments_update_noalloc: column %d values null or not 16B aligned", c);
    TXF_REQUIRE(dtype[c] == TXF_F32 || dtype[c] == TXF_I64, "moments_update_noalloc: column %d bad dtype", c);
    v.values[c] = values[c];
    v.valid[c] = valid ? valid[c] : nullptr;
    v.dtype[c] = dtype[c];
  }
  double* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int gx = moments_grid(n, ncols);
  // every partial the fold reads is written by the launch before it, so nothing needs clearing
  moments_partial_kernel<<<dim3(gx, ncols), kMomBlock, 0, s>>>(v, n, scratch);
  TXF_LAUNCH_CHECK();
  moments_fold_kernel<<<ncols, 32, 0, s>>>(scratch, gx, ncols, state_dev);
  TXF_LAUNCH_CHECK();
  count_launch(2);
  return TXF_OK;
}

int txf_moments_merge(const double* states_dev, int nstates, int ncols, double* state_out_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS && nstates >= 1, "moments_merge: bad sizes");
  TXF_REQUIRE(states_dev && state_out_dev, "moments_merge: null pointer");
  moments_merge_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(states_dev, nstates, ncols, state_out_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_moments_finalize(const double* state_dev, int ncols, float* mean_var_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_MAX_COLS, "moments_finalize: bad ncols");
  TXF_REQUIRE(state_dev && mean_var_dev, "moments_finalize: null pointer");
  moments_finalize_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(state_dev, ncols, mean_var_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
