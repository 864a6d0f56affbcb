// api.cu -- error plumbing, version, launch counter and the Fingerprint64 test hook.
#include <atomic>
#include <cstdarg>
#include <cstring>

#include "common.cuh"
#include "farmhash.cuh"

namespace txf {

static thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return TXF_OK;
  set_error("CUDA error %s (%d) at %s", cudaGetErrorString(e), static_cast<int>(e), what);
  return TXF_ERR_CUDA;
}

void count_launch(int k) { g_launches.fetch_add(k, std::memory_order_relaxed); }

__global__ void fingerprint_kernel(const int32_t* __restrict__ offsets, const uint8_t* __restrict__ pool,
                                   int64_t n, uint64_t* __restrict__ out) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  int32_t b = offsets[i], e = offsets[i + 1];
  out[i] = fingerprint64(pool + b, static_cast<uint32_t>(e - b));
}

}  // namespace txf

extern "C" {

const char* txf_last_error(void) { return txf::g_err; }
int txf_abi_version(void) { return TXF_ABI_VERSION; }
int64_t txf_launch_count(void) { return txf::g_launches.load(); }

int txf_fingerprint64_batch(const int32_t* offsets_dev, const uint8_t* pool_dev, int64_t n, uint64_t* out_dev,
                            txf_stream_t stream) {
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "fingerprint64_batch: n out of range");
  if (n == 0) return TXF_OK;
  TXF_REQUIRE(offsets_dev && pool_dev && out_dev, "fingerprint64_batch: null pointer");
  int block = 256;
  int grid = static_cast<int>((n + block - 1) / block);
  txf::fingerprint_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(offsets_dev, pool_dev, n, out_dev);
  txf::count_launch(1);
  TXF_LAUNCH_CHECK();
  return TXF_OK;
}

}  // extern "C"
