// copy_probe.cu -- which access pattern saturates HBM on this GPU?  int64 copy of N elements
// (N * 8 bytes in, N * 8 bytes out) with several kernel shapes, next to cudaMemcpyAsync D2D.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/copy_probe tools/probes/copy_probe.cu && /tmp/copy_probe
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

// A: persistent, tiles round-robin, U quads (16 B) in flight per thread
template <int U>
__global__ void __launch_bounds__(1024, 1) copy_persistent(const uint4* __restrict__ in, uint4* __restrict__ out, int64_t nv) {
  const int64_t tile = 1024LL * U;
  for (int64_t base = blockIdx.x * tile; base < nv; base += gridDim.x * tile) {
    uint4 r[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { int64_t i = base + threadIdx.x + u * 1024LL; if (i < nv) r[u] = __ldg(in + i); }
#pragma unroll
    for (int u = 0; u < U; ++u) { int64_t i = base + threadIdx.x + u * 1024LL; if (i < nv) __stcs(out + i, r[u]); }
  }
}
// B: one-shot grid, 256 threads, U vectors per thread
template <int U>
__global__ void __launch_bounds__(256) copy_oneshot(const uint4* __restrict__ in, uint4* __restrict__ out, int64_t nv) {
  const int64_t base = blockIdx.x * (256LL * U);
  uint4 r[U];
#pragma unroll
  for (int u = 0; u < U; ++u) { int64_t i = base + threadIdx.x + u * 256LL; if (i < nv) r[u] = __ldg(in + i); }
#pragma unroll
  for (int u = 0; u < U; ++u) { int64_t i = base + threadIdx.x + u * 256LL; if (i < nv) __stcs(out + i, r[u]); }
}
// C: persistent with 256-thread CTAs, several per SM
template <int U>
__global__ void __launch_bounds__(256) copy_persistent256(const uint4* __restrict__ in, uint4* __restrict__ out, int64_t nv) {
  const int64_t tile = 256LL * U;
  for (int64_t base = blockIdx.x * tile; base < nv; base += gridDim.x * tile) {
    uint4 r[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { int64_t i = base + threadIdx.x + u * 256LL; if (i < nv) r[u] = __ldg(in + i); }
#pragma unroll
    for (int u = 0; u < U; ++u) { int64_t i = base + threadIdx.x + u * 256LL; if (i < nv) __stcs(out + i, r[u]); }
  }
}
// D: 32-byte vectors (two adjacent uint4 per thread per access)
template <int U>
__global__ void __launch_bounds__(256) copy_oneshot32(const uint4* __restrict__ in, uint4* __restrict__ out, int64_t nv) {
  const int64_t base = blockIdx.x * (512LL * U);
  uint4 r[2 * U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    int64_t i = base + 2 * (threadIdx.x + u * 256LL);
    if (i + 1 < nv) { r[2 * u] = __ldg(in + i); r[2 * u + 1] = __ldg(in + i + 1); }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    int64_t i = base + 2 * (threadIdx.x + u * 256LL);
    if (i + 1 < nv) { __stcs(out + i, r[2 * u]); __stcs(out + i + 1, r[2 * u + 1]); }
  }
}

template <class F>
float timed(F f, int reps = 7) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  std::vector<float> t;
  for (int i = 0; i < reps; ++i) {
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); t.push_back(ms);
  }
  std::sort(t.begin(), t.end());
  return t[t.size() / 2];
}

int main() {
  const int64_t n = 100000000;              // int64 elements, like one pass-through column
  const int64_t bytes = n * 8, nv = bytes / 16;
  uint4 *in, *out;
  CK(cudaMalloc(&in, bytes)); CK(cudaMalloc(&out, bytes));
  CK(cudaMemset(in, 1, bytes));
  auto report = [&](const char* name, float ms) { printf("%-34s %7.3f ms  %7.1f GB/s\n", name, ms, 2.0 * bytes / ms / 1e6); };
  report("cudaMemcpyAsync D2D", timed([&] { cudaMemcpyAsync(out, in, bytes, cudaMemcpyDeviceToDevice); }));
  report("persistent 148x1024 U=4", timed([&] { copy_persistent<4><<<148, 1024>>>(in, out, nv); }));
  report("persistent 148x1024 U=2", timed([&] { copy_persistent<2><<<148, 1024>>>(in, out, nv); }));
  report("persistent 296x1024 U=4 (2 waves)", timed([&] { copy_persistent<4><<<296, 1024>>>(in, out, nv); }));
  report("oneshot 256 U=4", timed([&] { copy_oneshot<4><<<(unsigned)((nv + 1023) / 1024), 256>>>(in, out, nv); }));
  report("oneshot 256 U=8", timed([&] { copy_oneshot<8><<<(unsigned)((nv + 2047) / 2048), 256>>>(in, out, nv); }));
  report("oneshot 256 U=2", timed([&] { copy_oneshot<2><<<(unsigned)((nv + 511) / 512), 256>>>(in, out, nv); }));
  report("oneshot 256 U=1", timed([&] { copy_oneshot<1><<<(unsigned)((nv + 255) / 256), 256>>>(in, out, nv); }));
  report("persistent256 148*8 U=4", timed([&] { copy_persistent256<4><<<148 * 8, 256>>>(in, out, nv); }));
  report("persistent256 148*8 U=8", timed([&] { copy_persistent256<8><<<148 * 8, 256>>>(in, out, nv); }));
  report("persistent256 148*4 U=8", timed([&] { copy_persistent256<8><<<148 * 4, 256>>>(in, out, nv); }));
  report("oneshot32 256 U=2", timed([&] { copy_oneshot32<2><<<(unsigned)((nv + 1023) / 1024), 256>>>(in, out, nv); }));
  report("oneshot32 256 U=4", timed([&] { copy_oneshot32<4><<<(unsigned)((nv + 2047) / 2048), 256>>>(in, out, nv); }));
  CK(cudaDeviceSynchronize());
  return 0;
}
