// common.cuh -- shared device helpers for the taxi analyze+transform kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/taxi_xf.h"

namespace txf {

constexpr int kSMs = 148;  // B200

// ---- error plumbing (thread-local message, returned through txf_last_error) ---------------
void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
void count_launch(int k);
#define TXF_CUDA(expr)                                   \
  do {                                                   \
    int _rc = ::txf::check_cuda((expr), #expr);          \
    if (_rc) return _rc;                                 \
  } while (0)
#define TXF_REQUIRE(cond, ...)                           \
  do {                                                   \
    if (!(cond)) {                                       \
      ::txf::set_error(__VA_ARGS__);                     \
      return TXF_ERR_ARG;                                \
    }                                                    \
  } while (0)
#define TXF_LAUNCH_CHECK() TXF_CUDA(cudaGetLastError())

inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

// ---- column views passed by value to kernels ---------------------------------------------
struct NumColsView {
  const void* values[TXF_MAX_COLS];
  const uint8_t* valid[TXF_MAX_COLS];
  int32_t dtype[TXF_MAX_COLS];
};

// ---- validity -------------------------------------------------------------------------------
// Arrow bitmap, LSB first; nullptr == every row present.
__device__ __forceinline__ bool row_valid(const uint8_t* valid, int64_t i) {
  return valid == nullptr || ((__ldg(valid + (i >> 3)) >> (i & 7)) & 1);
}
// 4 validity bits of the quad starting at row q4 (q4 % 4 == 0)
__device__ __forceinline__ uint32_t quad_valid(const uint8_t* valid, int64_t q4) {
  if (valid == nullptr) return 0xFu;
  return (static_cast<uint32_t>(__ldg(valid + (q4 >> 3))) >> (q4 & 4)) & 0xFu;
}

// ---- canonical order-preserving key of an fp32 (NaN last, -0 == +0) -----------------------
__host__ __device__ __forceinline__ uint32_t f32_bits(float x) {
#ifdef __CUDA_ARCH__
  return __float_as_uint(x);
#else
  union { float f; uint32_t u; } c; c.f = x; return c.u;
#endif
}
__host__ __device__ __forceinline__ float bits_f32(uint32_t u) {
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
__host__ __device__ __forceinline__ uint32_t f32_key(float x) {
#ifdef __CUDA_ARCH__
  // same function in seven instructions: x + 0 turns -0 into +0, the sign is spread by an
  // arithmetic shift (negative: flip all bits, else set the top bit), NaNs are overridden last
  const uint32_t u = __float_as_uint(__fadd_rn(x, 0.0f));
  const uint32_t k = u ^ (static_cast<uint32_t>(static_cast<int32_t>(u) >> 31) | 0x80000000u);
  return (u & 0x7FFFFFFFu) > 0x7F800000u ? 0xFFFFFFFFu : k;
#else
  uint32_t u = f32_bits(x);
  if ((u & 0x7FFFFFFFu) > 0x7F800000u) return 0xFFFFFFFFu;  // any NaN
  if (u == 0x80000000u) u = 0u;                               // -0 -> +0
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
#endif
}
__host__ __device__ __forceinline__ float key_f32(uint32_t k) {
  return bits_f32((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}

// value of a numeric cell as the fp32 the tft ops see (int64 -> fp32 round-to-nearest-even)
__device__ __forceinline__ float i64_to_f32(long long v) { return __ll2float_rn(v); }

// ---- 16-byte vector access -----------------------------------------------------------------
__device__ __forceinline__ float4 ldg_f4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ int4 ldg_i4(const int* p) { return __ldg(reinterpret_cast<const int4*>(p)); }
__device__ __forceinline__ longlong2 ldg_l2(const long long* p) {
  return __ldg(reinterpret_cast<const longlong2*>(p));
}
// streaming (evict-first) stores for outputs that are never re-read by this pass
__device__ __forceinline__ void stcs_f4(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }
__device__ __forceinline__ void stcs_l2(long long* p, longlong2 v) { __stcs(reinterpret_cast<longlong2*>(p), v); }
__device__ __forceinline__ void stcs_i4(int* p, int4 v) { __stcs(reinterpret_cast<int4*>(p), v); }

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

// ---- explicit shared-window access ----------------------------------------------------------
// Hot loops address shared memory with 32-bit shared-window addresses and ld/st.shared.  Through
// ordinary pointers ptxas re-derives the window base (S2R SR_CgaCtaId + LEA) at every access site,
// which was a fifth of the instructions of the counting kernels.
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ unsigned long long lds_u64(uint32_t a) { unsigned long long v; asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"(static_cast<uint16_t>(v)) : "memory"); }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t atoms_add_u32(uint32_t a, uint32_t v) { uint32_t o; asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(o) : "r"(a), "r"(v) : "memory"); return o; }
__device__ __forceinline__ void reds_add_u32(uint32_t a, uint32_t v) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

}  // namespace txf
