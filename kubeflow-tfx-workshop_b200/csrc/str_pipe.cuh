// str_pipe.cuh -- double-buffered staging of string-column tiles into shared memory with the
// bulk async-copy engine (TMA 1-D: cp.async.bulk + mbarrier complete_tx; SASS UBLKCP).
//
// The term bytes of rows [r0, r1) of an Arrow string column are one contiguous span of the pool.
// One elected thread per CTA asks the copy engine to land that span (16-byte aligned, padded) in
// a shared-memory buffer and arms an mbarrier with the byte count; all threads later wait on the
// barrier and read their own terms from shared memory (an unaligned 8-byte fetch is two LDS
// instead of two scattered LDG).  Two buffers: tile k+1 streams in while tile k is processed, and
// the copy costs no issue slots of the compute warps.
#pragma once
#include "common.cuh"

namespace txf {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Shared-memory footprint: header (64 B) + 2 * cap bytes.  `cap` must be a multiple of 16.
struct StrTilePipe {
  uint64_t* bars;     // [2]
  long long* a0;      // [2] pool offset that buffer byte 0 corresponds to
  int* staged;        // [2] 1 = bytes are in the buffer, 0 = span too large: read the pool directly
  uint8_t* buf;       // 2 * cap bytes, 16-byte aligned
  int cap;
  const int32_t* offsets;
  const uint8_t* pool;
  bool pool_ok;       // pool base is 16-byte aligned (bulk copies need it)

  static __host__ __device__ constexpr int header_bytes() { return 64; }

  // all threads; `smem` 16-byte aligned
  __device__ __forceinline__ void init(uint8_t* smem, int cap_, const int32_t* off, const uint8_t* pool_) {
    bars = reinterpret_cast<uint64_t*>(smem);
    a0 = reinterpret_cast<long long*>(smem + 16);
    staged = reinterpret_cast<int*>(smem + 32);
    buf = smem + header_bytes();
    cap = cap_;
    offsets = off;
    pool = pool_;
    pool_ok = (reinterpret_cast<uintptr_t>(pool_) & 15) == 0;
    if (threadIdx.x == 0) {
      mbar_init(&bars[0], 1);
      mbar_init(&bars[1], 1);
      mbar_fence_init();
    }
    __syncthreads();
  }

  // ONE thread: start the copy of rows [r0, r1) into buffer b.  The buffer must no longer be read
  // by anyone (caller: __syncthreads() after the last consumer).
  __device__ __forceinline__ void issue(int64_t r0, int64_t r1, int b) {
    const long long b0 = __ldg(offsets + r0), e0 = __ldg(offsets + r1);
    const long long start = b0 & ~15LL;
    const long long need = ((e0 - start + 15) & ~15LL) + 16;  // fetch64 touches < 16 bytes past a term
    a0[b] = start;
    if (pool_ok && need <= cap) {
      staged[b] = 1;
      fence_proxy_async();
      mbar_arrive_expect_tx(&bars[b], static_cast<uint32_t>(need));
      bulk_copy_g2s(buf + static_cast<size_t>(b) * cap, pool + start, static_cast<uint32_t>(need), &bars[b]);
    } else {
      staged[b] = 0;
      mbar_arrive(&bars[b]);
    }
  }

  // all threads: wait for buffer b (use number `use` of that buffer: parity = use & 1); returns a
  // base pointer such that the term at pool offset o is at base + o
  __device__ __forceinline__ const uint8_t* acquire(int b, uint32_t use) {
    mbar_wait(&bars[b], use & 1u);
    return staged[b] ? buf + static_cast<size_t>(b) * cap - a0[b] : pool;
  }
};

}  // namespace txf
