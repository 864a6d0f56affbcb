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

// Shared-memory footprint: smem_bytes(cap, tile_rows).  `cap` must be a multiple of 16, `tile_rows`
// a multiple of 128.  Besides the term bytes, the tile's offsets (tile_rows + 4 int32) and its
// validity bytes (tile_rows / 8) are staged by the same barrier, so the compute warps touch no
// global memory at all on a full tile (the ragged last tile reads offsets/validity directly).
struct StrTilePipe {
  uint64_t* bars;     // [2]
  long long* a0;      // [2] pool offset that buffer byte 0 corresponds to
  int* staged;        // [2] 1 = bytes are in the buffer, 0 = span too large: read the pool directly
  int* meta;          // [2] 1 = offsets (+ validity) of the tile are staged too
  uint8_t* buf;       // 2 * cap bytes, 16-byte aligned
  int32_t* s_off;     // 2 * (tile_rows + 4)
  uint8_t* s_valid;   // 2 * tile_rows / 8
  int cap, tile_rows;
  const int32_t* offsets;
  const uint8_t* pool;
  const uint8_t* valid;
  int64_t n;
  bool pool_ok, meta_ok;  // 16-byte aligned bases (bulk copies need them)

  static __host__ __device__ constexpr int header_bytes() { return 64; }
  static __host__ __device__ constexpr int smem_bytes(int cap, int tile_rows) {
    return header_bytes() + 2 * cap + 2 * (tile_rows + 4) * 4 + 2 * (tile_rows / 8);
  }

  // all threads; `smem` 16-byte aligned
  __device__ __forceinline__ void init(uint8_t* smem, int cap_, int tile_rows_, const int32_t* off, const uint8_t* pool_,
                                       const uint8_t* valid_, int64_t n_) {
    bars = reinterpret_cast<uint64_t*>(smem);
    a0 = reinterpret_cast<long long*>(smem + 16);
    staged = reinterpret_cast<int*>(smem + 32);
    meta = reinterpret_cast<int*>(smem + 40);
    buf = smem + header_bytes();
    cap = cap_;
    tile_rows = tile_rows_;
    s_off = reinterpret_cast<int32_t*>(buf + 2 * static_cast<size_t>(cap));
    s_valid = reinterpret_cast<uint8_t*>(s_off + 2 * (tile_rows + 4));
    offsets = off;
    pool = pool_;
    valid = valid_;
    n = n_;
    pool_ok = (reinterpret_cast<uintptr_t>(pool_) & 15) == 0;
    meta_ok = (reinterpret_cast<uintptr_t>(off) & 15) == 0 && (reinterpret_cast<uintptr_t>(valid_) & 15) == 0;
    if (threadIdx.x == 0) {
      mbar_init(&bars[0], 1);
      mbar_init(&bars[1], 1);
      mbar_fence_init();
    }
    __syncthreads();
  }

  // ONE thread: start the copies of tile rows [r0, r1) into buffer b.  The buffer must no longer
  // be read by anyone (caller: __syncthreads() after the last consumer).
  __device__ __forceinline__ void issue(int64_t r0, int64_t r1, int b) {
    const long long b0 = __ldg(offsets + r0), e0 = __ldg(offsets + r1);
    const long long start = b0 & ~15LL;
    const long long need = ((e0 - start + 15) & ~15LL) + 32;  // window loads touch < 24 bytes past a term
    a0[b] = start;
    const bool do_bytes = pool_ok && need <= cap;
    const bool do_meta = meta_ok && (r1 - r0) == tile_rows && r0 + tile_rows + 4 <= n + 1;
    staged[b] = do_bytes;
    meta[b] = do_meta;
    const uint32_t off_bytes = static_cast<uint32_t>(tile_rows + 4) * 4u, val_bytes = static_cast<uint32_t>(tile_rows / 8);
    const uint32_t total = (do_bytes ? static_cast<uint32_t>(need) : 0u) + (do_meta ? off_bytes + (valid ? val_bytes : 0u) : 0u);
    if (total == 0) {
      mbar_arrive(&bars[b]);
      return;
    }
    fence_proxy_async();
    mbar_arrive_expect_tx(&bars[b], total);
    if (do_bytes) bulk_copy_g2s(buf + static_cast<size_t>(b) * cap, pool + start, static_cast<uint32_t>(need), &bars[b]);
    if (do_meta) {
      bulk_copy_g2s(s_off + static_cast<size_t>(b) * (tile_rows + 4), offsets + r0, off_bytes, &bars[b]);
      if (valid) bulk_copy_g2s(s_valid + static_cast<size_t>(b) * (tile_rows / 8), valid + (r0 >> 3), val_bytes, &bars[b]);
    }
  }

  // all threads: wait for buffer b (use number `use` of that buffer: parity = use & 1); returns a
  // base pointer such that the term at pool offset o is at base + o
  __device__ __forceinline__ const uint8_t* acquire(int b, uint32_t use) {
    mbar_wait(&bars[b], use & 1u);
    return staged[b] ? buf + static_cast<size_t>(b) * cap - a0[b] : pool;
  }
};

// Row accessors of the tile in buffer b (META = offsets/validity staged in shared memory)
template <bool META>
struct TileRows {
  const int32_t* off;    // META: shared copy indexed by (row - r0); else the global array indexed by row
  const uint8_t* valid;  // META: shared copy (bit index row - r0); else global bitmap (bit index row); may be null
  int64_t r0;
  __device__ __forceinline__ int32_t begin(int64_t row) const { return META ? off[row - r0] : __ldg(off + row); }
  __device__ __forceinline__ int32_t end(int64_t row) const { return META ? off[row - r0 + 1] : __ldg(off + row + 1); }
  __device__ __forceinline__ bool present(int64_t row) const {
    if (valid == nullptr) return true;
    const int64_t i = META ? row - r0 : row;
    const uint32_t byte = META ? valid[i >> 3] : __ldg(valid + (i >> 3));
    return (byte >> (i & 7)) & 1u;
  }
};
template <bool META>
__device__ __forceinline__ TileRows<META> tile_rows_of(const StrTilePipe& p, int b, int64_t r0) {
  TileRows<META> t;
  t.r0 = r0;
  if (META) {
    t.off = p.s_off + static_cast<size_t>(b) * (p.tile_rows + 4);
    t.valid = p.valid ? p.s_valid + static_cast<size_t>(b) * (p.tile_rows / 8) : nullptr;
  } else {
    t.off = p.offsets;
    t.valid = p.valid;
  }
  return t;
}

}  // namespace txf
