// vocab.cuh -- term identity, table layouts and string-tile staging shared by the vocabulary
// analyzer (vocab.cu) and the fused transform's lookup (transform.cu).
#pragma once
#include "common.cuh"
#include "farmhash.cuh"

namespace txf {

// ---------------------------------------------------------------------------------------------
// Term identity.  Counting and looking up terms does not need FarmHash (only the OOV bucket of a
// term that is NOT in the vocabulary does, reference call site taxi_utils.py:123-126), so the
// tables are keyed by a cheaper 128-bit value (a, b):
//   * terms of <= 15 bytes: the bytes themselves (zero padded) with the length in b's top byte --
//     exact, no hashing at all (payment types, most categorical strings);
//   * longer terms: a FINGERPRINT -- two independent 64-bit multiply-fold hashes of all bytes, with
//     the length (saturated at 127) stored verbatim in bits 55..61 of b and the top bit of b marking
//     the class, so the two classes never meet.  In the global count table, the id lookup and the
//     decode table a fingerprint only pre-filters: equality of long terms is decided on the BYTES
//     (the reference, tft, counts byte strings), so a collision there costs a few extra probes,
//     never a merged count or a wrong id.  Only the per-CTA front table of the count pass (a cache in
//     front of the global table) trusts the full 119-bit fingerprint + exact length.
// a == 0 and b == 0 are reserved for "unclaimed slot": b is never 0 by construction, and a == 0
// is remapped to 1 with a flag bit in b.
// ---------------------------------------------------------------------------------------------
struct StrKey {
  unsigned long long a, b;
};

constexpr uint32_t kExactKeyMax = 15;  // terms up to this length are their own identity
constexpr uint32_t kKeyLenSat = 127;   // length field of a fingerprint saturates here
constexpr unsigned long long kFpBitsB = (1ULL << 55) - 1;  // hash bits kept in word b of a fingerprint

__host__ __device__ __forceinline__ bool key_is_fingerprint(unsigned long long b) { return (b >> 63) != 0; }

// Test hook (txf_debug_set_identity_bits): the hash bits of a fingerprint are ANDed with these masks,
// so a test can make fingerprints collide at will and check that results stay byte-exact.
#ifdef __CUDACC__
static __constant__ unsigned long long c_idmask[2] = {~0ULL, kFpBitsB};
static inline cudaError_t set_identity_mask_this_tu(unsigned long long ma, unsigned long long mb) {
  const unsigned long long m[2] = {ma, mb & kFpBitsB};
  return cudaMemcpyToSymbol(c_idmask, m, sizeof(m));
}
#endif
__host__ __device__ __forceinline__ unsigned long long identity_mask(int i) {
#ifdef __CUDA_ARCH__
  return c_idmask[i];
#else
  return i ? kFpBitsB : ~0ULL;
#endif
}

__host__ __device__ __forceinline__ StrKey finish_key(uint64_t k0, uint64_t k1) {
  if (k0 == 0) {  // a == 0 marks an unclaimed slot
    k0 = 1;
    k1 |= 0x4000000000000000ULL;
  }
  return StrKey{k0, k1};
}

// Mixing of a long term.  Like FarmHash, terms of 16..32 bytes are covered by four (overlapping)
// 8-byte windows -- first 16 and last 16 bytes -- so there is no loop, no tail and no masking;
// 33..64 bytes take eight windows (first 32 and last 32 bytes), longer terms a loop over 32-byte
// blocks plus the last 32 bytes.
// The mixer is two independent 64-bit lanes of the multiply-fold kind: each lane keeps two 32-bit
// words, XORs one half of the window into each (x, y) and replaces them by the low and high half of
// x * y + (x << 32 | y) -- one wide multiply-add.  The added pair keeps a zero factor from wiping
// the lane: with x == 0 the state still carries y and vice versa (a bare product would map every
// prefix that reaches x == 0 to the same state).  Three instructions per lane and window on the GPU
// (two 3-input XORs, one IMAD.WIDE) where a 64-bit multiply-rotate chain costs about ten.
// The lanes use different constants, different starting words and opposite halves of the window.
__host__ __device__ __forceinline__ uint64_t fold_mul(uint32_t x, uint32_t y) {
  return static_cast<uint64_t>(x) * y + ((static_cast<uint64_t>(x) << 32) | y);
}
struct KeyMixer {
  uint32_t a1, b1, a2, b2, n;
  __host__ __device__ __forceinline__ explicit KeyMixer(uint32_t n_)
      : a1(n_ * 0x9E3779B1u + 0x7F4A7C15u), b1(n_ ^ 0x632BE5ABu), a2((n_ ^ 0xD6E8FEB8u) * 0xC2B2AE3Du), b2(n_ + 0x27D4EB2Fu), n(n_) {}
  __host__ __device__ __forceinline__ void add(uint32_t lo, uint32_t hi) {
    const uint64_t p = fold_mul(a1 ^ lo ^ 0x53C5CA59u, b1 ^ hi ^ 0x74743C1Bu);
    const uint64_t q = fold_mul(a2 ^ hi ^ 0xA0761D65u, b2 ^ lo ^ 0xE7037ED1u);
    a1 = static_cast<uint32_t>(p); b1 = static_cast<uint32_t>(p >> 32);
    a2 = static_cast<uint32_t>(q); b2 = static_cast<uint32_t>(q >> 32);
  }
  __host__ __device__ __forceinline__ void add(uint64_t w) { add(static_cast<uint32_t>(w), static_cast<uint32_t>(w >> 32)); }
  __host__ __device__ __forceinline__ StrKey finish() {
    // two data-free rounds so the words of the last window reach every output bit.  The low half of
    // a product is its weak half, so the low output word also takes the high half (lo ^ hi) and
    // the high output word is crossed with the previous round's low word.
    const uint64_t p = fold_mul(a1 ^ 0x8EBC6AF1u, b1 ^ 0x589965CDu);
    const uint64_t q = fold_mul(a2 ^ 0x1D8E4E27u, b2 ^ 0xEB44ACCBu);
    const uint32_t pl = static_cast<uint32_t>(p), ph = static_cast<uint32_t>(p >> 32);
    const uint32_t ql = static_cast<uint32_t>(q), qh = static_cast<uint32_t>(q >> 32);
    const uint64_t r = fold_mul(pl ^ 0x53C5CA59u, ph ^ 0x74743C1Bu);
    const uint64_t t = fold_mul(ql ^ 0xA0761D65u, qh ^ 0xE7037ED1u);
    const uint32_t rl = static_cast<uint32_t>(r), rh = static_cast<uint32_t>(r >> 32);
    const uint32_t tl = static_cast<uint32_t>(t), th = static_cast<uint32_t>(t >> 32);
    const uint64_t k0 = (static_cast<uint64_t>(rl ^ rh) | (static_cast<uint64_t>(rh ^ pl) << 32)) & identity_mask(0);
    const uint64_t k1 = (static_cast<uint64_t>(tl ^ th) | (static_cast<uint64_t>(th ^ ql) << 32)) & identity_mask(1);
    const uint64_t len7 = n < kKeyLenSat ? n : kKeyLenSat;
    return finish_key(k0, k1 | (len7 << 55) | 0x8000000000000000ULL);  // top byte 0x80..0xBF (0xC0.. with the a == 0 flag)
  }
};

// `fetch(off)` returns the 8 bytes at byte offset `off` of the term (little endian; bytes past the
// term may be anything) -- a functor so that the same code reads global or shared memory.
template <class Fetch>
__host__ __device__ __forceinline__ StrKey make_key_exact(Fetch fetch, uint32_t n) {  // n <= kExactKeyMax
  uint64_t k0 = 0, k1 = 0;
  if (n) {
    k0 = fetch(0);
    if (n < 8) k0 &= (1ULL << (8 * n)) - 1;
    if (n > 8) k1 = fetch(8) & ((1ULL << (8 * (n - 8))) - 1);
  }
  return finish_key(k0, k1 | (static_cast<uint64_t>(n + 1) << 56));  // top byte 1..16
}

template <class Fetch>
__host__ __device__ __forceinline__ StrKey make_key_hashed(Fetch fetch, uint32_t n) {  // n > kExactKeyMax
  KeyMixer m(n);
  if (n <= 32) {
    m.add(fetch(0)); m.add(fetch(8)); m.add(fetch(n - 16)); m.add(fetch(n - 8));
  } else if (n <= 64) {
    m.add(fetch(0)); m.add(fetch(8)); m.add(fetch(16)); m.add(fetch(24));
    m.add(fetch(n - 32)); m.add(fetch(n - 24)); m.add(fetch(n - 16)); m.add(fetch(n - 8));
  } else {
    uint32_t i = 0;
    for (; i + 32 <= n - 32; i += 32) { m.add(fetch(i)); m.add(fetch(i + 8)); m.add(fetch(i + 16)); m.add(fetch(i + 24)); }
    if (i < n - 32) { m.add(fetch(i)); m.add(fetch(i + 8)); m.add(fetch(i + 16)); m.add(fetch(i + 24)); }
    m.add(fetch(n - 32)); m.add(fetch(n - 24)); m.add(fetch(n - 16)); m.add(fetch(n - 8));
  }
  return m.finish();
}

struct PtrFetch {  // generic pointer (global memory, or host memory in the CPU-side unit test)
  const uint8_t* s;
  __host__ __device__ __forceinline__ uint64_t operator()(uint32_t off) const { return fetch64(s + off); }
};

__host__ __device__ inline StrKey make_key(const uint8_t* s, uint32_t n) {
  return n <= kExactKeyMax ? make_key_exact(PtrFetch{s}, n) : make_key_hashed(PtrFetch{s}, n);
}

// Byte equality of two terms of the same length n (generic pointers, any alignment; both buffers
// must be readable up to the next 8-byte boundary past the term plus 8, like every pool and arena).
__host__ __device__ inline bool term_bytes_equal(const uint8_t* x, const uint8_t* y, uint32_t n) {
  uint32_t i = 0;
  for (; i + 8 <= n; i += 8)
    if (fetch64(x + i) != fetch64(y + i)) return false;
  if (i < n) {
    const uint64_t m = (1ULL << (8 * (n - i))) - 1;
    if ((fetch64(x + i) ^ fetch64(y + i)) & m) return false;
  }
  return true;
}

#ifdef __CUDACC__
// ---- shared-memory fast paths (terms staged by StrTilePipe); same identities as above -------------
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
// the 16 bytes at shared address `addr` (any alignment) as four little-endian words
__device__ __forceinline__ void smem_words16(uint32_t addr, uint32_t (&w)[4]) {
  const uint32_t al = addr & ~3u, sh = (addr & 3u) << 3;
  const uint32_t x0 = lds32(al), x1 = lds32(al + 4), x2 = lds32(al + 8), x3 = lds32(al + 12), x4 = lds32(al + 16);
  w[0] = __funnelshift_r(x0, x1, sh); w[1] = __funnelshift_r(x1, x2, sh);
  w[2] = __funnelshift_r(x2, x3, sh); w[3] = __funnelshift_r(x3, x4, sh);
}
struct SmemFetch {  // 8 bytes at byte offset `off` of a staged term
  uint32_t addr;
  __device__ __forceinline__ uint64_t operator()(uint32_t off) const {
    const uint32_t a = addr + off, al = a & ~3u, sh = (a & 3u) << 3;
    const uint32_t x0 = lds32(al), x1 = lds32(al + 4), x2 = lds32(al + 8);
    return static_cast<uint64_t>(__funnelshift_r(x0, x1, sh)) | (static_cast<uint64_t>(__funnelshift_r(x1, x2, sh)) << 32);
  }
};
// byte masks of a short term: s_mask[n] keeps the first n of 16 bytes (n = 0..15), a 256-byte table
__device__ __forceinline__ void fill_short_masks(uint4* s_mask) {
  if (threadIdx.x < 16) {
    const int n = threadIdx.x;
    auto m = [&](int j) { int k = n - 4 * j; return k >= 4 ? 0xFFFFFFFFu : (k <= 0 ? 0u : (1u << (8 * k)) - 1u); };
    s_mask[n] = make_uint4(m(0), m(1), m(2), m(3));
  }
}
__device__ __forceinline__ StrKey make_key_exact_smem(uint32_t addr, uint32_t n, const uint4* s_mask) {
  uint32_t w[4] = {0u, 0u, 0u, 0u};
  if (n) {
    smem_words16(addr, w);
    const uint4 m = s_mask[n];
    w[0] &= m.x; w[1] &= m.y; w[2] &= m.z; w[3] &= m.w;
  }
  w[3] |= (n + 1) << 24;
  return finish_key(static_cast<uint64_t>(w[0]) | (static_cast<uint64_t>(w[1]) << 32),
                    static_cast<uint64_t>(w[2]) | (static_cast<uint64_t>(w[3]) << 32));
}
__device__ __forceinline__ StrKey make_key_hashed_smem(uint32_t addr, uint32_t n) {
  if (n <= 32) {  // two 16-byte windows, five word loads each
    uint32_t f[4], l[4];
    smem_words16(addr, f);
    smem_words16(addr + n - 16, l);
    KeyMixer m(n);
    m.add(f[0], f[1]); m.add(f[2], f[3]); m.add(l[0], l[1]); m.add(l[2], l[3]);
    return m.finish();
  }
  return make_key_hashed(SmemFetch{addr}, n);
}
#endif

// home slot of an identity: four 32-bit multiplies and a murmur-style fold (tables hold < 2^32 slots)
__host__ __device__ __forceinline__ uint64_t home_slot(uint64_t a, uint64_t b, uint64_t mask) {
  uint32_t h = static_cast<uint32_t>(a) * 0x9E3779B1u ^ static_cast<uint32_t>(a >> 32) * 0x85EBCA77u ^
               static_cast<uint32_t>(b) * 0xC2B2AE3Du ^ static_cast<uint32_t>(b >> 32) * 0x27D4EB2Fu;
  h ^= h >> 16;
  h *= 0x85EBCA6Bu;
  h ^= h >> 13;
  return static_cast<uint64_t>(h) & mask;
}

// Count table slot: 32 bytes = one DRAM sector.
struct CountSlot {
  unsigned long long a, b, count, strref;  // strref = (arena offset << 24) | length
};
static_assert(sizeof(CountSlot) == 32, "CountSlot must be one 32-byte sector");
constexpr uint32_t kMaxTermLen = (1u << 24) - 1;

// Lookup table slot (built once per vocabulary, probed per row): 32 bytes = one DRAM sector.
// strref = (byte offset << 24) | length of the term in the byte pool the table was built over (the
// count table's arena, or the pool of an explicit vocabulary): a fingerprint hit is confirmed there.
struct LookupSlot {
  unsigned long long a, b;  // b == 0: empty
  unsigned long long strref;
  int32_t id;
  int32_t pad;
};
static_assert(sizeof(LookupSlot) == 32, "LookupSlot must be 32 bytes");

}  // namespace txf
