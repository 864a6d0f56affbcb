// vocab.cuh -- term identity, table layouts and string-tile staging shared by the vocabulary
// analyzer (vocab.cu) and the fused transform's lookup (transform.cu).
#pragma once
#include "common.cuh"
#include "farmhash.cuh"

namespace txf {

// ---------------------------------------------------------------------------------------------
// Term identity.  Counting and looking up terms does not need FarmHash (only the OOV bucket of a
// term that is NOT in the vocabulary does, reference call site taxi_utils.py:123-126), so the
// tables are keyed by a cheaper 128-bit identity (a, b):
//   * terms of <= 15 bytes: the bytes themselves (zero padded) with the length in b's top byte --
//     exact, no hashing at all (payment types, most categorical strings);
//   * longer terms: two independent 64-bit multiply-rotate hashes of all bytes and the length
//     (126 effective bits; the top byte of b marks the class so the two classes never meet).
// a == 0 and b == 0 are reserved for "unclaimed slot": b is never 0 by construction, and a == 0
// is remapped to 1 with a flag bit in b.
// ---------------------------------------------------------------------------------------------
struct StrKey {
  unsigned long long a, b;
};

__host__ __device__ __forceinline__ uint64_t rotl64(uint64_t v, int s) { return (v << s) | (v >> (64 - s)); }

constexpr uint32_t kExactKeyMax = 15;  // terms up to this length are their own identity

__host__ __device__ __forceinline__ StrKey finish_key(uint64_t k0, uint64_t k1) {
  if (k0 == 0) {  // a == 0 marks an unclaimed slot
    k0 = 1;
    k1 |= 0x4000000000000000ULL;
  }
  return StrKey{k0, k1};
}

// Mixing of a long term.  Like FarmHash, terms of 16..32 bytes are covered by four (overlapping)
// 8-byte windows -- first 16 and last 16 bytes -- so there is no loop, no tail and no masking;
// 33..64 bytes take eight windows, longer terms a loop over 32-byte blocks plus the last 32 bytes.
struct KeyMixer {
  uint64_t h1, h2;
  __host__ __device__ __forceinline__ explicit KeyMixer(uint32_t n)
      : h1(static_cast<uint64_t>(n) * 0x9E3779B97F4A7C15ULL + 0x7F4A7C159E3779B9ULL),
        h2((static_cast<uint64_t>(n) ^ 0xD6E8FEB86659FD93ULL) * 0xC2B2AE3D27D4EB4FULL) {}
  __host__ __device__ __forceinline__ void add(uint64_t w) {
    h1 = (rotl64(h1, 27) ^ w) * 0xFF51AFD7ED558CCDULL;
    h2 = (rotl64(h2, 31) + w) * 0xC4CEB9FE1A85EC53ULL;
  }
  __host__ __device__ __forceinline__ StrKey finish() {
    h1 ^= h1 >> 32; h1 *= 0x9FB21C651E98DF25ULL; h1 ^= h1 >> 29;
    h2 ^= h2 >> 31; h2 *= 0xD6E8FEB86659FD93ULL; h2 ^= h2 >> 32;
    return finish_key(h1, (h2 & 0x3FFFFFFFFFFFFFFFULL) | 0x8000000000000000ULL);  // top byte 0x80..0xBF
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
  } else {
    uint32_t i = 0;
    for (; i + 32 <= n - 32; i += 32) { m.add(fetch(i)); m.add(fetch(i + 8)); m.add(fetch(i + 16)); m.add(fetch(i + 24)); }
    if (n <= 64 || i < n - 32) { m.add(fetch(i)); m.add(fetch(i + 8)); m.add(fetch(i + 16)); m.add(fetch(i + 24)); }
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
    m.add(static_cast<uint64_t>(f[0]) | (static_cast<uint64_t>(f[1]) << 32));
    m.add(static_cast<uint64_t>(f[2]) | (static_cast<uint64_t>(f[3]) << 32));
    m.add(static_cast<uint64_t>(l[0]) | (static_cast<uint64_t>(l[1]) << 32));
    m.add(static_cast<uint64_t>(l[2]) | (static_cast<uint64_t>(l[3]) << 32));
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

// Lookup table slot (built once per vocabulary, probed per row): 24 bytes.
struct LookupSlot {
  unsigned long long a, b;  // b == 0: empty
  int32_t id;
  int32_t pad;
};
static_assert(sizeof(LookupSlot) == 24, "LookupSlot must be 24 bytes");

}  // namespace txf
