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

__host__ __device__ inline StrKey make_key(const uint8_t* s, uint32_t n) {
  uint64_t k0, k1;
  if (n <= 15) {
    k0 = 0;
    k1 = 0;
    if (n) {
      k0 = fetch64(s);
      if (n < 8) k0 &= (1ULL << (8 * n)) - 1;
      if (n > 8) k1 = fetch64(s + 8) & ((1ULL << (8 * (n - 8))) - 1);
    }
    k1 |= static_cast<uint64_t>(n + 1) << 56;  // top byte 1..16
  } else {
    uint64_t h1 = static_cast<uint64_t>(n) * 0x9E3779B97F4A7C15ULL + 0x7F4A7C159E3779B9ULL;
    uint64_t h2 = (static_cast<uint64_t>(n) ^ 0xD6E8FEB86659FD93ULL) * 0xC2B2AE3D27D4EB4FULL;
    uint32_t i = 0;
    for (; i + 8 <= n; i += 8) {
      uint64_t w = fetch64(s + i);
      h1 = (rotl64(h1, 27) ^ w) * 0xFF51AFD7ED558CCDULL;
      h2 = (h2 + w) * 0xC4CEB9FE1A85EC53ULL;
      h2 ^= h2 >> 29;
    }
    if (i < n) {  // the last (overlapping) 8 bytes cover the tail
      uint64_t w = fetch64(s + n - 8);
      h1 = (rotl64(h1, 27) ^ w) * 0xFF51AFD7ED558CCDULL;
      h2 = (h2 + w) * 0xC4CEB9FE1A85EC53ULL;
      h2 ^= h2 >> 29;
    }
    h1 ^= h1 >> 32; h1 *= 0x9FB21C651E98DF25ULL; h1 ^= h1 >> 29;
    h2 ^= h2 >> 31; h2 *= 0xD6E8FEB86659FD93ULL; h2 ^= h2 >> 32;
    k0 = h1;
    k1 = (h2 & 0x3FFFFFFFFFFFFFFFULL) | 0x8000000000000000ULL;  // top byte 0x80..0xBF
  }
  if (k0 == 0) {
    k0 = 1;
    k1 |= 0x4000000000000000ULL;
  }
  return StrKey{k0, k1};
}

__host__ __device__ __forceinline__ uint64_t home_slot(uint64_t a, uint64_t b, uint64_t mask) {
  uint64_t h = (a ^ rotl64(b, 29)) * 0x9E3779B97F4A7C15ULL;
  return (h >> 24) & mask;
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
