// farmhash.cuh -- device FarmHash Fingerprint64 (farmhashna::Hash64).
//
// Replaces TF's StringToHashBucketFast on the OOV branch of
// tft.compute_and_apply_vocabulary (reference call site airflow-dags/taxi_utils.py:123-126).
// Written from the published algorithm.  It is only evaluated for terms
// that are NOT in the vocabulary; table identities use the cheaper key of vocab.cuh.
//
// All fetches are unaligned little-endian 8-byte reads assembled from two ALIGNED 8-byte
// loads, so the byte buffer must be readable up to the next 8-byte boundary past the last
// string byte plus 8 (TXF_POOL_SLACK).  Works on generic pointers (global or shared).
#pragma once
#include <stdint.h>

namespace txf {

constexpr uint64_t kK0 = 0xc3a5c85c97cb3127ULL;
constexpr uint64_t kK1 = 0xb492b66fbe98f273ULL;
constexpr uint64_t kK2 = 0x9ae16a3b2f90404fULL;

__host__ __device__ __forceinline__ uint64_t fetch64(const uint8_t* p) {
  uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const uint64_t* q = reinterpret_cast<const uint64_t*>(a & ~static_cast<uintptr_t>(7));
  uint32_t sh = static_cast<uint32_t>(a & 7) * 8;
  uint64_t lo = q[0], hi = q[1];
  return (lo >> sh) | ((hi << 1) << (63 - sh));  // sh == 0 -> second term is 0
}
__host__ __device__ __forceinline__ uint64_t fetch32(const uint8_t* p) { return fetch64(p) & 0xFFFFFFFFULL; }
__host__ __device__ __forceinline__ uint64_t rotr64(uint64_t v, int s) { return (v >> s) | (v << (64 - s)); }
__host__ __device__ __forceinline__ uint64_t shift_mix(uint64_t v) { return v ^ (v >> 47); }
__host__ __device__ __forceinline__ uint64_t hash_len16(uint64_t u, uint64_t v, uint64_t m) {
  uint64_t a = (u ^ v) * m;
  a ^= a >> 47;
  uint64_t b = (v ^ a) * m;
  b ^= b >> 47;
  return b * m;
}
__host__ __device__ __forceinline__ void weak_hash32(const uint8_t* p, uint64_t a, uint64_t b, uint64_t& o0,
                                                     uint64_t& o1) {
  uint64_t w = fetch64(p), x = fetch64(p + 8), y = fetch64(p + 16), z = fetch64(p + 24);
  a += w;
  b = rotr64(b + a + z, 21);
  uint64_t c = a;
  a += x;
  a += y;
  b += rotr64(a, 44);
  o0 = a + z;
  o1 = b + c;
}

// Fingerprint64 of s[0..n).
__host__ __device__ inline uint64_t fingerprint64(const uint8_t* s, uint32_t n) {
  uint64_t u, v, m;
  if (n <= 16) {
    if (n >= 8) {
      m = kK2 + static_cast<uint64_t>(n) * 2;
      uint64_t a = fetch64(s) + kK2;
      uint64_t b = fetch64(s + n - 8);
      u = rotr64(b, 37) * m + a;
      v = (rotr64(a, 25) + b) * m;
    } else if (n >= 4) {
      m = kK2 + static_cast<uint64_t>(n) * 2;
      uint64_t a = fetch32(s);
      u = n + (a << 3);
      v = fetch32(s + n - 4);
    } else if (n > 0) {
      uint32_t a = s[0], b = s[n >> 1], c = s[n - 1];
      uint32_t y = a + (b << 8);
      uint32_t z = n + (c << 2);
      return shift_mix(static_cast<uint64_t>(y) * kK2 ^ static_cast<uint64_t>(z) * kK0) * kK2;
    } else {
      return kK2;
    }
  } else if (n <= 32) {
    m = kK2 + static_cast<uint64_t>(n) * 2;
    uint64_t a = fetch64(s) * kK1;
    uint64_t b = fetch64(s + 8);
    uint64_t c = fetch64(s + n - 8) * m;
    uint64_t d = fetch64(s + n - 16) * kK2;
    u = rotr64(a + b, 43) + rotr64(c, 30) + d;
    v = a + rotr64(b + kK2, 18) + c;
  } else if (n <= 64) {
    m = kK2 + static_cast<uint64_t>(n) * 2;
    uint64_t a = fetch64(s) * kK2;
    uint64_t b = fetch64(s + 8);
    uint64_t c = fetch64(s + n - 8) * m;
    uint64_t d = fetch64(s + n - 16) * kK2;
    uint64_t y = rotr64(a + b, 43) + rotr64(c, 30) + d;
    uint64_t z = hash_len16(y, a + rotr64(b + kK2, 18) + c, m);
    uint64_t e = fetch64(s + 16) * m;
    uint64_t f = fetch64(s + 24);
    uint64_t g = (y + fetch64(s + n - 32)) * m;
    uint64_t h = (z + fetch64(s + n - 24)) * m;
    u = rotr64(e + f, 43) + rotr64(g, 30) + h;
    v = e + rotr64(f + a, 18) + g;
  } else {
    uint64_t x = 81, y = 81 * kK1 + 113, z = shift_mix(y * kK2 + 113) * kK2;
    uint64_t v0 = 0, v1 = 0, w0 = 0, w1 = 0;
    x = x * kK2 + fetch64(s);
    const uint8_t* end = s + ((n - 1) / 64) * 64;
    const uint8_t* last64 = s + n - 64;
    do {
      x = rotr64(x + y + v0 + fetch64(s + 8), 37) * kK1;
      y = rotr64(y + v1 + fetch64(s + 48), 42) * kK1;
      x ^= w1;
      y += v0 + fetch64(s + 40);
      z = rotr64(z + w0, 33) * kK1;
      weak_hash32(s, v1 * kK1, x + w0, v0, v1);
      weak_hash32(s + 32, z + w1, y + fetch64(s + 16), w0, w1);
      uint64_t t = z; z = x; x = t;
      s += 64;
    } while (s != end);
    m = kK1 + ((z & 0xff) << 1);
    s = last64;
    w0 += (n - 1) & 63;
    v0 += w0;
    w0 += v0;
    x = rotr64(x + y + v0 + fetch64(s + 8), 37) * m;
    y = rotr64(y + v1 + fetch64(s + 48), 42) * m;
    x ^= w1 * 9;
    y += v0 * 9 + fetch64(s + 40);
    z = rotr64(z + w0, 33) * m;
    weak_hash32(s, v1 * m, x + w0, v0, v1);
    weak_hash32(s + 32, z + w1, y + fetch64(s + 16), w0, w1);
    uint64_t t = z; z = x; x = t;
    u = hash_len16(v0, w0, m) + shift_mix(y) * kK0 + z;
    v = hash_len16(v1, w1, m) + x;
  }
  return hash_len16(u, v, m);
}

}  // namespace txf
