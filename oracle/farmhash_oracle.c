/*
 * oracle/farmhash_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C) of FarmHash Fingerprint64 (= farmhashna::Hash64), the hash
 * TensorFlow's StringToHashBucketFast applies to out-of-vocabulary terms inside
 * tft.compute_and_apply_vocabulary (reference call site:
 * airflow-dags/taxi_utils.py:123-126, num_oov_buckets=_OOV_SIZE at :54).
 *
 * The algorithm lives in a third-party dependency that is NOT vendored under
 * /root/reference (TensorFlow 1.13 -> farmhash, pinned via tensorflow-transform==0.13.0 at
 * notebooks/03_TensorFlow_Transform.ipynb:47). This file restates the published algorithm
 * (SURVEY.md Appendix A.3). Pinning:
 *   - len 0..16: TensorFlow's own string_to_hash_bucket_fast test expectations
 *     (tests/golden/farmhash_vectors.json, "tf_kat");
 *   - len 0..32: abseil's compiled CityHash64 (CityHash v1.1 shares the <=32-byte code paths
 *     with farmhashna; its 33..64 path is a different function) exported by pyarrow's
 *     libarrow_compute, see tests/golden/make_farmhash_golden.py;
 *   - len > 32: PARITY UNPINNED (no independent implementation is available offline; the
 *     only cross-check is SURVEY.md Appendix A.3's separately written restatement).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library.
 */
#include <stdint.h>
#include <string.h>
#include <stddef.h>

#define K0 0xc3a5c85c97cb3127ULL
#define K1 0xb492b66fbe98f273ULL
#define K2 0x9ae16a3b2f90404fULL

static inline uint64_t ld64(const uint8_t *p) { uint64_t v; memcpy(&v, p, 8); return v; }
static inline uint64_t ld32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }
static inline uint64_t ror(uint64_t v, int s) { return s == 0 ? v : (v >> s) | (v << (64 - s)); }
static inline uint64_t smix(uint64_t v) { return v ^ (v >> 47); }

static inline uint64_t h16(uint64_t u, uint64_t v, uint64_t m) {
  uint64_t a = (u ^ v) * m;
  a ^= a >> 47;
  uint64_t b = (v ^ a) * m;
  b ^= b >> 47;
  return b * m;
}

static uint64_t len0to16(const uint8_t *s, size_t n) {
  if (n >= 8) {
    uint64_t m = K2 + n * 2;
    uint64_t a = ld64(s) + K2;
    uint64_t b = ld64(s + n - 8);
    uint64_t c = ror(b, 37) * m + a;
    uint64_t d = (ror(a, 25) + b) * m;
    return h16(c, d, m);
  }
  if (n >= 4) {
    uint64_t m = K2 + n * 2;
    uint64_t a = ld32(s);
    return h16(n + (a << 3), ld32(s + n - 4), m);
  }
  if (n > 0) {
    uint8_t a = s[0], b = s[n >> 1], c = s[n - 1];
    uint32_t y = (uint32_t)a + ((uint32_t)b << 8);
    uint32_t z = (uint32_t)n + ((uint32_t)c << 2);
    return smix(y * K2 ^ z * K0) * K2;
  }
  return K2;
}

static uint64_t len17to32(const uint8_t *s, size_t n) {
  uint64_t m = K2 + n * 2;
  uint64_t a = ld64(s) * K1;
  uint64_t b = ld64(s + 8);
  uint64_t c = ld64(s + n - 8) * m;
  uint64_t d = ld64(s + n - 16) * K2;
  return h16(ror(a + b, 43) + ror(c, 30) + d, a + ror(b + K2, 18) + c, m);
}

static uint64_t len33to64(const uint8_t *s, size_t n) {
  uint64_t m = K2 + n * 2;
  uint64_t a = ld64(s) * K2;
  uint64_t b = ld64(s + 8);
  uint64_t c = ld64(s + n - 8) * m;
  uint64_t d = ld64(s + n - 16) * K2;
  uint64_t y = ror(a + b, 43) + ror(c, 30) + d;
  uint64_t z = h16(y, a + ror(b + K2, 18) + c, m);
  uint64_t e = ld64(s + 16) * m;
  uint64_t f = ld64(s + 24);
  uint64_t g = (y + ld64(s + n - 32)) * m;
  uint64_t h = (z + ld64(s + n - 24)) * m;
  return h16(ror(e + f, 43) + ror(g, 30) + h, e + ror(f + a, 18) + g, m);
}

static inline void weak32(const uint8_t *p, uint64_t a, uint64_t b, uint64_t *o0, uint64_t *o1) {
  uint64_t w = ld64(p), x = ld64(p + 8), y = ld64(p + 16), z = ld64(p + 24);
  a += w;
  b = ror(b + a + z, 21);
  uint64_t c = a;
  a += x;
  a += y;
  b += ror(a, 44);
  *o0 = a + z;
  *o1 = b + c;
}

uint64_t oracle_fingerprint64(const uint8_t *s, size_t n) {
  if (n <= 16) return len0to16(s, n);
  if (n <= 32) return len17to32(s, n);
  if (n <= 64) return len33to64(s, n);
  const uint64_t seed = 81;
  uint64_t x = seed, y = seed * K1 + 113, z = smix(y * K2 + 113) * K2;
  uint64_t v0 = 0, v1 = 0, w0 = 0, w1 = 0, t;
  x = x * K2 + ld64(s);
  const uint8_t *end = s + ((n - 1) / 64) * 64;
  const uint8_t *last64 = end + ((n - 1) & 63) - 63;
  do {
    x = ror(x + y + v0 + ld64(s + 8), 37) * K1;
    y = ror(y + v1 + ld64(s + 48), 42) * K1;
    x ^= w1;
    y += v0 + ld64(s + 40);
    z = ror(z + w0, 33) * K1;
    weak32(s, v1 * K1, x + w0, &v0, &v1);
    weak32(s + 32, z + w1, y + ld64(s + 16), &w0, &w1);
    t = z; z = x; x = t;
    s += 64;
  } while (s != end);
  uint64_t m = K1 + ((z & 0xff) << 1);
  s = last64;
  w0 += ((n - 1) & 63);
  v0 += w0;
  w0 += v0;
  x = ror(x + y + v0 + ld64(s + 8), 37) * m;
  y = ror(y + v1 + ld64(s + 48), 42) * m;
  x ^= w1 * 9;
  y += v0 * 9 + ld64(s + 40);
  z = ror(z + w0, 33) * m;
  weak32(s, v1 * m, x + w0, &v0, &v1);
  weak32(s + 32, z + w1, y + ld64(s + 16), &w0, &w1);
  t = z; z = x; x = t;
  return h16(h16(v0, w0, m) + smix(y) * K0 + z, h16(v1, w1, m) + x, m);
}

/* Bulk form over an Arrow-style string column: offsets[n+1] (int32) into pool. */
void oracle_fingerprint64_bulk(const int32_t *offsets, const uint8_t *pool, int64_t n,
                               uint64_t *out) {
  for (int64_t i = 0; i < n; ++i)
    out[i] = oracle_fingerprint64(pool + offsets[i], (size_t)(offsets[i + 1] - offsets[i]));
}
