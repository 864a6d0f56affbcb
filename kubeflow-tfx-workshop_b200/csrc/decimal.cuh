// decimal.cuh -- exact decimal -> double conversion used by the CSV ingest (csv.cu); __host__
// __device__ so that tests/test_host.py can check it against strtod on the host.
#pragma once
#include <math.h>
#include <stdint.h>

namespace txf {

#ifdef __CUDA_ARCH__
__device__ __forceinline__ int clz64(unsigned long long x) { return __clzll(x); }
__device__ __forceinline__ unsigned long long mulhi64(unsigned long long a, unsigned long long b) { return __umul64hi(a, b); }
#else
inline int clz64(unsigned long long x) { return x ? __builtin_clzll(x) : 64; }
inline unsigned long long mulhi64(unsigned long long a, unsigned long long b) {
  return static_cast<unsigned long long>((static_cast<unsigned __int128>(a) * b) >> 64);
}
#endif

// ---- exact decimal -> double for literals the one-operation path cannot take ------------------
// value = m * 10^e with m < 2^64 and |e| <= kBigPow10Max, evaluated in 320-bit integers: a product
// (e >= 0) or a 64-bit quotient with a sticky remainder (e < 0), then one round-half-even to 53 bits.
constexpr int kBigWords = 5;
constexpr int kBigPow10Max = 77;  // 10^77 < 2^256: the scaled numerator (divisor bits + 63) and m * 10^77 both fit 320 bits
struct UBig {
  unsigned long long w[kBigWords];  // little endian
};
__host__ __device__ inline int ubig_bitlen(const UBig& a) {
  for (int i = kBigWords - 1; i >= 0; --i)
    if (a.w[i]) return 64 * (i + 1) - clz64(a.w[i]);
  return 0;
}
__host__ __device__ inline void ubig_mul_small(UBig& a, unsigned long long k) {
  unsigned long long carry = 0;
  for (int i = 0; i < kBigWords; ++i) {
    const unsigned long long lo = a.w[i] * k, hi = mulhi64(a.w[i], k);
    a.w[i] = lo + carry;
    carry = hi + (a.w[i] < lo ? 1ULL : 0ULL);
  }
}
__host__ __device__ inline bool ubig_ge(const UBig& a, const UBig& b) {
  for (int i = kBigWords - 1; i > 0; --i)
    if (a.w[i] != b.w[i]) return a.w[i] > b.w[i];
  return a.w[0] >= b.w[0];
}
__host__ __device__ inline void ubig_sub(UBig& a, const UBig& b) {  // a >= b
  unsigned long long borrow = 0;
  for (int i = 0; i < kBigWords; ++i) {
    const unsigned long long t = a.w[i] - b.w[i];
    const unsigned long long nb = (a.w[i] < b.w[i] || t < borrow) ? 1ULL : 0ULL;
    a.w[i] = t - borrow;
    borrow = nb;
  }
}
__host__ __device__ inline void ubig_shl1(UBig& a, unsigned long long bit) {
  for (int i = kBigWords - 1; i > 0; --i) a.w[i] = (a.w[i] << 1) | (a.w[i - 1] >> 63);
  a.w[0] = (a.w[0] << 1) | bit;
}
// bits [pos, pos + 64) of a (bits beyond the top are zero)
__host__ __device__ inline unsigned long long ubig_window(const UBig& a, int pos) {
  unsigned long long r = 0;
  for (int i = 0; i < kBigWords; ++i) {
    const int sh = pos - 64 * i;  // word i holds bits [64 i, 64 i + 64)
    if (sh > -64 && sh < 64) r |= sh >= 0 ? (a.w[i] >> sh) : (a.w[i] << (-sh));
  }
  return r;
}
// q has its top bit set; value = (q + fraction) * 2^e2 with fraction in [0, 1), sticky = fraction != 0
__host__ __device__ inline double round_to_double(unsigned long long q, bool sticky, int e2) {
  unsigned long long mant = q >> 11;
  const unsigned long long rest = q & 0x7FFULL;
  if (rest > 0x400ULL || (rest == 0x400ULL && (sticky || (mant & 1ULL)))) ++mant;  // half to even
  return ldexp(static_cast<double>(mant), e2 + 11);  // mant <= 2^53: exact
}
__host__ __device__ inline double decimal_to_double_exact(unsigned long long m, int e10) {  // m != 0, |e10| <= kBigPow10Max
  UBig p{};
  p.w[0] = 1ULL;
  for (int i = 0, k = e10 < 0 ? -e10 : e10; i < k; ++i) ubig_mul_small(p, 10ULL);
  if (e10 >= 0) {
    ubig_mul_small(p, m);
    const int bl = ubig_bitlen(p);
    if (bl <= 64) return static_cast<double>(p.w[0]);  // one correctly rounded conversion
    const int drop = bl - 64;
    bool sticky = false;
    for (int i = 0; i < kBigWords; ++i) {  // any set bit below `drop`
      const int lo = 64 * i;
      if (drop >= lo + 64) sticky = sticky || p.w[i] != 0;
      else if (drop > lo) sticky = sticky || (p.w[i] & ((1ULL << (drop - lo)) - 1ULL)) != 0;
    }
    return round_to_double(ubig_window(p, drop), sticky, drop);
  }
  // m / 10^k: scale the numerator so the quotient has 63 or 64 bits
  const int ld = ubig_bitlen(p), ln = 64 - clz64(m);
  const int s = ld - ln + 63;  // numerator = m << s has ld + 63 bits
  UBig num{};
  {
    const int ws = s >> 6, bs = s & 63;
    num.w[ws] = m << bs;
    if (bs && ws + 1 < kBigWords) num.w[ws + 1] = m >> (64 - bs);
  }
  UBig r{};  // numerator >> 64: ld - 1 bits, below the divisor
  for (int i = 0; i + 1 < kBigWords; ++i) r.w[i] = num.w[i + 1];
  unsigned long long q = 0;
  for (int i = 63; i >= 0; --i) {
    ubig_shl1(r, (num.w[0] >> i) & 1ULL);
    q <<= 1;
    if (ubig_ge(r, p)) { ubig_sub(r, p); q |= 1ULL; }
  }
  int e2 = -s;
  if (!(q >> 63)) {  // 63-bit quotient: one more division step
    ubig_shl1(r, 0ULL);
    q <<= 1;
    if (ubig_ge(r, p)) { ubig_sub(r, p); q |= 1ULL; }
    --e2;
  }
  bool sticky = false;
  for (int i = 0; i < kBigWords; ++i) sticky = sticky || r.w[i] != 0;
  return round_to_double(q, sticky, e2);
}

}  // namespace txf
