// vocab.cuh -- table layouts shared by the vocabulary analyzer (vocab.cu) and the fused
// transform's lookup (transform.cu).
#pragma once
#include "common.cuh"
#include "farmhash.cuh"

namespace txf {

// Count table slot: 32 bytes = one DRAM sector.  fp == 0 / tag == 0 mean "unclaimed"; real
// zeros are remapped to 1 (table identity only -- OOV hashing uses the true fingerprint).
struct CountSlot {
  unsigned long long fp, tag, count, strref;  // strref = (arena offset << 24) | length
};
static_assert(sizeof(CountSlot) == 32, "CountSlot must be one 32-byte sector");
constexpr uint32_t kMaxTermLen = (1u << 24) - 1;

// Lookup table slot (built once per vocabulary, probed per row): 16 bytes.
struct LookupSlot {
  unsigned long long fp;  // 0 = empty
  uint32_t tag32;
  int32_t id;
};
static_assert(sizeof(LookupSlot) == 16, "LookupSlot must be 16 bytes");

__host__ __device__ __forceinline__ void table_identity(uint64_t fp, uint64_t tag, unsigned long long& kfp,
                                                        unsigned long long& ktag) {
  kfp = fp ? fp : 1ULL;
  ktag = tag ? tag : 1ULL;
}
__host__ __device__ __forceinline__ uint64_t home_slot(uint64_t kfp, uint64_t mask) {
  return (kfp ^ (kfp >> 29)) & mask;
}

}  // namespace txf
