// example.cu -- tf.Example / TFRecord encode of a table on the GPU (egress half).
// The reference materialises transformed rows as tf.Example protos in TFRecord files
// (notebooks/03a_TensorFlow_Transform_Advanced.ipynb:445-455 ExampleProtoCoder + WriteToTFRecord;
// read back at airflow-dags/taxi_utils.py:79-83,260-281).  Here every row of a device table becomes
// one framed record without leaving the GPU:
//
//   record  = u64 length | u32 masked_crc32c(length) | Example bytes | u32 masked_crc32c(Example bytes)
//   Example = 0x0A len Features;  Features = repeated (0x0A len MapEntry), entries in feature-name order
//   MapEntry = 0x0A len name | 0x12 len Feature
//   Feature  = 0x0A len BytesList{0x0A len bytes} | 0x12 len FloatList{0x0A 4 <f32 LE>} | 0x1A len Int64List{0x0A len varint}
//   a missing cell is the kind's list with no value (an empty list, like the reference's typed CSV loop)
//
// which is byte for byte what protobuf's deterministic serialisation of the same Example gives
// (tests/test_gpu_egress.py checks that against the protobuf runtime).  Two kernels, one thread per
// row: sizes (the caller prefix-sums them), then the bytes -- written into shared memory, 128 records
// per CTA, and copied out with coalesced 16-byte stores.
#include "common.cuh"

namespace txf {

struct ExFeat {
  const void* values;     // float* / int64* / int32 offsets*
  const uint8_t* pool;    // strings
  const uint8_t* valid;
  int32_t kind;           // TXF_F32 / TXF_I64 / TXF_EX_BYTES
  int32_t name_off, name_len;
};
struct ExFeats {
  ExFeat f[TXF_EX_MAX_FEATURES];
  int32_t n;
};

__device__ __forceinline__ uint32_t varint_len(unsigned long long v) {
  uint32_t n = 1;
  while (v >= 0x80ULL) { v >>= 7; ++n; }
  return n;
}

// length of the Feature message of one cell
__device__ __forceinline__ uint32_t feature_len(const ExFeat& f, int64_t row, uint32_t& inner) {
  const bool has = row_valid(f.valid, row);
  if (f.kind == TXF_F32) {
    inner = has ? 6u : 0u;                        // 0x0A 0x04 + 4 bytes
  } else if (f.kind == TXF_I64) {
    inner = has ? 2u + varint_len(static_cast<unsigned long long>(static_cast<const long long*>(f.values)[row])) : 0u;
  } else {
    const int32_t* off = static_cast<const int32_t*>(f.values);
    const uint32_t sl = static_cast<uint32_t>(off[row + 1] - off[row]);
    inner = has ? 1u + varint_len(sl) + sl : 0u;  // 0x0A len bytes
  }
  return 1u + varint_len(inner) + inner;          // list tag, list length, list payload
}

__device__ __forceinline__ uint32_t example_len(const ExFeats& fs, int64_t row, uint32_t& features_len) {
  uint32_t total = 0;
  for (int i = 0; i < fs.n; ++i) {
    uint32_t inner;
    const uint32_t fl = feature_len(fs.f[i], row, inner);
    const uint32_t entry = 1u + varint_len(fs.f[i].name_len) + fs.f[i].name_len + 1u + varint_len(fl) + fl;
    total += 1u + varint_len(entry) + entry;
  }
  features_len = total;
  return 1u + varint_len(total) + total;
}

__global__ void example_sizes_kernel(ExFeats fs, int64_t n, int64_t* __restrict__ sizes) {
  const int64_t row = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (row >= n) return;
  uint32_t fl;
  sizes[row] = 16 + static_cast<int64_t>(example_len(fs, row, fl));
}

// CRC-32C (Castagnoli), reflected, one byte per step from a table in shared memory; the running
// value is updated as each byte is written, so nothing is read back.
struct Emitter {
  uint8_t* p;
  uint32_t crc;
  const uint32_t* tab;
  __device__ __forceinline__ void put(uint32_t b) {
    *p++ = static_cast<uint8_t>(b);
    crc = tab[(crc ^ b) & 0xFFu] ^ (crc >> 8);
  }
  __device__ __forceinline__ void varint(unsigned long long v) {
    while (v >= 0x80ULL) { put((static_cast<uint32_t>(v) & 0x7Fu) | 0x80u); v >>= 7; }
    put(static_cast<uint32_t>(v));
  }
};
__device__ __forceinline__ uint32_t mask_crc(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }

constexpr int kExRows = 128;               // rows (= threads) per CTA
constexpr int kExStageBytes = 56 * 1024;   // records of one CTA staged in shared memory, then copied out coalesced

__global__ void __launch_bounds__(kExRows) example_encode_kernel(ExFeats fs, const uint8_t* __restrict__ names, int64_t n,
                                                                 const int64_t* __restrict__ offsets, uint8_t* __restrict__ out) {
  extern __shared__ __align__(16) uint8_t s_stage[];
  __shared__ uint32_t tab[256];
  for (int t = threadIdx.x; t < 256; t += kExRows) {
    uint32_t c = t;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ 0x82F63B78u : (c >> 1);
    tab[t] = c;
  }
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * kExRows;
  const int64_t r1 = r0 + kExRows < n ? r0 + kExRows : n;
  const int64_t g0 = offsets[r0], g1 = offsets[r1];
  const uint32_t skew = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(out + g0) & 15u);  // same 16-byte phase in both memories
  const bool staged = g1 - g0 + skew <= kExStageBytes;
  __syncthreads();
  const int64_t row = r0 + threadIdx.x;
  if (row < r1) {
    uint8_t* rec = staged ? s_stage + skew + (offsets[row] - g0) : out + offsets[row];
    uint32_t features_len;
    const uint32_t elen = example_len(fs, row, features_len);
    Emitter e{rec, 0xFFFFFFFFu, tab};
    for (int i = 0; i < 8; ++i) e.put(i < 4 ? (elen >> (8 * i)) & 0xFFu : 0u);  // u64 length (records are < 4 GiB)
    const uint32_t hcrc = mask_crc(~e.crc);
    for (int i = 0; i < 4; ++i) e.put((hcrc >> (8 * i)) & 0xFFu);
    e.crc = 0xFFFFFFFFu;
    e.put(0x0A);
    e.varint(features_len);
    for (int i = 0; i < fs.n; ++i) {
      const ExFeat& f = fs.f[i];
      uint32_t inner;
      const uint32_t fl = feature_len(f, row, inner);
      const uint32_t entry = 1u + varint_len(f.name_len) + f.name_len + 1u + varint_len(fl) + fl;
      e.put(0x0A);
      e.varint(entry);
      e.put(0x0A);
      e.varint(f.name_len);
      for (int k = 0; k < f.name_len; ++k) e.put(names[f.name_off + k]);
      e.put(0x12);
      e.varint(fl);
      const bool has = row_valid(f.valid, row);
      if (f.kind == TXF_F32) {
        e.put(0x12);
        e.varint(inner);
        if (has) {
          const uint32_t u = __float_as_uint(static_cast<const float*>(f.values)[row]);
          e.put(0x0A); e.put(4);
          for (int k = 0; k < 4; ++k) e.put((u >> (8 * k)) & 0xFFu);
        }
      } else if (f.kind == TXF_I64) {
        e.put(0x1A);
        e.varint(inner);
        if (has) {
          const unsigned long long v = static_cast<unsigned long long>(static_cast<const long long*>(f.values)[row]);
          e.put(0x0A);
          e.varint(varint_len(v));
          e.varint(v);
        }
      } else {
        e.put(0x0A);
        e.varint(inner);
        if (has) {
          const int32_t* off = static_cast<const int32_t*>(f.values);
          const uint32_t sl = static_cast<uint32_t>(off[row + 1] - off[row]);
          e.put(0x0A);
          e.varint(sl);
          const uint8_t* src = f.pool + off[row];
          for (uint32_t k = 0; k < sl; ++k) e.put(src[k]);
        }
      }
    }
    const uint32_t dcrc = mask_crc(~e.crc);
    for (int i = 0; i < 4; ++i) *e.p++ = static_cast<uint8_t>(dcrc >> (8 * i));
  }
  if (!staged) return;
  __syncthreads();
  // coalesced copy-out: ragged head up to the first 16-byte boundary, 16-byte body, ragged tail
  uint8_t* dst = out + g0 - skew;  // 16-byte aligned; byte j of the stage goes to dst[j]
  const int64_t lo = skew, hi = skew + (g1 - g0);
  const int64_t body0 = (lo + 15) & ~15LL, body1 = hi & ~15LL;
  if (body0 >= body1) {
    for (int64_t j = lo + threadIdx.x; j < hi; j += kExRows) dst[j] = s_stage[j];
    return;
  }
  for (int64_t j = lo + threadIdx.x; j < body0; j += kExRows) dst[j] = s_stage[j];
  for (int64_t j = body0 + threadIdx.x * 16LL; j < body1; j += kExRows * 16LL)
    __stcs(reinterpret_cast<uint4*>(dst + j), *reinterpret_cast<const uint4*>(s_stage + j));
  for (int64_t j = body1 + threadIdx.x; j < hi; j += kExRows) dst[j] = s_stage[j];
}

static int fill_feats(const txf_example_feature* feats, int nfeat, ExFeats* d) {
  TXF_REQUIRE(feats != nullptr && nfeat >= 1 && nfeat <= TXF_EX_MAX_FEATURES, "example: nfeat=%d out of [1,%d]", nfeat, TXF_EX_MAX_FEATURES);
  d->n = nfeat;
  for (int i = 0; i < nfeat; ++i) {
    const txf_example_feature& s = feats[i];
    TXF_REQUIRE(s.kind == TXF_F32 || s.kind == TXF_I64 || s.kind == TXF_EX_BYTES, "example: feature %d has bad kind %d", i, s.kind);
    TXF_REQUIRE(s.values != nullptr, "example: feature %d has no values", i);
    TXF_REQUIRE(s.kind != TXF_EX_BYTES || s.pool != nullptr, "example: string feature %d has no pool", i);
    TXF_REQUIRE(s.name_len >= 0 && s.name_off >= 0, "example: feature %d has a bad name range", i);
    d->f[i] = ExFeat{s.values, static_cast<const uint8_t*>(s.pool), static_cast<const uint8_t*>(s.valid), s.kind, s.name_off, s.name_len};
  }
  return TXF_OK;
}

}  // namespace txf

extern "C" {

int txf_example_sizes(const txf_example_feature* feats, int nfeat, int64_t n, int64_t* sizes_dev, txf_stream_t stream) {
  using namespace txf;
  ExFeats d{};
  int rc = fill_feats(feats, nfeat, &d);
  if (rc) return rc;
  TXF_REQUIRE(sizes_dev != nullptr && n >= 0 && n < (1LL << 31), "example_sizes: bad arguments");
  if (n == 0) return TXF_OK;
  example_sizes_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d, n, sizes_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_example_encode(const txf_example_feature* feats, int nfeat, const uint8_t* names_dev, int64_t n,
                       const int64_t* offsets_dev, uint8_t* out_dev, txf_stream_t stream) {
  using namespace txf;
  ExFeats d{};
  int rc = fill_feats(feats, nfeat, &d);
  if (rc) return rc;
  TXF_REQUIRE(names_dev && offsets_dev && out_dev && n >= 0 && n < (1LL << 31), "example_encode: bad arguments");
  if (n == 0) return TXF_OK;
  TXF_CUDA(cudaFuncSetAttribute(example_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kExStageBytes));
  example_encode_kernel<<<static_cast<unsigned>((n + kExRows - 1) / kExRows), kExRows, kExStageBytes, static_cast<cudaStream_t>(stream)>>>(
      d, names_dev, n, offsets_dev, out_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
