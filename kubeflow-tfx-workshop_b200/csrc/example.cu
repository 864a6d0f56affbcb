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
// row: sizes (the caller prefix-sums them), then the bytes.
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
__device__ __forceinline__ uint8_t* put_varint(uint8_t* p, unsigned long long v) {
  while (v >= 0x80ULL) { *p++ = static_cast<uint8_t>(v) | 0x80u; v >>= 7; }
  *p++ = static_cast<uint8_t>(v);
  return p;
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

// CRC-32C (Castagnoli), reflected, one byte per step from a table in shared memory
__device__ __forceinline__ uint32_t crc32c_bytes(const uint32_t* tab, const uint8_t* p, uint32_t n) {
  uint32_t c = 0xFFFFFFFFu;
  for (uint32_t i = 0; i < n; ++i) c = tab[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
  return ~c;
}
__device__ __forceinline__ uint32_t mask_crc(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }

__global__ void __launch_bounds__(256) example_encode_kernel(ExFeats fs, const uint8_t* __restrict__ names, int64_t n,
                                                             const int64_t* __restrict__ offsets, uint8_t* __restrict__ out) {
  __shared__ uint32_t tab[256];
  {
    uint32_t c = threadIdx.x;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ 0x82F63B78u : (c >> 1);
    tab[threadIdx.x] = c;
  }
  __syncthreads();
  const int64_t row = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (row >= n) return;
  uint8_t* rec = out + offsets[row];
  uint32_t features_len;
  const uint32_t elen = example_len(fs, row, features_len);
  // frame header
  const unsigned long long len64 = elen;
  for (int i = 0; i < 8; ++i) rec[i] = static_cast<uint8_t>(len64 >> (8 * i));
  const uint32_t hcrc = mask_crc(crc32c_bytes(tab, rec, 8));
  for (int i = 0; i < 4; ++i) rec[8 + i] = static_cast<uint8_t>(hcrc >> (8 * i));
  uint8_t* data = rec + 12;
  uint8_t* p = data;
  *p++ = 0x0A;
  p = put_varint(p, features_len);
  for (int i = 0; i < fs.n; ++i) {
    const ExFeat& f = fs.f[i];
    uint32_t inner;
    const uint32_t fl = feature_len(f, row, inner);
    const uint32_t entry = 1u + varint_len(f.name_len) + f.name_len + 1u + varint_len(fl) + fl;
    *p++ = 0x0A;
    p = put_varint(p, entry);
    *p++ = 0x0A;
    p = put_varint(p, f.name_len);
    for (int k = 0; k < f.name_len; ++k) *p++ = names[f.name_off + k];
    *p++ = 0x12;
    p = put_varint(p, fl);
    const bool has = row_valid(f.valid, row);
    if (f.kind == TXF_F32) {
      *p++ = 0x12;
      p = put_varint(p, inner);
      if (has) {
        const uint32_t u = __float_as_uint(static_cast<const float*>(f.values)[row]);
        *p++ = 0x0A; *p++ = 4;
        for (int k = 0; k < 4; ++k) *p++ = static_cast<uint8_t>(u >> (8 * k));
      }
    } else if (f.kind == TXF_I64) {
      *p++ = 0x1A;
      p = put_varint(p, inner);
      if (has) {
        const unsigned long long v = static_cast<unsigned long long>(static_cast<const long long*>(f.values)[row]);
        *p++ = 0x0A;
        p = put_varint(p, varint_len(v));
        p = put_varint(p, v);
      }
    } else {
      *p++ = 0x0A;
      p = put_varint(p, inner);
      if (has) {
        const int32_t* off = static_cast<const int32_t*>(f.values);
        const uint32_t sl = static_cast<uint32_t>(off[row + 1] - off[row]);
        *p++ = 0x0A;
        p = put_varint(p, sl);
        const uint8_t* src = f.pool + off[row];
        for (uint32_t k = 0; k < sl; ++k) *p++ = src[k];
      }
    }
  }
  const uint32_t dcrc = mask_crc(crc32c_bytes(tab, data, elen));
  for (int i = 0; i < 4; ++i) p[i] = static_cast<uint8_t>(dcrc >> (8 * i));
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
  example_encode_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d, names_dev, n, offsets_dev,
                                                                                                            out_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
