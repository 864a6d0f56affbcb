// example_decode.cu -- TFRecord / tf.Example DECODE on the GPU (ingest half).
// A TFX Transform executor is fed TFRecord files of tf.Example protos
// (airflow-dags/taxi_utils.py:79-83 names the gzip TFRecord reader, :260-281 reads them for training;
// airflow-dags/taxi_pipeline.py:70-89 wires CsvExampleGen's Examples into Transform).  Here the
// uncompressed record stream (gunzip stays on the host) is copied to the device once and becomes
// the Arrow-layout columns of columns.py without leaving the GPU:
//
//   1. RECORD INDEX without a sequential walk.  A record is  u64 length | u32 masked_crc32c(length) |
//      data | u32 masked_crc32c(data),  and the next one starts right after it, so a plain reader hops
//      through the file.  Here every byte position is tested as a record start (length field < 2^31 and
//      its CRC matches: a 2^-32 coincidence elsewhere); candidates are written in order; a candidate is a
//      true start iff it is position 0 or some candidate's record ends exactly where it begins (a
//      coincidental candidate inside a record's payload is never pointed at).  The caller checks that
//      the kept starts chain up to the end of the stream.
//   2. one thread per record walks  Example{1: Features{1: map<string, Feature>}}  with the CTA's records
//      staged in shared memory; map keys are matched against the requested feature names; a Feature is
//      bytes_list (1) / float_list (2) / int64_list (3), values packed or not.  No value = a missing
//      cell (validity 0), exactly one value = the cell, more than one = an error (the reference's
//      `_fill_in_missing` requires at most one value per row, taxi_utils.py:92-93).
//   3. string cells are gathered into the column's byte pool once their offsets are known.
#include "common.cuh"

namespace txf {

__device__ __forceinline__ uint32_t crc32c_byte(uint32_t crc, uint32_t b, const uint32_t* tab) { return tab[(crc ^ b) & 0xFFu] ^ (crc >> 8); }
__device__ __forceinline__ uint32_t mask_crc32c(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }
__device__ __forceinline__ void build_crc_table(uint32_t* tab, int nthreads) {
  for (int t = threadIdx.x; t < 256; t += nthreads) {
    uint32_t c = t;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ 0x82F63B78u : (c >> 1);
    tab[t] = c;
  }
}

constexpr int kTfrBlockBytes = 16384;  // byte positions tested by one CTA
constexpr int kTfrThreads = 256;

// is position p a record start?  data must be readable up to p + 12
__device__ __forceinline__ bool tfr_candidate(const uint8_t* __restrict__ data, int64_t p, int64_t nbytes, const uint32_t* tab,
                                              uint32_t* len_out) {
  if (p + 16 > nbytes) return false;  // header + trailer of an empty record
  // records are far below 2^31 bytes: bytes 4..7 of the length field are zero (rejects almost every position cheaply)
  if ((data[p + 4] | data[p + 5] | data[p + 6] | data[p + 7]) != 0) return false;
  uint32_t crc = 0xFFFFFFFFu;
  uint32_t lo = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint32_t b = data[p + i];
    if (i < 4) lo |= b << (8 * i);
    crc = crc32c_byte(crc, b, tab);
  }
  if (lo >= 0x80000000u) return false;
  const uint32_t stored = static_cast<uint32_t>(data[p + 8]) | (static_cast<uint32_t>(data[p + 9]) << 8) |
                          (static_cast<uint32_t>(data[p + 10]) << 16) | (static_cast<uint32_t>(data[p + 11]) << 24);
  if (mask_crc32c(~crc) != stored) return false;
  if (p + 16 + static_cast<int64_t>(lo) > nbytes) return false;
  *len_out = lo;
  return true;
}

// pass 0: candidates per block; pass 1: their positions, in order (block_base = exclusive prefix of the counts)
template <int PASS>
__global__ void __launch_bounds__(kTfrThreads) tfr_candidates_kernel(const uint8_t* __restrict__ data, int64_t nbytes,
                                                                     uint32_t* __restrict__ block_counts,
                                                                     const int64_t* __restrict__ block_base, int64_t* __restrict__ cand) {
  __shared__ uint32_t tab[256];
  __shared__ uint32_t s_warp[kTfrThreads / 32];
  __shared__ uint32_t s_run;
  build_crc_table(tab, kTfrThreads);
  if (threadIdx.x == 0) s_run = 0;
  __syncthreads();
  const int64_t b0 = static_cast<int64_t>(blockIdx.x) * kTfrBlockBytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int it = 0; it < kTfrBlockBytes / kTfrThreads; ++it) {
    const int64_t p = b0 + it * kTfrThreads + threadIdx.x;
    uint32_t len;
    const bool is = p < nbytes && tfr_candidate(data, p, nbytes, tab, &len);
    const unsigned m = __ballot_sync(0xffffffffu, is);
    if (lane == 0) s_warp[warp] = __popc(m);
    __syncthreads();
    uint32_t before = s_run;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    if (PASS == 1 && is) cand[block_base[blockIdx.x] + before + __popc(m & ((1u << lane) - 1u))] = p;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t t = 0;
      for (int w = 0; w < kTfrThreads / 32; ++w) t += s_warp[w];
      s_run += t;
    }
    __syncthreads();
  }
  if (PASS == 0 && threadIdx.x == 0) block_counts[blockIdx.x] = s_run;
}

// mark[j] = 1 iff candidate j is position 0 or the record of some candidate ends where j begins
__global__ void tfr_link_kernel(const uint8_t* __restrict__ data, const int64_t* __restrict__ cand, int64_t ncand, uint8_t* __restrict__ mark) {
  const int64_t j = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (j >= ncand) return;
  const int64_t p = cand[j];
  if (p == 0) mark[j] = 1;
  const uint32_t len = static_cast<uint32_t>(data[p]) | (static_cast<uint32_t>(data[p + 1]) << 8) | (static_cast<uint32_t>(data[p + 2]) << 16) |
                       (static_cast<uint32_t>(data[p + 3]) << 24);
  const int64_t nxt = p + 16 + len;
  int64_t lo = j + 1, hi = ncand;  // candidates are sorted: binary search for nxt
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (cand[mid] < nxt) lo = mid + 1; else hi = mid;
  }
  if (lo < ncand && cand[lo] == nxt) mark[lo] = 1;
}

// ---- Example parse ---------------------------------------------------------------------------
struct DecFeat {
  void* values;        // float* / int64* / uint32* position of the bytes in the stream (strings)
  int32_t* lens;       // strings
  uint32_t* valid;     // bitmap words (row0 % 32 == 0)
  int32_t kind, name_off, name_len;
};
struct DecFeats {
  DecFeat f[TXF_EX_MAX_FEATURES];
  int32_t n;
};
enum { kDecErrTruncated = 1, kDecErrWire = 2, kDecErrKind = 4, kDecErrMulti = 8, kDecErrCrc = 16 };

struct Reader {  // bounds-checked cursor over one record's payload
  const uint8_t* p;
  const uint8_t* end;
  bool ok;
  __device__ __forceinline__ bool more() const { return ok && p < end; }
  __device__ __forceinline__ uint32_t byte() {
    if (p >= end) { ok = false; return 0; }
    return *p++;
  }
  __device__ __forceinline__ unsigned long long varint() {
    unsigned long long v = 0;
    for (int s = 0; s < 70; s += 7) {
      const uint32_t b = byte();
      v |= static_cast<unsigned long long>(b & 0x7Fu) << s;
      if (!(b & 0x80u)) return v;
    }
    ok = false;
    return v;
  }
  __device__ __forceinline__ Reader sub(unsigned long long len) {
    Reader r{p, p + len, ok};
    if (len > static_cast<unsigned long long>(end - p)) { r.end = end; r.ok = false; ok = false; }
    p = r.end;
    return r;
  }
  __device__ __forceinline__ void skip(uint32_t wire) {
    if (wire == 0) varint();
    else if (wire == 1) sub(8);
    else if (wire == 2) sub(varint());
    else if (wire == 5) sub(4);
    else ok = false;
  }
};

constexpr int kDecRows = 128;
constexpr int kDecStageBytes = 64 * 1024;

__global__ void __launch_bounds__(kDecRows) example_decode_kernel(DecFeats fs, const uint8_t* __restrict__ names,
                                                                  const uint8_t* __restrict__ data, const int64_t* __restrict__ starts,
                                                                  int64_t n, int verify_crc, unsigned long long* __restrict__ status) {
  extern __shared__ __align__(16) uint8_t s_stage[];
  __shared__ uint32_t tab[256];
  build_crc_table(tab, kDecRows);
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * kDecRows;
  const int64_t r1 = r0 + kDecRows < n ? r0 + kDecRows : n;
  const int64_t g0 = starts[r0], g1 = starts[r1];
  const int64_t a0 = g0 & ~15LL;  // the stage holds [a0, g1): 16-byte loads from an aligned source
  const bool staged = g1 - a0 <= kDecStageBytes;
  if (staged)
    for (int64_t j = threadIdx.x * 16LL; j < g1 - a0; j += kDecRows * 16LL)
      *reinterpret_cast<uint4*>(s_stage + j) = __ldg(reinterpret_cast<const uint4*>(data + a0 + j));
  __syncthreads();
  const int64_t row = r0 + threadIdx.x;
  const bool live = row < r1;
  uint32_t err = 0;
  uint32_t have = 0, have_hi = 0;  // features seen with exactly one value (bit i), for the validity words
  if (live) {
    const int64_t s = starts[row], e = starts[row + 1];
    const uint8_t* rec = staged ? s_stage + (s - a0) : data + s;
    const uint32_t len = static_cast<uint32_t>(rec[0]) | (static_cast<uint32_t>(rec[1]) << 8) | (static_cast<uint32_t>(rec[2]) << 16) |
                         (static_cast<uint32_t>(rec[3]) << 24);
    if (static_cast<int64_t>(len) + 16 != e - s) err |= kDecErrTruncated;
    const uint8_t* payload = rec + 12;
    if (!err && verify_crc) {
      uint32_t crc = 0xFFFFFFFFu;
      for (uint32_t i = 0; i < len; ++i) crc = crc32c_byte(crc, payload[i], tab);
      const uint8_t* t = payload + len;
      const uint32_t stored = static_cast<uint32_t>(t[0]) | (static_cast<uint32_t>(t[1]) << 8) | (static_cast<uint32_t>(t[2]) << 16) |
                              (static_cast<uint32_t>(t[3]) << 24);
      if (mask_crc32c(~crc) != stored) err |= kDecErrCrc;
    }
    Reader ex{payload, payload + (err & kDecErrTruncated ? 0 : len), true};
    while (ex.more()) {  // Example
      const uint32_t tag = static_cast<uint32_t>(ex.varint());
      if (tag != 0x0A) { ex.skip(tag & 7); continue; }
      Reader feats = ex.sub(ex.varint());
      while (feats.more()) {  // Features: repeated map entries
        const uint32_t t2 = static_cast<uint32_t>(feats.varint());
        if (t2 != 0x0A) { feats.skip(t2 & 7); continue; }
        Reader entry = feats.sub(feats.varint());
        const uint8_t* key = nullptr;
        uint32_t key_len = 0;
        Reader feature{nullptr, nullptr, true};
        bool has_feature = false;
        while (entry.more()) {
          const uint32_t t3 = static_cast<uint32_t>(entry.varint());
          if (t3 == 0x0A) { const unsigned long long l = entry.varint(); key = entry.p; key_len = static_cast<uint32_t>(l); entry.sub(l); }
          else if (t3 == 0x12) { feature = entry.sub(entry.varint()); has_feature = true; }
          else entry.skip(t3 & 7);
        }
        if (!entry.ok) { feats.ok = false; break; }
        int fi = -1;  // which requested feature is this?
        for (int i = 0; i < fs.n && fi < 0; ++i) {
          if (static_cast<uint32_t>(fs.f[i].name_len) != key_len) continue;
          bool same = true;
          for (uint32_t k = 0; k < key_len && same; ++k) same = names[fs.f[i].name_off + k] == key[k];
          if (same) fi = i;
        }
        if (fi < 0 || !has_feature) continue;
        const DecFeat& f = fs.f[fi];
        int nvals = 0;
        while (feature.more()) {  // Feature: oneof kind
          const uint32_t t4 = static_cast<uint32_t>(feature.varint());
          if ((t4 & 7) != 2) { feature.skip(t4 & 7); continue; }
          Reader list = feature.sub(feature.varint());
          const int field = static_cast<int>(t4 >> 3);
          const int want = f.kind == TXF_EX_BYTES ? 1 : f.kind == TXF_F32 ? 2 : 3;
          if (field < 1 || field > 3) continue;
          if (field != want) { if (list.more()) err |= kDecErrKind; continue; }
          while (list.more()) {  // the list's repeated field 1
            const uint32_t t5 = static_cast<uint32_t>(list.varint());
            if ((t5 >> 3) != 1) { list.skip(t5 & 7); continue; }
            if (f.kind == TXF_EX_BYTES) {
              if ((t5 & 7) != 2) { err |= kDecErrWire; break; }
              const unsigned long long l = list.varint();
              const uint8_t* at = list.p;
              list.sub(l);
              if (nvals++ == 0 && list.ok) {
                static_cast<uint32_t*>(f.values)[row] = static_cast<uint32_t>(s + 12 + (at - payload));  // position in the stream
                f.lens[row] = static_cast<int32_t>(l);
              }
            } else if (f.kind == TXF_F32) {
              Reader vals = (t5 & 7) == 2 ? list.sub(list.varint()) : Reader{list.p, list.p + 4, list.ok};
              if ((t5 & 7) == 5) list.sub(4); else if ((t5 & 7) != 2) { err |= kDecErrWire; break; }
              while (vals.more()) {
                uint32_t u = vals.byte(); u |= vals.byte() << 8; u |= vals.byte() << 16; u |= vals.byte() << 24;
                if (nvals++ == 0 && vals.ok) static_cast<float*>(f.values)[row] = __uint_as_float(u);
              }
              if (!vals.ok) list.ok = false;
            } else {
              if ((t5 & 7) == 2) {
                Reader vals = list.sub(list.varint());
                while (vals.more()) {
                  const unsigned long long v = vals.varint();
                  if (nvals++ == 0 && vals.ok) static_cast<long long*>(f.values)[row] = static_cast<long long>(v);
                }
                if (!vals.ok) list.ok = false;
              } else if ((t5 & 7) == 0) {
                const unsigned long long v = list.varint();
                if (nvals++ == 0 && list.ok) static_cast<long long*>(f.values)[row] = static_cast<long long>(v);
              } else { err |= kDecErrWire; break; }
            }
          }
          if (!list.ok) feature.ok = false;
        }
        if (!feature.ok) { feats.ok = false; break; }
        if (nvals > 1) err |= kDecErrMulti;
        if (nvals == 1) { if (fi < 32) have |= 1u << fi; else have_hi |= 1u << (fi - 32); }
      }
      if (!feats.ok) ex.ok = false;
    }
    if (!ex.ok) err |= kDecErrWire;
  }
  // validity words: a warp covers 32 consecutive rows; absent cells get value 0 / length 0
  for (int i = 0; i < fs.n; ++i) {
    const bool present = live && (((i < 32 ? have >> i : have_hi >> (i - 32)) & 1u) != 0);
    const unsigned m = __ballot_sync(0xffffffffu, present);
    if ((threadIdx.x & 31) == 0 && r0 + threadIdx.x < n) fs.f[i].valid[(r0 + threadIdx.x) >> 5] = m;
    if (live && !present) {
      const DecFeat& f = fs.f[i];
      if (f.kind == TXF_F32) static_cast<float*>(f.values)[row] = 0.f;
      else if (f.kind == TXF_I64) static_cast<long long*>(f.values)[row] = 0;
      else { static_cast<uint32_t*>(f.values)[row] = 0; f.lens[row] = 0; }
    }
  }
  if (err) {
    atomicOr(&status[0], static_cast<unsigned long long>(err));
    atomicMin(&status[1], static_cast<unsigned long long>(row));
  }
}

__global__ void gather_bytes_kernel(const uint8_t* __restrict__ data, const uint32_t* __restrict__ pos, const int32_t* __restrict__ offsets,
                                    int64_t n, uint8_t* __restrict__ pool) {
  const int64_t row = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (row >= n) return;
  const int32_t b = offsets[row], e = offsets[row + 1];
  const uint8_t* src = data + pos[row];
  for (int32_t k = 0; k < e - b; ++k) pool[b + k] = src[k];
}

}  // namespace txf

extern "C" {

int64_t txf_tfrecord_block_bytes(void) { return txf::kTfrBlockBytes; }

int txf_tfrecord_count_candidates(const uint8_t* data_dev, int64_t nbytes, uint32_t* block_counts_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(data_dev && block_counts_dev && nbytes >= 0, "tfrecord_count_candidates: bad arguments");
  if (nbytes == 0) return TXF_OK;
  const unsigned blocks = static_cast<unsigned>((nbytes + kTfrBlockBytes - 1) / kTfrBlockBytes);
  tfr_candidates_kernel<0><<<blocks, kTfrThreads, 0, static_cast<cudaStream_t>(stream)>>>(data_dev, nbytes, block_counts_dev, nullptr, nullptr);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_tfrecord_candidates(const uint8_t* data_dev, int64_t nbytes, const int64_t* block_base_dev, int64_t* cand_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(data_dev && block_base_dev && cand_dev && nbytes >= 0, "tfrecord_candidates: bad arguments");
  if (nbytes == 0) return TXF_OK;
  const unsigned blocks = static_cast<unsigned>((nbytes + kTfrBlockBytes - 1) / kTfrBlockBytes);
  tfr_candidates_kernel<1><<<blocks, kTfrThreads, 0, static_cast<cudaStream_t>(stream)>>>(data_dev, nbytes, nullptr, block_base_dev, cand_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_tfrecord_link(const uint8_t* data_dev, const int64_t* cand_dev, int64_t ncand, uint8_t* mark_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ncand >= 0, "tfrecord_link: negative count");
  if (ncand == 0) return TXF_OK;
  TXF_REQUIRE(data_dev && cand_dev && mark_dev, "tfrecord_link: null pointer");
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(mark_dev, 0, static_cast<size_t>(ncand), cs));
  tfr_link_kernel<<<static_cast<unsigned>((ncand + 255) / 256), 256, 0, cs>>>(data_dev, cand_dev, ncand, mark_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_example_decode(const txf_example_column* feats, int nfeat, const uint8_t* names_dev, const uint8_t* data_dev,
                       const int64_t* record_start_dev, int64_t n, int verify_crc, uint64_t* status_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(feats != nullptr && nfeat >= 1 && nfeat <= TXF_EX_MAX_FEATURES, "example_decode: nfeat=%d out of [1,%d]", nfeat, TXF_EX_MAX_FEATURES);
  TXF_REQUIRE(names_dev && data_dev && record_start_dev && status_dev && n >= 0 && n < (1LL << 31), "example_decode: bad arguments");
  TXF_REQUIRE(aligned(data_dev, 16), "example_decode: the record stream must be 16-byte aligned");
  DecFeats d{};
  d.n = nfeat;
  for (int i = 0; i < nfeat; ++i) {
    const txf_example_column& s = feats[i];
    TXF_REQUIRE(s.kind == TXF_F32 || s.kind == TXF_I64 || s.kind == TXF_EX_BYTES, "example_decode: feature %d has bad kind %d", i, s.kind);
    TXF_REQUIRE(s.values != nullptr && s.valid != nullptr && aligned(s.valid, 4), "example_decode: feature %d needs values and a 4-byte aligned bitmap", i);
    TXF_REQUIRE(s.kind != TXF_EX_BYTES || s.lens != nullptr, "example_decode: string feature %d needs lens", i);
    d.f[i] = DecFeat{s.values, s.lens, static_cast<uint32_t*>(s.valid), s.kind, s.name_off, s.name_len};
  }
  if (n == 0) return TXF_OK;
  TXF_CUDA(cudaFuncSetAttribute(example_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDecStageBytes));
  example_decode_kernel<<<static_cast<unsigned>((n + kDecRows - 1) / kDecRows), kDecRows, kDecStageBytes, static_cast<cudaStream_t>(stream)>>>(
      d, names_dev, data_dev, record_start_dev, n, verify_crc, reinterpret_cast<unsigned long long*>(status_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_gather_bytes(const uint8_t* data_dev, const uint32_t* pos_dev, const int32_t* offsets_dev, int64_t n, uint8_t* pool_dev,
                     txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "gather_bytes: n out of range");
  if (n == 0) return TXF_OK;
  TXF_REQUIRE(data_dev && pos_dev && offsets_dev && pool_dev, "gather_bytes: null pointer");
  gather_bytes_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(data_dev, pos_dev, offsets_dev, n, pool_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
