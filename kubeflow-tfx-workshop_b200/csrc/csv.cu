// csv.cu -- CSV ingest on the GPU (SURVEY.md 8 f2 / north_star subsystem 1): the text of a CSV file
// already resident in HBM becomes typed device columns (int64 / float32 values + validity bitmaps,
// string positions) without touching the host again.  Reference path: TFX CsvExampleGen
// (airflow-dags/taxi_pipeline.py:70-73) and the schema-typed CSV -> Example loop of
// notebooks/04_TensorFlow_Model_Analysis.ipynb:222-240 ("empty field -> no value",
// int(value) / float(value) / bytes).
//
// Records are lines (the reference reads the file with a line reader, so a quoted field cannot
// hold a line break either); fields may be quoted with "" as the escaped quote.
//   1. csv_count_lines:  one CTA per 64 KB block counts '\n'           (16-byte loads)
//   2. csv_line_starts:  same blocks, ordered write of each line's first byte offset
//   3. csv_scan_rows:    256 rows per CTA; the CTA's byte span is staged in shared memory with
//                        coalesced 16-byte loads, then one thread walks one row.  MODE 0 classifies
//                        fields for type inference, MODE 1 parses them into the columns.
//   4. csv_copy_strings: bytes of one string column into its pool (un-escaping "" on the way)
// Decimal text -> float32 goes through the double the reference's float(value) produces: up to 19
// significant digits and a decimal exponent within +-22 give that double with ONE correctly
// rounded multiply or divide (both operands exact); anything else is reported, not approximated.
#include "common.cuh"
#include "decimal.cuh"

namespace txf {

constexpr int kCsvBlockBytes = 64 * 1024;
constexpr int kCsvThreads = 256;
constexpr int kCsvRowsPerCta = 256;
constexpr int kCsvStageBytes = 48 * 1024;  // rows of one CTA (256 x ~130 B for the taxi file); 4 CTAs per SM; longer spans are read from global

__constant__ double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                  1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

__global__ void __launch_bounds__(kCsvThreads) csv_count_lines_kernel(const uint8_t* __restrict__ text, int64_t nbytes,
                                                                       uint32_t* __restrict__ block_lines) {
  const int64_t b0 = static_cast<int64_t>(blockIdx.x) * kCsvBlockBytes;
  const int64_t b1 = b0 + kCsvBlockBytes < nbytes ? b0 + kCsvBlockBytes : nbytes;
  uint32_t cnt = 0;
  // text is 16-byte aligned and blocks are multiples of 16 bytes
  for (int64_t i = b0 + threadIdx.x * 16LL; i < b1; i += kCsvThreads * 16LL) {
    if (i + 16 <= b1) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(text + i));
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) cnt += __popc(__vcmpeq4(w[k], 0x0A0A0A0Au)) >> 3;
    } else {
      for (int64_t j = i; j < b1; ++j) cnt += text[j] == '\n';
    }
  }
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  __shared__ uint32_t s[kCsvThreads / 32];
  if (lane_id() == 0) s[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < kCsvThreads / 32; ++w) t += s[w];
    block_lines[blockIdx.x] = t;
  }
}

// line_start[k] = offset of the first byte of line k; line_start[0] = 0 is written by block 0.
// block_base[b] = number of '\n' before block b (exclusive prefix sum of block_lines).
// 16 bytes per thread and step: a 16-bit newline mask per thread, a block-wide exclusive scan of
// the mask populations, then every thread writes the starts that follow its own newlines.
__device__ __forceinline__ uint32_t newline_mask4(uint32_t w) {
  const uint32_t m = __vcmpeq4(w, 0x0A0A0A0Au) & 0x80808080u;  // bit 7 of every matching byte
  return ((m >> 7) | (m >> 14) | (m >> 21) | (m >> 28)) & 0xFu;
}
__global__ void __launch_bounds__(kCsvThreads) csv_line_starts_kernel(const uint8_t* __restrict__ text, int64_t nbytes,
                                                                       const int64_t* __restrict__ block_base,
                                                                       int64_t* __restrict__ line_start) {
  const int64_t b0 = static_cast<int64_t>(blockIdx.x) * kCsvBlockBytes;
  const int64_t b1 = b0 + kCsvBlockBytes < nbytes ? b0 + kCsvBlockBytes : nbytes;
  __shared__ uint32_t s_warp[kCsvThreads / 32];
  if (threadIdx.x == 0 && blockIdx.x == 0) line_start[0] = 0;
  int64_t run = block_base[blockIdx.x] + 1;  // index of the next line start to write
  for (int64_t c0 = b0; c0 < b1; c0 += kCsvThreads * 16LL) {
    const int64_t i = c0 + threadIdx.x * 16LL;
    uint32_t mask = 0;
    if (i + 16 <= b1) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(text + i));
      mask = newline_mask4(v.x) | (newline_mask4(v.y) << 4) | (newline_mask4(v.z) << 8) | (newline_mask4(v.w) << 12);
    } else {
      for (int k = 0; k < 16 && i + k < b1; ++k) mask |= (text[i + k] == '\n' ? 1u : 0u) << k;
    }
    const uint32_t cnt = __popc(mask);
    uint32_t incl = cnt;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (static_cast<int>(lane_id()) >= o) incl += t;
    }
    if (lane_id() == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t before = incl - cnt, total = 0;
    for (int w = 0; w < kCsvThreads / 32; ++w) {
      if (w < static_cast<int>(threadIdx.x >> 5)) before += s_warp[w];
      total += s_warp[w];
    }
    int64_t at = run + before;
    while (mask) {
      const int k = __ffs(mask) - 1;
      mask &= mask - 1;
      line_start[at++] = i + k + 1;
    }
    run += total;
    __syncthreads();
  }
}

struct CsvCols {
  void* values[TXF_CSV_MAX_COLS];     // int64* / float* (numbers); uint32* field position (strings)
  int32_t* lens[TXF_CSV_MAX_COLS];    // strings: un-escaped field length
  uint8_t* valid[TXF_CSV_MAX_COLS];   // validity bitmaps (bit = field not empty)
  int32_t kind[TXF_CSV_MAX_COLS];     // TXF_CSV_SKIP / I64 / F32 / BYTES
};

struct Field {
  uint32_t pos, raw_len;  // content (inside the quotes when quoted), relative to the staged base
  uint32_t escapes;       // number of "" pairs inside
  bool quoted;
};

// number grammar: [+-] digits [. digits] [(e|E) [+-] digits] with at least one digit in the mantissa
// returns 0 = not a number, 1 = integer text, 2 = decimal text
template <class Get>
__device__ __forceinline__ int number_class(Get get, uint32_t n) {
  uint32_t i = 0;
  if (i < n && (get(i) == '+' || get(i) == '-')) ++i;
  uint32_t digits = 0;
  while (i < n && get(i) >= '0' && get(i) <= '9') { ++i; ++digits; }
  if (i == n) return digits ? 1 : 0;
  bool frac = false;
  if (get(i) == '.') {
    ++i;
    frac = true;
    while (i < n && get(i) >= '0' && get(i) <= '9') { ++i; ++digits; }
  }
  if (!digits) return 0;
  if (i < n && (get(i) == 'e' || get(i) == 'E')) {
    ++i;
    if (i < n && (get(i) == '+' || get(i) == '-')) ++i;
    uint32_t ed = 0;
    while (i < n && get(i) >= '0' && get(i) <= '9') { ++i; ++ed; }
    if (!ed) return 0;
    frac = true;
  }
  return i == n ? (frac ? 2 : 1) : 0;
}

// err bits (TXF_CSV_ERR_*)
template <class Get>
__device__ __forceinline__ long long parse_i64(Get get, uint32_t n, uint32_t& err) {
  uint32_t i = 0;
  bool neg = false;
  if (i < n && (get(i) == '+' || get(i) == '-')) { neg = get(i) == '-'; ++i; }
  unsigned long long m = 0;
  uint32_t digits = 0;
  bool over = false;
  for (; i < n; ++i) {
    const uint32_t c = get(i) - '0';
    if (c > 9u) { err |= TXF_CSV_ERR_NOT_A_NUMBER; return 0; }
    if (m > (0xFFFFFFFFFFFFFFFFULL - c) / 10ULL) over = true;
    m = m * 10ULL + c;
    ++digits;
  }
  if (!digits) { err |= TXF_CSV_ERR_NOT_A_NUMBER; return 0; }
  if (over || m > (neg ? 0x8000000000000000ULL : 0x7FFFFFFFFFFFFFFFULL)) { err |= TXF_CSV_ERR_RANGE; return 0; }
  return neg ? static_cast<long long>(0ULL - m) : static_cast<long long>(m);
}

template <class Get>
__device__ __forceinline__ float parse_f32(Get get, uint32_t n, uint32_t& err) {
  uint32_t i = 0;
  bool neg = false;
  if (i < n && (get(i) == '+' || get(i) == '-')) { neg = get(i) == '-'; ++i; }
  unsigned long long m = 0;
  int sig = 0, e10 = 0;
  uint32_t digits = 0;
  bool dropped = false, seen_dot = false;
  for (; i < n; ++i) {
    const uint32_t ch = get(i);
    if (ch == '.') {
      if (seen_dot) { err |= TXF_CSV_ERR_NOT_A_NUMBER; return 0.f; }
      seen_dot = true;
      continue;
    }
    const uint32_t c = ch - '0';
    if (c > 9u) break;
    ++digits;
    if (sig < 19) {
      if (m != 0 || c != 0) { m = m * 10ULL + c; ++sig; }
      if (seen_dot) --e10;
    } else {
      if (c != 0) dropped = true;
      if (!seen_dot) ++e10;
    }
  }
  if (!digits) { err |= TXF_CSV_ERR_NOT_A_NUMBER; return 0.f; }
  if (i < n) {
    if (get(i) != 'e' && get(i) != 'E') { err |= TXF_CSV_ERR_NOT_A_NUMBER; return 0.f; }
    ++i;
    bool eneg = false;
    if (i < n && (get(i) == '+' || get(i) == '-')) { eneg = get(i) == '-'; ++i; }
    int ex = 0;
    uint32_t ed = 0;
    for (; i < n; ++i) {
      const uint32_t c = get(i) - '0';
      if (c > 9u) { err |= TXF_CSV_ERR_NOT_A_NUMBER; return 0.f; }
      if (ex < 100000) ex = ex * 10 + static_cast<int>(c);
      ++ed;
    }
    if (!ed) { err |= TXF_CSV_ERR_NOT_A_NUMBER; return 0.f; }
    e10 += eneg ? -ex : ex;
  }
  double d;
  if (m == 0) {
    d = 0.0;
  } else {
    if (dropped) { err |= TXF_CSV_ERR_PRECISION; return 0.f; }  // more than 19 significant digits
    // float32 saturates long before double does: magnitudes beyond 10^40 are inf, below 10^-50 are 0
    if (e10 + sig > 40) return neg ? -__int_as_float(0x7f800000) : __int_as_float(0x7f800000);
    if (e10 + sig < -50) return neg ? -0.f : 0.f;
    while (e10 > 0 && m <= 0x1FFFFFFFFFFFFFULL / 10ULL) { m *= 10ULL; --e10; }  // fold positive exponents into m while exact
    if (m <= 0x1FFFFFFFFFFFFFULL && e10 >= -22 && e10 <= 22) {
      // both operands exact in double: ONE correctly rounded operation gives the nearest double
      d = static_cast<double>(m);
      d = e10 < 0 ? __ddiv_rn(d, kPow10[-e10]) : __dmul_rn(d, kPow10[e10]);
    } else if (e10 >= -kBigPow10Max && e10 <= kBigPow10Max) {
      d = decimal_to_double_exact(m, e10);
    } else {
      err |= TXF_CSV_ERR_PRECISION;
      return 0.f;
    }
  }
  const float f = __double2float_rn(d);
  return neg ? -f : f;
}

// MODE 0: classify (type inference), MODE 1: parse
template <int MODE>
__global__ void __launch_bounds__(kCsvThreads) csv_scan_rows_kernel(const uint8_t* __restrict__ text,
                                                                    const int64_t* __restrict__ line_start, int64_t first_line,
                                                                    int64_t nrows, int ncols, CsvCols cols,
                                                                    uint32_t* __restrict__ col_flags,
                                                                    unsigned long long* __restrict__ status) {
  extern __shared__ __align__(16) uint8_t s_text[];
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * kCsvRowsPerCta;
  const int64_t r1 = r0 + kCsvRowsPerCta < nrows ? r0 + kCsvRowsPerCta : nrows;
  const int64_t t0 = line_start[first_line + r0], t1 = line_start[first_line + r1];
  const int64_t a0 = t0 & ~15LL;
  const bool staged = t1 - a0 <= kCsvStageBytes;
  if (staged) {
    for (int64_t i = a0 + threadIdx.x * 16LL; i < t1; i += kCsvThreads * 16LL)
      *reinterpret_cast<uint4*>(s_text + (i - a0)) = __ldg(reinterpret_cast<const uint4*>(text + i));  // reads < 16 B past t1: caller pads
  }
  __syncthreads();
  const int64_t row = r0 + threadIdx.x;
  const bool live = row < r1;
  uint32_t err = 0;
  int64_t p = 0, end = 0;
  if (live) {
    p = line_start[first_line + row];
    end = line_start[first_line + row + 1];
    // a line ends at its '\n' (the last one may have none) and an optional '\r' before it
    auto raw = [&](int64_t i) -> uint32_t { return staged ? s_text[i - a0] : text[i]; };
    if (end > p && raw(end - 1) == '\n') --end;
    if (end > p && raw(end - 1) == '\r') --end;
  }
  auto at = [&](int64_t i) -> uint32_t { return staged ? s_text[i - a0] : text[i]; };
  for (int c = 0; c < ncols; ++c) {
    // ---- field boundaries -------------------------------------------------------------------
    int64_t f0 = p, f1 = p;
    uint32_t escapes = 0;
    bool quoted = false, present_field = live && (c == 0 || p <= end);
    if (live) {
      if (c > 0 && p > end) {
        err |= TXF_CSV_ERR_FIELD_COUNT;  // fewer fields than columns
      } else if (p < end && at(p) == '"') {
        quoted = true;
        f0 = p + 1;
        int64_t i = f0;
        for (;;) {
          if (i >= end) { err |= TXF_CSV_ERR_QUOTE; f1 = end; p = end + 1; break; }
          if (at(i) == '"') {
            if (i + 1 < end && at(i + 1) == '"') { ++escapes; i += 2; continue; }
            f1 = i;
            if (i + 1 < end && at(i + 1) != ',') err |= TXF_CSV_ERR_QUOTE;
            p = i + 2;
            break;
          }
          ++i;
        }
      } else {
        int64_t i = p;
        while (i < end && at(i) != ',') ++i;
        f1 = i;
        p = i + 1;
      }
    }
    const uint32_t n = static_cast<uint32_t>(f1 - f0);
    const bool has = present_field && n > 0;  // an empty field is a missing value, quoted or not (the reference tests value == '')
    auto get = [&](uint32_t k) -> uint32_t { return at(f0 + k); };
    if (MODE == 0) {
      uint32_t fl = 0;
      if (has) {
        const int cls = (quoted || escapes) ? 0 : number_class(get, n);
        fl = TXF_CSV_SEEN_VALUE | (cls == 1 ? TXF_CSV_SEEN_INT : cls == 2 ? TXF_CSV_SEEN_DECIMAL : TXF_CSV_SEEN_TEXT);
      }
      for (int o = 16; o > 0; o >>= 1) fl |= __shfl_xor_sync(0xffffffffu, fl, o);
      if (lane_id() == 0 && fl) atomicOr(&col_flags[c], fl);
    } else {
      const int kind = cols.kind[c];
      if (kind == TXF_CSV_SKIP) continue;
      const unsigned vm = __ballot_sync(0xffffffffu, has);
      if (lane_id() == 0 && r0 + (threadIdx.x & ~31) < r1)
        reinterpret_cast<uint32_t*>(cols.valid[c])[(r0 + threadIdx.x) >> 5] = vm;  // rows of a CTA start at a multiple of 256
      if (!live) continue;
      if (kind == TXF_CSV_I64) {
        static_cast<long long*>(cols.values[c])[row] = has ? parse_i64(get, n, err) : 0LL;
      } else if (kind == TXF_CSV_F32) {
        static_cast<float*>(cols.values[c])[row] = has ? parse_f32(get, n, err) : 0.f;
      } else {
        static_cast<uint32_t*>(cols.values[c])[row] = static_cast<uint32_t>(f0) | (escapes ? 0x80000000u : 0u);
        cols.lens[c][row] = has ? static_cast<int32_t>(n - escapes) : 0;
      }
    }
  }
  if (live && p <= end) err |= TXF_CSV_ERR_FIELD_COUNT;  // more fields than columns
  if (err) {
    atomicOr(&status[0], static_cast<unsigned long long>(err));
    atomicMin(&status[1], static_cast<unsigned long long>(first_line + row));
  }
}

// pool[offsets[r] .. offsets[r+1]) = bytes of the field at text[pos[r]...] (bit 31 of pos: has "" pairs)
__global__ void csv_copy_strings_kernel(const uint8_t* __restrict__ text, const uint32_t* __restrict__ pos,
                                        const int32_t* __restrict__ offsets, int64_t nrows, uint8_t* __restrict__ pool) {
  const int64_t r = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (r >= nrows) return;
  const uint32_t ps = pos[r];
  const uint8_t* src = text + (ps & 0x7FFFFFFFu);
  uint8_t* dst = pool + offsets[r];
  const int32_t n = offsets[r + 1] - offsets[r];
  if (ps & 0x80000000u) {
    for (int32_t i = 0, j = 0; i < n; ++i, ++j) {
      dst[i] = src[j];
      if (src[j] == '"') ++j;  // "" -> "
    }
  } else {
    for (int32_t i = 0; i < n; ++i) dst[i] = src[i];
  }
}

}  // namespace txf

extern "C" {

int64_t txf_csv_block_bytes(void) { return txf::kCsvBlockBytes; }

int txf_csv_count_lines(const uint8_t* text_dev, int64_t nbytes, uint32_t* block_lines_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(text_dev != nullptr && block_lines_dev != nullptr, "csv_count_lines: null argument");
  TXF_REQUIRE(aligned(text_dev, 16), "csv_count_lines: text not 16B aligned");
  TXF_REQUIRE(nbytes >= 0 && nbytes < (1LL << 31), "csv_count_lines: at most 2 GiB of text per call");
  if (nbytes == 0) return TXF_OK;
  const unsigned grid = static_cast<unsigned>((nbytes + kCsvBlockBytes - 1) / kCsvBlockBytes);
  csv_count_lines_kernel<<<grid, kCsvThreads, 0, static_cast<cudaStream_t>(stream)>>>(text_dev, nbytes, block_lines_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_csv_line_starts(const uint8_t* text_dev, int64_t nbytes, const int64_t* block_base_dev, int64_t* line_start_dev,
                        txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(text_dev && block_base_dev && line_start_dev, "csv_line_starts: null argument");
  TXF_REQUIRE(nbytes >= 0 && nbytes < (1LL << 31), "csv_line_starts: at most 2 GiB of text per call");
  if (nbytes == 0) return TXF_OK;
  const unsigned grid = static_cast<unsigned>((nbytes + kCsvBlockBytes - 1) / kCsvBlockBytes);
  csv_line_starts_kernel<<<grid, kCsvThreads, 0, static_cast<cudaStream_t>(stream)>>>(text_dev, nbytes, block_base_dev, line_start_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

static int csv_scan(int mode, const uint8_t* text_dev, const int64_t* line_start_dev, int64_t first_line, int64_t nrows, int ncols,
                    const txf_csv_column* columns, uint32_t* col_flags_dev, uint64_t* status_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(text_dev && line_start_dev && status_dev, "csv: null argument");
  TXF_REQUIRE(aligned(text_dev, 16), "csv: text not 16B aligned");
  TXF_REQUIRE(ncols >= 1 && ncols <= TXF_CSV_MAX_COLS, "csv: ncols=%d out of [1,%d]", ncols, TXF_CSV_MAX_COLS);
  TXF_REQUIRE(nrows >= 0 && nrows < (1LL << 31) && first_line >= 0, "csv: row range out of bounds");
  if (nrows == 0) return TXF_OK;
  CsvCols c{};
  if (mode == 1) {
    TXF_REQUIRE(columns != nullptr, "csv_parse: null column table");
    for (int i = 0; i < ncols; ++i) {
      const txf_csv_column& k = columns[i];
      TXF_REQUIRE(k.kind >= TXF_CSV_SKIP && k.kind <= TXF_CSV_BYTES, "csv_parse: column %d has bad kind %d", i, k.kind);
      c.kind[i] = k.kind;
      if (k.kind == TXF_CSV_SKIP) continue;
      TXF_REQUIRE(k.values != nullptr && k.valid != nullptr && aligned(k.valid, 4), "csv_parse: column %d needs values and a 4B-aligned bitmap", i);
      TXF_REQUIRE(k.kind != TXF_CSV_BYTES || k.lens != nullptr, "csv_parse: string column %d needs lens", i);
      c.values[i] = k.values;
      c.lens[i] = k.lens;
      c.valid[i] = static_cast<uint8_t*>(k.valid);
    }
  } else {
    TXF_REQUIRE(col_flags_dev != nullptr, "csv_classify: null flags");
  }
  const unsigned grid = static_cast<unsigned>((nrows + kCsvRowsPerCta - 1) / kCsvRowsPerCta);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (mode == 0) {
    TXF_CUDA(cudaFuncSetAttribute(csv_scan_rows_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCsvStageBytes));
    csv_scan_rows_kernel<0><<<grid, kCsvThreads, kCsvStageBytes, s>>>(text_dev, line_start_dev, first_line, nrows, ncols, c, col_flags_dev,
                                                                     reinterpret_cast<unsigned long long*>(status_dev));
  } else {
    TXF_CUDA(cudaFuncSetAttribute(csv_scan_rows_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCsvStageBytes));
    csv_scan_rows_kernel<1><<<grid, kCsvThreads, kCsvStageBytes, s>>>(text_dev, line_start_dev, first_line, nrows, ncols, c, nullptr,
                                                                     reinterpret_cast<unsigned long long*>(status_dev));
  }
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

int txf_csv_classify(const uint8_t* text_dev, const int64_t* line_start_dev, int64_t first_line, int64_t nrows, int ncols,
                     uint32_t* col_flags_dev, uint64_t* status_dev, txf_stream_t stream) {
  return csv_scan(0, text_dev, line_start_dev, first_line, nrows, ncols, nullptr, col_flags_dev, status_dev, stream);
}

int txf_csv_parse(const uint8_t* text_dev, const int64_t* line_start_dev, int64_t first_line, int64_t nrows, int ncols,
                  const txf_csv_column* columns, uint64_t* status_dev, txf_stream_t stream) {
  return csv_scan(1, text_dev, line_start_dev, first_line, nrows, ncols, columns, nullptr, status_dev, stream);
}

int txf_csv_copy_strings(const uint8_t* text_dev, const uint32_t* pos_dev, const int32_t* offsets_dev, int64_t nrows,
                         uint8_t* pool_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(text_dev && pos_dev && offsets_dev && pool_dev, "csv_copy_strings: null argument");
  TXF_REQUIRE(nrows >= 0 && nrows < (1LL << 31), "csv_copy_strings: nrows out of range");
  if (nrows == 0) return TXF_OK;
  csv_copy_strings_kernel<<<static_cast<unsigned>((nrows + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      text_dev, pos_dev, offsets_dev, nrows, pool_dev);
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
