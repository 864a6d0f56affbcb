// stream_ops.cu -- the ops of preprocessing_fn that need NO analyzer result (categorical
// pass-through airflow-dags/taxi_utils.py:132-133 with the fill rule of :86-103, and the tips
// label :135-143) as a small-footprint streaming kernel that runs UNDER the analyzers.
//
// The counting analyzers (vocab_count_kernel, qcount_kernel) are bound by instruction issue and
// latency and leave three quarters of the HBM bandwidth idle, while half of the transform's bytes
// (six int64 pass-throughs, the string offsets, the label) do not depend on them.  This kernel moves
// those bytes while the analyzers run: one CTA of 128 threads per SM, ~42 KB of shared memory, few
// registers -- sized to sit next to an analyzer CTA on the same SM (the analyzers leave that room,
// see txf_vocab_count's cta_warps).  Bytes in flight do not come from threads but from the bulk-copy
// engine: a tile of 8 KB is landed in a shared-memory stage (cp.async.bulk, completion on the
// stage's mbarrier), fixed up in place if the op needs it (missing cells -> 0, offsets rebased, the
// label computed), and written back with a bulk store.  Five stages per CTA, three loads ahead.
// Tiles are handed out dynamically (chunks of an atomic counter), so SMs that are busier with an
// analyzer simply take fewer.
//
// Algorithmic bytes: every input cell read once, every output cell written once (SURVEY.md 8d:
// 56 + 4 + 8 B read and 56 + 4 B written per row on the taxi schema for these ops).
#include "common.cuh"
#include "str_pipe.cuh"

namespace txf {

constexpr int kSsThreads = 128;
constexpr int kSsConsumers = kSsThreads - 32;  // warp 0 produces, warps 1..3 consume
constexpr int kSsTile = 8192;    // data bytes of a stage: 1024 int64 cells / 2048 4-byte cells / 1024 rows of (fare, tips)
constexpr int kSsValid = 256;    // validity bytes of a stage (the label stages two bitmaps of 128 bytes)
constexpr int kSsStage = kSsTile + kSsValid;
constexpr int kSsStages = 5;
constexpr int kSsHeader = 512;   // mbarriers, job descriptors, per-op offsets base
constexpr int kSsChunk = 8;      // tiles claimed per atomic (the claim for the next chunk is issued a chunk ahead)
constexpr int kSsSmemBytes = kSsHeader + kSsStages * kSsStage;

struct SsOp {
  int32_t kind, per_tile;        // cells (rows) per tile
  const uint8_t* in0;
  const uint8_t* valid0;
  const uint8_t* in1;
  const uint8_t* valid1;
  uint8_t* out;
  long long first_job, ncell;    // ncell: rows, or rows + 1 for an offsets column
  float f0;
  int32_t in_size;               // bytes per input cell
};
struct SsProgram {
  SsOp ops[TXF_MAX_OPS];
  int32_t nops, pad;
  long long njobs;
};

__device__ __forceinline__ void bulk_copy_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void consumer_bar(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(kSsConsumers) : "memory"); }

__device__ __forceinline__ long long ss_label(float fare, float tips, float ratio) {
  const float thr = __fmul_rn(fare, ratio);  // fp32 product, never contracted (taxi_utils.py:143: 0.2 is a float32 constant)
  return (fare != fare) ? 0LL : (tips > thr ? 1LL : 0LL);
}

// job descriptor: op index in bits 0..7, bit 8 = partial tile (processed from global memory, nothing staged),
// bit 9 = validity not staged (misaligned bitmap: read it from global memory), tile index from bit 16; -1 = no more work
constexpr long long kSsPartial = 1 << 8, kSsValidGlobal = 1 << 9;

// does a full tile of this op need the threads at all?  (otherwise it is a pure bulk load + bulk store)
__device__ __forceinline__ bool ss_needs_fix(const SsOp& op, int32_t base) {
  if (op.kind == TXF_OP_RATIO_LABEL) return true;
  if (op.kind == TXF_OP_FILL_STR) return base != 0;
  return op.valid0 != nullptr;
}

// In-place fix-up of a staged tile by the CTA's consumer threads (t = 0..kSsConsumers-1; 16-byte accesses).
__device__ __forceinline__ void ss_fix_tile(const SsOp& op, long long desc, uint8_t* stage, int32_t base, int t) {
  const long long tile = desc >> 16;
  const long long c0 = tile * op.per_tile;
  if (desc & kSsPartial) {  // ragged last tile: plain loads and stores, cell by cell
    for (long long i = c0 + t; i < op.ncell; i += kSsConsumers) {
      switch (op.kind) {
        case TXF_OP_FILL_I64:
          reinterpret_cast<long long*>(op.out)[i] = row_valid(op.valid0, i) ? reinterpret_cast<const long long*>(op.in0)[i] : 0LL;
          break;
        case TXF_OP_FILL_F32:
          reinterpret_cast<float*>(op.out)[i] = row_valid(op.valid0, i) ? reinterpret_cast<const float*>(op.in0)[i] : 0.f;
          break;
        case TXF_OP_FILL_STR:
          reinterpret_cast<int32_t*>(op.out)[i] = reinterpret_cast<const int32_t*>(op.in0)[i] - base;
          break;
        default: {
          const float fare = row_valid(op.valid0, i) ? reinterpret_cast<const float*>(op.in0)[i] : 0.f;
          const float tips = row_valid(op.valid1, i) ? reinterpret_cast<const float*>(op.in1)[i] : 0.f;
          reinterpret_cast<long long*>(op.out)[i] = ss_label(fare, tips, op.f0);
        }
      }
    }
    return;
  }
  const bool vglobal = (desc & kSsValidGlobal) != 0;
  const uint8_t* v0 = vglobal ? op.valid0 + (c0 >> 3) : stage + kSsTile;
  uint4* d = reinterpret_cast<uint4*>(stage);
  constexpr int kVec = kSsTile / 16;  // 512 x 16 bytes
  switch (op.kind) {
    case TXF_OP_FILL_I64: {
      for (int idx = t; idx < kVec; idx += kSsConsumers) {  // two cells each
        const uint32_t bits = (v0[idx >> 2] >> ((idx & 3) * 2)) & 3u;
        if (bits != 3u) {
          uint4 x = d[idx];
          if (!(bits & 1u)) { x.x = 0; x.y = 0; }
          if (!(bits & 2u)) { x.z = 0; x.w = 0; }
          d[idx] = x;
        }
      }
      break;
    }
    case TXF_OP_FILL_F32: {
      for (int idx = t; idx < kVec; idx += kSsConsumers) {  // four cells each
        const uint32_t bits = (v0[idx >> 1] >> ((idx & 1) * 4)) & 15u;
        if (bits != 15u) {
          uint4 x = d[idx];
          if (!(bits & 1u)) x.x = 0;
          if (!(bits & 2u)) x.y = 0;
          if (!(bits & 4u)) x.z = 0;
          if (!(bits & 8u)) x.w = 0;
          d[idx] = x;
        }
      }
      break;
    }
    case TXF_OP_FILL_STR: {
      for (int idx = t; idx < kVec; idx += kSsConsumers) {
        uint4 x = d[idx];
        x.x -= static_cast<uint32_t>(base); x.y -= static_cast<uint32_t>(base);
        x.z -= static_cast<uint32_t>(base); x.w -= static_cast<uint32_t>(base);
        d[idx] = x;
      }
      break;
    }
    default: {  // TXF_OP_RATIO_LABEL: fare in the first 4 KB, tips in the second; 1024 int64 labels overwrite both
      const uint8_t* v1 = vglobal ? op.valid1 + (c0 >> 3) : stage + kSsTile + 128;
      constexpr int kPer = (256 + kSsConsumers - 1) / kSsConsumers;  // quads of 4 rows per thread
      long long r[kPer][4];
#pragma unroll
      for (int j = 0; j < kPer; ++j) {
        const int q = t + kSsConsumers * j;
        if (q < 256) {
          float4 fare = reinterpret_cast<const float4*>(stage)[q];
          float4 tips = reinterpret_cast<const float4*>(stage + 4096)[q];
          const uint32_t bf = op.valid0 ? (v0[q >> 1] >> ((q & 1) * 4)) & 15u : 15u;
          const uint32_t bt = op.valid1 ? (v1[q >> 1] >> ((q & 1) * 4)) & 15u : 15u;
          if (!(bf & 1u)) fare.x = 0.f; if (!(bf & 2u)) fare.y = 0.f; if (!(bf & 4u)) fare.z = 0.f; if (!(bf & 8u)) fare.w = 0.f;
          if (!(bt & 1u)) tips.x = 0.f; if (!(bt & 2u)) tips.y = 0.f; if (!(bt & 4u)) tips.z = 0.f; if (!(bt & 8u)) tips.w = 0.f;
          r[j][0] = ss_label(fare.x, tips.x, op.f0); r[j][1] = ss_label(fare.y, tips.y, op.f0);
          r[j][2] = ss_label(fare.z, tips.z, op.f0); r[j][3] = ss_label(fare.w, tips.w, op.f0);
        }
      }
      consumer_bar(2);  // every thread has read its inputs: the labels may overwrite them
#pragma unroll
      for (int j = 0; j < kPer; ++j) {
        const int q = t + kSsConsumers * j;
        if (q < 256) {
          reinterpret_cast<longlong2*>(stage)[2 * q] = make_longlong2(r[j][0], r[j][1]);
          reinterpret_cast<longlong2*>(stage)[2 * q + 1] = make_longlong2(r[j][2], r[j][3]);
        }
      }
    }
  }
}

// Warp 0 is the producer (lane 0 claims tiles and starts their bulk loads as soon as a stage is free);
// warps 1..3 are the consumers: they wait for a tile to land, fix it up in place if the op needs it,
// and one of them starts the bulk store and hands the stage back once the store has read it.  The two
// sides only meet at the stages' full / empty mbarriers (empty: one arrival per consumer warp), so a fix-up never
// delays the next load.
__global__ void __launch_bounds__(kSsThreads, 6) stream_ops_kernel(const __grid_constant__ SsProgram prog,
                                                                   unsigned long long* __restrict__ counter) {
  extern __shared__ __align__(128) uint8_t ss_smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(ss_smem);                          // [kSsStages]
  uint64_t* empty = reinterpret_cast<uint64_t*>(ss_smem + 64);                    // [kSsStages]
  long long* s_job = reinterpret_cast<long long*>(ss_smem + 128);                 // [kSsStages]
  int32_t* s_base = reinterpret_cast<int32_t*>(ss_smem + 256);                    // [TXF_MAX_OPS] first offset of an offsets column
  uint8_t* stages = ss_smem + kSsHeader;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < prog.nops)
    s_base[threadIdx.x] = prog.ops[threadIdx.x].kind == TXF_OP_FILL_STR ? __ldg(reinterpret_cast<const int32_t*>(prog.ops[threadIdx.x].in0)) : 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kSsStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kSsConsumers / 32); }
    mbar_fence_init();
  }
  __syncthreads();

  if (warp == 0) {
    if (lane != 0) return;
    // ---- producer ----
    constexpr unsigned long long kCh = kSsChunk;
    unsigned long long cur = atomicAdd(counter, kCh), nxt = atomicAdd(counter, kCh);  // the next chunk is always claimed one chunk ahead
    int in_chunk = 0, o_run = 0;  // jobs are claimed in increasing order: the op index only moves forward
    for (uint32_t k = 0;; ++k) {
      const int s = static_cast<int>(k % kSsStages);
      mbar_wait(&empty[s], ((k / kSsStages) & 1u) ^ 1u);  // first use of a stage passes at once
      const long long j = static_cast<long long>(cur) + in_chunk;
      if (++in_chunk == kSsChunk) { cur = nxt; nxt = atomicAdd(counter, kCh); in_chunk = 0; }
      if (j >= prog.njobs) {
        s_job[s] = -1;
        mbar_arrive(&full[s]);
        break;
      }
      while (o_run + 1 < prog.nops && j >= prog.ops[o_run + 1].first_job) ++o_run;
      const SsOp& op = prog.ops[o_run];
      const long long tile = j - op.first_job, c0 = tile * op.per_tile;
      long long desc = (tile << 16) | o_run;
      uint8_t* st = stages + s * kSsStage;
      if (c0 + op.per_tile > op.ncell) {  // ragged last tile: nothing staged
        s_job[s] = desc | kSsPartial;
        mbar_arrive(&full[s]);
        continue;
      }
      const bool label = op.kind == TXF_OP_RATIO_LABEL;
      const uint32_t vbytes = static_cast<uint32_t>(op.per_tile >> 3);
      const uint8_t* v0 = op.valid0 ? op.valid0 + (c0 >> 3) : nullptr;
      const uint8_t* v1 = (label && op.valid1) ? op.valid1 + (c0 >> 3) : nullptr;
      const bool v_ok = ((reinterpret_cast<uintptr_t>(v0) | reinterpret_cast<uintptr_t>(v1)) & 15) == 0;
      if (!v_ok) desc |= kSsValidGlobal;
      s_job[s] = desc;
      uint32_t total = kSsTile;
      if (v_ok && v0) total += vbytes;
      if (v_ok && v1) total += vbytes;
      fence_proxy_async();  // the stage was last written by the consumers' ordinary stores
      mbar_arrive_expect_tx(&full[s], total);
      if (label) {
        bulk_copy_g2s(st, op.in0 + c0 * 4, 4096, &full[s]);
        bulk_copy_g2s(st + 4096, op.in1 + c0 * 4, 4096, &full[s]);
      } else {
        bulk_copy_g2s(st, op.in0 + c0 * op.in_size, kSsTile, &full[s]);
      }
      if (v_ok && v0) bulk_copy_g2s(st + kSsTile, v0, vbytes, &full[s]);
      if (v_ok && v1) bulk_copy_g2s(st + kSsTile + 128, v1, vbytes, &full[s]);
    }
    return;
  }

  // ---- consumers ----
  const int t = threadIdx.x - 32;
  const bool elected = t == 0;
  int prev_stage = -1;  // elected thread: stage whose store is the most recent bulk group
  for (uint32_t k = 0;; ++k) {
    const int s = static_cast<int>(k % kSsStages);
    mbar_wait(&full[s], (k / kSsStages) & 1u);
    const long long desc = s_job[s];
    __syncwarp();  // every lane has seen this phase of full[s] and read the job before a lane of the warp moves on
    if (desc < 0) break;
    const int o = static_cast<int>(desc & 0xFF);
    const SsOp& op = prog.ops[o];
    const bool partial = (desc & kSsPartial) != 0;
    uint8_t* st = stages + s * kSsStage;
    if (partial || ss_needs_fix(op, s_base[o])) {
      ss_fix_tile(op, desc, st, s_base[o], t);
      fence_proxy_async();  // the bulk store reads what the threads just wrote
      consumer_bar(1);
    }
    // A stage goes back to the producer once EVERY consumer warp is done with it: the elected thread arrives when the
    // store has read the stage (one iteration later), the other warps' first lanes here -- so no warp can fall a
    // whole round of stages behind and miss a phase of full[s] (pure-copy tiles have no other meeting point).
    if (lane == 0 && !elected) mbar_arrive(&empty[s]);
    if (elected) {
      if (!partial) {
        const long long c0 = (desc >> 16) * op.per_tile;
        const int out_size = op.kind == TXF_OP_FILL_I64 || op.kind == TXF_OP_RATIO_LABEL ? 8 : 4;
        bulk_copy_s2g(op.out + c0 * out_size, st, kSsTile);
      }
      bulk_commit();  // one group per tile (possibly empty)
      if (prev_stage >= 0) {
        bulk_wait_read_1();  // the previous tile's store has read its stage: hand it back to the producer
        mbar_arrive(&empty[prev_stage]);
      }
      prev_stage = s;
    }
  }
  if (elected) bulk_wait_all();
}

}  // namespace txf

extern "C" {

size_t txf_transform_stream_workspace_bytes(void) { return 256; }

int txf_transform_stream(const txf_op* ops, int nops, int64_t n, void* workspace_dev, txf_stream_t stream) {
  using namespace txf;
  TXF_REQUIRE(ops != nullptr && nops >= 1 && nops <= TXF_MAX_OPS, "transform_stream: nops=%d out of [1,%d]", nops, TXF_MAX_OPS);
  TXF_REQUIRE(n >= 0 && n < (1LL << 31), "transform_stream: n out of range");
  TXF_REQUIRE(workspace_dev != nullptr && aligned(workspace_dev, 8), "transform_stream: workspace null or misaligned");
  if (n == 0) return TXF_OK;
  SsProgram prog{};
  prog.nops = nops;
  long long jobs = 0;
  for (int o = 0; o < nops; ++o) {
    const txf_op& s = ops[o];
    SsOp& d = prog.ops[o];
    TXF_REQUIRE(s.in0 != nullptr && s.out != nullptr, "transform_stream: op %d has a null input/output", o);
    TXF_REQUIRE(aligned(s.in0, 16) && aligned(s.out, 16), "transform_stream: op %d input/output not 16B aligned", o);
    d.kind = s.kind;
    d.in0 = static_cast<const uint8_t*>(s.in0); d.valid0 = s.valid0;
    d.in1 = static_cast<const uint8_t*>(s.in1); d.valid1 = s.valid1;
    d.out = static_cast<uint8_t*>(s.out);
    d.f0 = s.f0;
    d.ncell = n;
    switch (s.kind) {
      case TXF_OP_FILL_I64: d.in_size = 8; d.per_tile = kSsTile / 8; break;
      case TXF_OP_FILL_F32: d.in_size = 4; d.per_tile = kSsTile / 4; break;
      case TXF_OP_FILL_STR: d.in_size = 4; d.per_tile = kSsTile / 4; d.ncell = n + 1; d.valid0 = nullptr; break;
      case TXF_OP_RATIO_LABEL:
        TXF_REQUIRE(s.in1 != nullptr && aligned(s.in1, 16), "transform_stream: op %d (label) needs in1 (16B aligned)", o);
        TXF_REQUIRE(s.in_dtype == TXF_F32, "transform_stream: op %d (label) takes fp32 columns", o);
        d.in_size = 4; d.per_tile = kSsTile / 8;
        break;
      default:
        TXF_REQUIRE(false, "transform_stream: op %d (kind %d) needs analyzer state; it belongs to txf_transform", o, s.kind);
    }
    d.first_job = jobs;
    jobs += (d.ncell + d.per_tile - 1) / d.per_tile;
  }
  prog.njobs = jobs;
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  TXF_CUDA(cudaMemsetAsync(workspace_dev, 0, sizeof(unsigned long long), cs));
  TXF_CUDA(cudaFuncSetAttribute(stream_ops_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSsSmemBytes));
  // the SM's L1/shared split is only changed when the SM is idle: ask for the largest shared-memory split up front, or an
  // analyzer CTA that needs 185 KB could not join this kernel's CTA on the SM before it has finished
  TXF_CUDA(cudaFuncSetAttribute(stream_ops_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  const long long want = (jobs + kSsChunk - 1) / kSsChunk;
  const int grid = static_cast<int>(want < kSMs ? want : kSMs);
  stream_ops_kernel<<<grid, kSsThreads, kSsSmemBytes, cs>>>(prog, static_cast<unsigned long long*>(workspace_dev));
  TXF_LAUNCH_CHECK();
  count_launch(1);
  return TXF_OK;
}

}  // extern "C"
