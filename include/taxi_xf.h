/*
 * taxi_xf.h -- C ABI of libtaxi_xf.so: the B200 (sm_100a) analyze+transform kernels behind
 * `preprocessing_fn` of the Chicago-taxi TFX pipeline.
 *
 * The reference (pablomendes/kubeflow-tfx-workshop) has no FFI of its own: the hot path is the
 * Python callback airflow-dags/taxi_utils.py:106-145, whose arithmetic runs inside
 * tensorflow-transform / TensorFlow / Beam.  Each entry point below therefore cites the
 * reference CALL SITE (file:line under /root/reference) whose behaviour it replaces.
 *
 * Conventions
 *  - every function returns 0 (TXF_OK) or a negative TXF_ERR_*; txf_last_error() returns the
 *    message of the calling thread's last failure;
 *  - plain pointers and sizes only.  Pointers named *_dev / marked "device" are CUDA device
 *    addresses owned by the caller (e.g. torch tensors' data_ptr()); nothing here allocates or
 *    frees device memory, and nothing synchronises the stream unless its name ends in _sync;
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *  - a chunk of rows is at most 2^31-1 rows, string pools at most 2^31-1 bytes (int32
 *    offsets, Arrow layout).  Larger shards are fed chunk by chunk; analyzer state accumulates;
 *  - numeric values pointers must be 16-byte aligned; validity is an Arrow bitmap (LSB first,
 *    1 = present) or NULL when every row is present; string pools must be readable for
 *    TXF_POOL_SLACK bytes past their last byte.
 */
#ifndef TAXI_XF_H_
#define TAXI_XF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TXF_ABI_VERSION 1
#define TXF_MAX_COLS 16   /* columns handled by one analyzer launch */
#define TXF_MAX_OPS 32    /* column ops in one fused transform program */
#define TXF_POOL_SLACK 64 /* readable bytes required past the end of a string pool */
#define TXF_QMAX_TARGETS 64 /* bracket-count fast path: at most this many boundaries */

enum { TXF_OK = 0, TXF_ERR_ARG = -1, TXF_ERR_CUDA = -2, TXF_ERR_CAPACITY = -3 };
enum { TXF_F32 = 0, TXF_I64 = 1 };

typedef void* txf_stream_t;

#if defined(__GNUC__)
#define TXF_API __attribute__((visibility("default")))
#else
#define TXF_API
#endif

TXF_API const char* txf_last_error(void);
TXF_API int txf_abi_version(void);

/* ------------------------------------------------------------------------------------------
 * Moments analyzer  -- tft.scale_to_z_score's mean/variance combiner
 * (taxi_utils.py:116-119).  State per column: double[3] = (count, mean, M2) over the
 * zero-filled column (_fill_in_missing, taxi_utils.py:86-103); int64 cells are rounded to fp32
 * first, as tft does.
 * ------------------------------------------------------------------------------------------ */
TXF_API size_t txf_moments_workspace_bytes(int ncols);
/* accumulate one chunk into state_dev[ncols*3] (zero-initialised by the caller) */
TXF_API int txf_moments_update(const void* const* values, const uint8_t* const* valid, const int32_t* dtype,
                       int ncols, int64_t n, double* state_dev, void* workspace_dev, txf_stream_t stream);
/* Same accumulation, for callers that cannot hand over a workspace.  (No counterpart in the
 * reference: this entry point exists for the hosts described here.)  Some hosts embed this library
 * as a plug-in of a larger runtime -- a Beam/TFX worker, a serving process with a fixed device-memory
 * budget -- which passes input and output buffers in but gives the plug-in no device allocator, and
 * where a cudaMalloc would stall every stream of the process and take memory outside that budget.
 * For them the per-block partials live in a fixed array inside the library instead of
 * workspace_dev: nothing is allocated or freed by this call, ever.
 *  - limit: at most TXF_NOALLOC_MAX_COLS columns per call (more -> TXF_ERR_ARG with a message);
 *    the array (about 111 KB) is held by every GPU that loads the library until the process exits;
 *  - the first call on a device looks up the array's device address; make that first call
 *    before any stream capture begins on that device;
 *  - all calls on a device share the array (it keeps whatever the last call left there; each
 *    call overwrites every partial it later reads, so no clearing is needed): two calls whose
 *    kernels may overlap on the GPU -- different streams, no ordering between them -- race. */
#define TXF_NOALLOC_MAX_COLS 4
TXF_API int txf_moments_update_noalloc(const void* const* values, const uint8_t* const* valid, const int32_t* dtype,
                               int ncols, int64_t n, double* state_dev, txf_stream_t stream);
/* Chan-merge nstates states laid out [nstates][ncols][3] in index order (rank order) */
TXF_API int txf_moments_merge(const double* states_dev, int nstates, int ncols, double* state_out_dev,
                      txf_stream_t stream);
/* state -> float[ncols*2] = (mean_f32, population variance_f32) */
TXF_API int txf_moments_finalize(const double* state_dev, int ncols, float* mean_var_dev, txf_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Exact multi-rank selection on 4-byte keys (radix select, 4 passes x 8 bits).  The general
 * engine behind tft.bucketize's quantile analyzer (taxi_utils.py:128-130) and behind the
 * count cut-off of the vocabulary top-k (taxi_utils.py:123-126).  Histograms are plain integer
 * sums, so shards/ranks merge by addition (all-reduce) between _hist and _resolve.
 *
 * state layout (bytes): u64 hist[ncols][R][256] first (txf_msel_hist_bytes), then private.
 * kind: 0 = numeric column cells (zero-fill, ->fp32, order key); 1 = raw uint32 keys.
 * ranks are 1-based positions in ascending order, ascending per column; 0 = inactive entry.
 * ------------------------------------------------------------------------------------------ */
TXF_API size_t txf_msel_state_bytes(int ncols, int R);
TXF_API size_t txf_msel_hist_bytes(int ncols, int R);
TXF_API int txf_msel_begin(void* state_dev, int ncols, int R, const uint64_t* ranks_dev, txf_stream_t stream);
/* n_dev (device, per column) overrides n when non-NULL; n then only bounds the grid */
TXF_API int txf_msel_hist(void* state_dev, int ncols, int R, int pass, const void* const* data,
                  const uint8_t* const* valid, const int32_t* dtype, int kind, int64_t n,
                  const uint64_t* n_dev, txf_stream_t stream);
TXF_API int txf_msel_resolve(void* state_dev, int ncols, int R, int pass, txf_stream_t stream);
/* after pass 3: keys_out_dev[ncols*R] = selected keys (inactive entries -> 0) */
TXF_API int txf_msel_result(const void* state_dev, int ncols, int R, uint32_t* keys_out_dev, txf_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Quantile boundaries, single-pass fast path (tft.bucketize analyzer, taxi_utils.py:128-130):
 * sample -> brackets (via txf_msel_* on the sample) -> ONE counting pass over the column ->
 * exact order statistics from the few cells that straddle each target rank.
 * ------------------------------------------------------------------------------------------ */
/* gather M stratified sample keys per column: sample_dev[ncols][M] */
TXF_API int txf_qsample(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols,
                int64_t n, int64_t M, uint64_t seed, uint32_t* sample_dev, txf_stream_t stream);
/* edges_dev[ncols][2T] ascending keys (lo0,hi0,lo1,hi1,...).  Accumulates
 * cells_dev[ncols][4T+1] (u64; cell 2p = strictly between edge p-1 and p, cell 2p+1 = equal to
 * edge p) and appends keys strictly inside a bracket to cand_dev[ncols][cand_cap]
 * (cand_count_dev[ncols], u64, keeps counting past cand_cap so overflow is detectable). */
TXF_API int txf_qcount(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols,
               int64_t n, int T, const uint32_t* edges_dev, uint64_t* cells_dev, uint32_t* cand_dev,
               uint64_t* cand_count_dev, int64_t cand_cap, txf_stream_t stream);
/* per target: answer directly (direct_key, cand_rank = 0) or give its 1-based rank inside the
 * column's candidate set; status_dev[col] != 0 when a target fell outside its bracket or the
 * candidate buffer overflowed (caller then uses the general txf_msel_* path on the column) */
TXF_API int txf_qresolve(const uint64_t* cells_dev, const uint32_t* edges_dev, const uint64_t* cand_count_dev,
                 int64_t cand_cap, const uint64_t* target_ranks_dev, int ncols, int T,
                 uint32_t* direct_key_dev, uint64_t* cand_rank_dev, int32_t* status_dev,
                 txf_stream_t stream);
/* Selection among the candidates in ONE histogram pass (instead of txf_msel_*'s four): the candidates of bracket p
 * are counted by (key - lower edge - 1) into B bins of hist_dev (u32 [ncols][T][B], then [ncols] overflow words;
 * txf_bsel_hist zeroes it).  Shards / ranks merge by adding hist_dev.  txf_bsel_resolve then writes the selected
 * keys; a column whose bracket spans more than B keys gets status bit 4 (use txf_msel_* on its candidates instead). */
TXF_API size_t txf_bsel_hist_bytes(int ncols, int T, int B);
TXF_API int txf_bsel_hist(const uint32_t* cand_dev, int64_t cand_cap, const uint64_t* cand_count_dev, const uint32_t* edges_dev,
                          int ncols, int T, int B, uint32_t* hist_dev, txf_stream_t stream);
TXF_API int txf_bsel_resolve(const uint32_t* hist_dev, const uint64_t* cells_dev, const uint32_t* edges_dev,
                             const uint64_t* cand_rank_dev, int ncols, int T, int B, uint32_t* selected_key_dev,
                             int32_t* status_dev, txf_stream_t stream);
/* keys (direct where cand_rank == 0, else selected) -> distinct fp32 boundaries
 * boundaries_dev[ncols][T], nbound_dev[ncols] */
TXF_API int txf_qfinal(const uint32_t* direct_key_dev, const uint64_t* cand_rank_dev, const uint32_t* selected_key_dev,
               int ncols, int T, float* boundaries_dev, int32_t* nbound_dev, txf_stream_t stream);

/* Quantile boundaries for MANY targets (num_buckets up to 1000 and more), two passes over the column
 * (tft.bucketize analyzer, taxi_utils.py:128-130, at the sizes of BASELINE configs[4]):
 *   edges_dev[ncols][E]: E <= 4096 sorted splitter keys (equi-depth order statistics of a sample, via txf_msel_*);
 *   1. txf_qsplit_count   accumulates cells_dev[ncols][2E+1] (u64; cell 2p = strictly between splitter p-1 and p,
 *                          2p+1 = equal to splitter p); shards / ranks merge by addition;
 *   2. txf_qsplit_resolve  per target rank (ascending, 1-based; T <= 1024): the splitter it falls on (direct_key,
 *                          cand_rank = 0) or its cell -- flagged in flags_dev[ncols][txf_qsplit_flag_words(E)] -- and
 *                          its 1-based rank among all keys pass 2 will collect for the column; cand_total_dev[ncols];
 *   3. txf_qsplit_collect  appends the keys of flagged cells to cand_dev[ncols][cand_cap] (cand_count_dev[ncols] keeps
 *                          counting past cand_cap so that overflow is detectable);
 *   4. txf_msel_* on the collected keys with cand_rank, then txf_qfinal as for the single-pass engine. */
/* edges_dev[ncols][S] from sample_dev[ncols][M] (uint32 keys): S a power of two in [16, 4096], M >= 8 S */
TXF_API int txf_qsplit_edges(const uint32_t* sample_dev, int ncols, int64_t M, int S, uint32_t* edges_dev, txf_stream_t stream);
TXF_API size_t txf_qsplit_scratch_bytes(int ncols, int E);
TXF_API size_t txf_qsplit_flag_words(int E);
TXF_API int txf_qsplit_count(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols,
                             int64_t n, int E, const uint32_t* edges_dev, uint64_t* cells_dev, txf_stream_t stream);
TXF_API int txf_qsplit_resolve(const uint64_t* cells_dev, const uint32_t* edges_dev, const uint64_t* target_ranks_dev,
                               int ncols, int E, int T, uint32_t* direct_key_dev, uint64_t* cand_rank_dev,
                               uint32_t* flags_dev, uint64_t* cand_total_dev, void* scratch_dev, txf_stream_t stream);
TXF_API int txf_qsplit_collect(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols,
                               int64_t n, int E, const uint32_t* edges_dev, const uint32_t* flags_dev, uint32_t* cand_dev,
                               uint64_t* cand_count_dev, int64_t cand_cap, txf_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Vocabulary analyzer -- tft.compute_and_apply_vocabulary(top_k, num_oov_buckets)
 * (taxi_utils.py:121-126): exact occurrence counts in an open-addressing table keyed by the
 * term identity, first-seen bytes kept in an arena, then the top_k terms by (count, bytes)
 * descending.  Fingerprint64 is evaluated only for out-of-vocabulary terms.
 *
 * table: capacity (power of two) slots of 32 bytes {u64 a, u64 b, u64 count, u64 strref}.  (a, b) is
 * the term's key: the bytes themselves up to 15 bytes; beyond that a fingerprint (119 hash bits + the
 * exact length) that only pre-filters -- in this table, in the id lookup and in the decode table every
 * match of a longer term is confirmed on its bytes (tft counts byte strings), so a fingerprint
 * collision there never merges two terms.  (The count pass keeps a per-CTA cache in shared memory in
 * front of this table; the cache alone trusts the full fingerprint.);
 * ctrl_dev: u64[8] = {arena_used, slots_used, flags, kept, total, sel, dropped, 0}; flags: 1 = arena
 * full, 2 = table full, 4 = a fixed-layout exchange section overflowed, 16 = a count exceeds 2^32-1
 * (the top-k cut-off is selected on 32-bit counts), 32 = internal wait limit, 64 = some rank's local
 * pass overflowed (its flags in bits 8..13).  Table and ctrl must be zero-initialised by the caller
 * before the first count.  arena_dev must be readable for TXF_POOL_SLACK bytes past arena_cap.
 * ------------------------------------------------------------------------------------------ */
TXF_API size_t txf_vocab_table_bytes(int64_t capacity);
/* codes_dev (optional, uint32[n]) receives the dictionary code of every row's term and
 * local2global_dev (uint32[txf_vocab_local_codes()]) the table slot behind each CTA-local code;
 * txf_vocab_final_codes turns them into a decode table the fused transform gathers from, so
 * analyze+transform of the SAME rows never re-reads or re-hashes the term bytes. */
/* cta_warps: 0 or 32 = the count CTA takes a whole SM (fastest alone); 24 = it leaves room for a
 * txf_transform_stream CTA on the same SM (use when that kernel runs concurrently on another stream). */
TXF_API int64_t txf_vocab_local_codes(void);
TXF_API int txf_vocab_count(const int32_t* offsets_dev, const uint8_t* pool_dev, const uint8_t* valid_dev, int64_t n,
                    void* table_dev, int64_t capacity, uint8_t* arena_dev, int64_t arena_cap,
                    uint64_t* ctrl_dev, uint32_t* codes_dev, uint32_t* local2global_dev, int cta_warps,
                    txf_stream_t stream);
/* final_code_dev: int64[txf_vocab_local_codes() + capacity]; lookup_dev: the vocabulary's lookup
 * table (txf_vocab_build_lookup / _lookup_from_terms) and lookup_pool_dev the bytes it was built
 * over; semantics of the value = TXF_OP_VOCAB's output for that term */
TXF_API int txf_vocab_final_codes(const void* table_dev, int64_t capacity, const uint8_t* arena_dev,
                          const uint32_t* local2global_dev, const void* lookup_dev, int64_t lookup_cap,
                          const uint8_t* lookup_pool_dev, int64_t vocab_size, int64_t num_oov_buckets,
                          int64_t default_value, int64_t* final_code_dev, txf_stream_t stream);
/* dense list of occupied slots: ent_dev[capacity] (u32 slot ids), ctrl[4] = total distinct.
 * Terms tft never admits ('' and anything containing \n or \r) are listed last and not
 * counted in ctrl[3] = kept; counts_dev[i] = count of ent_dev[i] saturated to u32. */
TXF_API int txf_vocab_compact(const void* table_dev, int64_t capacity, const uint8_t* arena_dev, uint32_t* ent_dev,
                      uint32_t* counts_dev, uint64_t* ctrl_dev, txf_stream_t stream);
/* serialise m listed entries for exchange between ranks: fp/tag/count/len arrays, then (after
 * the caller turned len into exclusive offsets) the packed bytes */
TXF_API int txf_vocab_export_meta(const void* table_dev, const uint32_t* ent_dev, int64_t m, uint64_t* fp_dev,
                          uint64_t* tag_dev, uint64_t* count_dev, int32_t* len_dev, txf_stream_t stream);
TXF_API int txf_vocab_export_bytes(const void* table_dev, const uint8_t* arena_dev, const uint32_t* ent_dev, int64_t m,
                           const int64_t* out_offsets_dev, uint8_t* out_pool_dev, txf_stream_t stream);
/* fixed-layout exchange of a vocabulary between ranks without host-side sizes: a section is
 * u64 words [count | a[K] | b[K] | count[K] | len[K] | K byte slots of S bytes].  export fills this
 * rank's section from the first ctrl[3] listed entries; insert re-counts the sections of all ranks
 * (rows of an all-gathered buffer) into one table and sets ctrl[2] bit 2 if any section overflowed
 * (more than K admitted terms or a term longer than S) -- the caller then uses the sized path. */
TXF_API int txf_vocab_export_blob(const void* table_dev, const uint8_t* arena_dev, const uint32_t* ent_dev,
                          const uint64_t* ctrl_dev, int64_t K, int64_t S, uint64_t* section_dev, txf_stream_t stream);
TXF_API int txf_vocab_insert_blob(const uint64_t* blob_dev, int64_t stride_words, int64_t section_offset_words, int nranks,
                          int64_t K, int64_t S, void* table_dev, int64_t capacity, uint8_t* arena_dev,
                          int64_t arena_cap, uint64_t* ctrl_dev, txf_stream_t stream);
/* add m serialised entries (the ranks' exports) into a table.  owners > 1: only the entries whose key
 * hashes to `owner` (mod owners) are added -- every rank sees all entries and re-counts its own share,
 * which yields exact global counts per term at 1/owners of the insert work (the Beam CombinePerKey the
 * reference relies on, SURVEY.md 3.1, partitions by key the same way); owners = 1 adds everything. */
TXF_API int txf_vocab_insert(const uint64_t* fp_dev, const uint64_t* tag_dev, const uint64_t* count_dev,
                     const int64_t* offsets_dev, const uint8_t* pool_dev, int64_t m, int owners, int owner,
                     void* table_dev, int64_t capacity, uint8_t* arena_dev, int64_t arena_cap,
                     uint64_t* ctrl_dev, txf_stream_t stream);
/* keep entries with count >= *threshold_key_dev: ent_out_dev, ctrl[5] = how many */
TXF_API int txf_vocab_filter_ge(const uint32_t* ent_dev, const uint32_t* counts_dev, int64_t m,
                        const uint32_t* threshold_key_dev, uint32_t* ent_out_dev, uint64_t* ctrl_dev,
                        txf_stream_t stream);
/* sort m entries by (count, bytes) descending, in place (scratch: u32[next_pow2(m)]) */
TXF_API size_t txf_vocab_sort_scratch_bytes(int64_t m);
TXF_API int txf_vocab_sort(const void* table_dev, const uint8_t* arena_dev, uint32_t* ent_dev, int64_t m,
                   void* scratch_dev, txf_stream_t stream);
/* first k sorted entries -> id lookup table of txf_vocab_lookup_bytes(lookup_cap) bytes (lookup_cap:
 * power of two, >= 2k; 32-byte slots {a, b, strref, id}).  strref points into the byte pool the
 * table was built over -- arena_dev of the count table here, voc_pool_dev for _lookup_from_terms --
 * and that pool must stay alive and be passed wherever the table is used (TXF_OP_VOCAB p1,
 * txf_vocab_final_codes).  The vocabulary's bytes in id order come from txf_vocab_export_* on the same list */
TXF_API size_t txf_vocab_lookup_bytes(int64_t lookup_cap);
TXF_API int txf_vocab_build_lookup(const void* table_dev, const uint32_t* ent_sorted_dev, int64_t k, void* lookup_dev,
                           int64_t lookup_cap, txf_stream_t stream);
/* lookup table from an explicit vocabulary (saved transform state; tft.apply_vocabulary) */
TXF_API int txf_vocab_lookup_from_terms(const int32_t* voc_offsets_dev, const uint8_t* voc_pool_dev, int64_t k,
                                void* lookup_dev, int64_t lookup_cap, txf_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Fused transform -- the per-row phase of preprocessing_fn (taxi_utils.py:115-145): one
 * kernel runs a program of column ops over row tiles.
 * ------------------------------------------------------------------------------------------ */
enum {
  TXF_OP_ZSCORE = 0,     /* :116-119  out f32 = (f32(fill(x)) - mean) / sqrt(var); p0 = float[2] {mean,var} */
  TXF_OP_VOCAB = 1,      /* :121-126  in0 = int32 offsets, pool; p0 = lookup table, p1 = the byte pool the
                                       lookup table was built over, a0 = lookup_cap, a1 = vocab size,
                                       a2 = num_oov_buckets, a3 = default_value; out i64 */
  TXF_OP_BUCKETIZE = 2,  /* :128-130  p0 = float boundaries, p1 = int32* count (device); out i64 */
  TXF_OP_FILL_I64 = 3,   /* :132-133  out i64 = fill(x) */
  TXF_OP_FILL_F32 = 4,   /* :132-133  out f32 = fill(x) */
  TXF_OP_FILL_STR = 5,   /* :132-133  out int32 offsets rebased to 0 (bytes are copied by the caller) */
  TXF_OP_RATIO_LABEL = 6, /* :135-143  out i64 = isnan(in0) ? 0 : (in1 > in0 * f0); in0 = fare, in1 = tips */
  TXF_OP_VOCAB_CODED = 7  /* :121-126  in0 = uint32 codes from txf_vocab_count, p0 = int64 decode table
                                       (txf_vocab_final_codes); out i64 = table[code] */
};

typedef struct txf_op {
  int32_t kind;
  int32_t in_dtype; /* TXF_F32 / TXF_I64 where the op accepts both */
  const void* in0;
  const uint8_t* valid0;
  const void* in1;
  const uint8_t* valid1;
  const uint8_t* pool;
  void* out;
  const void* p0;
  const void* p1;
  int64_t a0, a1, a2, a3;
  float f0;
  int32_t reserved;
} txf_op;

TXF_API int txf_transform(const txf_op* ops, int nops, int64_t n, txf_stream_t stream);
/* Bytes of a pass-through string feature (TXF_OP_FILL_STR rebases the offsets to 0): out_pool[0 .. m) =
 * pool[offsets[0] .. offsets[n]), m = offsets[n] - offsets[0] <= max_bytes (the caller's bound on the span,
 * which only sizes the grid).  out_pool_dev: 8-byte aligned, m bytes + TXF_POOL_SLACK. */
TXF_API int txf_copy_string_bytes(const int32_t* offsets_dev, const uint8_t* pool_dev, int64_t n, int64_t max_bytes,
                                  uint8_t* out_pool_dev, txf_stream_t stream);
/* The ops of a program that need no analyzer result (TXF_OP_FILL_I64 / _F32 / _STR, TXF_OP_RATIO_LABEL;
 * taxi_utils.py:132-143) as a small-footprint streaming kernel meant to run on a second stream WHILE the
 * analyzers of the same rows run (tft_beam.AnalyzeAndTransformDataset makes one pass of both,
 * notebooks/06_Airflow_Feature_Analysis.ipynb:236-262): one 128-thread CTA per SM, 42 KB of shared
 * memory, tiles moved by the bulk-copy engine.  Same results as txf_transform for these ops; any other
 * kind is refused.  workspace_dev: txf_transform_stream_workspace_bytes() bytes, used by one call at a time. */
TXF_API size_t txf_transform_stream_workspace_bytes(void);
TXF_API int txf_transform_stream(const txf_op* ops, int nops, int64_t n, void* workspace_dev, txf_stream_t stream);
/* name of the transform kernel and launch geometry of the last call (for profiling notes) */
TXF_API int txf_transform_launch_info(int* grid, int* block, int* smem_bytes);

/* ------------------------------------------------------------------------------------------
 * CSV ingest on the GPU (north_star subsystem 1).  Replaces, for this path, TFX
 * CsvExampleGen (airflow-dags/taxi_pipeline.py:70-73) and the typed CSV -> Example loop of
 * notebooks/04_TensorFlow_Model_Analysis.ipynb:222-240: the text of the file, copied to the device
 * once, becomes int64 / float32 / string columns with validity bitmaps.  Records are lines; fields
 * may be quoted ("" = escaped quote); an empty field, quoted or not, is a missing value.
 *   text_dev: 16-byte aligned, with at least 16 readable bytes past nbytes; nbytes < 2^31.
 *   1. txf_csv_count_lines -> block_lines_dev[ceil(nbytes / txf_csv_block_bytes())] ('\n' per block)
 *   2. caller: exclusive prefix sum -> block_base_dev (int64); nlines = sum (+1 if the text does
 *      not end with '\n')
 *   3. txf_csv_line_starts -> line_start_dev[0 .. sum]; the caller sets line_start[nlines] = nbytes
 *   4. txf_csv_classify (optional, type inference): ORs TXF_CSV_SEEN_* into col_flags_dev[ncols]
 *   5. txf_csv_parse: rows [first_line, first_line + nrows) into the columns below
 *   6. per string column: caller prefix-sums lens into int32 offsets, then txf_csv_copy_strings
 * status_dev[0] = OR of TXF_CSV_ERR_* over all rows, status_dev[1] = smallest offending line
 * (the caller initialises them to 0 and UINT64_MAX).  Decimal text becomes the float32 nearest to
 * the double Python's float() gives, exactly, for up to 19 significant digits and a decimal
 * exponent within +-22 after folding; longer literals raise TXF_CSV_ERR_PRECISION.
 * ------------------------------------------------------------------------------------------ */
#define TXF_CSV_MAX_COLS 64
enum { TXF_CSV_SKIP = 0, TXF_CSV_I64 = 1, TXF_CSV_F32 = 2, TXF_CSV_BYTES = 3 };
enum { TXF_CSV_SEEN_VALUE = 1, TXF_CSV_SEEN_INT = 2, TXF_CSV_SEEN_DECIMAL = 4, TXF_CSV_SEEN_TEXT = 8 };
enum {
  TXF_CSV_ERR_FIELD_COUNT = 1,   /* a line has more or fewer fields than ncols */
  TXF_CSV_ERR_QUOTE = 2,         /* unterminated quote, or text after a closing quote */
  TXF_CSV_ERR_NOT_A_NUMBER = 4,  /* a field of a numeric column is not a number */
  TXF_CSV_ERR_RANGE = 8,         /* integer outside int64 */
  TXF_CSV_ERR_PRECISION = 16     /* decimal literal beyond the exactly convertible range */
};
typedef struct txf_csv_column {
  int32_t kind;     /* TXF_CSV_* */
  int32_t reserved;
  void* values;     /* int64[nrows] | float[nrows] | uint32 field position[nrows] (strings) */
  int32_t* lens;    /* strings: field length after un-escaping; else NULL */
  void* valid;      /* bitmap, ceil(nrows / 32) * 4 bytes, 4-byte aligned; nrows' first row % 256 == 0 */
} txf_csv_column;
TXF_API int64_t txf_csv_block_bytes(void);
TXF_API int txf_csv_count_lines(const uint8_t* text_dev, int64_t nbytes, uint32_t* block_lines_dev, txf_stream_t stream);
TXF_API int txf_csv_line_starts(const uint8_t* text_dev, int64_t nbytes, const int64_t* block_base_dev,
                                int64_t* line_start_dev, txf_stream_t stream);
TXF_API int txf_csv_classify(const uint8_t* text_dev, const int64_t* line_start_dev, int64_t first_line, int64_t nrows,
                             int ncols, uint32_t* col_flags_dev, uint64_t* status_dev, txf_stream_t stream);
TXF_API int txf_csv_parse(const uint8_t* text_dev, const int64_t* line_start_dev, int64_t first_line, int64_t nrows,
                          int ncols, const txf_csv_column* columns, uint64_t* status_dev, txf_stream_t stream);
TXF_API int txf_csv_copy_strings(const uint8_t* text_dev, const uint32_t* pos_dev, const int32_t* offsets_dev,
                                 int64_t nrows, uint8_t* pool_dev, txf_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * tf.Example / TFRecord encode on the GPU (egress).  Replaces, for this path,
 * ExampleProtoCoder.encode + WriteToTFRecord (notebooks/03a_TensorFlow_Transform_Advanced.ipynb:445-455;
 * the files are read back at airflow-dags/taxi_utils.py:79-83,260-281).  Row i of the given columns
 * becomes one framed record: u64 length, u32 masked CRC-32C of the length, the serialized Example
 * (map entries in the order of `feats`, which the caller sorts by name -- protobuf's deterministic
 * order), u32 masked CRC-32C of the data.  A missing cell is an empty value list.
 *   1. txf_example_sizes -> sizes_dev[n] (bytes of each framed record)
 *   2. caller: exclusive prefix sum -> offsets_dev[n + 1]
 *   3. txf_example_encode -> out_dev[offsets[n]]
 * names_dev holds the UTF-8 feature names back to back; name_off/name_len index it.
 * ------------------------------------------------------------------------------------------ */
#define TXF_EX_MAX_FEATURES 64
#define TXF_EX_BYTES 2 /* kind of a string feature (numbers use TXF_F32 / TXF_I64) */
typedef struct txf_example_feature {
  int32_t kind;       /* TXF_F32 | TXF_I64 | TXF_EX_BYTES */
  int32_t name_off;
  int32_t name_len;
  int32_t reserved;
  const void* values; /* float[n] | int64[n] | int32 offsets[n + 1] */
  const void* pool;   /* strings: byte pool */
  const void* valid;  /* validity bitmap or NULL */
} txf_example_feature;
TXF_API int txf_example_sizes(const txf_example_feature* feats, int nfeat, int64_t n, int64_t* sizes_dev,
                              txf_stream_t stream);
TXF_API int txf_example_encode(const txf_example_feature* feats, int nfeat, const uint8_t* names_dev, int64_t n,
                               const int64_t* offsets_dev, uint8_t* out_dev, txf_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * TFRecord / tf.Example DECODE on the GPU (ingest): what a TFX Transform executor is fed
 * (airflow-dags/taxi_pipeline.py:70-89; the reader is named at airflow-dags/taxi_utils.py:79-83 and used at
 * :260-281).  data_dev: the UNCOMPRESSED record stream (gunzip stays on the host), 16-byte aligned, with at
 * least 16 readable bytes past nbytes; nbytes < 2^31.
 *   1. txf_tfrecord_count_candidates -> block_counts_dev[ceil(nbytes / txf_tfrecord_block_bytes())]: positions
 *      that look like a record start (length field < 2^31 whose masked CRC-32C matches)
 *   2. caller: exclusive prefix sum -> block_base_dev (int64); ncand = sum
 *   3. txf_tfrecord_candidates -> cand_dev[ncand], ascending
 *   4. txf_tfrecord_link -> mark_dev[ncand]: 1 for position 0 and for every candidate some candidate's record
 *      ends at.  The marked candidates, in order, are the records; the caller checks that each one ends where
 *      the next begins and the last one at nbytes (anything else is a corrupt file) and appends nbytes:
 *      record_start_dev[n + 1]
 *   5. txf_example_decode: row i of every requested feature from record i.  A feature that is absent or has
 *      an empty list is a missing cell (validity 0, value 0 / length 0); one value is the cell; more than one
 *      sets TXF_EX_ERR_MULTI (the reference requires at most one value per row, taxi_utils.py:92-93).
 *      status_dev[0] = OR of TXF_EX_ERR_*, status_dev[1] = smallest offending record (caller: 0, UINT64_MAX)
 *   6. per string feature: caller prefix-sums lens into int32 offsets, then txf_gather_bytes
 * ------------------------------------------------------------------------------------------ */
enum {
  TXF_EX_ERR_TRUNCATED = 1, /* the length field does not match the record's extent */
  TXF_EX_ERR_WIRE = 2,      /* malformed protobuf wire data */
  TXF_EX_ERR_KIND = 4,      /* a feature holds a list of another kind than requested */
  TXF_EX_ERR_MULTI = 8,     /* more than one value in a cell */
  TXF_EX_ERR_CRC = 16       /* data CRC mismatch (verify_crc) */
};
typedef struct txf_example_column {
  int32_t kind;     /* TXF_F32 | TXF_I64 | TXF_EX_BYTES */
  int32_t name_off;
  int32_t name_len;
  int32_t reserved;
  void* values;     /* float[n] | int64[n] | uint32 position of the bytes in the stream[n] (strings) */
  int32_t* lens;    /* strings: byte length[n]; else NULL */
  void* valid;      /* bitmap, ceil(n / 32) * 4 bytes, 4-byte aligned */
} txf_example_column;
TXF_API int64_t txf_tfrecord_block_bytes(void);
TXF_API int txf_tfrecord_count_candidates(const uint8_t* data_dev, int64_t nbytes, uint32_t* block_counts_dev, txf_stream_t stream);
TXF_API int txf_tfrecord_candidates(const uint8_t* data_dev, int64_t nbytes, const int64_t* block_base_dev, int64_t* cand_dev,
                                    txf_stream_t stream);
TXF_API int txf_tfrecord_link(const uint8_t* data_dev, const int64_t* cand_dev, int64_t ncand, uint8_t* mark_dev, txf_stream_t stream);
TXF_API int txf_example_decode(const txf_example_column* feats, int nfeat, const uint8_t* names_dev, const uint8_t* data_dev,
                               const int64_t* record_start_dev, int64_t n, int verify_crc, uint64_t* status_dev,
                               txf_stream_t stream);
TXF_API int txf_gather_bytes(const uint8_t* data_dev, const uint32_t* pos_dev, const int32_t* offsets_dev, int64_t n,
                             uint8_t* pool_dev, txf_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Sibling analyzers / mappers of the workshop notebooks.
 *   min/max analyzer + tft.scale_to_0_1 / scale_by_min_max:
 *     notebooks/03_TensorFlow_Transform.ipynb:148, 03a_TensorFlow_Transform_Advanced.ipynb:323,333
 *   fp32 arithmetic between columns and analyzer scalars (x - tft.mean(x), a * b):
 *     notebooks/03_TensorFlow_Transform.ipynb:147,150
 * min/max state: float[ncols][2] = (min, max) of the FILLED fp32 values (absent cell = 0; int64
 * cells are converted to fp32 first; NaN cells are skipped).  txf_minmax_begin sets (+inf, -inf);
 * update folds one chunk into the state; merge folds `nstates` states laid out [state][col][2].
 * ------------------------------------------------------------------------------------------ */
TXF_API size_t txf_minmax_workspace_bytes(int ncols);
TXF_API int txf_minmax_begin(float* state_dev, int ncols, txf_stream_t stream);
TXF_API int txf_minmax_update(const void* const* values, const uint8_t* const* valid, const int32_t* dtype, int ncols,
                              int64_t n, float* state_dev, void* workspace_dev, txf_stream_t stream);
TXF_API int txf_minmax_merge(const float* states_dev, int nstates, int ncols, float* state_out_dev, txf_stream_t stream);

/* One operand of an elementwise fp32 op: a column (fp32 or int64, fill-missing applied, int64
 * converted to fp32) when col != NULL, else the device scalar *scalar. */
typedef struct txf_operand {
  const void* col;
  const void* valid;
  const float* scalar;
  int32_t dtype; /* TXF_F32 | TXF_I64 (columns only) */
  int32_t reserved;
} txf_operand;
enum { TXF_ARITH_ADD = 0, TXF_ARITH_SUB = 1, TXF_ARITH_MUL = 2, TXF_ARITH_DIV = 3 };
/* out[i] = a[i] (op) b[i]: one correctly rounded fp32 operation per row. */
TXF_API int txf_map_binary(int opcode, const txf_operand* a, const txf_operand* b, float* out_dev, int64_t n,
                           txf_stream_t stream);
/* tf.to_int64 / tf.to_float of a filled column (kubeflow-pipelines/taxi/preprocessing.py:92):
 * to_dtype TXF_I64 truncates toward zero, TXF_F32 rounds to nearest even. */
TXF_API int txf_map_cast(const txf_operand* x, int to_dtype, void* out_dev, int64_t n, txf_stream_t stream);
/* tf.as_string on an int64 feature (kubeflow-pipelines/taxi/preprocessing.py:82), two steps so the
 * caller owns the prefix sum: lengths (absent cell = "0"), then digits at offsets[i]..offsets[i+1]. */
TXF_API int txf_i64_format_lengths(const int64_t* values_dev, const uint8_t* valid_dev, int64_t n,
                                   int32_t* len_out_dev, txf_stream_t stream);
TXF_API int txf_i64_format_write(const int64_t* values_dev, const uint8_t* valid_dev, int64_t n,
                                 const int32_t* offsets_dev, uint8_t* pool_out_dev, txf_stream_t stream);
/* out[i] = (min < max ? (x[i] - min) / (max - min) : 0.5) * (out_max - out_min) + out_min. */
TXF_API int txf_map_scale_minmax(const txf_operand* x, const float* minmax_dev, float out_min, float out_max,
                                 float* out_dev, int64_t n, txf_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Test hook: Fingerprint64 of every string of a column (TF StringToHashBucketFast's hash).
 * ------------------------------------------------------------------------------------------ */
TXF_API int txf_fingerprint64_batch(const int32_t* offsets_dev, const uint8_t* pool_dev, int64_t n, uint64_t* out_dev,
                            txf_stream_t stream);

/* Test hook: keep only the low `bits` (1..119) hash bits of the fingerprints of terms longer than
 * 15 bytes, on the current device, so that fingerprints collide at will; results must not change,
 * because every fingerprint match in the global table, the lookup and the decode table is confirmed
 * on the bytes (the count pass bypasses its shared-memory cache for long terms while bits < 119).
 * 119 restores the default. */
TXF_API int txf_debug_set_identity_bits(int bits);

/* number of kernels launched by this library in the calling process (bench.py gpu_launches) */
TXF_API int64_t txf_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TAXI_XF_H_ */
