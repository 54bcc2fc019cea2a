/* openbnsl_b200.h -- C ABI of the B200 (sm_100a) contingency-counting engine.
 *
 * This is the drop-in boundary for OpenBNSL's data-parallel hot path.  The
 * reference has no FFI layer of its own: its only accelerator slot is
 * src/template_cuda.cu (an unbound matmul demo, include/template.h:6-11), and
 * its hot path is the per-tuple pair
 *     ContingencyTable<true>(var_ids, df)          include/base/contingency_table.h:34-56
 *  -> calculate_local_score<double,true>(...)      include/score/local_score.h:22-62
 *  |  citest<true>(...)                            include/citest/citest.h:41-148
 * called from exhaustive_search.cpp:55-72 and pc.cpp:38-39 one tuple at a
 * time.  The functions below are what those call sites bind to instead: the
 * same computations over BATCHES of tuples.  Plain pointers and sizes only;
 * no C++ or torch types.  See INTEGRATION.md for the reference-side stubs.
 *
 * Conventions
 *  - Every function returns an int status (OBN_OK == 0).  obn_last_error()
 *    returns a thread-local message for the last failing call.
 *  - All pointer arguments are HOST pointers unless the name ends in `_dev`.
 *    The library copies; it never keeps a host pointer after returning.
 *  - Calls are synchronous: outputs are valid on return.
 *  - A ctx is bound to ONE CUDA device (one process per GPU) and must be used
 *    by one host thread at a time.  A multi-GPU job gives every rank's ctx a
 *    communicator (obn_comm_init, NCCL); the sharded calls (obn_pc_round with
 *    shard != 0, obn_comm_allgather_inplace) are then COLLECTIVE: every rank
 *    makes the same call with the same batch.
 *  - The library reads no environment variables; tuning goes through
 *    obn_ctx_set_option.
 *  - There is no CPU fallback.  Without a usable CUDA device obn_ctx_create
 *    fails with OBN_ERR_CUDA and nothing else can be called.
 */
#ifndef OPENBNSL_B200_H
#define OPENBNSL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define OBN_API __attribute__((visibility("default")))
#else
#define OBN_API
#endif

enum {
  OBN_OK = 0,
  OBN_ERR_INVALID = 1,     /* bad argument (maps to std::invalid_argument) */
  OBN_ERR_CUDA = 2,        /* CUDA runtime error / no device               */
  OBN_ERR_OOM = 3,         /* device or host allocation failed             */
  OBN_ERR_UNSUPPORTED = 4, /* e.g. a dense table with more than 2^31 cells */
  OBN_ERR_RANGE = 5,       /* index out of range (maps to std::out_of_range) */
  OBN_ERR_COMM = 6         /* NCCL error / no communicator                 */
};

enum { OBN_CI_CHI2 = 0, OBN_CI_G2 = 1 };

typedef struct obn_ctx obn_ctx;
typedef struct obn_dataset obn_dataset;

OBN_API const char* obn_last_error(void);
OBN_API const char* obn_version(void);

/* ---- context ------------------------------------------------------------ */
OBN_API int obn_ctx_create(int device, obn_ctx** out);
OBN_API void obn_ctx_destroy(obn_ctx* ctx);
/* The CUDA stream (cudaStream_t) all work of this ctx is issued on. */
OBN_API void* obn_ctx_stream(obn_ctx* ctx);
/* Tuning knobs for A/B runs and tests (defaults are what the published numbers use):
 *   "threads_per_sm" (768), "force_threads" (0), "force_splits" (0), "streams" (4)   counting-kernel launch policy
 *   "pack4" (1)                          keep a second copy of the matrix with two rows per byte for the variables of <= 16 states
 *                                        and read that one whenever every column of a table has it (half the L2 traffic)
 *   "rep" (1), "chunk" (1), "min_waves" (0)   replicated shared-memory tables on / off; tables beyond shared memory counted
 *                                        chunk-wise in shared memory instead of by L2 reductions; row-split small launches to this many waves
 *   "largest_first" (0)                  launch order of a batch's classes (A/B)
 *   "min_class" (0), "lut_max_mb" (2048)  smallest table-size class used (fewer, larger launches per batch); cap of the
 *                                        exact-order lgamma tables kept on the device (evicted and rebuilt beyond it)
 *   "lattice_budget_gb" (24), "lattice_max_l" (20), "lattice_bottom" (1), "lattice_reorder" (1),
 *   "lattice_graph" (1), "lattice_lanes" (2), "lattice_no_lut" (0), "debug" (0)      marginalisation lattice
 * Unknown names fail with OBN_ERR_INVALID. */
OBN_API int obn_ctx_set_option(obn_ctx* ctx, const char* name, double value);

/* ---- communicator (multi-GPU jobs, one process per GPU) --------------------
 * NCCL over NVLink, owned by the ctx, bound at run time (a process that already
 * holds libnccl.so.2 shares it).  Rank 0 calls obn_comm_unique_id and hands the
 * OBN_COMM_ID_BYTES bytes to every rank by any host-side means; every rank then
 * calls obn_comm_init (collective).  The exchanges of the path run on device
 * buffers of the library: all-reduce(min) of the first passing candidate per
 * edge inside obn_pc_round, in-place all-gathers of the exhaustive-search
 * buffers through obn_comm_allgather_inplace. */
#define OBN_COMM_ID_BYTES 128
OBN_API int obn_comm_unique_id(uint8_t* id_out);
OBN_API int obn_comm_init(obn_ctx* ctx, const uint8_t* id_bytes, int rank, int world);
OBN_API void obn_comm_destroy(obn_ctx* ctx);
OBN_API int obn_comm_rank(const obn_ctx* ctx);
OBN_API int obn_comm_world(const obn_ctx* ctx);
/* dev: DEVICE buffer of world * bytes_per_rank bytes, rank r contributes slice r; synchronous. */
OBN_API int obn_comm_allgather_inplace(obn_ctx* ctx, void* dev, size_t bytes_per_rank);
/* All-gather of small HOST payloads (recv: world * bytes_per_rank bytes, rank order), staged through the device. */
OBN_API int obn_comm_allgather_host(obn_ctx* ctx, const void* send, void* recv, size_t bytes_per_rank);
/* In-place elementwise MIN over ranks of a HOST int32 array (staged through the device). */
OBN_API int obn_comm_allreduce_min_i32(obn_ctx* ctx, int32_t* host, size_t count);

/* ---- dataset: replaces DataframeWrapper's two vector<vector<uint8_t>>
 * (include/base/dataframe_wrapper.h:42-45) by ONE padded column-major uint8
 * matrix in HBM: column v, row r lives at v * stride + r, stride a multiple of
 * 128 bytes.  `cards[v]` is DataframeWrapper::num_of_values[v] (<= 256). */
OBN_API int obn_dataset_upload(obn_ctx* ctx, const uint8_t* col_major,
                               size_t n_vars, size_t n_rows, size_t row_stride,
                               const uint32_t* cards, obn_dataset** out);
OBN_API void obn_dataset_destroy(obn_dataset* ds);
/* Page-lock / unlock a caller-owned host buffer (e.g. the column-major matrix a
 * DataframeWrapper keeps, dataframe_wrapper.h:42) so that obn_dataset_upload
 * runs at full PCIe rate.  Optional. */
OBN_API int obn_host_pin(obn_ctx* ctx, const void* p, size_t bytes);
OBN_API int obn_host_unpin(obn_ctx* ctx, const void* p);
OBN_API size_t obn_dataset_n_vars(const obn_dataset* ds);
OBN_API size_t obn_dataset_n_rows(const obn_dataset* ds);

/* ---- batched hot path ---------------------------------------------------
 * A batch is a CSR list of tuples: tuple t covers the variables
 * var_ids[offsets[t] .. offsets[t+1]), ASCENDING (contingency_table.h:37-38:
 * unsorted -> OBN_ERR_INVALID "var_ids must be sorted"). */

/* Replaces the ContingencyTable ctor (contingency_table.h:34-56).  Dense int32
 * counts per tuple in REFERENCE KEY ORDER (first variable most significant,
 * :76-85), tuple t written at counts_out + out_offsets[t]; zero cells are the
 * keys the reference's std::map would not contain. */
OBN_API int obn_count_tables(obn_ctx* ctx, const obn_dataset* ds,
                             const uint32_t* var_ids, const uint64_t* offsets,
                             size_t n_tuples, int32_t* counts_out,
                             const uint64_t* out_offsets);

/* ONE table, no allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code
 * under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * The same counts as obn_count_tables for one tuple -- dense int32, reference
 * key order -- but every buffer is the CALLER's device memory and the global
 * scratch table the kernel merges its CTAs into is a file-scope __device__
 * array of the library (256 KiB + 256 bytes, held by every GPU that loads the
 * library until the process exits), so the call issues no cudaMalloc / cudaFree:
 * one cudaMemsetAsync of that array and one kernel launch on `stream`
 * (a cudaStream_t; NULL = the default stream).  Needs no ctx.
 *   data_dev        column-major uint8 matrix on `device`: column v, row r at v * row_stride + r
 *   cols, cards     HOST arrays: the n_cols <= 16 columns of the table, ASCENDING, and their cardinalities
 *   counts_out_dev  DEVICE buffer of n_cells = prod(cards) int32; n_cells > OBN_STATIC_MAX_CELLS
 *                   fails with OBN_ERR_UNSUPPORTED (use obn_count_tables)
 * ASYNCHRONOUS: the counts are valid once `stream` has reached this point.  All
 * calls on a device share the one array (cleared at the start of every call):
 * calls that overlap on the GPU -- different streams, no dependency -- race.  The
 * first call per device resolves the array's address; make it outside capture. */
#define OBN_STATIC_MAX_CELLS 65536
OBN_API int obn_count_table_static(int device, void* stream, const uint8_t* data_dev,
                                   size_t n_rows, size_t row_stride,
                                   const uint32_t* cols, const uint32_t* cards, size_t n_cols,
                                   int32_t* counts_out_dev, size_t n_cells);

/* Replaces ContingencyTable + calculate_local_score (local_score.h:22-62) for
 * BDeu: score of child[t] given the other variables of tuple t.
 * exact_order != 0 reproduces the reference's floating-point term order with
 * host-libm lgamma tables (bit-identical scores; slower).  */
OBN_API int obn_bdeu_scores(obn_ctx* ctx, const obn_dataset* ds,
                            const uint32_t* var_ids, const uint64_t* offsets,
                            size_t n_tuples, const uint32_t* child, double ess,
                            int exact_order, double* scores_out);

/* The reference's own sharing in exhaustive_search.cpp:55-72: ONE table per
 * subset W serves |W| scores.  scores_out[offsets[t] + i] is the score of
 * child var_ids[offsets[t]+i] given W \ child. */
OBN_API int obn_bdeu_subset_scores(obn_ctx* ctx, const obn_dataset* ds,
                                   const uint32_t* var_ids,
                                   const uint64_t* offsets, size_t n_tuples,
                                   double ess, int exact_order,
                                   double* scores_out);

/* Replaces ContingencyTable + citest (citest.h:41-148) for chi-square / G2:
 * tuple t must contain x[t] and y[t]; all its other variables condition.
 * Outputs the statistic, the degrees of freedom and p = pchisq(stat, dof)
 * (citest.h:17-21); the decision p >= level stays with the caller.  Any of
 * the three output pointers may be NULL.
 * Tests of one batch that are asked over the SAME variable set share one
 * table on the device (the reference's table does not depend on which two of
 * its variables are tested, citest.h:63-70; a PC frontier asks every set
 * |Z|+1 times or more): it is counted once and read once per test. */
OBN_API int obn_ci_tests(obn_ctx* ctx, const obn_dataset* ds,
                         const uint32_t* var_ids, const uint64_t* offsets,
                         size_t n_tuples, const uint32_t* x, const uint32_t* y,
                         int kind, double* stat_out, uint64_t* dof_out,
                         double* p_out);

/* One round of the PC skeleton search (src/structure_learning/pc.cpp:27-47):
 * the batch holds candidate tests of n_slots edges, test t belonging to edge
 * slot_of_test[t], in the reference's enumeration order within every edge.
 * The decision p >= level (citest.h:147) is taken on the device and
 * first_out[s] = smallest t of edge s that passed -- the candidate the serial
 * reference stops at -- or INT32_MAX when none did.  Only that array leaves
 * the device.  shard != 0 and a communicator on the ctx: COLLECTIVE -- every
 * rank passes the same batch, evaluates its share of the tables and the
 * per-edge minima are merged by ncclAllReduce(min) on the device buffer. */
OBN_API int obn_pc_round(obn_ctx* ctx, const obn_dataset* ds,
                         const uint32_t* var_ids, const uint64_t* offsets, size_t n_tests,
                         const uint32_t* x, const uint32_t* y, const uint32_t* slot_of_test,
                         size_t n_slots, int kind, double level, int shard, int32_t* first_out);

/* obn_ci_tests / obn_pc_round in two halves, so that the caller's host thread can prepare its next
 * batch while the device works on this one.  submit validates, packs, uploads and enqueues everything
 * and returns; the input arrays may be reused as soon as it has returned.  collect waits for the OLDEST
 * submitted batch and writes its results (any pointer may be NULL; p_out only if want_p was set).
 * Up to TWO batches may be in flight per ctx (they run in submission order); a third submit, and any
 * other batch call on the ctx while one is in flight, returns OBN_ERR_INVALID.  collect without a
 * submitted batch (or after submitting zero tuples) is a no-op that leaves the output arrays
 * untouched: a caller of obn_pc_round_collect pre-fills first_out with INT32_MAX.
 * Chained PC rounds (chain != 0): the batch continues the PC round submitted just before it -- the same
 * n_slots edges in the same numbering, that round still in flight or the last one collected.  Tests of
 * edges the previous round decided (first != none) are skipped on the device: not counted, not
 * evaluated, never reported -- what the serial reference does by stopping at an edge's first
 * independent candidate (pc.cpp:40-42).  The caller can therefore enqueue round r+1 of a depth before
 * it has seen round r's result, and the device never idles between the two.  The profile counters
 * account a chained round net of what was skipped. */
OBN_API int obn_ci_tests_submit(obn_ctx* ctx, const obn_dataset* ds,
                                const uint32_t* var_ids, const uint64_t* offsets,
                                size_t n_tuples, const uint32_t* x, const uint32_t* y,
                                int kind, int want_p);
OBN_API int obn_ci_tests_collect(obn_ctx* ctx, double* stat_out, uint64_t* dof_out, double* p_out);
OBN_API int obn_pc_round_submit(obn_ctx* ctx, const obn_dataset* ds,
                                const uint32_t* var_ids, const uint64_t* offsets, size_t n_tests,
                                const uint32_t* x, const uint32_t* y, const uint32_t* slot_of_test,
                                size_t n_slots, int kind, double level, int shard, int chain);
OBN_API int obn_pc_round_collect(obn_ctx* ctx, int32_t* first_out);

/* Diagnostics (host only, needs no device): the table sharing obn_ci_tests / obn_pc_round use.
 * Groups in CSR form (group_var_ids needs room for offsets[n_tests] ids, group_offsets for
 * n_tests + 1), the group of every test and its digits (px | py << 8: positions of x < y in the
 * group's ascending variable list).  rank / world: the share of a sharded round that rank evaluates
 * (0 / 1: everything) -- every variable set goes to exactly one rank, group_of_test is 0xffffffff
 * for the tests of the other ranks.  Any output pointer but n_groups may be NULL. */
OBN_API int obn_ci_group_plan(const uint32_t* cards, size_t n_vars, const uint32_t* var_ids,
                              const uint64_t* offsets, size_t n_tests, const uint32_t* x, const uint32_t* y,
                              int rank, int world,
                              uint32_t* group_var_ids, uint64_t* group_offsets, uint32_t* group_of_test,
                              uint32_t* digits_of_test, size_t* n_groups);

/* ---- exhaustive search on the device --------------------------------------
 * Level-synchronous Silander-Myllymaki DP with the reference's tie-breaking
 * (src/structure_learning/exhaustive_search.cpp:47-148).  The DP tables
 * (best_ls_tbl: n * 2^n (score, parent set); best_gs_tbl / best_ch_tbl: 2^n,
 * :33-40) live in HBM for the whole search.  Per lattice level the caller
 *   1. obn_es_level_begin   hands over the level's subsets (bit masks, all of
 *                           size k, in the reference's Gosper order, :51,:129);
 *                           masks may be NULL for a level that needs no scores
 *                           (k = 0, k > max_parents + 1, or after obn_es_lattice):
 *                           n_subsets must then be C(n, k) and the DP walks
 *                           level lists built on the device,
 *   2. obn_es_level_scores  has the |W| local scores of the subsets
 *                           [begin, end) computed (one table per subset, :55-72)
 *                           into the level buffer -- a multi-GPU caller gives
 *                           each rank one range and all-gathers the buffer
 *                           (obn_es_level_buffer) in place,
 *   3. obn_es_level_dp      runs :59-116 for every subset of the level.
 * obn_es_finish backtracks (:133-148): parent_masks[v] = bit mask of v's parents. */
typedef struct obn_es obn_es;
OBN_API int obn_es_begin(obn_ctx* ctx, const obn_dataset* ds, double ess, size_t max_parents,
                         int exact_order, obn_es** out);
OBN_API int obn_es_level_begin(obn_es* es, const uint64_t* masks, size_t n_subsets, int k,
                               size_t buffer_subsets);
OBN_API int obn_es_level_scores(obn_es* es, size_t begin, size_t end);
OBN_API int obn_es_level_buffer(obn_es* es, void** scores_dev, size_t* n_doubles);
OBN_API int obn_es_level_dp(obn_es* es);
/* Multi-GPU, after obn_es_lattice: the DP sharded by child.  The best-parent-set entry of (W, c)
 * (:75-87) depends only on entries of the same child c, so rank r of `world` keeps the DP columns
 * of children [n r / world, n (r+1) / world) and never needs the others.  Only the sink choice
 * (:89-94) needs every child: obn_es_level_dp_part runs the level for this rank's children and
 * writes (best sum, child) per subset into slice `rank` of a device buffer of world * bytes_per_rank
 * bytes, the caller all-gathers that buffer in place, obn_es_level_dp_commit takes the reference's
 * choice (largest sum, earliest child) and writes best_gs / best_ch on every rank.  Once a level
 * has been done this way the rest of the search must be too; obn_es_finish then fills only this
 * rank's children in parent_masks (zero elsewhere: OR the ranks' arrays). */
OBN_API int obn_es_level_dp_part(obn_es* es, int rank, int world, void** parts_dev, size_t* bytes_per_rank);
/* The same level in ONE call when the ctx owns the communicator (obn_comm_init): this rank's children, then
 * max over ranks of the sum and min over ranks of the child among those that reach it (two all-reduces of 8
 * and 4 bytes per subset instead of gathering 16 bytes per subset and rank), then the commit -- all enqueued
 * on the ctx's stream, no host synchronisation per level.  Collective. */
OBN_API int obn_es_level_dp_sharded(obn_es* es);
OBN_API int obn_es_level_dp_commit(obn_es* es, int world);
/* Fast mode for deep searches: instead of steps 2 of every level, compute
 *   F(S) = sum over the cells of S of lgamma(N_c + ess/cells(S)) - lgamma(ess/cells(S))
 * for ALL subsets at once by a marginalisation lattice over the joint table
 * (one pass over the data; see csrc/obn_lattice.cuh); the DP then uses
 * score(child | pa) = F(pa + child) - F(pa)  (local_score.h:38-61 regrouped; 1e-9
 * relative, not exact-order).  Rank r of `world` computes its share of the
 * blocks; obn_es_f_buffer exposes F for the in-place all-gather
 * (world * doubles_per_rank doubles, rank r owns slice r). */
OBN_API int obn_es_lattice(obn_es* es, int rank, int world);
OBN_API int obn_es_f_buffer(obn_es* es, void** f_dev, size_t* doubles_per_rank);
/* Diagnostics (host only, needs no device): the per-thread program of the lattice's bottom-tier
 * kernel (csrc/obn_lattice.cuh).  geometry[6] = {bottom variables, threads, quads per chunk, quad
 * buffer size, upper variables, phase-2 quads per thread}; words[upper variables][threads]: source
 * quad | log2(distance to the second source) << 10 | destination quad << 13 for levels
 * upper-1 .. 0; patterns[slots * threads]: upper pattern of quad q (0xffff = none).  words /
 * patterns may be NULL to query the geometry only.  The CPU test-suite replays the program in
 * numpy against directly computed marginals. */
OBN_API int obn_lattice_bottom_program(uint32_t* words, size_t n_words, uint16_t* patterns, size_t n_patterns,
                                       int* geometry);

OBN_API int obn_es_finish(obn_es* es, uint64_t* parent_masks, double* best_score);
OBN_API void obn_es_end(obn_es* es);

/* ---- instrumentation -----------------------------------------------------
 * Device time (CUDA events on the ctx stream) and launch count of the
 * counting kernels since the last reset; used by bench.py for the roofline. */
typedef struct obn_profile {
  double count_kernel_ms;    /* sum of counting-kernel durations            */
  uint64_t count_launches;   /* number of counting-kernel launches          */
  uint64_t total_launches;   /* all kernels launched by this ctx            */
  double algorithmic_bytes;  /* sum over tuples (CI: tests) of k * n_rows   */
  uint64_t tuples;           /* tuples (CI: tests) evaluated by this ctx    */
  double h2d_bytes;          /* bytes copied host -> device (batches + datasets) */
  double d2h_bytes;          /* bytes copied device -> host (results)       */
  double host_prep_ms;       /* host: validate + classify + pack the batch  */
  double host_wait_ms;       /* host: enqueue copies/launches + wait for the stream */
  double scanned_bytes;      /* sum over TABLES of k * n_rows: what the kernels read (tests sharing a table count once) */
  double increments;         /* sum over tables of n_rows: shared-memory / L2 increments applied */
  uint64_t tables;           /* contingency tables counted                  */
  uint64_t collectives;      /* NCCL collectives issued                     */
  double collective_ms;      /* host wall time of the synchronous collectives (all-gathers) */
  uint64_t exact_fallbacks;  /* exact_order == 2 batches that fell back to fast mode (lgamma tables over the cap) */
} obn_profile;
OBN_API int obn_profile_enable(obn_ctx* ctx, int on);
OBN_API int obn_profile_read(obn_ctx* ctx, obn_profile* out, int reset);

#ifdef __cplusplus
}
#endif
#endif /* OPENBNSL_B200_H */
