/*
 * mcb200.h -- C ABI of libmcb200.so: the B200 (sm_100a) implementation of the metacells per-pile
 * k-NN similarity-graph hot path.  Plain pointers and sizes only.
 *
 * Every entry point replaces one function of the reference's pybind11 module
 * `metacells.extensions` (reference file:line cited per function; paths relative to
 * tanaylab/metacells v0.9.5).  Two forms exist for each operation:
 *
 *   mcb_host_<op>(...)          pointers are HOST memory (numpy buffers).  The call copies to the
 *                               GPU, runs the kernels, copies back and returns when done: the exact
 *                               contract of the reference function it replaces.
 *   mcb_dev_<op>(..., stream)   pointers are DEVICE memory; work is enqueued on `stream`
 *                               (a cudaStream_t passed as void*).  Used by the device-resident
 *                               pipeline so a pile never leaves HBM between stages.
 *
 * All functions return 0 on success; otherwise a non-zero code, with a message available from
 * mcb_last_error() (thread local).  Where the reference calls std::exit(1) on a failed assertion
 * (extensions.h:59-72) this library returns MCB_ERR_ARG instead.
 *
 * Type suffixes follow the reference's symbol names: f32 = float32_t, f64 = float64_t,
 * i32 = int32_t.  Only the instantiations the hot path uses are provided.
 */
#ifndef MCB200_H
#define MCB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------------------- */

const char* mcb_last_error(void);
int mcb_version(void);
int mcb_device_count(void);
/* Select the CUDA device of the calling thread; fails unless it is compute capability 10.x. */
int mcb_set_device(int device);
/* The device of the calling thread.  The current device is per host thread and a new thread starts on device 0, so
 * worker threads of a process that drives another GPU call mcb_set_device before anything else. */
int mcb_current_device(int* device);
int mcb_synchronize(void* stream);
/* Streaming multiprocessors of the calling thread's device: how many one-CTA optimisations (mcb_host_choose_seeds,
 * mcb_host_optimize_partitions, each on its caller thread's own stream) run side by side. */
int mcb_device_sm_count(int* count);
/* Grows the device's scratch pool (stream-ordered allocator, kept between calls) to at least `bytes` free now. */
int mcb_reserve_scratch(long long bytes);
/* Number of kernels this library has launched in the calling process so far. */
long long mcb_launch_count(void);
/* Optional per-kernel timing: when enabled every launch is bracketed by CUDA events on its own
 * stream; mcb_profile_read sums the recorded milliseconds of one kernel name (synchronises). */
int mcb_profile_enable(int enable);
int mcb_profile_read(const char* name, double* total_ms, long long* count);
int mcb_profile_names(char* buffer, long long size);
int mcb_profile_reset(void);

/* ---- downsample.cpp ------------------------------------------------------------------------- */

/* downsample_compressed<float32,int32,float32>  (downsample.cpp:171-192).
 * data/out: nnz values; indptr: bands+1; samples: bands.  Rows with samples < 0 or no elements
 * leave `out` untouched, as the reference does (:86-88). */
int mcb_host_downsample_compressed_f32_i32_f32(const float* data, const int32_t* indptr, float* out,
                                               const int32_t* samples, int64_t bands, uint64_t random_seed);
/* max_band_size: an upper bound of indptr[b+1]-indptr[b] (the host form computes it itself). */
int mcb_dev_downsample_compressed_f32_i32_f32(const float* data, const int32_t* indptr, float* out,
                                              const int32_t* samples, int64_t bands, int64_t max_band_size,
                                              uint64_t random_seed, void* stream);

/* downsample_dense<float32,float32>  (downsample.cpp:131-151): rows x columns, C contiguous. */
int mcb_host_downsample_dense_f32_f32(const float* input, float* output, const int32_t* samples, int64_t rows,
                                      int64_t columns, uint64_t random_seed);
int mcb_dev_downsample_dense_f32_f32(const float* input, float* output, const int32_t* samples, int64_t rows,
                                     int64_t columns, uint64_t random_seed, void* stream);

/* downsample_array<float32,float32>  (downsample.cpp:116-128): the seed is used as given. */
int mcb_host_downsample_array_f32_f32(const float* input, float* output, int64_t size, int32_t samples,
                                      uint64_t random_seed);

/* ---- correlate.cpp --------------------------------------------------------------------------- */

/* correlate_dense<F>  (correlate.cpp:350-431): all-pairs Pearson r between the rows of
 * input[rows, columns] (row stride in elements >= columns).  output: float32 [rows, rows],
 * symmetric, unit diagonal, 0 for constant rows, clamped to [-1, 1]; NaN results are stored as 0
 * (the reference's caller does that at computation.py:465).
 * float32 input runs on the tensor cores (tcgen05, 3xTF32 split); float64 input in float64 FMA. */
int mcb_host_correlate_dense_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, float* output);
int mcb_host_correlate_dense_f64(const double* input, int64_t rows, int64_t columns, int64_t row_stride, float* output);
/* Device form; absolute != 0 stores |r| (fuses tools/similarity.py:214-215 "abs_pearson"). */
int mcb_dev_correlate_dense_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, float* output,
                                int64_t out_stride, int absolute, void* stream);

/* cross_correlate_dense<F>  (correlate.cpp:433-494): output float32 [first_rows, second_rows]. */
int mcb_host_cross_correlate_dense_f32(const float* first, int64_t first_rows, int64_t first_stride, const float* second,
                                       int64_t second_rows, int64_t second_stride, int64_t columns, float* output);
int mcb_host_cross_correlate_dense_f64(const double* first, int64_t first_rows, int64_t first_stride, const double* second,
                                       int64_t second_rows, int64_t second_stride, int64_t columns, float* output);
int mcb_dev_cross_correlate_dense_f32(const float* first, int64_t first_rows, const float* second, int64_t second_rows,
                                      int64_t columns, float* output, int64_t out_stride, void* stream);

/* One block of rows of the similarity matrix of `all_rows` (tools/similarity.py:210-215 + computation.py:465-466):
 * output[i, j] = r(rows_block[i, :], all_rows[j, :]) (|r| when absolute != 0), NaN -> 0, and 1 where
 * j == first_row + i.  rows_block is all_rows + first_row * columns on the rank that owns the block; used by the
 * row-block sharded similarity of an oversized pile.  Device pointers. */
int mcb_dev_similarity_block_f32(const float* rows_block, int64_t block_rows, int64_t first_row, const float* all_rows,
                                 int64_t all_count, int64_t columns, float* output, int64_t out_stride, int absolute,
                                 void* stream);

/* pairs_correlate_dense<F>  (correlate.cpp:496-527): output[i] = r(first[i, :], second[i, :]). */
int mcb_host_pairs_correlate_dense_f32(const float* first, int64_t first_stride, const float* second, int64_t second_stride,
                                       int64_t rows, int64_t columns, float* output);
int mcb_host_pairs_correlate_dense_f64(const double* first, int64_t first_stride, const double* second, int64_t second_stride,
                                       int64_t rows, int64_t columns, float* output);
int mcb_dev_pairs_correlate_dense_f32(const float* first, const float* second, int64_t rows, int64_t columns, float* output,
                                      void* stream);

/* ---- rank.cpp ----------------------------------------------------------------------------------- */

/* rank_matrix<float32>  (rank.cpp:38-87): in place, every row becomes its 1-based ranks
 * (ascending != 0: smallest value gets rank 1).  Equal values rank by position (lower first). */
int mcb_host_rank_matrix_f32(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending);
int mcb_dev_rank_matrix_f32(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, void* stream);

/* mcb_dev_rank_matrix_f32 for callers that cannot allocate.  No counterpart in the reference: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same
 * kernels and the same ranks as mcb_dev_rank_matrix_f32; the one piece of scratch (the list of rows handed from the
 * bucket sort to the merge sort) lives in a file-scope device array of the library, so the call makes no cudaMalloc /
 * cudaFree / pool allocation.  Limits: rows <= MCB_RANK_MATRIX_NOALLOC_MAX_ROWS, columns <=
 * MCB_RANK_MATRIX_NOALLOC_MAX_COLUMNS; beyond them the call fails with MCB_ERR_ARG and a message.  Every GPU that loads
 * the library holds the array (257 KiB) until exit.  All calls share it: calls that overlap on one GPU (two streams,
 * two host threads) race; calls on one stream, one after the other, are fine (every call clears what it needs). */
#define MCB_RANK_MATRIX_NOALLOC_MAX_ROWS 65536
#define MCB_RANK_MATRIX_NOALLOC_MAX_COLUMNS 24576
int mcb_dev_rank_matrix_noalloc_f32(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, void* stream);

/* rank_rows<float32>  (rank.cpp:5-36): output[row] = value of 0-based `rank` in ascending order. */
int mcb_host_rank_rows_f32(const float* matrix, int64_t rows, int64_t columns, int64_t row_stride, float* output, int64_t rank);
int mcb_dev_rank_rows_f32(const float* matrix, int64_t rows, int64_t columns, int64_t row_stride, float* output, int64_t rank,
                          void* stream);

/* ---- top_per.cpp -------------------------------------------------------------------------------- */

/* collect_top<float32>  (top_per.cpp:5-92): per row the `degree` largest values; out_indices
 * (ascending per row) and out_data have rows * degree entries; ranks != 0 stores the 1-based
 * ascending rank among the kept values instead of the values. */
int mcb_host_collect_top_f32(int64_t degree, const float* similarity, int64_t rows, int64_t columns, int64_t row_stride,
                             int32_t* out_indices, float* out_data, int ranks);
int mcb_dev_collect_top_f32(int64_t degree, const float* similarity, int64_t rows, int64_t columns, int64_t row_stride,
                            int32_t* out_indices, float* out_data, int ranks, void* stream);

/* ---- prune_per.cpp ------------------------------------------------------------------------------ */

/* collect_pruned<float32,int32,int32>  (prune_per.cpp:56-105): out_indptr has bands + 1 entries;
 * out_data / out_indices need room for sum(min(band size, degree)) entries. */
int mcb_host_collect_pruned_f32_i32_i32(int64_t degree, const float* data, const int32_t* indices, const int32_t* indptr,
                                        int64_t bands, float* out_data, int32_t* out_indices, int32_t* out_indptr);
int mcb_dev_collect_pruned_f32_i32_i32(int64_t degree, const float* data, const int32_t* indices, const int32_t* indptr,
                                       int64_t bands, float* out_data, int32_t* out_indices, int32_t* out_indptr, void* stream);

/* ---- relayout.cpp ------------------------------------------------------------------------------- */

/* collect_compressed<float32,int32,int32>  (relayout.cpp:73-108).  out_indptr_shifted has out_bands
 * entries: on entry the start of every output band, on return its end (the caller passes
 * output_indptr[1:], computation.py:268-292).  The order inside an output band is unspecified,
 * as in the reference; follow with sort_compressed_indices. */
int mcb_host_collect_compressed_f32_i32_i32(const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands,
                                            float* out_data, int32_t* out_indices, int32_t* out_indptr_shifted,
                                            int64_t out_bands);
/* sort_compressed_indices<float32,int32,int32>  (relayout.cpp:152-167): in place. */
int mcb_host_sort_compressed_indices_f32_i32_i32(float* data, int32_t* indices, const int32_t* indptr, int64_t bands);

/* ---- tools/knn_graph.py:203-407 (numpy/scipy glue of the reference, fused on the device) -------- */

/* A device-resident CSR graph: `weights` (obs_outgoing_weights) and `ranks` (obs_pruned_ranks)
 * share indices / indptr. */
typedef struct mcb_graph mcb_graph;

/* compute_obs_obs_knn_graph on a float32 similarity matrix [n, n]: rank (knn_graph.py:285-297),
 * balance (:300-340), prune (:343-392), weigh (:395-407), growing k until every row keeps
 * min_outgoing_degree edges (:251-268).  Free the result with mcb_graph_free. */
/* Test aid, computed on the HOST (no GPU needed): sums[r] = numpy's float32 np.add.reduceat over values[offsets[r] ..
 * offsets[r + 1]) -- the row totals scipy's csr sum(axis=1) yields for tools/knn_graph.py:399 (_weigh_edges) -- by the very
 * code the device runs for rows whose float32 total is not exact (csrc/numpy_sum.cuh).  Empty segments yield 0. */
int mcb_host_numpy_add_reduce_f32(const float* values, const int64_t* offsets, int64_t rows, float* sums);
int mcb_host_knn_graph_f32(const float* similarity, int64_t n, int64_t k, double balanced_ranks_factor,
                           double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                           mcb_graph** out);
int mcb_dev_knn_graph_f32(const float* similarity, int64_t n, int64_t k, double balanced_ranks_factor,
                          double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                          void* stream, mcb_graph** out);
int64_t mcb_graph_size(const mcb_graph* graph);
int64_t mcb_graph_nnz(const mcb_graph* graph);
int64_t mcb_graph_k(const mcb_graph* graph); /* the k finally used */
/* Copy to host buffers (any may be NULL): weights/ranks/indices have nnz entries, indptr size + 1. */
int mcb_graph_read(const mcb_graph* graph, float* weights, float* ranks, int32_t* indices, int32_t* indptr);
/* The same copies started on a stream of the graph's own, without waiting: a worker that walks its piles (one after the
 * other, pipeline/divide_and_conquer.py:1264-1295) reads the graph of pile k while the kernels of pile k + 1 run.  The
 * host buffers (page-locked, or the copies are not asynchronous) hold the graph after mcb_graph_read_finish returns;
 * mcb_graph_free also waits for the copies. */
int mcb_graph_read_start(mcb_graph* graph, float* weights, float* ranks, int32_t* indices, int32_t* indptr);
int mcb_graph_read_finish(mcb_graph* graph);
int mcb_graph_device_arrays(const mcb_graph* graph, const float** weights, const float** ranks, const int32_t** indices,
                            const int32_t** indptr);
void mcb_graph_free(mcb_graph* graph);

/* ---- pipeline/direct.py:214-283: one pile, device resident -------------------------------------- */

/* A pile = the cells x genes UMI matrix (CSR float32/int32/int32) of compute_direct_metacells,
 * uploaded once (create_host) or borrowed from device memory (create_dev; the arrays must outlive
 * the pile).  All later calls run on the pile's stream and keep every intermediate in HBM.
 * create_host starts the upload in row chunks on a stream of its own and returns; with pinned host memory the
 * buffers must stay valid until a later call on the pile has synchronised (mcb_pile_similarity, _knn_graph,
 * _gene_totals, _read_downsampled).  mcb_pile_downsample then works chunk by chunk under the rest of the upload. */
typedef struct mcb_pile mcb_pile;
int mcb_pile_create_host(const float* data, const int32_t* indices, const int32_t* indptr, int64_t cells, int64_t genes,
                         mcb_pile** out);
/* The same pile from a COMPACT host copy: 16-bit gene indices (at most 65536 genes), 8-bit counts, 3 bytes per non-zero
 * instead of 8 on the host link, which is what bounds the end-to-end step.  A count of 255 stands
 * for "see the escape list": escape_at int64[escapes] (positions in the arrays, ascending) and escape_values
 * float32[escapes] hold the counts of 255 and more (and any count that is not a whole number).  The chunks are widened on
 * the device, on the copy stream, to the float32 / int32 arrays the path reads; every mcb_pile_* call then works as on a
 * pile of mcb_pile_create_host, bit for bit.  No counterpart in the reference (its callers hold scipy CSR float32 /
 * int32): for callers that keep an atlas compact on the host or on disk. */
int mcb_pile_create_host_compact(const uint16_t* indices, const uint8_t* counts, const int32_t* indptr, int64_t cells,
                                 int64_t genes, const int64_t* escape_at, const float* escape_values, int64_t escapes,
                                 mcb_pile** out);
/* Optional: the largest total of a cell (the pipeline has the totals on the host, tools/downsample.py:84-90).
 * Sizes the downsampling scratch without a pass over the data and a synchronisation; a cell above it is
 * reported by the next synchronising call on the pile. */
int mcb_pile_set_max_cell_total(mcb_pile* pile, uint64_t max_total);
/* Page-lock / release a host buffer (cudaHostRegister).  mcb_pile_create_host returns before the upload has finished
 * only for page-locked buffers; a caller that walks over many piles (ut.parallel_map over piles,
 * pipeline/divide_and_conquer.py:1264-1295) pins the arrays of the next pile and creates it before it asks for the graph
 * of the current one, so that the upload runs under the current pile's kernels. */
int mcb_host_pin(const void* buffer, uint64_t bytes);
int mcb_host_unpin(const void* buffer);
int mcb_pile_create_dev(const float* data, const int32_t* indices, const int32_t* indptr, int64_t cells, int64_t genes,
                        void* stream, mcb_pile** out);
int64_t mcb_pile_nnz(const mcb_pile* pile);
/* Returns when the upload mcb_pile_create_host started has arrived.  Host threads that walk piles side by side take
 * turns uploading (one upload at a time has the whole link, and falls under the other threads' kernels). */
int mcb_pile_wait_upload(mcb_pile* pile);
/* tools/downsample.py:99 -> downsample_compressed (downsample.cpp:171-192), every cell to `samples`. */
int mcb_pile_downsample(mcb_pile* pile, int32_t samples, uint64_t random_seed);
int mcb_pile_read_downsampled(mcb_pile* pile, float* data); /* nnz values, same structure as the input */
/* Per-gene totals of the (downsampled, if done) matrix to a host array of `genes` floats. */
int mcb_pile_gene_totals(mcb_pile* pile, float* totals);
/* The `count` genes with the largest (downsampled) totals, equal totals by lower gene index, ascending gene
 * indices: == np.sort(np.argsort(-totals, kind="stable")[:count]).  The list stays on the device for
 * mcb_pile_similarity / mcb_pile_knn_graph (pass feature_genes = NULL there); genes_host (nullable) receives a
 * copy.  A stand-in for pipeline/select.py on synthetic data, computed without leaving the GPU. */
int mcb_pile_top_genes(mcb_pile* pile, int64_t count, int32_t* genes_host);
int64_t mcb_pile_top_genes_count(const mcb_pile* pile); /* the length of the list on the device (0: none yet) */

/* The O(non-zeros) part of the feature-gene selection (pipeline/select.py:145-220 over tools/high.py:29-135, 246-303):
 * per gene of the downsampled pile the total (tools/high.py:73 sum_per), the sum of squares (computation.py:1526-1546
 * variance_per -> sum_squared_per) and the number of cells holding at least `min_top_value` UMIs of it (the
 * `rank_per(..., rank=n_obs - topN - 1) >= min_gene_topN` test of tools/high.py:121-124 is "at least topN + 1 cells hold
 * >= min_gene_topN": no dense [genes, cells] matrix is ranked).  Ask before mcb_pile_downsample: its kernel adds them up
 * where it writes every count (one 64-bit integer add per non-zero); read after.  totals float32[genes] (exact below
 * 2^24), squares uint64[genes] (exact), top_counts int32[genes].  At most 65535 cells. */
int mcb_pile_want_gene_statistics(mcb_pile* pile, float min_top_value);
int mcb_pile_gene_statistics(mcb_pile* pile, float* totals, uint64_t* squares, int32_t* top_counts);
/* The non-zero values of `count` genes of the (downsampled) pile in cell order, values[g * cells + k], sizes[g] of them:
 * the columns of the CSC matrix the reference sums (computation.py:1224-1268 on `layout="column_major"`).  For the few
 * genes whose float32 sums pass 2^24, where the order of numpy's pairwise summation shows: the host re-sums them with
 * numpy itself. */
int mcb_pile_gene_columns(mcb_pile* pile, const int32_t* genes, int64_t count, float* values, int32_t* sizes);

/* medians[i] = the median of values[i .. i + window) for i = 0 .. count - window (odd window <= 127; host arrays): the
 * `rolling(window).median()` of sliding_window_function (computation.py:1836-1897) over the genes ordered by their mean
 * expression, of which relative_variance_per subtracts the result (computation.py:1590-1595).  The median of an odd
 * window is one of its elements, so no arithmetic is involved; a window holding a NaN or an infinity yields NaN, as pandas'
 * rolling median does (it takes infinities for missing values). */
int mcb_host_sliding_median_f32(const float* values, int64_t count, int64_t window, float* medians);
/* The whole of sliding_window_function(values, "median", window, order_by) (computation.py:1836-1897) but its argsort:
 * `order` (int64[count], host) sorts the entries by `order_by`; medians[e] (float64[count], host) = the median of the
 * `window` entries nearest to entry e in that order, the windows at both ends padded as the reference pads them
 * (:1877-1883).  The gathers by `order`, the padding and the scatter back happen on the device. */
int mcb_host_sliding_median_ordered_f32(const float* values, const int64_t* order, int64_t count, int64_t window,
                                        double* medians);

/* direct.py:242-249 -> computation.py:679-741 (log_data): the dense float32 [cells, features] operand of the similarity,
 * dense_host[c, f] = log_table[count of gene feature_genes[f] in cell c] (log_table[c] = np.log2(float32(c) +
 * regularization), made by numpy on the host, so the result is bit-identical to the reference's `data += 1/7;
 * np.log2(data)`).  feature_genes as for mcb_pile_similarity.  A value that is not an integer count below table_size
 * is an error (MCB_ERR_ARG; *outside_table, nullable, is set to 1). */
int mcb_pile_log_fraction(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table,
                          int64_t table_size, float regularization, float* dense_host, int* outside_table);

/* direct.py:242-249 + tools/similarity.py:178-246: dense slice of `feature_genes` (host, distinct
 * gene indices; NULL = the list of mcb_pile_top_genes), log2(x + regularization) through `log_table` (host; log_table[c] = np.log2(c +
 * regularization) for c = 0 .. table_size-1), Pearson (absolute != 0: |r|) into similarity_dev
 * [cells, cells] (device).  Values the table misses are an error (see mcb_pile_log_fraction); *outside_table (host,
 * may be NULL) reports them. */
int mcb_pile_similarity(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table,
                        int64_t table_size, float regularization, int absolute, float* similarity_dev, int* outside_table);
/* ... followed by compute_obs_obs_knn_graph (see mcb_dev_knn_graph_f32). */
int mcb_pile_knn_graph(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table,
                       int64_t table_size, float regularization, int absolute, int64_t k, double balanced_ranks_factor,
                       double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                       int* outside_table, mcb_graph** out);
void mcb_pile_free(mcb_pile* pile);

/* ---- pipeline/divide_and_conquer.py:1250-1309: the atlas resident in HBM, piles gathered on the device ------------
 *
 * _compute_metacells_of_piles forces the whole cells x genes matrix to CSR once (:1286) and every pile of every phase
 * (preliminary, final, outlier levels: :990-1079) is `ut.slice(adata, obs=pile mask)` of it (:1267).  Here the atlas is
 * uploaded once and stays in HBM (float32 values, int32 gene indices, 64-bit row pointers: a million cells hold more
 * than 2^31 non-zeros; 8 bytes per non-zero, 29 GB for a million cells of 3 600 genes); a pile is a list of atlas
 * rows, gathered by a kernel, so that a pile of any phase costs its index list on the host link, not its UMIs. */
typedef struct mcb_atlas mcb_atlas;
/* Host arrays in (page-locked sources upload at the link's speed); returns when the atlas is resident. */
int mcb_atlas_create_host(const float* data, const int32_t* indices, const int64_t* indptr, int64_t cells, int64_t genes,
                          mcb_atlas** out);
/* Borrowed device arrays (they must outlive the atlas), e.g. blocks uploaded rank by rank and exchanged over NVLink;
 * the row pointers are given on the device and on the host. */
int mcb_atlas_create_dev(const float* data, const int32_t* indices, const int64_t* indptr_device, const int64_t* indptr_host,
                         int64_t cells, int64_t genes, mcb_atlas** out);
int64_t mcb_atlas_cells(const mcb_atlas* atlas);
int64_t mcb_atlas_nnz(const mcb_atlas* atlas);
/* Total UMIs of every cell (direct.py:214-218: the row sums, float32) to a host array of `cells` floats. */
int mcb_atlas_cell_totals(const mcb_atlas* atlas, float* totals_host);
/* sums_device[g, :] (device, float32 [groups, genes], overwritten) = the sum of the rows of the cells with
 * group_of_cells[c] == g (host int32 [cells]; negative: the cell is left out): the per-metacell totals of
 * collect_metacells (pipeline/collect.py:162-201 -> ut.sum_groups), which the grouping of the preliminary metacells
 * into the final piles starts from (divide_and_conquer.py:1153-1245). */
int mcb_atlas_sum_groups(const mcb_atlas* atlas, const int32_t* group_of_cells, int64_t groups, float* sums_device, void* stream);
void mcb_atlas_free(mcb_atlas* atlas);
/* The pile of the atlas rows `rows` (host int64 [count], in pile order, = ut.slice(adata, obs=...)): owned device
 * copies of the rows, on the calling thread's own stream; every mcb_pile_* call works on it as on an uploaded pile. */
int mcb_pile_create_from_atlas(const mcb_atlas* atlas, const int64_t* rows, int64_t count, mcb_pile** out);

/* The pile of a dense float32 [cells, genes] matrix on the device (the metacells of collect_metacells as the "cells"
 * of the grouping pile, divide_and_conquer.py:1153-1245): its non-zeros in column order, as scipy's csr_matrix(dense). */
int mcb_pile_create_dense_dev(const float* dense, int64_t cells, int64_t genes, mcb_pile** out);

/* sums float64 [groups, genes] (host) = the per-gene UMIs of the pile's cells with group_of_cells[c] == g (host int32
 * [cells]; negative: left out), of the pile's own (not downsampled) counts: the `data[cells, :].sum(axis=0)` and
 * `sum_per(data, per="column")` of tools/dissolve.py:101-102, :187-189.  Exact integers, so a sum below 2^24 is the
 * float32 value numpy computes in any order. */
int mcb_pile_sum_groups(mcb_pile* pile, const int32_t* group_of_cells, int64_t groups, double* sums);

/* collect_metacells for the cells of a pile (pipeline/collect.py:127-232): per metacell and gene the UMIs
 * (`umis_per_gene`, sums_device float32 [metacells, genes]), the cells left without the gene after each cell is downsampled
 * to its metacell's `zeros_downsample_umis` (`zeros_per_gene`, zeros_device int32 [metacells, genes]; the cells of a
 * metacell are sampled as one matrix: seed_index_of_cells is a cell's row in it, :160-171 -> downsample.cpp:148) and the
 * fractions (fractions_device float32 [metacells, genes]): geometric != 0 (`metacell_geo_mean`, the default): the weighted
 * geometric mean over the metacell's cells of umis / total + regularization, less that of the regularization, zeroed where
 * the gene has no UMIs or the result is negative, normalised to 1 (:196-225); geometric == 0: the downsampled UMIs' share
 * is NOT provided here, the plain UMIs' share is (:181-194 with the stored counts).  `scale` (per metacell) multiplies
 * the float32 fraction as `fraction *= total_umis` does for the metacell groups (:229-230); 1 leaves it.  Host arrays per
 * cell: metacell (0 .. metacells - 1, the cells in metacell order), weight (log2 of the cell's UMIs), regularization
 * (`metacell_umis_regularization` / UMIs), total UMIs (float64), samples, seed index; per metacell:
 * mean_log_regularization = sum(weight * log(regularization)) / sum(weight), weight_sum, scale, total_umis, cells.  The
 * sums are made in cell order without atomics (the same bits in every run and on every rank); the sums of UMIs and the
 * zeros are exact; the fractions are float64 arithmetic with the device's log / exp, within a few ulps of
 * numpy's (the parity tests hold them to 1e-6 relative). */
int mcb_pile_collect_metacells(mcb_pile* pile, const int32_t* metacell_of_cells, int64_t metacells, const double* weight_of_cells,
                               const double* regularization_of_cells, const double* total_of_cells,
                               const int32_t* samples_of_cells, const int32_t* seed_index_of_cells, uint64_t random_seed,
                               int geometric, const double* mean_log_regularization, const double* weight_sum, const double* scale,
                               const double* total_umis, const int32_t* cells_of_metacells, float* fractions_device,
                               float* sums_device, int32_t* zeros_device);

/* ---- gaps.cpp ------------------------------------------------------------------------------------ */

/* compute_cell_gaps (gaps.cpp:6-186; bound as xt.compute_cell_gaps, tools/deviants.py:309-326): per candidate metacell and
 * gene, the cells are ordered by the gene's log fraction (the reference's own std::sort, whose order of equal values
 * carries over from gene to gene, is reproduced) and the cells beyond a gap at either end of the order receive the gap's
 * size in max_gap_per_cell (float32[cells], read and written: the maximum with what is there).  The two matrices are
 * float32 [cells, genes], C contiguous; candidate_of_cells int32[cells]; deviant_per_cell / active_per_cell bool[cells];
 * min_gap_per_gene float32[genes].  The reference's third matrix (the fractions) is only checked for its shape there and
 * is not an argument here.  At most 8192 cells in a candidate, 65535 candidates. */
int mcb_host_compute_cell_gaps(const float* umis_per_gene_per_cell, const float* log_fraction_per_gene_per_cell, int64_t cells,
                               int64_t genes, const int32_t* candidate_of_cells, const uint8_t* deviant_per_cell,
                               const uint8_t* active_per_cell, const float* min_gap_per_gene, int64_t candidates_count,
                               int64_t gap_skip_cells, int64_t max_deviant_cells_count, float max_deviant_cells_fraction,
                               float min_compare_umis, float* max_gap_per_cell);
/* As above with the two matrices on the device (the vectors stay host arrays); synchronises the stream. */
int mcb_dev_compute_cell_gaps(const float* umis_per_gene_per_cell, const float* log_fraction_per_gene_per_cell, int64_t cells,
                              int64_t genes, const int32_t* candidate_of_cells, const uint8_t* deviant_per_cell,
                              const uint8_t* active_per_cell, const float* min_gap_per_gene, int64_t candidates_count,
                              int64_t gap_skip_cells, int64_t max_deviant_cells_count, float max_deviant_cells_fraction,
                              float min_compare_umis, float* max_gap_per_cell, void* stream);

/* The dense inputs of compute_cell_gaps for a pile, made on the device from its (not downsampled) UMIs
 * (tools/deviants.py:270-292: to_numpy_matrix, the fractions umis / total, log2(fraction + regularization)).  The division
 * and the addition are IEEE float32 operations; the logarithm is numpy's float32 log2, which is not correctly rounded, so
 * the host supplies it: log_table float32 [cells, table_counts] holds log2(count / total + regularization) of every cell
 * for the counts 0 .. table_counts - 1 (one vectorised numpy call), and the entries whose count the table does not hold
 * are handed back (mcb_gap_inputs_exceptions / _read_exceptions: their fraction + regularization) and patched with the
 * host's logarithms (mcb_gap_inputs_patch).  All host arrays. */
typedef struct mcb_gap_inputs mcb_gap_inputs;
int mcb_pile_gap_inputs_create(mcb_pile* pile, const float* total_umis_per_cell, const float* regularization_per_cell,
                               const float* log_table, int64_t table_counts, mcb_gap_inputs** out);
int64_t mcb_gap_inputs_exceptions(const mcb_gap_inputs* inputs);
int mcb_gap_inputs_read_exceptions(const mcb_gap_inputs* inputs, float* values);
int mcb_gap_inputs_patch(mcb_gap_inputs* inputs, const float* logs);
/* The two matrices to host arrays of cells x genes floats (either may be null): for tests and small piles. */
int mcb_gap_inputs_read(const mcb_gap_inputs* inputs, float* umis, float* log_fraction);
int mcb_gap_inputs_device_arrays(const mcb_gap_inputs* inputs, const float** umis, const float** log_fraction);
/* compute_cell_gaps on the resident matrices: one call per round of tools/deviants.py:299-356. */
int mcb_gap_inputs_cell_gaps(const mcb_gap_inputs* inputs, const int32_t* candidate_of_cells, const uint8_t* deviant_per_cell,
                             const uint8_t* active_per_cell, const float* min_gap_per_gene, int64_t candidates_count,
                             int64_t gap_skip_cells, int64_t max_deviant_cells_count, float max_deviant_cells_fraction,
                             float min_compare_umis, float* max_gap_per_cell);
void mcb_gap_inputs_free(mcb_gap_inputs* inputs);

/* ---- choose_seeds.cpp --------------------------------------------------------------------------- */

/* choose_seeds (choose_seeds.cpp:264-372; bound as xt.choose_seeds, tools/candidates.py:733-745).  The graph is
 * given twice, as outgoing CSR and incoming CSC (float32 data, sorted int32 indices, int32 indptr of nodes + 1
 * entries, the one the transpose of the other).  seed_of_nodes int32[nodes]: on entry the seeds already given
 * (-1 = none), on return the seed of every node; *seeds_count receives the value the reference returns.
 * Bit-identical to the reference for a given random_seed: std::minstd_rand, uniform_real_distribution and the
 * element order libstdc++'s std::nth_element / std::sort leave behind are reproduced.  Where the reference's
 * FastAssert would exit(1), returns MCB_ERR_ARG with the text in mcb_last_error().  At most 65535 nodes and
 * 16384 seeds. */
int mcb_host_choose_seeds(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                          const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                          int64_t nodes, uint64_t random_seed, int64_t max_seeds_count, float min_seed_size_quantile,
                          float max_seed_size_quantile, int32_t* seed_of_nodes, int64_t* seeds_count);
/* As above on device arrays (seeds_count stays a host pointer); synchronises the stream. */
int mcb_dev_choose_seeds(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                         const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                         int64_t nodes, uint64_t random_seed, int64_t max_seeds_count, float min_seed_size_quantile,
                         float max_seed_size_quantile, int32_t* seed_of_nodes, int64_t* seeds_count, void* stream);

/* The opening of compute_candidate_metacells on a device-resident k-NN graph (tools/candidates.py:128-196 and the
 * first _optimize_partitions of _reduce_communities, :351-368): the incoming layout is built on the device, choose_seeds
 * fills the partition, optimize_partitions (every community hot) improves it; only mass_of_nodes goes up and the
 * partition comes back.  Runs on the calling thread's own stream: graphs of different piles are handled side by side by
 * as many host threads, one SM each.  partition_of_nodes int32[nodes] (host, output); *seeds_count and *score as
 * mcb_host_choose_seeds / mcb_host_optimize_partitions return them.  Bit-identical to the reference's two calls. */
int mcb_graph_seed_partition(const mcb_graph* graph, const float* mass_of_nodes, uint64_t random_seed, int64_t max_seeds_count,
                             float min_seed_size_quantile, float max_seed_size_quantile, double low_mass, double target_mass,
                             double high_mass, double low_size, double target_size, double high_size, double cooldown_pass,
                             double cooldown_node, double cold_temperature, int32_t* partition_of_nodes, int64_t* seeds_count,
                             double* score);

/* The whole loop of compute_candidate_metacells (tools/candidates.py:128-289) on a device-resident k-NN graph.  A
 * session keeps what the loop's native calls share between them in HBM: the graph (borrowed: it must outlive the
 * session), its incoming layout (:135) and the node masses (cell_umis, :145-149).  Only the community vector (int32 per
 * node) and the hot mask travel per call.  Every call runs on the calling thread's own stream and returns when its
 * results are on the host; sessions of different piles are driven side by side by as many host threads. */
typedef struct mcb_candidates mcb_candidates;
int mcb_candidates_create(const mcb_graph* graph, const float* mass_of_nodes, mcb_candidates** out);
void mcb_candidates_free(mcb_candidates* session);
/* _choose_seeds (tools/candidates.py:720-749 -> xt.choose_seeds): seed_of_nodes int32[nodes] (host) holds the communities
 * already given (-1 = none) and receives a seed for every node; as mcb_host_choose_seeds. */
int mcb_candidates_choose_seeds(mcb_candidates* session, uint64_t random_seed, int64_t max_seeds_count,
                                float min_seed_size_quantile, float max_seed_size_quantile, int32_t* seed_of_nodes,
                                int64_t* seeds_count);
/* _optimize_partitions (tools/candidates.py:831-880 -> xt.optimize_partitions) with the hot-communities mask and the cold
 * temperature of the round; as mcb_host_optimize_partitions (stats nullable). */
int mcb_candidates_optimize_partitions(mcb_candidates* session, uint32_t random_seed, double low_mass, double target_mass,
                                       double high_mass, double low_size, double target_size, double high_size,
                                       double cooldown_pass, double cooldown_node, int32_t* partition_of_nodes,
                                       const int8_t* hot_partitions_mask, int64_t mask_count, double cold_temperature,
                                       int64_t* stats, double* score);
/* The symmetric sub-graph of every listed community (tools/candidates.py:545-546: outgoing[mask, :][:, mask] plus its
 * transpose): members int32[member_offsets[communities]] lists the nodes community by community (ascending inside one, as
 * np.where gives them; a node at most once), member_offsets int64[communities + 1]; dense (host) receives the
 * communities' [size, size] float32 blocks one after the other. */
int mcb_candidates_community_weights(mcb_candidates* session, const int32_t* members, const int64_t* member_offsets,
                                     int64_t communities, float* dense);
/* ut.min_cut of every listed community's symmetric sub-graph (utilities/computation.py:2368-2449, called at
 * tools/candidates.py:548), one CTA per community.  The reference hands the edges to igraph's Graph.mincut (absent from
 * this image); here Stoer & Wagner's algorithm in float64 with stated tie rules (each phase starts at the lowest live
 * vertex; the most tightly connected vertex next, the lowest index among equals; the first phase of the least weight
 * wins; its last vertex's side is cut.partition[0]) -- the same minimal weight, but among cuts of equal weight igraph's
 * choice is its own.  is_second int8[members]: 1 = the member is in cut.partition[1]; totals float64[communities, 3]: the
 * weight inside the first side, inside the second side and across the cut, added in float32 in the order :2414-2426
 * walks the edges (Python float += numpy float32 stays float32 under NumPy >= 2).  At most 4096 nodes in a community. */
int mcb_candidates_min_cuts(mcb_candidates* session, const int32_t* members, const int64_t* member_offsets,
                            int64_t communities, int8_t* is_second, double* totals);

/* ---- partitions.cpp ----------------------------------------------------------------------------- */

/* optimize_partitions (partitions.cpp:1234-1302; bound as xt.optimize_partitions, tools/candidates.py:853-873).
 * The graph is given twice, as outgoing CSR and incoming CSC (float32 data, sorted int32 indices, int32
 * indptr of nodes + 1 entries); mass_of_nodes float32[nodes]; partition_of_nodes int32[nodes] holds the
 * initial partition (every index >= 0, no empty partition) and receives the optimised one;
 * hot_partitions_mask int8[mask_count >= partitions], 0 = the nodes of that partition start at
 * cold_temperature.  *score receives the value the reference returns.  Bit-identical to the reference
 * for a given random_seed (std::minstd_rand / std::shuffle / glibc log are reproduced).  Where the
 * reference's FastAssert would exit(1), returns MCB_ERR_ARG with the assert's text in mcb_last_error().
 * stats (nullable, int64[16], host): passes, improve_node calls, moves, skipped visits, batches, 0, 0, 0,
 * then SM-cycle counters of the phases (diagnostics).
 * At most 65535 nodes. */
int mcb_host_optimize_partitions(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                                 const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                                 int64_t nodes, uint32_t random_seed, const float* mass_of_nodes, double low_mass,
                                 double target_mass, double high_mass, double low_size, double target_size, double high_size,
                                 double cooldown_pass, double cooldown_node, int32_t* partition_of_nodes,
                                 const int8_t* hot_partitions_mask, int64_t mask_count, double cold_temperature, int64_t* stats,
                                 double* score);
/* As above on device arrays (score and stats stay host pointers); synchronises the stream. */
int mcb_dev_optimize_partitions(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                                const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                                int64_t nodes, uint32_t random_seed, const float* mass_of_nodes, double low_mass,
                                double target_mass, double high_mass, double low_size, double target_size, double high_size,
                                double cooldown_pass, double cooldown_node, int32_t* partition_of_nodes,
                                const int8_t* hot_partitions_mask, int64_t mask_count, double cold_temperature, int64_t* stats,
                                double* score, void* stream);

/* score_partitions (partitions.cpp:1304-1338, the 16-argument C++ signature). */
int mcb_host_score_partitions(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                              const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                              int64_t nodes, const float* mass_of_nodes, double low_mass, double target_mass, double high_mass,
                              double low_size, double target_size, double high_size, const int32_t* partition_of_nodes,
                              const int8_t* hot_partitions_mask, int64_t mask_count, int with_orphans, double* score);
int mcb_dev_score_partitions(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                             const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                             int64_t nodes, const float* mass_of_nodes, double low_mass, double target_mass, double high_mass,
                             double low_size, double target_size, double high_size, const int32_t* partition_of_nodes,
                             const int8_t* hot_partitions_mask, int64_t mask_count, int with_orphans, double* score,
                             void* stream);

/* The double-precision log the scores are built from: glibc 2.39's log() (extensions.h:135-138 calls libm),
 * evaluated on the device bit for bit (csrc/glibc_log.cuh).  Host arrays; exported so the parity tests
 * can pin it against libm. */
int mcb_host_glibc_log_f64(const double* values, double* logs, int64_t count);

/* ---- projection front-end (BASELINE.json config 5) ----------------------------------------
 * Replaces, for tools/project.py:249-270 + :309-331, the pair `ut.cross_corrcoef_rows(query, atlas)`
 * (computation.py:478-524 -> cross_correlate_dense, correlate.cpp:433-494) and the per-row
 * `np.argsort(-correlations)` of which the candidates loop consumes a prefix: for every query row the `k` atlas rows
 * of the largest Pearson correlation, in descending order of the correlation, equal correlations by ascending atlas
 * index (a stable argsort of the negated row).  The [query, atlas] matrix is never materialised beyond one block of
 * query rows in HBM (MCB_PROJECT_BLOCK_BYTES, default 4 GiB).  out_indices / out_correlations: [query_rows, k].
 * The inputs must not hold NaNs (log fractions never do). */
int mcb_host_project_top_f32(const float* query, int64_t query_rows, int64_t query_stride, const float* atlas, int64_t atlas_rows,
                             int64_t atlas_stride, int64_t columns, int64_t k, int32_t* out_indices, float* out_correlations);
int mcb_dev_project_top_f32(const float* query, int64_t query_rows, const float* atlas, int64_t atlas_rows, int64_t columns,
                            int64_t k, int32_t* out_indices, float* out_correlations, void* stream);
/* The selection stage alone on a device matrix [rows, row_stride >= columns]: per row the `k` largest values in
 * descending order (ties by ascending column), 1 <= k <= min(columns, 4096). */
int mcb_dev_ordered_top_f32(const float* values, int64_t rows, int64_t columns, int64_t row_stride, int64_t k,
                            int32_t* out_indices, float* out_values, void* stream);

/* @@APPEND@@ */

#ifdef __cplusplus
}
#endif

#endif /* MCB200_H */
