// C-ABI entry points for the ordering / compressed-matrix functions and the fused k-NN graph builder.
#include <algorithm>
#include <cmath>
#include <memory>
#include <utility>

#include "numpy_sum.cuh"
#include "common.cuh"
#include "knn_graph.cuh"
#include "../../include/mcb200.h"

namespace mcb {
int rank_matrix_launch(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, cudaStream_t stream);
int rank_rows_launch(const float* matrix, int64_t rows, int64_t columns, int64_t row_stride, float* output, int64_t rank,
                     cudaStream_t stream);
int collect_top_launch(int64_t degree, const float* similarity, int64_t rows, int64_t columns, int64_t row_stride,
                       int32_t* out_indices, float* out_data, int ranks, cudaStream_t stream);
int collect_pruned_launch(int64_t degree, const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands,
                          float* out_data, int32_t* out_indices, int32_t* out_indptr, cudaStream_t stream);
int collect_compressed_launch(const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands, float* out_data,
                              int32_t* out_indices, int32_t* cursor, cudaStream_t stream);
int sort_compressed_indices_launch(float* data, int32_t* indices, const int32_t* indptr, int64_t bands, int64_t widest_band,
                                   cudaStream_t stream);
int rank_outgoing_launch(float* work, int64_t n, cudaStream_t stream, const uint32_t* known_min);

int weigh_edges_launch(const float* data, const int32_t* indptr, int64_t n, float* weights, int32_t* min_degree_host,
                       cudaStream_t stream);
}  // namespace mcb

using namespace mcb;

// A device-resident CSR graph (see mcb200.h).
struct mcb_graph {
    int64_t n = 0, nnz = 0, k = 0;
    Scratch weights, ranks, indices, indptr;
    cudaStream_t stream = nullptr;
    cudaStream_t read_stream = nullptr;  // mcb_graph_read_start: the copies to the host run beside the next pile's kernels
    cudaEvent_t complete = nullptr;      // recorded on `stream` behind the last kernel that writes the graph
    ~mcb_graph() {
        if (complete)
            cudaEventDestroy(complete);
        if (read_stream) {
            cudaStreamSynchronize(read_stream);
            cudaStreamDestroy(read_stream);
        }
    }
};

namespace {

template <typename T>
int to_device(Scratch& device, const T* host, size_t count, cudaStream_t stream) {
    MCB_TRY(device.alloc(count * sizeof(T), stream));
    if (count)
        MCB_CUDA(cudaMemcpyAsync(device.ptr, host, count * sizeof(T), cudaMemcpyHostToDevice, stream));
    return MCB_OK;
}
template <typename T>
int to_host(T* host, const Scratch& device, size_t count, cudaStream_t stream) {
    if (count)
        MCB_CUDA(cudaMemcpyAsync(host, device.ptr, count * sizeof(T), cudaMemcpyDeviceToHost, stream));
    return MCB_OK;
}

// tools/knn_graph.py:251-275: rank, balance, prune; grow k until every row keeps min_outgoing_degree edges.
// consumed != nullptr: the caller gives its similarity buffer away (ranked in place, no copy) and may pass the
// matrix minimum it already knows.
int knn_graph_device(const float* similarity, int64_t n, int64_t k, double balanced_ranks_factor, double incoming_degree_factor,
                     double outgoing_degree_factor, int64_t min_outgoing_degree, cudaStream_t stream, mcb_graph* graph,
                     float* consumed = nullptr, const uint32_t* known_min = nullptr) {
    MCB_REQUIRE(n >= 2, "knn_graph: needs at least 2 elements (got %lld)", (long long)n);
    MCB_REQUIRE(k >= 1 && balanced_ranks_factor > 0 && incoming_degree_factor > 0 && outgoing_degree_factor > 0 &&
                    min_outgoing_degree >= 1,
                "knn_graph: bad parameters");  // knn_graph.py:216-220
    trace_lap("knn: enter", stream);
    Scratch copy;
    float* ranks = consumed;
    if (ranks == nullptr) {
        MCB_TRY(copy.alloc((size_t)n * n * sizeof(float), stream));
        MCB_CUDA(cudaMemcpyAsync(copy.ptr, similarity, (size_t)n * n * sizeof(float), cudaMemcpyDeviceToDevice, stream));
        ranks = copy.as<float>();
    }
    // The ranks do not depend on k: computed once (the reference recomputes them every round, knn_graph.py:252).
    trace_lap("knn: copy similarity", stream);
    MCB_TRY(rank_outgoing_launch(ranks, n, stream, known_min));
    trace_lap("knn: rank_outgoing", stream);
    std::unique_ptr<PrunedRanks> pruned_holder;
    for (;;) {
        const double max_rank = (double)k * balanced_ranks_factor;                   // knn_graph.py:313
        const float shift = (float)(max_rank + 1.0 - 2097152.0);                     // knn_graph.py:325
        int64_t incoming = (int64_t)std::nearbyint((double)k * incoming_degree_factor);  // Python round(): half to even
        int64_t outgoing = (int64_t)std::nearbyint((double)k * outgoing_degree_factor);
        incoming = incoming < n - 1 ? incoming : n - 1;
        outgoing = outgoing < n - 1 ? outgoing : n - 1;
        BalancedRanks balanced;
        MCB_TRY(balance_ranks_launch(ranks, n, shift, &balanced, stream));
        trace_lap("knn: balance_ranks", stream);
        pruned_holder.reset(new PrunedRanks());
        PrunedRanks& pruned = *pruned_holder;
        MCB_TRY(prune_ranks_launch(balanced, n, incoming, outgoing, &pruned, stream));
        trace_lap("knn: prune_ranks", stream);
        int32_t min_degree = 0;
        MCB_TRY(weigh_edges_launch(pruned.data.as<float>(), pruned.indptr.as<int32_t>(), n, nullptr, &min_degree, stream));
        if (min_degree >= min_outgoing_degree)
            break;
        k += (k / 10) > 1 ? (k / 10) : 1;  // knn_graph.py:268
    }
    PrunedRanks& pruned = *pruned_holder;
    graph->n = n;
    graph->nnz = pruned.nnz;
    graph->k = k;
    graph->stream = stream;
    MCB_TRY(graph->weights.alloc((size_t)pruned.nnz * sizeof(float), stream));
    MCB_TRY(weigh_edges_launch(pruned.data.as<float>(), pruned.indptr.as<int32_t>(), n, graph->weights.as<float>(), nullptr, stream));
    trace_lap("knn: weigh_edges", stream);
    std::swap(graph->ranks.ptr, pruned.data.ptr);
    std::swap(graph->ranks.stream, pruned.data.stream);
    std::swap(graph->indices.ptr, pruned.indices.ptr);
    std::swap(graph->indices.stream, pruned.indices.stream);
    std::swap(graph->indptr.ptr, pruned.indptr.ptr);
    std::swap(graph->indptr.stream, pruned.indptr.stream);
    // Other streams (mcb_graph_seed_partition on a worker thread's stream, mcb_graph_read_start) order themselves
    // behind this event: the weights are still being written when this function returns.
    MCB_CUDA(cudaEventCreateWithFlags(&graph->complete, cudaEventDisableTiming));
    MCB_CUDA(cudaEventRecord(graph->complete, stream));
    return MCB_OK;
}

}  // namespace

namespace mcb {
// Orders `stream` behind the kernels that write the graph (a no-op on the graph's own stream).
int graph_wait_complete(const mcb_graph* graph, cudaStream_t stream) {
    if (graph->complete != nullptr && stream != graph->stream)
        MCB_CUDA(cudaStreamWaitEvent(stream, graph->complete, 0));
    return MCB_OK;
}

// The pile path: `similarity` (device, n x n) is consumed; known_min as in rank_outgoing_launch.
int knn_graph_consuming(float* similarity, const uint32_t* known_min, int64_t n, int64_t k, double balanced_ranks_factor,
                        double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                        cudaStream_t stream, mcb_graph** out) {
    MCB_REQUIRE(similarity != nullptr && out != nullptr, "knn_graph: null argument");
    std::unique_ptr<mcb_graph> graph(new mcb_graph());
    MCB_TRY(knn_graph_device(similarity, n, k, balanced_ranks_factor, incoming_degree_factor, outgoing_degree_factor,
                             min_outgoing_degree, stream, graph.get(), similarity, known_min));
    *out = graph.release();
    return MCB_OK;
}
}  // namespace mcb

extern "C" {

/* ---- rank.cpp / top_per.cpp ------------------------------------------------------------------ */

int mcb_dev_rank_matrix_f32(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, void* stream) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "rank_matrix: bad shape");
    return rank_matrix_launch(matrix, rows, columns, row_stride, ascending, (cudaStream_t)stream);
}

int mcb_host_rank_matrix_f32(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "rank_matrix: bad shape");
    if (rows == 0 || columns == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch device;
    MCB_TRY(device.alloc((size_t)rows * columns * sizeof(float), stream));
    MCB_CUDA(cudaMemcpy2DAsync(device.ptr, columns * sizeof(float), matrix, row_stride * sizeof(float), columns * sizeof(float), rows,
                               cudaMemcpyHostToDevice, stream));
    MCB_TRY(rank_matrix_launch(device.as<float>(), rows, columns, columns, ascending, stream));
    MCB_CUDA(cudaMemcpy2DAsync(matrix, row_stride * sizeof(float), device.ptr, columns * sizeof(float), columns * sizeof(float), rows,
                               cudaMemcpyDeviceToHost, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_rank_rows_f32(const float* matrix, int64_t rows, int64_t columns, int64_t row_stride, float* output, int64_t rank) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "rank_rows: bad shape");
    MCB_REQUIRE(rank >= 0 && rank < columns, "rank_rows: rank %lld out of range", (long long)rank);  // rank.cpp:33
    if (rows == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch device, result;
    MCB_TRY(device.alloc((size_t)rows * columns * sizeof(float), stream));
    MCB_CUDA(cudaMemcpy2DAsync(device.ptr, columns * sizeof(float), matrix, row_stride * sizeof(float), columns * sizeof(float), rows,
                               cudaMemcpyHostToDevice, stream));
    MCB_TRY(result.alloc((size_t)rows * sizeof(float), stream));
    MCB_TRY(rank_rows_launch(device.as<float>(), rows, columns, columns, result.as<float>(), rank, stream));
    MCB_TRY(to_host(output, result, (size_t)rows, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_dev_rank_rows_f32(const float* matrix, int64_t rows, int64_t columns, int64_t row_stride, float* output, int64_t rank,
                          void* stream) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "rank_rows: bad shape");
    return rank_rows_launch(matrix, rows, columns, row_stride, output, rank, (cudaStream_t)stream);
}

int mcb_dev_collect_top_f32(int64_t degree, const float* similarity, int64_t rows, int64_t columns, int64_t row_stride,
                            int32_t* out_indices, float* out_data, int ranks, void* stream) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "collect_top: bad shape");
    return collect_top_launch(degree, similarity, rows, columns, row_stride, out_indices, out_data, ranks, (cudaStream_t)stream);
}

int mcb_host_collect_top_f32(int64_t degree, const float* similarity, int64_t rows, int64_t columns, int64_t row_stride,
                             int32_t* out_indices, float* out_data, int ranks) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "collect_top: bad shape");
    MCB_REQUIRE(degree > 0 && degree < columns, "collect_top: degree %lld must be in 1..columns-1", (long long)degree);
    if (rows == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch device, indices, data;
    MCB_TRY(device.alloc((size_t)rows * columns * sizeof(float), stream));
    MCB_CUDA(cudaMemcpy2DAsync(device.ptr, columns * sizeof(float), similarity, row_stride * sizeof(float), columns * sizeof(float),
                               rows, cudaMemcpyHostToDevice, stream));
    MCB_TRY(indices.alloc((size_t)rows * degree * sizeof(int32_t), stream));
    MCB_TRY(data.alloc((size_t)rows * degree * sizeof(float), stream));
    MCB_TRY(collect_top_launch(degree, device.as<float>(), rows, columns, columns, indices.as<int32_t>(), data.as<float>(), ranks, stream));
    MCB_TRY(to_host(out_indices, indices, (size_t)rows * degree, stream));
    MCB_TRY(to_host(out_data, data, (size_t)rows * degree, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

/* ---- prune_per.cpp / relayout.cpp -------------------------------------------------------------- */

int mcb_dev_collect_pruned_f32_i32_i32(int64_t degree, const float* data, const int32_t* indices, const int32_t* indptr,
                                       int64_t bands, float* out_data, int32_t* out_indices, int32_t* out_indptr, void* stream) {
    MCB_REQUIRE(bands >= 0, "collect_pruned: negative bands count");
    return collect_pruned_launch(degree, data, indices, indptr, bands, out_data, out_indices, out_indptr, (cudaStream_t)stream);
}

int mcb_host_collect_pruned_f32_i32_i32(int64_t degree, const float* data, const int32_t* indices, const int32_t* indptr,
                                        int64_t bands, float* out_data, int32_t* out_indices, int32_t* out_indptr) {
    MCB_REQUIRE(bands >= 0 && degree >= 0, "collect_pruned: bad arguments");
    cudaStream_t stream = nullptr;
    const int64_t nnz = bands > 0 ? indptr[bands] : 0;
    Scratch d_data, d_indices, d_indptr, o_data, o_indices, o_indptr;
    MCB_TRY(to_device(d_data, data, (size_t)nnz, stream));
    MCB_TRY(to_device(d_indices, indices, (size_t)nnz, stream));
    MCB_TRY(to_device(d_indptr, indptr, (size_t)bands + 1, stream));
    MCB_TRY(o_data.alloc((size_t)nnz * sizeof(float), stream));
    MCB_TRY(o_indices.alloc((size_t)nnz * sizeof(int32_t), stream));
    MCB_TRY(o_indptr.alloc((size_t)(bands + 1) * sizeof(int32_t), stream));
    MCB_TRY(collect_pruned_launch(degree, d_data.as<float>(), d_indices.as<int32_t>(), d_indptr.as<int32_t>(), bands,
                                  o_data.as<float>(), o_indices.as<int32_t>(), o_indptr.as<int32_t>(), stream));
    MCB_TRY(to_host(out_indptr, o_indptr, (size_t)bands + 1, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    const int64_t out_nnz = out_indptr[bands];
    MCB_TRY(to_host(out_data, o_data, (size_t)out_nnz, stream));
    MCB_TRY(to_host(out_indices, o_indices, (size_t)out_nnz, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_collect_compressed_f32_i32_i32(const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands,
                                            float* out_data, int32_t* out_indices, int32_t* out_indptr_shifted,
                                            int64_t out_bands) {
    MCB_REQUIRE(bands >= 0 && out_bands >= 0, "collect_compressed: bad arguments");
    if (bands == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    const int64_t nnz = indptr[bands];
    for (int64_t e = 0; e < nnz; ++e)
        MCB_REQUIRE(indices[e] >= 0 && indices[e] < out_bands, "collect_compressed: index %d out of range at %lld", indices[e], (long long)e);
    Scratch d_data, d_indices, d_indptr, o_data, o_indices, cursor;
    MCB_TRY(to_device(d_data, data, (size_t)nnz, stream));
    MCB_TRY(to_device(d_indices, indices, (size_t)nnz, stream));
    MCB_TRY(to_device(d_indptr, indptr, (size_t)bands + 1, stream));
    MCB_TRY(to_device(cursor, out_indptr_shifted, (size_t)out_bands, stream));
    MCB_TRY(o_data.alloc((size_t)nnz * sizeof(float), stream));
    MCB_TRY(o_indices.alloc((size_t)nnz * sizeof(int32_t), stream));
    MCB_TRY(collect_compressed_launch(d_data.as<float>(), d_indices.as<int32_t>(), d_indptr.as<int32_t>(), bands, o_data.as<float>(),
                                      o_indices.as<int32_t>(), cursor.as<int32_t>(), stream));
    MCB_TRY(to_host(out_data, o_data, (size_t)nnz, stream));
    MCB_TRY(to_host(out_indices, o_indices, (size_t)nnz, stream));
    MCB_TRY(to_host(out_indptr_shifted, cursor, (size_t)out_bands, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_sort_compressed_indices_f32_i32_i32(float* data, int32_t* indices, const int32_t* indptr, int64_t bands) {
    MCB_REQUIRE(bands >= 0, "sort_compressed_indices: negative bands count");
    if (bands == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    const int64_t nnz = indptr[bands];
    int64_t widest = 0;
    for (int64_t b = 0; b < bands; ++b) widest = std::max<int64_t>(widest, indptr[b + 1] - indptr[b]);
    Scratch d_data, d_indices, d_indptr;
    MCB_TRY(to_device(d_data, data, (size_t)nnz, stream));
    MCB_TRY(to_device(d_indices, indices, (size_t)nnz, stream));
    MCB_TRY(to_device(d_indptr, indptr, (size_t)bands + 1, stream));
    MCB_TRY(sort_compressed_indices_launch(d_data.as<float>(), d_indices.as<int32_t>(), d_indptr.as<int32_t>(), bands, widest, stream));
    MCB_TRY(to_host(data, d_data, (size_t)nnz, stream));
    MCB_TRY(to_host(indices, d_indices, (size_t)nnz, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

/* ---- tools/knn_graph.py:203-407 -------------------------------------------------------------------- */

int mcb_dev_knn_graph_f32(const float* similarity, int64_t n, int64_t k, double balanced_ranks_factor,
                          double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                          void* stream, mcb_graph** out) {
    MCB_REQUIRE(out != nullptr, "knn_graph: null result pointer");
    std::unique_ptr<mcb_graph> graph(new mcb_graph());
    MCB_TRY(knn_graph_device(similarity, n, k, balanced_ranks_factor, incoming_degree_factor, outgoing_degree_factor,
                             min_outgoing_degree, (cudaStream_t)stream, graph.get()));
    *out = graph.release();
    return MCB_OK;
}

int mcb_host_numpy_add_reduce_f32(const float* values, const int64_t* offsets, int64_t rows, float* sums) {
    MCB_REQUIRE(offsets != nullptr && sums != nullptr && rows >= 0, "numpy_add_reduce: bad arguments");
    for (int64_t row = 0; row < rows; ++row) {
        const int64_t count = offsets[row + 1] - offsets[row];
        MCB_REQUIRE(count >= 0 && count <= 0x7fffffff, "numpy_add_reduce: offsets are not monotone at %lld", (long long)row);
        sums[row] = count == 0 ? 0.0f : mcb::numpy_add_reduce_f32(values + offsets[row], (int)count);
    }
    return MCB_OK;
}

int mcb_host_knn_graph_f32(const float* similarity, int64_t n, int64_t k, double balanced_ranks_factor,
                           double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                           mcb_graph** out) {
    MCB_REQUIRE(out != nullptr && n >= 2, "knn_graph: bad arguments");
    cudaStream_t stream = nullptr;
    Scratch device;
    MCB_TRY(to_device(device, similarity, (size_t)n * n, stream));
    return mcb_dev_knn_graph_f32(device.as<float>(), n, k, balanced_ranks_factor, incoming_degree_factor, outgoing_degree_factor,
                                 min_outgoing_degree, stream, out);
}

int64_t mcb_graph_size(const mcb_graph* graph) { return graph ? graph->n : 0; }
int64_t mcb_graph_nnz(const mcb_graph* graph) { return graph ? graph->nnz : 0; }
int64_t mcb_graph_k(const mcb_graph* graph) { return graph ? graph->k : 0; }

int mcb_graph_device_arrays(const mcb_graph* graph, const float** weights, const float** ranks, const int32_t** indices,
                            const int32_t** indptr) {
    MCB_REQUIRE(graph != nullptr, "graph: null handle");
    if (weights) *weights = graph->weights.as<float>();
    if (ranks) *ranks = graph->ranks.as<float>();
    if (indices) *indices = graph->indices.as<int32_t>();
    if (indptr) *indptr = graph->indptr.as<int32_t>();
    return MCB_OK;
}

int mcb_graph_read(const mcb_graph* graph, float* weights, float* ranks, int32_t* indices, int32_t* indptr) {
    MCB_REQUIRE(graph != nullptr, "graph: null handle");
    cudaStream_t stream = graph->stream;
    if (weights) MCB_TRY(to_host(weights, graph->weights, (size_t)graph->nnz, stream));
    if (ranks) MCB_TRY(to_host(ranks, graph->ranks, (size_t)graph->nnz, stream));
    if (indices) MCB_TRY(to_host(indices, graph->indices, (size_t)graph->nnz, stream));
    if (indptr) MCB_TRY(to_host(indptr, graph->indptr, (size_t)graph->n + 1, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_graph_read_start(mcb_graph* graph, float* weights, float* ranks, int32_t* indices, int32_t* indptr) {
    MCB_REQUIRE(graph != nullptr, "graph: null handle");
    if (graph->read_stream == nullptr)
        MCB_CUDA(cudaStreamCreateWithFlags(&graph->read_stream, cudaStreamNonBlocking));
    MCB_TRY(graph_wait_complete(graph, graph->read_stream));
    cudaStream_t stream = graph->read_stream;
    if (weights) MCB_TRY(to_host(weights, graph->weights, (size_t)graph->nnz, stream));
    if (ranks) MCB_TRY(to_host(ranks, graph->ranks, (size_t)graph->nnz, stream));
    if (indices) MCB_TRY(to_host(indices, graph->indices, (size_t)graph->nnz, stream));
    if (indptr) MCB_TRY(to_host(indptr, graph->indptr, (size_t)graph->n + 1, stream));
    return MCB_OK;
}

int mcb_graph_read_finish(mcb_graph* graph) {
    MCB_REQUIRE(graph != nullptr, "graph: null handle");
    if (graph->read_stream)
        MCB_CUDA(cudaStreamSynchronize(graph->read_stream));
    return MCB_OK;
}

void mcb_graph_free(mcb_graph* graph) { delete graph; }

}  // extern "C"
