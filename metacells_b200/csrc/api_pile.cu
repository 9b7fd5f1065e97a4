// C-ABI of the device-resident pile (see pile.cu and mcb200.h).
#include <memory>
#include <vector>

#include "common.cuh"
#include "../../include/mcb200.h"

namespace mcb {
int downsample_launch(const float* input, float* output, const int32_t* indptr, const int32_t* samples,
                      int32_t samples_all, int64_t rows, int64_t columns, int64_t max_row_size, uint64_t seed,
                      int use_row_seed, cudaStream_t stream);
int column_totals_launch(const float* data, const int32_t* indices, int64_t nnz, int64_t columns, float* totals,
                         cudaStream_t stream);
int log_fraction_dense_launch(const float* data, const int32_t* indices, const int32_t* indptr, int64_t rows,
                              const int32_t* slot_of_gene, int64_t features, const float* table, int64_t table_size,
                              float regularization, float* dense, int* outside_table, cudaStream_t stream);
}  // namespace mcb

using namespace mcb;

struct mcb_pile {
    int64_t cells = 0, genes = 0, nnz = 0, widest = 0;
    cudaStream_t stream = nullptr;
    // the UMI matrix (CSR float32 / int32 / int32): owned copies, or borrowed device pointers
    Scratch own_data, own_indices, own_indptr;
    const float* data = nullptr;
    const int32_t* indices = nullptr;
    const int32_t* indptr = nullptr;
    Scratch downsampled;  // same sparsity structure as `data`
    bool has_downsampled = false;
};

namespace {

int finish_create(mcb_pile* pile, const int32_t* host_indptr) {
    int64_t widest = 0;
    for (int64_t c = 0; c < pile->cells; ++c) {
        MCB_REQUIRE(host_indptr[c + 1] >= host_indptr[c], "pile: indptr is not monotone at cell %lld", (long long)c);
        widest = std::max<int64_t>(widest, host_indptr[c + 1] - host_indptr[c]);
    }
    pile->widest = widest;
    pile->nnz = host_indptr[pile->cells];
    return MCB_OK;
}

}  // namespace

extern "C" {

int mcb_pile_create_host(const float* data, const int32_t* indices, const int32_t* indptr, int64_t cells, int64_t genes,
                         mcb_pile** out) {
    MCB_REQUIRE(out && cells >= 1 && genes >= 1, "pile: bad arguments");
    std::unique_ptr<mcb_pile> pile(new mcb_pile());
    pile->cells = cells;
    pile->genes = genes;
    MCB_TRY(finish_create(pile.get(), indptr));
    cudaStream_t stream = pile->stream;
    MCB_TRY(pile->own_data.alloc((size_t)pile->nnz * sizeof(float), stream));
    MCB_TRY(pile->own_indices.alloc((size_t)pile->nnz * sizeof(int32_t), stream));
    MCB_TRY(pile->own_indptr.alloc((size_t)(cells + 1) * sizeof(int32_t), stream));
    MCB_CUDA(cudaMemcpyAsync(pile->own_data.ptr, data, (size_t)pile->nnz * sizeof(float), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(pile->own_indices.ptr, indices, (size_t)pile->nnz * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(pile->own_indptr.ptr, indptr, (size_t)(cells + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    pile->data = pile->own_data.as<float>();
    pile->indices = pile->own_indices.as<int32_t>();
    pile->indptr = pile->own_indptr.as<int32_t>();
    *out = pile.release();
    return MCB_OK;
}

int mcb_pile_create_dev(const float* data, const int32_t* indices, const int32_t* indptr, int64_t cells, int64_t genes,
                        void* stream, mcb_pile** out) {
    MCB_REQUIRE(out && cells >= 1 && genes >= 1, "pile: bad arguments");
    std::unique_ptr<mcb_pile> pile(new mcb_pile());
    pile->cells = cells;
    pile->genes = genes;
    pile->stream = (cudaStream_t)stream;
    std::vector<int32_t> host_indptr((size_t)cells + 1);
    MCB_CUDA(cudaMemcpyAsync(host_indptr.data(), indptr, (size_t)(cells + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    MCB_TRY(finish_create(pile.get(), host_indptr.data()));
    pile->data = data;
    pile->indices = indices;
    pile->indptr = indptr;
    *out = pile.release();
    return MCB_OK;
}

int mcb_pile_downsample(mcb_pile* pile, int32_t samples, uint64_t random_seed) {
    MCB_REQUIRE(pile != nullptr, "pile: null handle");
    MCB_REQUIRE(samples >= 0, "pile: negative samples");
    if (!pile->downsampled.ptr)
        MCB_TRY(pile->downsampled.alloc((size_t)pile->nnz * sizeof(float), pile->stream));
    MCB_TRY(downsample_launch(pile->data, pile->downsampled.as<float>(), pile->indptr, nullptr, samples, pile->cells, 0,
                              pile->widest, random_seed, 1, pile->stream));
    pile->has_downsampled = true;
    return MCB_OK;
}

int mcb_pile_gene_totals(mcb_pile* pile, float* totals) {
    MCB_REQUIRE(pile != nullptr && totals != nullptr, "pile: null argument");
    const float* values = pile->has_downsampled ? pile->downsampled.as<float>() : pile->data;
    Scratch device;
    MCB_TRY(device.alloc((size_t)pile->genes * sizeof(float), pile->stream));
    MCB_TRY(column_totals_launch(values, pile->indices, pile->nnz, pile->genes, device.as<float>(), pile->stream));
    MCB_CUDA(cudaMemcpyAsync(totals, device.ptr, (size_t)pile->genes * sizeof(float), cudaMemcpyDeviceToHost, pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    return MCB_OK;
}

int mcb_pile_read_downsampled(mcb_pile* pile, float* data) {
    MCB_REQUIRE(pile != nullptr && pile->has_downsampled, "pile: not downsampled yet");
    MCB_CUDA(cudaMemcpyAsync(data, pile->downsampled.ptr, (size_t)pile->nnz * sizeof(float), cudaMemcpyDeviceToHost, pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    return MCB_OK;
}

int mcb_pile_similarity(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table,
                        int64_t table_size, float regularization, int absolute, float* similarity_dev, int* outside_table) {
    MCB_REQUIRE(pile != nullptr && feature_genes && log_table && similarity_dev, "pile: null argument");
    MCB_REQUIRE(features >= 1 && features <= pile->genes, "pile: bad feature genes count %lld", (long long)features);
    cudaStream_t stream = pile->stream;
    std::vector<int32_t> slot((size_t)pile->genes, -1);
    for (int64_t f = 0; f < features; ++f) {
        MCB_REQUIRE(feature_genes[f] >= 0 && feature_genes[f] < pile->genes, "pile: feature gene %d out of range", feature_genes[f]);
        MCB_REQUIRE(slot[feature_genes[f]] < 0, "pile: feature gene %d listed twice", feature_genes[f]);
        slot[feature_genes[f]] = (int32_t)f;
    }
    trace_lap("similarity: enter", stream);
    Scratch d_slot, d_table, d_dense, d_flag;
    MCB_TRY(d_slot.alloc((size_t)pile->genes * sizeof(int32_t), stream));
    MCB_TRY(d_table.alloc((size_t)table_size * sizeof(float), stream));
    MCB_TRY(d_dense.alloc((size_t)pile->cells * features * sizeof(float), stream));
    MCB_TRY(d_flag.alloc(sizeof(int), stream));
    MCB_CUDA(cudaMemcpyAsync(d_slot.ptr, slot.data(), (size_t)pile->genes * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_table.ptr, log_table, (size_t)table_size * sizeof(float), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemsetAsync(d_flag.ptr, 0, sizeof(int), stream));
    const float* values = pile->has_downsampled ? pile->downsampled.as<float>() : pile->data;
    MCB_TRY(log_fraction_dense_launch(values, pile->indices, pile->indptr, pile->cells, d_slot.as<int32_t>(), features,
                                      d_table.as<float>(), table_size, regularization, d_dense.as<float>(), d_flag.as<int>(), stream));
    trace_lap("similarity: log_fraction", stream);
    MCB_TRY(mcb_dev_correlate_dense_f32(d_dense.as<float>(), pile->cells, features, features, similarity_dev, pile->cells, absolute,
                                        stream));
    trace_lap("similarity: correlate", stream);
    int flag = 0;
    MCB_CUDA(cudaMemcpyAsync(&flag, d_flag.ptr, sizeof(int), cudaMemcpyDeviceToHost, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    if (outside_table)
        *outside_table = flag;
    return MCB_OK;
}

int mcb_pile_knn_graph(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table,
                       int64_t table_size, float regularization, int absolute, int64_t k, double balanced_ranks_factor,
                       double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                       int* outside_table, mcb_graph** out) {
    MCB_REQUIRE(pile != nullptr && out != nullptr, "pile: null argument");
    Scratch similarity;
    MCB_TRY(similarity.alloc((size_t)pile->cells * pile->cells * sizeof(float), pile->stream));
    MCB_TRY(mcb_pile_similarity(pile, feature_genes, features, log_table, table_size, regularization, absolute,
                                similarity.as<float>(), outside_table));
    return mcb_dev_knn_graph_f32(similarity.as<float>(), pile->cells, k, balanced_ranks_factor, incoming_degree_factor,
                                 outgoing_degree_factor, min_outgoing_degree, pile->stream, out);
}

int64_t mcb_pile_nnz(const mcb_pile* pile) { return pile ? pile->nnz : 0; }

void mcb_pile_free(mcb_pile* pile) { delete pile; }

}  // extern "C"
