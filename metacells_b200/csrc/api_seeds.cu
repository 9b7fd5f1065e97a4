// C-ABI entry points of the seed chooser (seeds.cu).
#include <algorithm>

#include "common.cuh"
#include "../../include/mcb200.h"

namespace mcb {
int seeds_launch(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w, const int32_t* in_idx,
                 const int32_t* in_ptr, int64_t nodes, uint64_t random_seed, int64_t max_seeds_count,
                 float min_seed_size_quantile, float max_seed_size_quantile, int32_t* seed_of_nodes, int64_t* seeds_count,
                 cudaStream_t stream);
int partitions_launch(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w,
                      const int32_t* in_idx, const int32_t* in_ptr, int64_t nodes, uint32_t random_seed,
                      const float* mass_of_nodes, const double* bounds6, double cooldown_pass, double cooldown_node,
                      int32_t* partition_of_nodes, const int8_t* hot_mask, int64_t mask_count, double cold_temperature,
                      int optimize, int with_orphans, double* score, int64_t* stats, cudaStream_t stream);
int relayout_launch(const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands_in, int64_t bands_out,
                    int64_t nnz, float* out_data, int32_t* out_indices, int32_t* out_indptr, cudaStream_t stream);
int graph_wait_complete(const mcb_graph* graph, cudaStream_t stream);
}  // namespace mcb

using namespace mcb;

namespace {

template <typename T>
int upload(Scratch& device, const T* host, size_t count, cudaStream_t stream) {
    MCB_TRY(device.alloc(count * sizeof(T), stream));
    if (count)
        MCB_CUDA(cudaMemcpyAsync(device.ptr, host, count * sizeof(T), cudaMemcpyHostToDevice, stream));
    return MCB_OK;
}

}  // namespace

extern "C" {

int mcb_host_choose_seeds(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                          const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                          int64_t nodes, uint64_t random_seed, int64_t max_seeds_count, float min_seed_size_quantile,
                          float max_seed_size_quantile, int32_t* seed_of_nodes, int64_t* seeds_count) {
    MCB_REQUIRE(nodes > 0 && outgoing_indptr && incoming_indptr && seed_of_nodes && seeds_count,
                "choose_seeds: null argument or no nodes");
    MCB_REQUIRE(outgoing_indptr[0] == 0 && incoming_indptr[0] == 0, "choose_seeds: indptr[0] must be 0");
    const size_t out_edges = size_t(outgoing_indptr[nodes]), in_edges = size_t(incoming_indptr[nodes]);
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch d_out_w, d_out_idx, d_out_ptr, d_in_w, d_in_idx, d_in_ptr, d_seeds;
    MCB_TRY(upload(d_out_w, outgoing_weights, out_edges, stream));
    MCB_TRY(upload(d_out_idx, outgoing_indices, out_edges, stream));
    MCB_TRY(upload(d_out_ptr, outgoing_indptr, size_t(nodes) + 1, stream));
    MCB_TRY(upload(d_in_w, incoming_weights, in_edges, stream));
    MCB_TRY(upload(d_in_idx, incoming_indices, in_edges, stream));
    MCB_TRY(upload(d_in_ptr, incoming_indptr, size_t(nodes) + 1, stream));
    MCB_TRY(upload(d_seeds, seed_of_nodes, size_t(nodes), stream));
    MCB_TRY(seeds_launch(d_out_w.as<float>(), d_out_idx.as<int32_t>(), d_out_ptr.as<int32_t>(), d_in_w.as<float>(),
                         d_in_idx.as<int32_t>(), d_in_ptr.as<int32_t>(), nodes, random_seed, max_seeds_count,
                         min_seed_size_quantile, max_seed_size_quantile, d_seeds.as<int32_t>(), seeds_count, stream));
    MCB_TRY(copy_to_host(seed_of_nodes, d_seeds.ptr, size_t(nodes) * 4, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_dev_choose_seeds(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                         const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                         int64_t nodes, uint64_t random_seed, int64_t max_seeds_count, float min_seed_size_quantile,
                         float max_seed_size_quantile, int32_t* seed_of_nodes, int64_t* seeds_count, void* stream) {
    return seeds_launch(outgoing_weights, outgoing_indices, outgoing_indptr, incoming_weights, incoming_indices, incoming_indptr,
                        nodes, random_seed, max_seeds_count, min_seed_size_quantile, max_seed_size_quantile, seed_of_nodes,
                        seeds_count, static_cast<cudaStream_t>(stream));
}

int mcb_graph_seed_partition(const mcb_graph* graph, const float* mass_of_nodes, uint64_t random_seed, int64_t max_seeds_count,
                             float min_seed_size_quantile, float max_seed_size_quantile, double low_mass, double target_mass,
                             double high_mass, double low_size, double target_size, double high_size, double cooldown_pass,
                             double cooldown_node, double cold_temperature, int32_t* partition_of_nodes, int64_t* seeds_count,
                             double* score) {
    MCB_REQUIRE(graph && mass_of_nodes && partition_of_nodes && seeds_count && score, "graph_seed_partition: null argument");
    const float* weights = nullptr;
    const int32_t* indices = nullptr;
    const int32_t* indptr = nullptr;
    MCB_TRY(mcb_graph_device_arrays(graph, &weights, nullptr, &indices, &indptr));
    const int64_t nodes = mcb_graph_size(graph), edges = mcb_graph_nnz(graph);
    MCB_REQUIRE(nodes > 0, "graph_seed_partition: empty graph");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    MCB_TRY(graph_wait_complete(graph, stream));  // the graph's last kernels may still be queued on the stream that built it
    // the incoming layout (tools/candidates.py:135: to_layout(..., "column_major")), on the device
    Scratch in_w, in_idx, in_ptr, d_mass, d_partition, d_mask;
    MCB_TRY(in_w.alloc(size_t(edges) * sizeof(float), stream));
    MCB_TRY(in_idx.alloc(size_t(edges) * sizeof(int32_t), stream));
    MCB_TRY(in_ptr.alloc(size_t(nodes + 1) * sizeof(int32_t), stream));
    MCB_TRY(relayout_launch(weights, indices, indptr, nodes, nodes, edges, in_w.as<float>(), in_idx.as<int32_t>(),
                            in_ptr.as<int32_t>(), stream));
    MCB_TRY(upload(d_mass, mass_of_nodes, size_t(nodes), stream));
    MCB_TRY(d_partition.alloc(size_t(nodes) * sizeof(int32_t), stream));
    MCB_CUDA(cudaMemsetAsync(d_partition.ptr, 0xff, size_t(nodes) * sizeof(int32_t), stream));  // -1: no seed yet
    MCB_TRY(seeds_launch(weights, indices, indptr, in_w.as<float>(), in_idx.as<int32_t>(), in_ptr.as<int32_t>(), nodes, random_seed,
                         max_seeds_count, min_seed_size_quantile, max_seed_size_quantile, d_partition.as<int32_t>(), seeds_count,
                         stream));
    const int64_t mask_count = std::max<int64_t>(*seeds_count, 2);  // every community hot (tools/candidates.py:189)
    MCB_TRY(d_mask.alloc(size_t(mask_count), stream));
    MCB_CUDA(cudaMemsetAsync(d_mask.ptr, 1, size_t(mask_count), stream));
    const double bounds[6] = {low_mass, target_mass, high_mass, low_size, target_size, high_size};
    MCB_TRY(partitions_launch(weights, indices, indptr, in_w.as<float>(), in_idx.as<int32_t>(), in_ptr.as<int32_t>(), nodes,
                              uint32_t(random_seed), d_mass.as<float>(), bounds, cooldown_pass, cooldown_node,
                              d_partition.as<int32_t>(), d_mask.as<int8_t>(), mask_count, cold_temperature, 1, 1, score, nullptr,
                              stream));
    MCB_TRY(copy_to_host(partition_of_nodes, d_partition.ptr, size_t(nodes) * sizeof(int32_t), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

}  // extern "C"
