// C-ABI entry points of the seed chooser (seeds.cu).
#include "common.cuh"
#include "../../include/mcb200.h"

namespace mcb {
int seeds_launch(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w, const int32_t* in_idx,
                 const int32_t* in_ptr, int64_t nodes, uint64_t random_seed, int64_t max_seeds_count,
                 float min_seed_size_quantile, float max_seed_size_quantile, int32_t* seed_of_nodes, int64_t* seeds_count,
                 cudaStream_t stream);
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

}  // extern "C"
