// C-ABI entry points of the partition optimiser (partitions.cu) and of the device log it scores with.
#include "common.cuh"
#include "glibc_log.cuh"
#include "../../include/mcb200.h"

namespace mcb {
int partitions_launch(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w,
                      const int32_t* in_idx, const int32_t* in_ptr, int64_t nodes, uint32_t random_seed,
                      const float* mass_of_nodes, const double* bounds6, double cooldown_pass, double cooldown_node,
                      int32_t* partition_of_nodes, const int8_t* hot_mask, int64_t mask_count, double cold_temperature,
                      int optimize, int with_orphans, double* score, int64_t* stats, cudaStream_t stream);
}  // namespace mcb

using namespace mcb;

namespace {

__global__ void glibc_log_kernel(const double* __restrict__ x, double* __restrict__ y, int64_t count) {
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < count; i += int64_t(gridDim.x) * blockDim.x)
        y[i] = glibc_log::log(x[i]);
}

template <typename T>
int upload(Scratch& device, const T* host, size_t count, cudaStream_t stream) {
    MCB_TRY(device.alloc(count * sizeof(T), stream));
    if (count)
        MCB_CUDA(cudaMemcpyAsync(device.ptr, host, count * sizeof(T), cudaMemcpyHostToDevice, stream));
    return MCB_OK;
}

// Host arrays in, partition_of_nodes (when optimising) and the score out.
int partitions_host(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w, const int32_t* in_idx,
                    const int32_t* in_ptr, int64_t nodes, uint32_t random_seed, const float* mass_of_nodes, const double* bounds6,
                    double cooldown_pass, double cooldown_node, int32_t* partition_of_nodes, const int8_t* hot_mask,
                    int64_t mask_count, double cold_temperature, int optimize, int with_orphans, double* score, int64_t* stats) {
    MCB_REQUIRE(nodes > 0 && out_ptr && in_ptr && partition_of_nodes && hot_mask && mass_of_nodes && score,
                "optimize_partitions: null argument or no nodes");
    MCB_REQUIRE(out_ptr[0] == 0 && in_ptr[0] == 0, "optimize_partitions: indptr[0] must be 0");
    const size_t out_edges = size_t(out_ptr[nodes]), in_edges = size_t(in_ptr[nodes]);
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch d_out_w, d_out_idx, d_out_ptr, d_in_w, d_in_idx, d_in_ptr, d_mass, d_partition, d_mask;
    MCB_TRY(upload(d_out_w, out_w, out_edges, stream));
    MCB_TRY(upload(d_out_idx, out_idx, out_edges, stream));
    MCB_TRY(upload(d_out_ptr, out_ptr, size_t(nodes) + 1, stream));
    MCB_TRY(upload(d_in_w, in_w, in_edges, stream));
    MCB_TRY(upload(d_in_idx, in_idx, in_edges, stream));
    MCB_TRY(upload(d_in_ptr, in_ptr, size_t(nodes) + 1, stream));
    MCB_TRY(upload(d_mass, mass_of_nodes, size_t(nodes), stream));
    MCB_TRY(upload(d_partition, partition_of_nodes, size_t(nodes), stream));
    MCB_TRY(upload(d_mask, hot_mask, size_t(mask_count), stream));
    MCB_TRY(partitions_launch(d_out_w.as<float>(), d_out_idx.as<int32_t>(), d_out_ptr.as<int32_t>(), d_in_w.as<float>(),
                              d_in_idx.as<int32_t>(), d_in_ptr.as<int32_t>(), nodes, random_seed, d_mass.as<float>(), bounds6,
                              cooldown_pass, cooldown_node, d_partition.as<int32_t>(), d_mask.as<int8_t>(), mask_count,
                              cold_temperature, optimize, with_orphans, score, stats, stream));
    if (optimize) {
        MCB_TRY(copy_to_host(partition_of_nodes, d_partition.ptr, size_t(nodes) * 4, stream));
        MCB_CUDA(cudaStreamSynchronize(stream));
    }
    return MCB_OK;
}

}  // namespace

extern "C" {

int mcb_host_optimize_partitions(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                                 const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                                 int64_t nodes, uint32_t random_seed, const float* mass_of_nodes, double low_mass,
                                 double target_mass, double high_mass, double low_size, double target_size, double high_size,
                                 double cooldown_pass, double cooldown_node, int32_t* partition_of_nodes,
                                 const int8_t* hot_partitions_mask, int64_t mask_count, double cold_temperature, int64_t* stats,
                                 double* score) {
    const double bounds[6] = {low_mass, target_mass, high_mass, low_size, target_size, high_size};
    return partitions_host(outgoing_weights, outgoing_indices, outgoing_indptr, incoming_weights, incoming_indices,
                           incoming_indptr, nodes, random_seed, mass_of_nodes, bounds, cooldown_pass, cooldown_node,
                           partition_of_nodes, hot_partitions_mask, mask_count, cold_temperature, 1, 1, score, stats);
}

int mcb_dev_optimize_partitions(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                                const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                                int64_t nodes, uint32_t random_seed, const float* mass_of_nodes, double low_mass,
                                double target_mass, double high_mass, double low_size, double target_size, double high_size,
                                double cooldown_pass, double cooldown_node, int32_t* partition_of_nodes,
                                const int8_t* hot_partitions_mask, int64_t mask_count, double cold_temperature, int64_t* stats,
                                double* score, void* stream) {
    const double bounds[6] = {low_mass, target_mass, high_mass, low_size, target_size, high_size};
    return partitions_launch(outgoing_weights, outgoing_indices, outgoing_indptr, incoming_weights, incoming_indices,
                             incoming_indptr, nodes, random_seed, mass_of_nodes, bounds, cooldown_pass, cooldown_node,
                             partition_of_nodes, hot_partitions_mask, mask_count, cold_temperature, 1, 1, score, stats,
                             static_cast<cudaStream_t>(stream));
}

int mcb_host_score_partitions(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                              const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                              int64_t nodes, const float* mass_of_nodes, double low_mass, double target_mass, double high_mass,
                              double low_size, double target_size, double high_size, const int32_t* partition_of_nodes,
                              const int8_t* hot_partitions_mask, int64_t mask_count, int with_orphans, double* score) {
    const double bounds[6] = {low_mass, target_mass, high_mass, low_size, target_size, high_size};
    return partitions_host(outgoing_weights, outgoing_indices, outgoing_indptr, incoming_weights, incoming_indices,
                           incoming_indptr, nodes, 0, mass_of_nodes, bounds, 0.0, 0.0, const_cast<int32_t*>(partition_of_nodes),
                           hot_partitions_mask, mask_count, 0.0, 0, with_orphans, score, nullptr);
}

int mcb_dev_score_partitions(const float* outgoing_weights, const int32_t* outgoing_indices, const int32_t* outgoing_indptr,
                             const float* incoming_weights, const int32_t* incoming_indices, const int32_t* incoming_indptr,
                             int64_t nodes, const float* mass_of_nodes, double low_mass, double target_mass, double high_mass,
                             double low_size, double target_size, double high_size, const int32_t* partition_of_nodes,
                             const int8_t* hot_partitions_mask, int64_t mask_count, int with_orphans, double* score,
                             void* stream) {
    const double bounds[6] = {low_mass, target_mass, high_mass, low_size, target_size, high_size};
    return partitions_launch(outgoing_weights, outgoing_indices, outgoing_indptr, incoming_weights, incoming_indices,
                             incoming_indptr, nodes, 0, mass_of_nodes, bounds, 0.0, 0.0, const_cast<int32_t*>(partition_of_nodes),
                             hot_partitions_mask, mask_count, 0.0, 0, with_orphans, score, nullptr,
                             static_cast<cudaStream_t>(stream));
}

int mcb_host_glibc_log_f64(const double* values, double* logs, int64_t count) {
    MCB_REQUIRE(count >= 0 && (count == 0 || (values && logs)), "glibc_log: null argument");
    cudaStream_t stream = nullptr;
    Scratch input, output;
    MCB_TRY(upload(input, values, size_t(count), stream));
    MCB_TRY(output.alloc(size_t(count) * 8, stream));
    if (count) {
        glibc_log_kernel<<<int(std::min<int64_t>(ceil_div(count, 256), 4096)), 256, 0, stream>>>(input.as<double>(),
                                                                                                 output.as<double>(), count);
        MCB_LAUNCH_CHECK();
        MCB_TRY(copy_to_host(logs, output.ptr, size_t(count) * 8, stream));
    }
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

}  // extern "C"
