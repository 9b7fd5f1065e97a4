// The candidate-metacells loop on a device-resident k-NN graph (tools/candidates.py:128-289): the graph, its incoming
// layout and the node masses stay in HBM between the loop's calls of choose_seeds and optimize_partitions; the
// communities' symmetric sub-graphs (:545-546) and their minimal cuts (:548, utilities/computation.py:2368-2449) are
// computed for all communities of a round at once, one CTA per community.
#include <algorithm>
#include <vector>

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

struct mcb_candidates {
    const mcb_graph* graph = nullptr;
    int64_t nodes = 0, edges = 0;
    const float* out_w = nullptr;
    const int32_t* out_idx = nullptr;
    const int32_t* out_ptr = nullptr;
    Scratch in_w, in_idx, in_ptr, mass;
};

namespace {

constexpr int64_t MAX_CUT_COMMUNITY = 4096;  // the cut kernel's work matrix is doubles of the community's size squared
constexpr int CUT_THREADS = 128;

template <typename T>
int upload(Scratch& device, const T* host, size_t count, cudaStream_t stream) {
    MCB_TRY(device.alloc(count * sizeof(T), stream));
    if (count)
        MCB_CUDA(cudaMemcpyAsync(device.ptr, host, count * sizeof(T), cudaMemcpyHostToDevice, stream));
    return MCB_OK;
}

// One warp per listed node: its outgoing edges that stay inside its community are added to both mirrored entries of the
// community's dense block, which so holds outgoing[mask, :][:, mask] + its transpose (two float32 terms per entry: the sum
// does not depend on their order).
__global__ void community_weights_kernel(const float* __restrict__ out_w, const int32_t* __restrict__ out_idx,
                                         const int32_t* __restrict__ out_ptr, const int32_t* __restrict__ members,
                                         int64_t listed, const int32_t* __restrict__ community_of_node,
                                         const int32_t* __restrict__ slot_of_node, const int64_t* __restrict__ member_offsets,
                                         const int64_t* __restrict__ block_offsets, float* __restrict__ dense) {
    const int64_t item = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (item >= listed) return;
    const int node = members[item];
    const int community = community_of_node[node];
    const int64_t size = member_offsets[community + 1] - member_offsets[community];
    const int64_t row = slot_of_node[node];
    float* block = dense + block_offsets[community];
    for (int edge = out_ptr[node] + lane; edge < out_ptr[node + 1]; edge += 32) {
        const int other = out_idx[edge];
        if (community_of_node[other] != community) continue;
        const int64_t column = slot_of_node[other];
        const float weight = out_w[edge];
        atomicAdd(block + row * size + column, weight);
        atomicAdd(block + column * size + row, weight);
    }
}

// The minimal cut of every listed community's symmetric sub-graph (Stoer & Wagner), one CTA per community, in float64 as
// igraph's capacities are.  Each phase grows a set from the lowest live vertex by the most tightly connected vertex (ties:
// the lowest index); the last vertex added, with everything merged into it so far, against the rest is the phase's cut; the
// first phase with the least weight wins and its last vertex's side is the first partition.  Then the weights of the
// edges inside each side and across the cut are added up (in float32) in the order computation.py:2414-2426 walks them (source
// ascending, target ascending below the source).
__global__ void __launch_bounds__(CUT_THREADS)
    min_cut_kernel(const float* __restrict__ dense, const int64_t* __restrict__ member_offsets,
                   const int64_t* __restrict__ block_offsets, double* __restrict__ work, double* __restrict__ key_of,
                   int32_t* __restrict__ root_of, uint8_t* __restrict__ state_of, int8_t* __restrict__ is_second,
                   double* __restrict__ totals) {
    const int community = blockIdx.x;
    const int n = int(member_offsets[community + 1] - member_offsets[community]);
    const int64_t first_member = member_offsets[community];
    const float* D = dense + block_offsets[community];
    double* W = work + block_offsets[community];
    double* key = key_of + first_member;
    int32_t* root = root_of + first_member;
    uint8_t* state = state_of + first_member;  // bit 0: the vertex is live (not merged away); bit 1: in the phase's set
    int8_t* second = is_second + first_member;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ double warp_key[CUT_THREADS / 32];
    __shared__ int warp_vertex[CUT_THREADS / 32];

    for (int64_t at = tid; at < int64_t(n) * n; at += CUT_THREADS) W[at] = double(D[at]);
    for (int v = tid; v < n; v += CUT_THREADS) {
        root[v] = v;
        state[v] = 1;
        second[v] = 0;
    }
    __syncthreads();
    double best = 1e300;
    for (int phase = 0; phase + 1 < n; ++phase) {
        for (int v = tid; v < n; v += CUT_THREADS) {
            key[v] = 0.0;
            state[v] &= 1;
        }
        __syncthreads();
        int last = -1, previous = -1;
        double last_key = 0.0;
        for (int step = 0; step < n - phase; ++step) {
            double my_key = -1.0;
            int my_vertex = 0x7fffffff;
            for (int v = tid; v < n; v += CUT_THREADS)
                if (state[v] == 1 && key[v] > my_key) {  // ascending v: the lowest index among equal keys
                    my_key = key[v];
                    my_vertex = v;
                }
#pragma unroll
            for (int offset = 16; offset > 0; offset >>= 1) {
                const double other_key = __shfl_xor_sync(0xffffffffu, my_key, offset);
                const int other_vertex = __shfl_xor_sync(0xffffffffu, my_vertex, offset);
                if (other_key > my_key || (other_key == my_key && other_vertex < my_vertex)) {
                    my_key = other_key;
                    my_vertex = other_vertex;
                }
            }
            if (lane == 0) {
                warp_key[warp] = my_key;
                warp_vertex[warp] = my_vertex;
            }
            __syncthreads();
            my_key = warp_key[0];
            my_vertex = warp_vertex[0];
#pragma unroll
            for (int w = 1; w < CUT_THREADS / 32; ++w)
                if (warp_key[w] > my_key || (warp_key[w] == my_key && warp_vertex[w] < my_vertex)) {
                    my_key = warp_key[w];
                    my_vertex = warp_vertex[w];
                }
            previous = last;
            last = my_vertex;
            last_key = my_key;
            const double* row = W + int64_t(last) * n;
            for (int v = tid; v < n; v += CUT_THREADS) {
                if (v == last)
                    state[v] |= 2;
                else if (state[v] == 1)
                    key[v] += row[v];
            }
            __syncthreads();
        }
        if (last_key < best) {
            best = last_key;
            for (int v = tid; v < n; v += CUT_THREADS) second[v] = root[v] != last;
        }
        double* kept = W + int64_t(previous) * n;
        const double* gone = W + int64_t(last) * n;
        for (int v = tid; v < n; v += CUT_THREADS) {
            if ((state[v] & 1) && v != last && v != previous) {
                const double sum = kept[v] + gone[v];
                kept[v] = sum;
                W[int64_t(v) * n + previous] = sum;
            }
            if (root[v] == last) root[v] = previous;
        }
        __syncthreads();
        if (tid == 0) state[last] = 0;
        __syncthreads();
    }
    if (tid == 0) {
        // float32 sums: `total += weight` of a Python float and a numpy float32 stays float32 (NumPy >= 2 promotion)
        float first_total = 0.0f, second_total = 0.0f, cut_total = 0.0f;
        for (int source = 0; source < n; ++source)
            for (int target = 0; target < source; ++target) {
                const float weight = D[int64_t(source) * n + target];
                if (!(weight > 0.0f)) continue;
                if (second[source] != second[target])
                    cut_total += weight;
                else if (second[source])
                    second_total += weight;
                else
                    first_total += weight;
            }
        totals[3 * int64_t(community) + 0] = double(first_total);
        totals[3 * int64_t(community) + 1] = double(second_total);
        totals[3 * int64_t(community) + 2] = double(cut_total);
    }
}

// The listed communities on the device: members, offsets, and for every node of the graph its community and its position
// in it (-1: not listed).
struct Listed {
    Scratch members, member_offsets, block_offsets, community_of_node, slot_of_node, dense;
    int64_t listed = 0, entries = 0;
};

int upload_listed(const mcb_candidates* session, const int32_t* members, const int64_t* member_offsets, int64_t communities,
                  int64_t max_size, Listed& device, cudaStream_t stream) {
    MCB_REQUIRE(member_offsets[0] == 0, "candidates: member_offsets[0] must be 0");
    const int64_t listed = member_offsets[communities];
    std::vector<int32_t> community_of_node(size_t(session->nodes), -1), slot_of_node(size_t(session->nodes), -1);
    std::vector<int64_t> block_offsets(size_t(communities) + 1, 0);
    for (int64_t community = 0; community < communities; ++community) {
        const int64_t size = member_offsets[community + 1] - member_offsets[community];
        MCB_REQUIRE(size >= 0, "candidates: member_offsets must not decrease");
        MCB_REQUIRE(max_size <= 0 || size <= max_size, "candidates: a community of %lld nodes is beyond the %lld the cut supports",
                    (long long)size, (long long)max_size);
        block_offsets[size_t(community) + 1] = block_offsets[size_t(community)] + size * size;
        for (int64_t slot = 0; slot < size; ++slot) {
            const int32_t node = members[member_offsets[community] + slot];
            MCB_REQUIRE(node >= 0 && node < session->nodes, "candidates: member %d is not a node of the graph", node);
            MCB_REQUIRE(community_of_node[size_t(node)] < 0, "candidates: node %d is listed twice", node);
            community_of_node[size_t(node)] = int32_t(community);
            slot_of_node[size_t(node)] = int32_t(slot);
        }
    }
    device.listed = listed;
    device.entries = block_offsets[size_t(communities)];
    MCB_TRY(upload(device.members, members, size_t(listed), stream));
    MCB_TRY(upload(device.member_offsets, member_offsets, size_t(communities) + 1, stream));
    MCB_TRY(upload(device.block_offsets, block_offsets.data(), size_t(communities) + 1, stream));
    MCB_TRY(upload(device.community_of_node, community_of_node.data(), size_t(session->nodes), stream));
    MCB_TRY(upload(device.slot_of_node, slot_of_node.data(), size_t(session->nodes), stream));
    MCB_TRY(device.dense.alloc(size_t(device.entries) * sizeof(float), stream));
    MCB_CUDA(cudaMemsetAsync(device.dense.ptr, 0, size_t(device.entries) * sizeof(float), stream));
    if (listed > 0) {
        const int64_t blocks = ceil_div(listed * 32, 256);
        MCB_PROFILED("community_weights", stream,
                     community_weights_kernel<<<unsigned(blocks), 256, 0, stream>>>(
                         session->out_w, session->out_idx, session->out_ptr, device.members.as<int32_t>(), listed,
                         device.community_of_node.as<int32_t>(), device.slot_of_node.as<int32_t>(),
                         device.member_offsets.as<int64_t>(), device.block_offsets.as<int64_t>(), device.dense.as<float>()));
        MCB_LAUNCH_CHECK();
    }
    // the host vectors above are pageable: their copies have been staged by the time cudaMemcpyAsync returns
    return MCB_OK;
}

}  // namespace

extern "C" {

int mcb_candidates_create(const mcb_graph* graph, const float* mass_of_nodes, mcb_candidates** out) {
    MCB_REQUIRE(graph && mass_of_nodes && out, "candidates_create: null argument");
    *out = nullptr;
    const float* weights = nullptr;
    const int32_t* indices = nullptr;
    const int32_t* indptr = nullptr;
    MCB_TRY(mcb_graph_device_arrays(graph, &weights, nullptr, &indices, &indptr));
    const int64_t nodes = mcb_graph_size(graph), edges = mcb_graph_nnz(graph);
    MCB_REQUIRE(nodes > 0, "candidates_create: empty graph");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    MCB_TRY(graph_wait_complete(graph, stream));
    mcb_candidates* session = new mcb_candidates();
    session->graph = graph;
    session->nodes = nodes;
    session->edges = edges;
    session->out_w = weights;
    session->out_idx = indices;
    session->out_ptr = indptr;
    int code = MCB_OK;
    // tools/candidates.py:135: the incoming (column-major) layout of the same edges
    if ((code = session->in_w.alloc(size_t(edges) * sizeof(float), stream)) == MCB_OK &&
        (code = session->in_idx.alloc(size_t(edges) * sizeof(int32_t), stream)) == MCB_OK &&
        (code = session->in_ptr.alloc(size_t(nodes + 1) * sizeof(int32_t), stream)) == MCB_OK &&
        (code = relayout_launch(weights, indices, indptr, nodes, nodes, edges, session->in_w.as<float>(),
                                session->in_idx.as<int32_t>(), session->in_ptr.as<int32_t>(), stream)) == MCB_OK &&
        (code = upload(session->mass, mass_of_nodes, size_t(nodes), stream)) == MCB_OK) {
        if (cudaStreamSynchronize(stream) != cudaSuccess)
            code = fail(MCB_ERR_CUDA, "candidates_create: the incoming layout failed");
    }
    if (code != MCB_OK) {
        delete session;
        return code;
    }
    *out = session;
    return MCB_OK;
}

void mcb_candidates_free(mcb_candidates* session) { delete session; }

int mcb_candidates_choose_seeds(mcb_candidates* session, uint64_t random_seed, int64_t max_seeds_count,
                                float min_seed_size_quantile, float max_seed_size_quantile, int32_t* seed_of_nodes,
                                int64_t* seeds_count) {
    MCB_REQUIRE(session && seed_of_nodes && seeds_count, "candidates_choose_seeds: null argument");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch seeds;
    MCB_TRY(upload(seeds, seed_of_nodes, size_t(session->nodes), stream));
    MCB_TRY(seeds_launch(session->out_w, session->out_idx, session->out_ptr, session->in_w.as<float>(),
                         session->in_idx.as<int32_t>(), session->in_ptr.as<int32_t>(), session->nodes, random_seed,
                         max_seeds_count, min_seed_size_quantile, max_seed_size_quantile, seeds.as<int32_t>(), seeds_count,
                         stream));
    MCB_TRY(copy_to_host(seed_of_nodes, seeds.ptr, size_t(session->nodes) * sizeof(int32_t), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_candidates_optimize_partitions(mcb_candidates* session, uint32_t random_seed, double low_mass, double target_mass,
                                       double high_mass, double low_size, double target_size, double high_size,
                                       double cooldown_pass, double cooldown_node, int32_t* partition_of_nodes,
                                       const int8_t* hot_partitions_mask, int64_t mask_count, double cold_temperature,
                                       int64_t* stats, double* score) {
    MCB_REQUIRE(session && partition_of_nodes && hot_partitions_mask && score, "candidates_optimize_partitions: null argument");
    MCB_REQUIRE(mask_count > 0, "candidates_optimize_partitions: empty hot mask");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch partition, mask;
    MCB_TRY(upload(partition, partition_of_nodes, size_t(session->nodes), stream));
    MCB_TRY(upload(mask, hot_partitions_mask, size_t(mask_count), stream));
    const double bounds[6] = {low_mass, target_mass, high_mass, low_size, target_size, high_size};
    MCB_TRY(partitions_launch(session->out_w, session->out_idx, session->out_ptr, session->in_w.as<float>(),
                              session->in_idx.as<int32_t>(), session->in_ptr.as<int32_t>(), session->nodes, random_seed,
                              session->mass.as<float>(), bounds, cooldown_pass, cooldown_node, partition.as<int32_t>(),
                              mask.as<int8_t>(), mask_count, cold_temperature, 1, 1, score, stats, stream));
    MCB_TRY(copy_to_host(partition_of_nodes, partition.ptr, size_t(session->nodes) * sizeof(int32_t), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_candidates_community_weights(mcb_candidates* session, const int32_t* members, const int64_t* member_offsets,
                                     int64_t communities, float* dense) {
    MCB_REQUIRE(session && member_offsets && communities >= 0, "candidates_community_weights: null argument");
    MCB_REQUIRE(members || member_offsets[communities] == 0, "candidates_community_weights: null members");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Listed device;
    MCB_TRY(upload_listed(session, members, member_offsets, communities, 0, device, stream));
    MCB_REQUIRE(dense || device.entries == 0, "candidates_community_weights: null output");
    MCB_TRY(copy_to_host(dense, device.dense.ptr, size_t(device.entries) * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_candidates_min_cuts(mcb_candidates* session, const int32_t* members, const int64_t* member_offsets,
                            int64_t communities, int8_t* is_second, double* totals) {
    MCB_REQUIRE(session && member_offsets && communities >= 0, "candidates_min_cuts: null argument");
    if (communities == 0) return MCB_OK;
    MCB_REQUIRE(members && is_second && totals, "candidates_min_cuts: null argument");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Listed device;
    MCB_TRY(upload_listed(session, members, member_offsets, communities, MAX_CUT_COMMUNITY, device, stream));
    Scratch work, key, root, state, second, sums;
    MCB_TRY(work.alloc(size_t(device.entries) * sizeof(double), stream));
    MCB_TRY(key.alloc(size_t(device.listed) * sizeof(double), stream));
    MCB_TRY(root.alloc(size_t(device.listed) * sizeof(int32_t), stream));
    MCB_TRY(state.alloc(size_t(device.listed), stream));
    MCB_TRY(second.alloc(size_t(device.listed), stream));
    MCB_TRY(sums.alloc(size_t(communities) * 3 * sizeof(double), stream));
    MCB_PROFILED("min_cut", stream,
                 min_cut_kernel<<<unsigned(communities), CUT_THREADS, 0, stream>>>(
                     device.dense.as<float>(), device.member_offsets.as<int64_t>(), device.block_offsets.as<int64_t>(),
                     work.as<double>(), key.as<double>(), root.as<int32_t>(), state.as<uint8_t>(), second.as<int8_t>(),
                     sums.as<double>()));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(is_second, second.ptr, size_t(device.listed), stream));
    MCB_TRY(copy_to_host(totals, sums.ptr, size_t(communities) * 3 * sizeof(double), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

}  // extern "C"
