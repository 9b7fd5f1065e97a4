// partitions.cu -- the candidate-metacell partition optimiser.
//
// Replaces metacells/partitions.cpp: optimize_partitions (:1234-1302), score_partitions (:1304-1338) and
// struct OptimizePartitions (:243-1232) of the reference.  The reference is a strictly sequential
// annealing loop (every accepted move changes the state the next node sees, and the random stream
// is consumed only by moves); results here are bit-identical to it for a given random_seed:
//   - the float64 operation order of every accumulated sum is kept (per-partition cold diffs in
//     merged-neighbour order, candidate sums in partition order, the running temperature product);
//   - log2 is glibc's log (glibc_log.cuh) times 1/log(2), as extensions.h:131-138;
//   - std::minstd_rand, libstdc++'s std::shuffle / uniform_int_distribution / generate_canonical
//     are restated;
//   - the reference's three candidate lists alias one vector (partitions.cpp:26-29 with
//     extensions.h:483-486), so every admissible partition is one candidate list in partition order.
//
// What is parallel: one persistent CTA runs the pass loop.  Within a pass the visiting order, the
// temperature of every visit and the set of skipped nodes are known up front, and a node that finds no
// candidate changes nothing but its own temperature, so the next W non-skipped nodes are evaluated
// SPECULATIVELY, one warp each, against the current state; the first of them (in visiting order) that
// has candidates draws from the generator and moves, the evaluations after it are discarded, the ones
// before it are committed as "unimproved".  About one visit in ten moves, so a batch retires ~8 nodes.
// Inside a node the O(P) partition scan, the O(degree) neighbour re-scoring (a log each) and the
// O(degree) state update of a move are data-parallel over lanes / the whole CTA.
//
// Layout in HBM: score table Entry[N][P] (node-major, 32 B entries: a node's row is scanned coalesced);
// self[N] = the entry of every node for its own partition plus that partition (one gather per
// neighbour instead of two dependent ones); the merged (union) neighbour list of every node, built
// once (u, outgoing weight or 0, incoming weight or 0) in ascending u.
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "common.cuh"
#include "glibc_log.cuh"

namespace mcb {

namespace {

constexpr double EPSILON = 1e-6;                 // extensions.h:131
constexpr int OPT_THREADS = 512;                 // 16 warps: the speculation depth the move rate supports
constexpr int OPT_MAX_WARPS = OPT_THREADS / 32;
constexpr unsigned FULL = 0xffffffffu;

// One merged neighbour of a node: the other node, the weight of node -> other (0 if absent) and of other -> node.
struct alignas(16) Neighbour {
    int32_t node;
    float outgoing;
    float incoming;
    int32_t unused;
};

struct alignas(32) Entry {
    double outgoing;
    double incoming;
    double score;
    long long partition;  // only meaningful in self[]
};

__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double ddiv(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ double clamp0(double x) { return x < 0.0 ? 0.0 : x; }  // std::max(x, 0.0)

// The glibc log tables, staged in shared memory (lanes index them divergently).
struct LogTables {
    const unsigned long long* tab;
    const unsigned long long* poly;
    const unsigned long long* poly1;
};
constexpr int LOG_TABLE_WORDS = 256 + 8 + 16;

__device__ __forceinline__ void stage_log_tables(unsigned long long* dst) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) dst[i] = glibc_log_tab[i];
    if (threadIdx.x < 5) dst[256 + threadIdx.x] = glibc_log_poly[threadIdx.x];
    if (threadIdx.x < 11) dst[264 + threadIdx.x] = glibc_log_poly1[threadIdx.x];
}
__device__ __forceinline__ LogTables log_tables(const unsigned long long* staged) {
    return LogTables{staged, staged + 256, staged + 264};
}

// extensions.h:132,135-138: log(x) * (1.0 / log(2.0)); 1/log(2) = 0x3FF71547652B82FE with glibc's log(2).
__device__ __forceinline__ double mc_log2(double x, const LogTables& t) {
    return dmul(glibc_log::log(x, t.tab, t.poly, t.poly1), __longlong_as_double(0x3FF71547652B82FELL));
}

// partitions.cpp:5-19
__device__ __forceinline__ double mass_factor(double mass, double low, double target, double high, const LogTables& t) {
    const double inner_slope = 0.02, outer_slope = 0.5;
    const double inner = dsub(1.0, inner_slope);
    const double outer = dsub(ddiv(1.0, dsub(1.0, outer_slope)), 1.0);
    if (mass < low) return mc_log2(ddiv(dmul(inner, low), dadd(low, dmul(outer, dsub(low, mass)))), t);
    if (mass < target) return mc_log2(dsub(1.0, ddiv(dmul(inner_slope, dsub(target, mass)), dsub(target, low))), t);
    if (mass < high) return mc_log2(dsub(1.0, ddiv(dmul(inner_slope, dsub(mass, target)), dsub(high, target))), t);
    return mc_log2(ddiv(dmul(inner, low), dadd(low, dmul(outer, dsub(mass, high)))), t);
}

// NodeScore::rescore / connectivity, partitions.cpp:62,73-79
__device__ __forceinline__ double rescore(double outgoing, double incoming, const LogTables& t) {
    return dmul(mc_log2(dadd(EPSILON, dmul(outgoing, incoming)), t), 0.5);  // "/ 2.0" is exact either way
}
__device__ __forceinline__ int connectivity(double outgoing, double incoming) {
    return int(outgoing > EPSILON) + int(incoming > 0.0);
}

struct Bounds {
    double low_mass, target_mass, high_mass, low_size, target_size, high_size;
};

// ---------------------------------------------------------------------------------------------------------
// Construction (partitions.cpp:88-236, 264-320): all sums keep the reference's order.
// ---------------------------------------------------------------------------------------------------------

// flags[0] = max partition, flags[1] = 1 when a partition index is negative.
__global__ void partition_range_kernel(const int32_t* __restrict__ partition_of_nodes, int nodes, int* __restrict__ flags) {
    int most = -1, negative = 0;
    for (int node = blockIdx.x * blockDim.x + threadIdx.x; node < nodes; node += gridDim.x * blockDim.x) {
        const int partition = partition_of_nodes[node];
        most = max(most, partition);
        negative |= partition < 0;
    }
    most = __reduce_max_sync(FULL, most);
    negative = __any_sync(FULL, negative);
    if ((threadIdx.x & 31) == 0) {
        atomicMax(&flags[0], most);
        if (negative) atomicOr(&flags[1], 1);
    }
}

// Merged neighbour list of every node: ascending u over outgoing U incoming, a node on both lists once.
// Node v owns slots [out_ptr[v] + in_ptr[v], +union_count[v]).
__global__ void build_union_kernel(const float* __restrict__ out_w, const int32_t* __restrict__ out_idx,
                                   const int32_t* __restrict__ out_ptr, const float* __restrict__ in_w,
                                   const int32_t* __restrict__ in_idx, const int32_t* __restrict__ in_ptr, int nodes,
                                   Neighbour* __restrict__ neighbours, int32_t* __restrict__ union_count) {
    const int node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= nodes) return;
    int o = out_ptr[node], i = in_ptr[node];
    const int o_end = out_ptr[node + 1], i_end = in_ptr[node + 1];
    int64_t at = int64_t(o) + i;
    const int64_t begin = at;
    while (o < o_end || i < i_end) {
        const int o_node = o < o_end ? out_idx[o] : nodes;
        const int i_node = i < i_end ? in_idx[i] : nodes;
        const int other = min(o_node, i_node);
        const bool is_out = o_node == other, is_in = i_node == other;
        neighbours[at] = Neighbour{other, is_out ? out_w[o] : 0.0f, is_in ? in_w[i] : 0.0f, 0};
        ++at;
        o += is_out;
        i += is_in;
    }
    union_count[node] = int(at - begin);
}

__global__ void clear_table_kernel(Entry* __restrict__ table, int64_t entries) {
    for (int64_t at = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; at < entries; at += int64_t(gridDim.x) * blockDim.x)
        table[at] = Entry{0.0, 0.0, 0.0, 0};
}

// partitions.cpp:143-185: totals of node v towards every partition, edges in ascending order of the other node
// (the reference adds the incoming ones while visiting the source nodes in ascending order: the same order).
// status[0] |= 1: outgoing weights do not sum to 1 (:182-183); |= 2: no incoming weight (:123).
__global__ void accumulate_totals_kernel(const float* __restrict__ out_w, const int32_t* __restrict__ out_idx,
                                         const int32_t* __restrict__ out_ptr, const float* __restrict__ in_w,
                                         const int32_t* __restrict__ in_idx, const int32_t* __restrict__ in_ptr,
                                         const int32_t* __restrict__ partition_of_nodes, int nodes, int partitions,
                                         Entry* __restrict__ table, double* __restrict__ incoming_log2,
                                         int* __restrict__ status) {
    __shared__ unsigned long long staged[LOG_TABLE_WORDS];
    stage_log_tables(staged);
    __syncthreads();
    const LogTables t = log_tables(staged);
    const int node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= nodes) return;
    Entry* row = table + int64_t(node) * partitions;
    double total_outgoing = 0.0;
    for (int e = out_ptr[node]; e < out_ptr[node + 1]; ++e) {
        const double weight = out_w[e];
        total_outgoing = dadd(total_outgoing, weight);
        Entry& entry = row[partition_of_nodes[out_idx[e]]];
        entry.outgoing = clamp0(dadd(entry.outgoing, weight));
    }
    double total_incoming = 0.0;
    for (int e = in_ptr[node]; e < in_ptr[node + 1]; ++e) {
        const double weight = in_w[e];
        total_incoming = dadd(total_incoming, weight);
        Entry& entry = row[partition_of_nodes[in_idx[e]]];
        entry.incoming = clamp0(dadd(entry.incoming, weight));
    }
    int bad = 0;
    if (!(total_outgoing >= 1 - EPSILON) || !(total_outgoing < 1 + EPSILON)) bad |= 1;
    if (!(total_incoming > 0)) bad |= 2;
    if (bad) atomicOr(status, bad);
    incoming_log2[node] = total_incoming > 0 ? mc_log2(total_incoming, t) : 0.0;  // :124
}

// Scores of the totals, and `linked_bits`: bit p of node v's words is set iff v has any weight towards partition p, that is
// iff table[v][p] is not the default entry (0, 0, log2(EPSILON) / 2).  One lane per (node, partition padded to 32).
__global__ void rescore_table_kernel(Entry* __restrict__ table, int nodes, int partitions, int words,
                                     uint32_t* __restrict__ linked_bits) {
    __shared__ unsigned long long staged[LOG_TABLE_WORDS];
    stage_log_tables(staged);
    __syncthreads();
    const LogTables t = log_tables(staged);
    const double default_score = rescore(0.0, 0.0, t);  // NodeScore(): log2(EPSILON) / 2
    const int64_t items = int64_t(nodes) * words * 32;
    for (int64_t at = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; at < items; at += int64_t(gridDim.x) * blockDim.x) {
        const int64_t word = at >> 5;
        const int node = int(word / words);
        const int p = int(word - int64_t(node) * words) * 32 + int(at & 31);
        bool linked = false;
        if (p < partitions) {
            Entry* slot = table + int64_t(node) * partitions + p;
            Entry entry = *slot;
            linked = entry.outgoing != 0.0 || entry.incoming != 0.0;
            entry.score = linked ? rescore(entry.outgoing, entry.incoming, t) : default_score;
            *slot = entry;
        }
        const unsigned mask = __ballot_sync(FULL, linked);
        if ((at & 31) == 0) linked_bits[word] = mask;
    }
}

// Per partition, over the nodes in ascending order (partitions.cpp:97, 208-236): size, mass, score.
// status |= 4 when a partition is empty (:101).  One thread per partition; node data is broadcast.
__global__ void partition_totals_kernel(const int32_t* __restrict__ partition_of_nodes, const float* __restrict__ mass_of_nodes,
                                        const Entry* __restrict__ table, int nodes, int partitions,
                                        int32_t* __restrict__ sizes, double* __restrict__ masses,
                                        double* __restrict__ scores, int* __restrict__ status) {
    const int partition = blockIdx.x * blockDim.x + threadIdx.x;
    if (partition >= partitions) return;
    int size = 0;
    double mass = 0.0, score = 0.0;
    for (int node = 0; node < nodes; ++node) {
        if (partition_of_nodes[node] != partition) continue;
        ++size;
        mass = dadd(mass, double(mass_of_nodes[node]));
        score = dadd(score, table[int64_t(node) * partitions + partition].score);
    }
    sizes[partition] = size;
    masses[partition] = mass;
    scores[partition] = score;
    if (size == 0) atomicOr(status, 4);
}

// self[v] = table[v][partition(v)] + that partition; node temperatures and frozen flags (partitions.cpp:398-417).
__global__ void node_state_kernel(const int32_t* __restrict__ partition_of_nodes, const Entry* __restrict__ table,
                                  const int8_t* __restrict__ hot_mask, double cold_temperature, int nodes, int partitions,
                                  Entry* __restrict__ self, double* __restrict__ temperature_of_nodes,
                                  uint8_t* __restrict__ frozen, int32_t* __restrict__ order) {
    const int node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= nodes) return;
    const int partition = partition_of_nodes[node];
    Entry entry = table[int64_t(node) * partitions + partition];
    entry.partition = partition;
    self[node] = entry;
    const bool hot = hot_mask[partition] != 0;
    temperature_of_nodes[node] = hot ? 1.0 : cold_temperature;
    frozen[node] = hot ? 0 : 1;
    order[node] = node;
}

// partitions.cpp:322-339, by one thread (every sum sequential as there).  incoming_scale = sum over the nodes in
// ascending order of log2(total incoming weight) (:107-128).
__device__ double final_score(int nodes, int partitions, const double* incoming_log2, const int32_t* sizes,
                              const double* masses, const double* scores, const Bounds& bounds, bool with_orphans,
                              const LogTables& t) {
    double incoming_scale = 0.0;
    for (int node = 0; node < nodes; ++node) incoming_scale = dadd(incoming_scale, incoming_log2[node]);
    double total = dsub(dmul(double(nodes), mc_log2(double(nodes), t)), incoming_scale);
    long long orphans = nodes;
    for (int p = 0; p < partitions; ++p) {
        const double size = double(sizes[p]);
        total = dadd(total, dsub(scores[p], dmul(size, mc_log2(size, t))));
        total = dadd(total, dmul(size, mass_factor(masses[p], bounds.low_mass, bounds.target_mass, bounds.high_mass, t)));
        orphans -= sizes[p];
    }
    if (with_orphans) {
        total = dadd(total, dmul(double(orphans), rescore(0.0, 0.0, t)));
        return ddiv(total, double(nodes));
    }
    return ddiv(total, double(nodes - orphans));
}

__global__ void score_kernel(int nodes, int partitions, const double* incoming_log2, const int32_t* sizes, const double* masses,
                             const double* scores, Bounds bounds, int with_orphans, double* result) {
    __shared__ unsigned long long staged[LOG_TABLE_WORDS];
    stage_log_tables(staged);
    __syncthreads();
    if (threadIdx.x == 0)
        *result = final_score(nodes, partitions, incoming_log2, sizes, masses, scores, bounds, with_orphans != 0, log_tables(staged));
}

// ---------------------------------------------------------------------------------------------------------
// The optimiser (partitions.cpp:388-1231)
// ---------------------------------------------------------------------------------------------------------

struct OptParams {
    const Neighbour* neighbours;
    const int32_t* union_count;
    const int32_t* out_ptr;
    const int32_t* in_ptr;
    const float* mass_of_nodes;
    int32_t* partition_of_nodes;
    Entry* table;
    uint32_t* linked_bits;  // [nodes][words]: the entries of the table that are not the default one
    Entry* self;
    double* temperature_of_nodes;
    uint8_t* frozen;
    int32_t* order;
    double* visit_temperature;  // per pass: the temperature of every visit
    int32_t* active;            // per pass: positions that are not skipped, ascending
    const double* incoming_log2;
    int32_t* sizes;
    double* masses;
    double* scores;
    int nodes, partitions, padded, warps, words;
    int parallel_shuffle;  // the shared-memory plan has room for the raw generator stream (8 bytes per node)
    Bounds bounds;
    double cooldown_pass, cooldown_node;
    uint32_t seed;
    double* result;
    long long* stats;  // [0] passes [1] improve_node calls [2] moves [3] skipped visits [4] batches [5] error
                       // [8..13] SM cycles of thread 0: pass set-up, evaluation (to the barrier), choice, move, and of
                       // warp 0 inside evaluate_node: the partition scan + neighbours, the candidate scan
};

struct Control {
    double temperature;
    double temperature_end;
    double cold_from, cold_to;
    uint32_t rng;
    int did_improve, did_skip;
    int mover_node, mover_from, mover_to;
    int error;
    long long cycles_neighbours, cycles_candidates;
    long long cycles_fine[4];  // warp 0: the partition scan, gathers + re-scoring, the ordered sums, (unused)
    int flags[OPT_MAX_WARPS];
    int counts[OPT_MAX_WARPS];
    int scan[34];
    double reduce[OPT_MAX_WARPS];
};

// Shared state of all warps: the per-partition vectors of the reference plus values cached per partition
// that depend on nothing but the partition (so they change only when a move touches it).
struct PartitionState {
    int32_t* sizes;
    double* masses;
    double* scores;
    double* old_adjusted;     // score - size log2(size) + size min(mass_factor(mass), mass_factor(size))  (:892-896)
    double* grown_log;        // (size + 1) log2(size + 1)                                                (:914)
    double* grown_factor;     // mass_factor(size + 1) over the size bounds                                (:912)
};

struct WarpBuffers {
    double* cold;    // tmp_partition_cold_diffs
    double* total;   // tmp_partition_hot_diffs until the candidate scan, then tmp_partition_total_diffs
    int32_t* linked; // connectivity diff + 65536 * disconnected diff (both small integers in the reference's doubles)
    uint16_t* list;  // the (aliased) candidate list, in partition order
    double* staging; // one chunk of neighbour diffs, 32 doubles
    long long* cycles;  // warp 0 only: {scan + neighbours, candidates}
    long long* fine;    // warp 0 only: Control::cycles_fine
};
constexpr int EVAL_UNROLL = 4;

__device__ __forceinline__ uint32_t rng_next(uint32_t& state) {
    state = uint32_t((uint64_t(state) * 48271u) % 2147483647u);
    return state;
}

// uniform_int_distribution<size_t>{0, high} on minstd_rand: libstdc++ uniform_int_dist.h:284-345, "downscaling"
// fallback branch (high + 1 < 2^31 - 2 for every supported node count).
__device__ __forceinline__ uint32_t uniform_int(uint32_t& rng, uint32_t high) {
    const uint32_t range = 2147483645u;
    const uint32_t wanted = high + 1;
    const uint32_t scaling = range / wanted;
    const uint32_t past = wanted * scaling;
    uint32_t draw;
    do {
        draw = rng_next(rng) - 1u;
    } while (draw >= past);
    return draw / scaling;
}

// std::shuffle, libstdc++ stl_algo.h:3742-3804, over the visiting order staged in shared memory.
__device__ void shuffle_order(uint16_t* order, int nodes, uint32_t& rng) {
    if (nodes == 0) return;
    if (2147483645u / uint32_t(nodes) >= uint32_t(nodes)) {
        int i = 1;
        if ((nodes & 1) == 0) {
            const uint32_t j = uniform_int(rng, 1);
            const uint16_t tmp = order[i]; order[i] = order[j]; order[j] = tmp;
            ++i;
        }
        while (i != nodes) {
            const uint32_t swap_range = uint32_t(i) + 1;
            const uint32_t x = uniform_int(rng, swap_range * (swap_range + 1) - 1);
            const uint32_t first = x / (swap_range + 1), second = x - first * (swap_range + 1);
            uint16_t tmp = order[i]; order[i] = order[first]; order[first] = tmp;
            ++i;
            tmp = order[i]; order[i] = order[second]; order[second] = tmp;
            ++i;
        }
        return;
    }
    for (int i = 1; i != nodes; ++i) {
        const uint32_t j = uniform_int(rng, uint32_t(i));
        const uint16_t tmp = order[i]; order[i] = order[j]; order[j] = tmp;
    }
}

// x * y mod (2^31 - 1) for x, y < 2^31 - 1, by Mersenne folding.
__device__ __forceinline__ uint32_t mulmod_mersenne(uint32_t x, uint32_t y) {
    const uint64_t product = uint64_t(x) * uint64_t(y);
    uint32_t folded = uint32_t(product & 2147483647u) + uint32_t(product >> 31);
    folded = (folded & 2147483647u) + (folded >> 31);
    return folded >= 2147483647u ? folded - 2147483647u : folded;
}

// The same std::shuffle by the whole CTA (nodes <= 46340: the two-swaps-per-draw branch; `raw` and `scal` are shared
// memory of `nodes` and nodes / 2 + 2 words).  What is sequential in it is little: which raw generator values the
// rejection loops of uniform_int_distribution consume, and the swaps.  So (1) every thread writes its share of the raw
// stream x0 a^(n+1) (closed form; pow_thread = a^(tid+1), pow_block = a^blockDim) and the scaling of every draw;
// (2) warp 0 pairs draws with raw values, 32 at a time, restarting behind the first rejected value; (3) every thread turns
// its accepted values into the two swap targets; (4) one thread swaps.  Returns the new generator state to thread 0.
__device__ uint32_t shuffle_order_parallel(uint16_t* order, int nodes, uint32_t state, uint32_t* raw, uint32_t* scal,
                                           uint32_t pow_thread, uint32_t pow_block) {
    const int tid = threadIdx.x, lane = tid & 31;
    const int special = (nodes & 1) == 0 ? 1 : 0;      // an even count starts with one draw from {0, 1}
    const int draws = special + (nodes - 1 - special) / 2;
    const int capacity = nodes;                         // raw values kept; more are generated one by one (never, in practice)
    {
        uint32_t current = mulmod_mersenne(state, pow_thread);
        for (int n = tid; n < capacity; n += blockDim.x) {
            raw[n] = current;
            current = mulmod_mersenne(current, pow_block);
        }
        for (int d = tid; d < draws; d += blockDim.x) {
            const uint32_t i = d < special ? 1u : uint32_t(2 * (d - special) + 1 + special);
            const uint32_t wanted = d < special ? 2u : (i + 1u) * (i + 2u);
            scal[d] = 2147483645u / wanted;
        }
    }
    __syncthreads();
    uint32_t last = state;
    if (tid < 32) {  // (2)
        int d0 = 0, n0 = 0;
        while (d0 < draws) {
            const int d = d0 + lane, n = n0 + lane;
            bool have = d < draws, rejected = false;
            uint32_t value = 0;
            if (have && n < capacity) {
                const uint32_t i = d < special ? 1u : uint32_t(2 * (d - special) + 1 + special);
                const uint32_t wanted = d < special ? 2u : (i + 1u) * (i + 2u);
                value = raw[n] - 1u;
                rejected = value >= wanted * scal[d];
            }
            const unsigned beyond = __ballot_sync(FULL, have && n >= capacity);
            const unsigned rejects = __ballot_sync(FULL, rejected);
            const unsigned stop = beyond | rejects;
            const int taken = stop ? __ffs(stop) - 1 : min(32, draws - d0);  // lanes below the first rejected / missing value
            __syncwarp();
            if (lane < taken) raw[d] = value;  // d <= n: the accepted values are compacted in place
            __syncwarp();
            if (taken > 0) last = __shfl_sync(FULL, value, taken - 1) + 1u;
            d0 += taken;
            n0 += taken;
            if (stop && (rejects >> taken & 1u)) {  // the rejected value is consumed too
                last = __shfl_sync(FULL, value, taken) + 1u;
                n0 += 1;
            } else if (stop) {  // out of precomputed values: this draw one value at a time, as the serial loop does
                if (lane == 0) {
                    const uint32_t i = d0 < special ? 1u : uint32_t(2 * (d0 - special) + 1 + special);
                    const uint32_t wanted = d0 < special ? 2u : (i + 1u) * (i + 2u);
                    const uint32_t past = wanted * scal[d0];
                    uint32_t draw;
                    do {
                        draw = rng_next(last) - 1u;
                    } while (draw >= past);
                    raw[d0] = draw;
                }
                last = __shfl_sync(FULL, last, 0);
                d0 += 1;
                n0 = capacity;
            }
        }
    }
    __syncthreads();
    for (int d = tid; d < draws; d += blockDim.x) {  // (3)
        const uint32_t x = raw[d] / scal[d];
        if (d < special) {
            scal[d] = x;
        } else {
            const uint32_t i = uint32_t(2 * (d - special) + 1 + special);
            const uint32_t first = x / (i + 2u), second = x - first * (i + 2u);
            scal[d] = first | second << 16;
        }
    }
    __syncthreads();
    if (tid == 0) {  // (4) stl_algo.h:3771-3790
        int i = 1;
        if (special) {
            const uint32_t j = scal[0];
            const uint16_t tmp = order[i]; order[i] = order[j]; order[j] = tmp;
            ++i;
        }
        for (int d = special; d < draws; ++d) {
            const uint32_t packed = scal[d];
            const uint32_t first = packed & 0xffffu, second = packed >> 16;
            uint16_t tmp = order[i]; order[i] = order[first]; order[first] = tmp;
            ++i;
            tmp = order[i]; order[i] = order[second]; order[second] = tmp;
            ++i;
        }
    }
    return last;
}

// uniform_real_distribution<double>(0, high): generate_canonical<double, 53> (random.tcc:3349-3381), two draws.
__device__ __forceinline__ double uniform_real(uint32_t& rng, double high) {
    const double r = 2147483646.0;
    double sum = dadd(0.0, dmul(double(rng_next(rng) - 1u), 1.0));
    sum = dadd(sum, dmul(double(rng_next(rng) - 1u), r));
    double canonical = ddiv(sum, dmul(r, r));
    if (canonical >= 1.0) canonical = __longlong_as_double(0x3FEFFFFFFFFFFFFFLL);  // nextafter(1, 0)
    return dadd(dmul(canonical, dsub(high, 0.0)), 0.0);
}

__device__ __forceinline__ double min_like_std(double a, double b) { return b < a ? b : a; }  // std::min(a, b)

__device__ void refresh_partition(const PartitionState& s, int p, const Bounds& b, const LogTables& t) {
    const double size = double(s.sizes[p]);
    double adjusted = dsub(s.scores[p], dmul(size, mc_log2(size, t)));
    const double by_mass = mass_factor(s.masses[p], b.low_mass, b.target_mass, b.high_mass, t);
    const double by_size = mass_factor(size, b.low_size, b.target_size, b.high_size, t);
    s.old_adjusted[p] = dadd(adjusted, dmul(size, min_like_std(by_mass, by_size)));
    const double grown = double(s.sizes[p] + 1);
    s.grown_log[p] = dmul(grown, mc_log2(grown, t));
    s.grown_factor[p] = mass_factor(grown, b.low_size, b.target_size, b.high_size, t);
}

// improve_node up to (not including) the choice: partitions.cpp:568-608 + 647-1005, by one warp, read-only on
// the global state.  Returns the number of candidate partitions (0 = the node stays: "unimproved").
__device__ int evaluate_node(const OptParams& P, const PartitionState& S, const WarpBuffers& B, const LogTables& t,
                             int node, double temperature, int lane) {
    const int partitions = P.partitions;
    const long long entered = clock64();
    const Entry* row = P.table + int64_t(node) * partitions;
    const uint32_t* bits = P.linked_bits + int64_t(node) * P.words;
    uint32_t first_bits = lane < P.words ? bits[lane] : 0u;  // in flight beside the gather below
    const Entry current_entry = P.self[node];                // == row[current], with the partition: one gather
    const int current = int(current_entry.partition);
    if (S.sizes[current] < 2) return 0;  // :575
    const int current_connectivity = connectivity(current_entry.outgoing, current_entry.incoming);

    // :647-662 and :699-723.  Most of the row is the default entry (a node has weight towards a few dozen partitions):
    // every slot gets the default entry's values, then only the entries `linked_bits` marks are read (a 16 KB row scan by
    // each of 16 warps at once is bound by what one SM takes from L2; the listed entries are a tenth of it).
    {
        const double default_score = rescore(0.0, 0.0, t);
        const int unlinked = current_connectivity > 0 ? 65536 - current_connectivity : 0;
        for (int p = lane; p < partitions; p += 32) {
            B.cold[p] = default_score;
            B.total[p] = default_score;
            B.linked[p] = unlinked;
        }
        int listed = 0;
        for (int first = 0; first < P.words; first += 32) {
            uint32_t mine = first == 0 ? first_bits : (first + lane < P.words ? bits[first + lane] : 0u);
            int before = __popc(mine);
#pragma unroll
            for (int step = 1; step < 32; step <<= 1) {
                const int below = __shfl_up_sync(FULL, before, step);
                if (lane >= step) before += below;
            }
            const int total = __shfl_sync(FULL, before, 31);
            int at = listed + before - __popc(mine);
            while (mine) {
                B.list[at++] = uint16_t((first + lane) * 32 + __ffs(mine) - 1);
                mine &= mine - 1u;
            }
            listed += total;
        }
        __syncwarp();
#pragma unroll 2
        for (int i = lane; i < listed; i += 32) {
            const int p = B.list[i];
            if (p == current) continue;
            const Entry entry = row[p];
            const int other_connectivity = connectivity(entry.outgoing, entry.incoming);
            int disconnected = 0;
            if (current_connectivity > 0 && other_connectivity == 0) disconnected = 1;
            else if (current_connectivity == 0 && other_connectivity > 0) disconnected = -1;
            B.cold[p] = entry.score;
            B.total[p] = entry.score;
            B.linked[p] = (other_connectivity - current_connectivity) + disconnected * 65536;
        }
        if (lane == 0) {
            B.cold[current] = -current_entry.score;
            B.total[current] = -current_entry.score;
            B.linked[current] = 0;
        }
    }
    __syncwarp();
    long long lap = clock64();
    if (B.fine && lane == 0) B.fine[0] += lap - entered;

    // :725-802: every neighbour (merged list, ascending) re-scored as if the node had left `current` / joined the
    // neighbour's partition; the cold diffs of one partition are added in neighbour order.  EVAL_UNROLL chunks of 32
    // neighbours are loaded and re-scored together (the two dependent gathers and the log are latency, not work).
    const int64_t base = int64_t(P.out_ptr[node]) + P.in_ptr[node];
    const int count = P.union_count[node];
    Neighbour ahead[EVAL_UNROLL];
#pragma unroll
    for (int j = 0; j < EVAL_UNROLL; ++j) {
        const int e = j * 32 + lane;
        ahead[j] = P.neighbours[base + (e < count ? e : 0)];
    }
    for (int start = 0; start < count; start += 32 * EVAL_UNROLL) {
        float outgoing_weight[EVAL_UNROLL], incoming_weight[EVAL_UNROLL];
        Entry score[EVAL_UNROLL];
#pragma unroll
        for (int j = 0; j < EVAL_UNROLL; ++j) {
            score[j] = P.self[ahead[j].node];
            outgoing_weight[j] = ahead[j].outgoing;  // node -> other: other's incoming total
            incoming_weight[j] = ahead[j].incoming;  // other -> node: other's outgoing total
        }
        if (start + 32 * EVAL_UNROLL < count) {  // the next group's neighbours, in flight during this group's work
#pragma unroll
            for (int j = 0; j < EVAL_UNROLL; ++j) {
                const int e = start + 32 * EVAL_UNROLL + j * 32 + lane;
                ahead[j] = P.neighbours[base + (e < count ? e : 0)];
            }
        }
        int q[EVAL_UNROLL], linked[EVAL_UNROLL];
        double diff[EVAL_UNROLL];
#pragma unroll
        for (int j = 0; j < EVAL_UNROLL; ++j) {
            q[j] = int(score[j].partition);
            const double direction = q[j] == current ? -1.0 : 1.0;
            const int old_connectivity = connectivity(score[j].outgoing, score[j].incoming);
            const double incoming = clamp0(dadd(score[j].incoming, dmul(direction, double(outgoing_weight[j]))));
            const double outgoing = clamp0(dadd(score[j].outgoing, dmul(direction, double(incoming_weight[j]))));
            const int new_connectivity = connectivity(outgoing, incoming);
            diff[j] = dsub(rescore(outgoing, incoming, t), score[j].score);
            linked[j] = new_connectivity - old_connectivity;
            if (old_connectivity == 0 && new_connectivity > 0) linked[j] -= 65536;
            if (old_connectivity > 0 && new_connectivity == 0) linked[j] += 65536;
            if (start + j * 32 + lane >= count) {  // lanes past the end: a private group each, nothing to add
                q[j] = -1 - lane;
                linked[j] = 0;
            }
        }
        if (B.fine) {  // the re-scored terms are complete here (the sums below consume them)
            const long long now = clock64();
            if (lane == 0) B.fine[1] += now - lap + (long long)(diff[0] == 1e300) + (long long)(diff[EVAL_UNROLL - 1] == 1e300);
            lap = now;
        }
        // The groups of equal partitions of all four chunks first (independent of each other), and for every group a run of
        // consecutive staging slots, in lane order inside the run: the first lane of a group then adds exactly its group's
        // terms from addresses known before the chain starts (32 predicated steps over all the slots, or a walk over the
        // set bits whose loads depend on the bit found, were both measured slower).
        int leader[EVAL_UNROLL], members[EVAL_UNROLL], first[EVAL_UNROLL], slot[EVAL_UNROLL];
#pragma unroll
        for (int j = 0; j < EVAL_UNROLL; ++j) {
            const unsigned peers = __match_any_sync(FULL, q[j]);
            leader[j] = __ffs(peers) - 1;
            members[j] = __popc(peers);
            slot[j] = __popc(peers & ((1u << lane) - 1u));
            first[j] = leader[j] == lane ? members[j] : 0;
        }
#pragma unroll
        for (int step = 1; step < 32; step <<= 1) {
#pragma unroll
            for (int j = 0; j < EVAL_UNROLL; ++j) {
                const int below = __shfl_up_sync(FULL, first[j], step);
                if (lane >= step) first[j] += below;
            }
        }
#pragma unroll
        for (int j = 0; j < EVAL_UNROLL; ++j) {
            first[j] = __shfl_sync(FULL, first[j] - members[j], leader[j]);  // at a group's first lane: the slots before it
            slot[j] += first[j];
        }
#pragma unroll
        for (int j = 0; j < EVAL_UNROLL; ++j) {
            if (start + j * 32 >= count) break;
            if (linked[j] != 0) atomicAdd(&B.linked[q[j]], linked[j]);  // integers: any order; rarely non-zero
            B.staging[slot[j]] = diff[j];
            __syncwarp();
            if (q[j] >= 0 && lane == leader[j]) {
                double chain = B.cold[q[j]];
                const double* terms = B.staging + first[j];
#pragma unroll 4
                for (int i = 0; i < members[j]; ++i) chain = dadd(chain, terms[i]);
                B.cold[q[j]] = chain;
            }
            __syncwarp();
        }
        if (B.fine) {
            const long long now = clock64();
            if (lane == 0) B.fine[2] += now - lap;
            lap = now;
        }
    }

    const long long scanned = clock64();
    // :815-876, the partition the node would leave
    const double node_mass = P.mass_of_nodes[node];
    const double current_hot_diff = B.total[current];
    const double current_cold_diff = B.cold[current];
    const int min_connectivity_diff = current_connectivity == 0 ? 1 : 0;
    double current_adjusted_cold_diff;
    {
        const double shrunk = double(S.sizes[current] - 1);
        const double by_mass = mass_factor(dsub(S.masses[current], node_mass), P.bounds.low_mass, P.bounds.target_mass, P.bounds.high_mass, t);
        const double by_size = mass_factor(shrunk, P.bounds.low_size, P.bounds.target_size, P.bounds.high_size, t);
        double adjusted = dsub(dadd(S.scores[current], current_cold_diff), dmul(shrunk, mc_log2(shrunk, t)));
        adjusted = dadd(adjusted, dmul(shrunk, min_like_std(by_mass, by_size)));
        current_adjusted_cold_diff = dsub(adjusted, S.old_adjusted[current]);
    }
    const int current_linked = B.linked[current];
    const int current_connectivity_diff = int16_t(current_linked & 0xffff);
    const int current_disconnected_diff = (current_linked - current_connectivity_diff) >> 16;
    const double cooled = dsub(1.0, temperature);
    __syncwarp();

    // :880-1002; all three push_back sites feed the one aliased list.  First the partitions that pass the two
    // connectivity filters (:950-957; few: a node is linked to a handful of partitions), in partition order ...
    // (four rounds of 32 partitions at a time: their loads and tests are independent, only the counts chain)
    int admissible = 0;
    for (int first = 0; first < partitions; first += 128) {
        bool passes[4];
        unsigned mask[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = first + 32 * u + lane;
            passes[u] = false;
            if (p < partitions && p != current) {
                const int linked = B.linked[p];
                const int connectivity_diff = int16_t(linked & 0xffff);
                const int disconnected_diff = (linked - connectivity_diff) >> 16;
                passes[u] = disconnected_diff + current_disconnected_diff <= 0 &&
                            connectivity_diff + current_connectivity_diff >= min_connectivity_diff;
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) mask[u] = __ballot_sync(FULL, passes[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (passes[u]) B.list[admissible + __popc(mask[u] & ((1u << lane) - 1u))] = uint16_t(first + 32 * u + lane);
            admissible += __popc(mask[u]);
        }
    }
    __syncwarp();
    // ... then their total diffs (a log each), keeping those that are candidates (compacted in place, in order).
    int candidates = 0;
    for (int first = 0; first < admissible; first += 32) {
        const int i = first + lane;
        bool candidate = false;
        int p = 0;
        if (i < admissible) {
            p = B.list[i];
            const double grown = double(S.sizes[p] + 1);
            const double by_mass = mass_factor(dadd(S.masses[p], node_mass), P.bounds.low_mass, P.bounds.target_mass, P.bounds.high_mass, t);
            double adjusted = dsub(dadd(S.scores[p], B.cold[p]), S.grown_log[p]);
            adjusted = dadd(adjusted, dmul(grown, min_like_std(by_mass, S.grown_factor[p])));
            const double adjusted_cold_diff = dsub(adjusted, S.old_adjusted[p]);
            const double total_diff = dadd(dmul(dadd(current_hot_diff, B.total[p]), temperature),
                                           dmul(dadd(current_adjusted_cold_diff, adjusted_cold_diff), cooled));
            B.total[p] = total_diff;
            candidate = total_diff >= EPSILON || min_connectivity_diff > 0;
        }
        const unsigned mask = __ballot_sync(FULL, candidate);
        __syncwarp();
        if (candidate) B.list[candidates + __popc(mask & ((1u << lane) - 1u))] = uint16_t(p);
        candidates += __popc(mask);
    }
    __syncwarp();
    if (B.cycles && lane == 0) {
        B.cycles[0] += scanned - entered;
        B.cycles[1] += clock64() - scanned;
    }
    return candidates;
}

// choose_target_partition / choose_biased_partition (:1007-1072) by the mover's warp; returns the chosen
// partition to lane 0 (-1: the reference's FastAssert(false) at :1071).
__device__ int choose_partition(const WarpBuffers& B, int candidates, uint32_t& rng, int lane) {
    double least = __longlong_as_double(0x7FF0000000000000LL);
    for (int i = lane; i < candidates; i += 32) least = fmin(least, B.total[B.list[i]]);
    for (int offset = 16; offset > 0; offset >>= 1) least = fmin(least, __shfl_xor_sync(FULL, least, offset));
    int chosen = -1;
    if (lane == 0) {
        double sum = 0.0;
        for (int i = 0; i < candidates; ++i) sum = dadd(sum, dsub(B.total[B.list[i]], least));
        double pick = uniform_real(rng, sum);
        for (int i = 0; i < candidates; ++i) {
            pick = dsub(pick, dsub(B.total[B.list[i]], least));
            if (pick <= EPSILON) {
                chosen = B.list[i];
                break;
            }
        }
    }
    return chosen;
}

// The accepted move (:625-636, 1074-1231) by the whole CTA: every neighbour's totals towards the two
// partitions, the self[] copies that mirror them, then the per-partition vectors.
__device__ void apply_move(const OptParams& P, const PartitionState& S, const LogTables& t, int node, int from, int to,
                           double cold_from, double cold_to) {
    const int partitions = P.partitions;
    const int64_t base = int64_t(P.out_ptr[node]) + P.in_ptr[node];
    const int count = P.union_count[node];
    // what the one thread of the tail needs, in flight during the loop (the node's own row is not touched by its move)
    Entry own{};
    double node_mass = 0.0;
    if (threadIdx.x == 0) {
        own = P.table[int64_t(node) * partitions + to];
        node_mass = P.mass_of_nodes[node];
    }
    // Up to MOVE_ITEMS neighbours per thread and round: every load of all of them (two entries, two bitmap words and the
    // neighbour's partition each: latency of HBM, the table is larger than L2) before the first store.  The neighbours are
    // distinct nodes, so their rows, words and self[] entries do not overlap.
    constexpr int MOVE_ITEMS = 1;
    for (int e0 = threadIdx.x; e0 < count; e0 += MOVE_ITEMS * blockDim.x) {
        Neighbour neighbour[MOVE_ITEMS];
        Entry left[MOVE_ITEMS], joined[MOVE_ITEMS];
        uint32_t word_from[MOVE_ITEMS], word_to[MOVE_ITEMS];
        int own_partition[MOVE_ITEMS];
        bool has[MOVE_ITEMS];
#pragma unroll
        for (int u = 0; u < MOVE_ITEMS; ++u) {
            has[u] = e0 + u * int(blockDim.x) < count;
            neighbour[u] = P.neighbours[base + (has[u] ? e0 + u * int(blockDim.x) : e0)];
        }
#pragma unroll
        for (int u = 0; u < MOVE_ITEMS; ++u) {
            if (!has[u]) continue;
            const Entry* row = P.table + int64_t(neighbour[u].node) * partitions;
            const uint32_t* words = P.linked_bits + int64_t(neighbour[u].node) * P.words;
            word_from[u] = words[from >> 5];
            word_to[u] = words[to >> 5];
            left[u] = row[from];
            joined[u] = row[to];
            own_partition[u] = int(P.self[neighbour[u].node].partition);
        }
#pragma unroll
        for (int u = 0; u < MOVE_ITEMS; ++u) {
            if (!has[u]) continue;
            const int other = neighbour[u].node;
            const double outgoing_weight = neighbour[u].outgoing;
            const double incoming_weight = neighbour[u].incoming;
            Entry* row = P.table + int64_t(other) * partitions;
            uint32_t* words = P.linked_bits + int64_t(other) * P.words;
            left[u].incoming = clamp0(dsub(left[u].incoming, outgoing_weight));
            left[u].outgoing = clamp0(dsub(left[u].outgoing, incoming_weight));
            left[u].score = rescore(left[u].outgoing, left[u].incoming, t);
            row[from] = left[u];
            joined[u].incoming = clamp0(dadd(joined[u].incoming, outgoing_weight));
            joined[u].outgoing = clamp0(dadd(joined[u].outgoing, incoming_weight));
            joined[u].score = rescore(joined[u].outgoing, joined[u].incoming, t);
            row[to] = joined[u];
            const bool left_empty = left[u].outgoing == 0.0 && left[u].incoming == 0.0;
            const bool joined_linked = joined[u].outgoing != 0.0 || joined[u].incoming != 0.0;
            if ((from >> 5) == (to >> 5)) {
                uint32_t word = word_from[u];
                if (left_empty) word &= ~(1u << (from & 31));
                if (joined_linked) word |= 1u << (to & 31);
                if (word != word_from[u]) words[from >> 5] = word;
            } else {
                if (left_empty) words[from >> 5] = word_from[u] & ~(1u << (from & 31));
                if (joined_linked && !(word_to[u] >> (to & 31) & 1u)) words[to >> 5] = word_to[u] | 1u << (to & 31);
            }
            if (own_partition[u] == from) {
                left[u].partition = from;
                P.self[other] = left[u];
            } else if (own_partition[u] == to) {
                joined[u].partition = to;
                P.self[other] = joined[u];
            }
        }
    }
    if (threadIdx.x == 0) {
        P.partition_of_nodes[node] = to;
        own.partition = to;
        P.self[node] = own;
        S.sizes[from] -= 1;
        S.sizes[to] += 1;
        S.scores[from] = dadd(S.scores[from], cold_from);
        S.scores[to] = dadd(S.scores[to], cold_to);
        S.masses[from] = dsub(S.masses[from], node_mass);
        S.masses[to] = dadd(S.masses[to], node_mass);
    }
    __syncthreads();
    if (threadIdx.x == 0) refresh_partition(S, from, P.bounds, t);
    if (threadIdx.x == 32) refresh_partition(S, to, P.bounds, t);
}

__host__ __device__ inline size_t align_up(size_t value, size_t to) { return (value + to - 1) / to * to; }

// Shared memory: Control | log tables | 6 per-partition vectors | max(warps x WarpBuffers, uint16 order[N]).
__host__ __device__ inline size_t partition_state_bytes(int padded) { return size_t(padded) * (4 + 5 * 8); }
__host__ __device__ inline size_t warp_buffer_bytes(int padded) { return size_t(padded) * (8 + 8 + 4 + 2) + 256; }
__host__ __device__ inline size_t fixed_shared_bytes(int padded) {
    return align_up(sizeof(Control), 16) + LOG_TABLE_WORDS * 8 + align_up(partition_state_bytes(padded), 16);
}

__global__ void __launch_bounds__(OPT_THREADS, 1) optimize_partitions_kernel(OptParams P) {
    extern __shared__ __align__(16) unsigned char shared[];
    Control* control = reinterpret_cast<Control*>(shared);
    unsigned long long* staged = reinterpret_cast<unsigned long long*>(shared + align_up(sizeof(Control), 16));
    unsigned char* cursor = reinterpret_cast<unsigned char*>(staged + LOG_TABLE_WORDS);
    PartitionState S;
    S.masses = reinterpret_cast<double*>(cursor);
    S.scores = S.masses + P.padded;
    S.old_adjusted = S.scores + P.padded;
    S.grown_log = S.old_adjusted + P.padded;
    S.grown_factor = S.grown_log + P.padded;
    S.sizes = reinterpret_cast<int32_t*>(S.grown_factor + P.padded);
    unsigned char* region = shared + fixed_shared_bytes(P.padded);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpBuffers B;
    {
        unsigned char* mine = region + size_t(warp < P.warps ? warp : 0) * warp_buffer_bytes(P.padded);
        B.cycles = warp == 0 ? &control->cycles_neighbours : nullptr;
        B.fine = warp == 0 ? control->cycles_fine : nullptr;
        B.staging = reinterpret_cast<double*>(mine);
        B.cold = B.staging + 32;
        B.total = B.cold + P.padded;
        B.linked = reinterpret_cast<int32_t*>(B.total + P.padded);
        B.list = reinterpret_cast<uint16_t*>(B.linked + P.padded);
    }
    uint16_t* staged_order = reinterpret_cast<uint16_t*>(region);

    // a^(tid+1) and a^blockDim of std::minstd_rand, for the parallel shuffle
    uint32_t pow_thread = 1, pow_block = 1;
    {
        uint32_t base = 48271u;
        for (unsigned e = threadIdx.x + 1; e; e >>= 1) {
            if (e & 1u) pow_thread = mulmod_mersenne(pow_thread, base);
            base = mulmod_mersenne(base, base);
        }
        base = 48271u;
        for (unsigned e = blockDim.x; e; e >>= 1) {
            if (e & 1u) pow_block = mulmod_mersenne(pow_block, base);
            base = mulmod_mersenne(base, base);
        }
    }
    stage_log_tables(staged);
    const LogTables t = log_tables(staged);
    for (int p = threadIdx.x; p < P.partitions; p += blockDim.x) {
        S.sizes[p] = P.sizes[p];
        S.masses[p] = P.masses[p];
        S.scores[p] = P.scores[p];
    }
    if (threadIdx.x == 0) {
        control->temperature = 1.0;
        control->temperature_end = 1.0;
        control->rng = P.seed % 2147483647u == 0 ? 1u : P.seed % 2147483647u;  // std::minstd_rand(seed)
        control->did_improve = 1;
        control->did_skip = 0;
        control->error = 0;
        control->cycles_neighbours = 0;
        control->cycles_candidates = 0;
        for (int i = 0; i < 4; ++i) control->cycles_fine[i] = 0;
    }
    __syncthreads();
    for (int p = threadIdx.x; p < P.partitions; p += blockDim.x) refresh_partition(S, p, P.bounds, t);

    const int nodes = P.nodes;
    const double cooldown_rate = dsub(1.0, ddiv(P.cooldown_pass, double(nodes)));  // :443
    const double node_cooling = dsub(1.0, P.cooldown_node);                        // :521
    long long passes = 0, visits = 0, moves = 0, skipped_visits = 0, batches = 0;
    long long cycles_setup = 0, cycles_evaluate = 0, cycles_choose = 0, cycles_move = 0;

    while (true) {
        __syncthreads();
        double temperature = control->temperature;
        int did_improve = control->did_improve, did_skip = control->did_skip;
        const int error = control->error;
        __syncthreads();
        if (error || !(temperature > 0 || did_improve)) break;  // :451
        if (did_improve) {  // :460-478
            did_improve = 0;
            did_skip = 0;
        } else if (did_skip) {
            did_skip = 0;
            double least = temperature;
            for (int node = threadIdx.x; node < nodes; node += blockDim.x)
                if (!P.frozen[node]) least = fmin(least, P.temperature_of_nodes[node]);
            for (int offset = 16; offset > 0; offset >>= 1) least = fmin(least, __shfl_xor_sync(FULL, least, offset));
            if (lane == 0) control->reduce[warp] = least;
            __syncthreads();
            for (int w = 0; w < OPT_MAX_WARPS; ++w) least = fmin(least, control->reduce[w]);
            temperature = least;
        } else {
            temperature = 0.0;
        }
        ++passes;
        long long mark = clock64();

        // :480 the visiting order (one thread: the swaps form a chain), and beside it the temperature of every
        // visit (:485, a chain of products).
        for (int i = threadIdx.x; i < nodes; i += blockDim.x) staged_order[i] = uint16_t(P.order[i]);
        __syncthreads();
        if (P.parallel_shuffle) {
            uint32_t* raw = reinterpret_cast<uint32_t*>(region + align_up(size_t(nodes) * 2, 16));
            uint32_t* scal = raw + nodes;
            const uint32_t rng = shuffle_order_parallel(staged_order, nodes, control->rng, raw, scal, pow_thread, pow_block);
            if (threadIdx.x == 0) control->rng = rng;
            if (threadIdx.x == 32) {  // beside thread 0's swaps: the temperature of every visit
                double running = temperature;
                for (int i = 0; i < nodes; ++i) {
                    running = dmul(running, cooldown_rate);
                    P.visit_temperature[i] = running;
                }
                control->temperature_end = running;
            }
        } else if (threadIdx.x == 0) {
            uint32_t rng = control->rng;
            shuffle_order(staged_order, nodes, rng);
            control->rng = rng;
        } else if (threadIdx.x == 32) {
            double running = temperature;
            for (int i = 0; i < nodes; ++i) {
                running = dmul(running, cooldown_rate);
                P.visit_temperature[i] = running;
            }
            control->temperature_end = running;
        }
        __syncthreads();

        // :493 which visits are skipped is known now: a node's own temperature changes only at its own visit.
        int active_count = 0;
        for (int first = 0; first < nodes; first += blockDim.x) {
            const int i = first + threadIdx.x;
            int keep = 0;
            if (i < nodes) {
                const int node = staged_order[i];
                P.order[i] = node;
                keep = !(P.temperature_of_nodes[node] < P.visit_temperature[i]);
            }
            int total;
            const int at = block_exclusive_scan(keep, control->scan, &total);
            if (keep) P.active[active_count + at] = i;
            active_count += total;
        }
        if (active_count < nodes) did_skip = 1;
        skipped_visits += nodes - active_count;
        __syncthreads();
        cycles_setup += clock64() - mark;

        for (int done = 0; done < active_count;) {
            const int limit = min(P.warps, active_count - done);
            mark = clock64();
            int node = -1;
            double visit = 0.0;
            if (warp < limit) {
                const int i = P.active[done + warp];
                node = P.order[i];
                visit = P.visit_temperature[i];
                // the table row and the neighbour list of the node this warp is likely to evaluate next (the one P.warps
                // visits on, unless a move restarts the batch earlier: then its rows are in the caches already), into L2
                if (done + P.warps + warp < active_count) {
                    const int next = P.order[P.active[done + P.warps + warp]];
                    const char* row = reinterpret_cast<const char*>(P.table + int64_t(next) * P.partitions);
                    for (int line = lane * 128; line < P.partitions * int(sizeof(Entry)); line += 32 * 128)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(row + line));
                    const char* list = reinterpret_cast<const char*>(P.neighbours + int64_t(P.out_ptr[next]) + P.in_ptr[next]);
                    for (int line = lane * 128; line < P.union_count[next] * int(sizeof(Neighbour)); line += 32 * 128)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(list + line));
                }
                const int candidates = evaluate_node(P, S, B, t, node, visit, lane);
                if (lane == 0) {
                    control->flags[warp] = candidates > 0;
                    control->counts[warp] = candidates;
                }
            }
            __syncthreads();
            cycles_evaluate += clock64() - mark;
            mark = clock64();
            int mover = limit;
            for (int w = 0; w < limit; ++w)
                if (control->flags[w]) {
                    mover = w;
                    break;
                }
            ++batches;
            if (warp < mover && warp < limit && lane == 0) {  // :518-527 unimproved
                P.temperature_of_nodes[node] = dmul(visit, node_cooling);
                P.frozen[node] = 0;
            }
            if (mover < limit) {
                if (warp == mover) {  // :609-644
                    uint32_t rng = control->rng;
                    const int chosen = choose_partition(B, control->counts[warp], rng, lane);
                    if (lane == 0) {
                        control->rng = rng;
                        control->mover_node = node;
                        control->mover_from = P.partition_of_nodes[node];
                        control->mover_to = chosen;
                        control->cold_from = B.cold[control->mover_from];
                        control->cold_to = chosen >= 0 ? B.cold[chosen] : 0.0;
                        P.frozen[node] = 0;
                        if (chosen < 0) control->error = 1;
                    }
                }
                __syncthreads();
                cycles_choose += clock64() - mark;
                mark = clock64();
                if (control->mover_to >= 0)
                    apply_move(P, S, t, control->mover_node, control->mover_from, control->mover_to, control->cold_from,
                               control->cold_to);
                did_improve = 1;
                ++moves;
                visits += mover + 1;
                done += mover + 1;
                if (control->mover_to < 0) done = active_count;
                __syncthreads();
                cycles_move += clock64() - mark;
            } else {
                visits += limit;
                done += limit;
            }
            __syncthreads();
        }

        if (threadIdx.x == 0) {
            control->temperature = control->temperature_end;
            control->did_improve = did_improve;
            control->did_skip = did_skip;
        }
    }

    for (int p = threadIdx.x; p < P.partitions; p += blockDim.x) {
        P.sizes[p] = S.sizes[p];
        P.masses[p] = S.masses[p];
        P.scores[p] = S.scores[p];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        *P.result = final_score(nodes, P.partitions, P.incoming_log2, S.sizes, S.masses, S.scores, P.bounds, true, t);
        if (P.stats) {
            P.stats[0] = passes; P.stats[1] = visits; P.stats[2] = moves; P.stats[3] = skipped_visits;
            P.stats[4] = batches; P.stats[5] = control->error;
            P.stats[8] = cycles_setup; P.stats[9] = cycles_evaluate; P.stats[10] = cycles_choose; P.stats[11] = cycles_move;
            P.stats[12] = control->cycles_neighbours; P.stats[13] = control->cycles_candidates;
            P.stats[6] = control->cycles_fine[0]; P.stats[7] = control->cycles_fine[1]; P.stats[14] = control->cycles_fine[2];
        }
    }
}

}  // namespace

// Device-side entry: every pointer is device memory except `score` / `stats` (host).  `optimize` = 0 only scores
// the given partition (score_partitions).  Synchronises `stream`.
int partitions_launch(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w,
                      const int32_t* in_idx, const int32_t* in_ptr, int64_t nodes, uint32_t random_seed,
                      const float* mass_of_nodes, const double* bounds6, double cooldown_pass, double cooldown_node,
                      int32_t* partition_of_nodes, const int8_t* hot_mask, int64_t mask_count, double cold_temperature,
                      int optimize, int with_orphans, double* score, int64_t* stats, cudaStream_t stream) {
    MCB_REQUIRE(nodes > 0, "optimize_partitions: no nodes");
    MCB_REQUIRE(nodes <= 65535, "optimize_partitions: %lld nodes; at most 65535 are supported", (long long)nodes);
    const Bounds bounds{bounds6[0], bounds6[1], bounds6[2], bounds6[3], bounds6[4], bounds6[5]};
    // partitions.cpp:314-319, 437-442
    MCB_REQUIRE(0.0 < bounds.low_mass && bounds.low_mass < bounds.target_mass && bounds.target_mass < bounds.high_mass,
                "failed assert: 0 < low_mass < target_mass < high_mass");
    MCB_REQUIRE(0.0 < bounds.low_size && bounds.low_size < bounds.target_size && bounds.target_size < bounds.high_size,
                "failed assert: 0 < low_size < target_size < high_size");
    if (optimize) {
        MCB_REQUIRE(cooldown_node >= 0.0 && cooldown_node <= 1.0, "failed assert: 0 <= cooldown_node <= 1");
        MCB_REQUIRE(cooldown_pass == 0.0 || (cooldown_pass > 0.0 && cooldown_pass < 1.0), "failed assert: 0 < cooldown_pass < 1");
    }
    const int n = int(nodes);

    Scratch flags_device;
    MCB_TRY(flags_device.alloc(8 * sizeof(int), stream));
    int* flags = flags_device.as<int>();
    const int flags_init[8] = {-1, 0, 0, 0, 0, 0, 0, 0};
    MCB_TRY(upload_small(flags, flags_init, sizeof(flags_init), stream));
    MCB_PROFILED("partition_range", stream, partition_range_kernel<<<std::min<int>(int(ceil_div(n, 256)), 1024), 256, 0, stream>>>(
                     partition_of_nodes, n, flags));
    MCB_LAUNCH_CHECK();
    int flags_host[8];
    int edges_host[2];
    MCB_TRY(copy_to_host(flags_host, flags, sizeof(flags_host), stream));
    MCB_TRY(copy_to_host(&edges_host[0], out_ptr + n, sizeof(int), stream));
    MCB_TRY(copy_to_host(&edges_host[1], in_ptr + n, sizeof(int), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    MCB_REQUIRE(flags_host[1] == 0, "failed assert: partition_index >= 0");  // partitions.cpp:93
    const int partitions = flags_host[0] + 1;
    MCB_REQUIRE(mask_count >= partitions, "hot_partitions_mask has %lld entries for %d partitions", (long long)mask_count, partitions);
    const int padded = int(ceil_div(partitions, 32) * 32);
    const int64_t union_capacity = int64_t(edges_host[0]) + edges_host[1];

    // How many nodes are evaluated side by side: one warp each, bounded by shared memory.
    const size_t shared_limit = 227 * 1024;
    const size_t fixed = fixed_shared_bytes(padded);
    MCB_REQUIRE(fixed + warp_buffer_bytes(padded) <= shared_limit && fixed + size_t(n) * 2 <= shared_limit,
                "optimize_partitions: %d partitions x %d nodes do not fit the shared-memory plan", partitions, n);
    const int warps = int(std::min<size_t>(OPT_MAX_WARPS, (shared_limit - fixed) / warp_buffer_bytes(padded)));
    // the visiting order (2 bytes per node) and, when it fits, the raw generator stream and the scaling of every draw of
    // the shuffle (shuffle_order_parallel: 4 + 2 bytes per node more)
    const size_t shuffle_bytes = align_up(size_t(n) * 2, 16) + size_t(n) * 4 + (size_t(n) / 2 + 2) * 4;
    const bool parallel_shuffle = n >= 64 && n <= 46340 && fixed + shuffle_bytes <= shared_limit && std::getenv("MCB_SERIAL_SHUFFLE") == nullptr;
    const size_t shared_bytes = fixed + std::max(size_t(warps) * warp_buffer_bytes(padded), parallel_shuffle ? shuffle_bytes : size_t(n) * 2);

    Scratch table, linked_bits, self, neighbours, union_count, incoming_log2, sizes, masses, scores, temperatures, frozen,
        order, visit_temperature, active, result, stats_device;
    const int64_t entries = int64_t(n) * partitions;
    MCB_TRY(table.alloc(size_t(entries) * sizeof(Entry), stream));
    const int words = int(ceil_div(partitions, 32));
    MCB_TRY(linked_bits.alloc(size_t(n) * words * 4, stream));
    MCB_TRY(self.alloc(size_t(n) * sizeof(Entry), stream));
    MCB_TRY(neighbours.alloc(size_t(union_capacity) * sizeof(Neighbour), stream));
    MCB_TRY(union_count.alloc(size_t(n) * 4, stream));
    MCB_TRY(incoming_log2.alloc(size_t(n) * 8, stream));
    MCB_TRY(sizes.alloc(size_t(padded) * 4, stream));
    MCB_TRY(masses.alloc(size_t(padded) * 8, stream));
    MCB_TRY(scores.alloc(size_t(padded) * 8, stream));
    MCB_TRY(temperatures.alloc(size_t(n) * 8, stream));
    MCB_TRY(frozen.alloc(size_t(n), stream));
    MCB_TRY(order.alloc(size_t(n) * 4, stream));
    MCB_TRY(visit_temperature.alloc(size_t(n) * 8, stream));
    MCB_TRY(active.alloc(size_t(n) * 4, stream));
    MCB_TRY(result.alloc(8, stream));
    MCB_TRY(stats_device.alloc(16 * sizeof(long long), stream));
    MCB_CUDA(cudaMemsetAsync(stats_device.ptr, 0, 16 * sizeof(long long), stream));

    const int node_blocks = int(ceil_div(n, 128));
    const int wide_blocks = int(std::min<int64_t>(ceil_div(entries, 256), int64_t(sm_count()) * 8));
    MCB_PROFILED("build_union", stream, build_union_kernel<<<node_blocks, 128, 0, stream>>>(
                     out_w, out_idx, out_ptr, in_w, in_idx, in_ptr, n, neighbours.as<Neighbour>(), union_count.as<int32_t>()));
    MCB_LAUNCH_CHECK();
    MCB_PROFILED("clear_table", stream, clear_table_kernel<<<wide_blocks, 256, 0, stream>>>(table.as<Entry>(), entries));
    MCB_LAUNCH_CHECK();
    MCB_PROFILED("accumulate_totals", stream, accumulate_totals_kernel<<<node_blocks, 128, 0, stream>>>(
                     out_w, out_idx, out_ptr, in_w, in_idx, in_ptr, partition_of_nodes, n, partitions, table.as<Entry>(),
                     incoming_log2.as<double>(), flags + 2));
    MCB_LAUNCH_CHECK();
    MCB_PROFILED("rescore_table", stream, rescore_table_kernel<<<wide_blocks, 256, 0, stream>>>(
                     table.as<Entry>(), n, partitions, words, linked_bits.as<uint32_t>()));
    MCB_LAUNCH_CHECK();
    MCB_PROFILED("partition_totals", stream, partition_totals_kernel<<<int(ceil_div(partitions, 32)), 32, 0, stream>>>(
                     partition_of_nodes, mass_of_nodes, table.as<Entry>(), n, partitions, sizes.as<int32_t>(),
                     masses.as<double>(), scores.as<double>(), flags + 2));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(flags_host, flags, sizeof(flags_host), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    MCB_REQUIRE((flags_host[2] & 1) == 0, "failed assert: the outgoing weights of every node sum to 1");  // :182-183
    MCB_REQUIRE((flags_host[2] & 2) == 0, "failed assert: total_incoming_weights > 0");                    // :123
    MCB_REQUIRE((flags_host[2] & 4) == 0, "failed assert: size_of_partitions[partition_index] > 0");      // :101

    if (optimize) {
        MCB_PROFILED("node_state", stream, node_state_kernel<<<node_blocks, 128, 0, stream>>>(
                         partition_of_nodes, table.as<Entry>(), hot_mask, cold_temperature, n, partitions, self.as<Entry>(),
                         temperatures.as<double>(), frozen.as<uint8_t>(), order.as<int32_t>()));
        MCB_LAUNCH_CHECK();
        OptParams params;
        params.neighbours = neighbours.as<Neighbour>();
        params.union_count = union_count.as<int32_t>();
        params.out_ptr = out_ptr;
        params.in_ptr = in_ptr;
        params.mass_of_nodes = mass_of_nodes;
        params.partition_of_nodes = partition_of_nodes;
        params.table = table.as<Entry>();
        params.linked_bits = linked_bits.as<uint32_t>();
        params.words = words;
        params.self = self.as<Entry>();
        params.temperature_of_nodes = temperatures.as<double>();
        params.frozen = frozen.as<uint8_t>();
        params.order = order.as<int32_t>();
        params.visit_temperature = visit_temperature.as<double>();
        params.active = active.as<int32_t>();
        params.incoming_log2 = incoming_log2.as<double>();
        params.sizes = sizes.as<int32_t>();
        params.masses = masses.as<double>();
        params.scores = scores.as<double>();
        params.nodes = n;
        params.partitions = partitions;
        params.padded = padded;
        params.warps = warps;
        params.parallel_shuffle = parallel_shuffle ? 1 : 0;
        params.bounds = bounds;
        params.cooldown_pass = cooldown_pass;
        params.cooldown_node = cooldown_node;
        params.seed = random_seed;
        params.result = result.as<double>();
        params.stats = stats_device.as<long long>();
        // the device's limit, once: optimisations of several host threads run side by side
        MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(optimize_partitions_kernel), std::max<size_t>(shared_bytes, shared_limit - 4096)));
        MCB_PROFILED("optimize_partitions", stream,
                     optimize_partitions_kernel<<<1, OPT_THREADS, shared_bytes, stream>>>(params));
        MCB_LAUNCH_CHECK();
    } else {
        MCB_PROFILED("score_partitions", stream, score_kernel<<<1, 256, 0, stream>>>(
                         n, partitions, incoming_log2.as<double>(), sizes.as<int32_t>(), masses.as<double>(),
                         scores.as<double>(), bounds, with_orphans, result.as<double>()));
        MCB_LAUNCH_CHECK();
    }
    long long stats_host[16];
    MCB_TRY(copy_to_host(score, result.ptr, 8, stream));
    MCB_TRY(copy_to_host(stats_host, stats_device.ptr, sizeof(stats_host), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    MCB_REQUIRE(stats_host[5] == 0, "failed assert: no candidate partition was chosen (partitions.cpp:1071)");
    if (stats)
        for (int i = 0; i < 16; ++i) stats[i] = stats_host[i];
    return MCB_OK;
}

}  // namespace mcb
