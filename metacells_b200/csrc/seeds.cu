// choose_seeds (choose_seeds.cpp:264-372) on the device.
//
// The reference is serial and its result depends on WHERE libstdc++'s std::nth_element leaves every element
// (choose_seeds.cpp:68-70 reads an element between two cuts), so the selection is libstdc++'s __introselect
// step for step, with its one data-parallel step spread over a CTA:
//
//   * Hoare partition (__unguarded_partition).  The sequential scan pairs the k-th element >= pivot from the
//     left with the k-th element <= pivot from the right and swaps them while the left one lies before the
//     right one.  Neither scan ever looks at a slot an earlier swap wrote, so the pairs are a function of the
//     unswapped range: two ranks from warp ballots + one scan over the warps give every "stopper" its k, the
//     pairs swap in parallel, and the cut is min(left[K], right[K-1]) for the first K that does not swap.
//   * The median-of-three, ranges of at most SERIAL_LIMIT elements, the depth-limit heap select and the closing
//     insertion sort are run by one thread (they are short or never happen).
//
// connected_nodes[node] of the reference is only ever asked for its size, so it is a counter per node here.
// The seeding loop runs in one CTA (compaction of the candidates, two selections, the seed's band, the
// counters of the neighbours of the nodes the seed took); do_complete_seeds is a chain through the generator
// and the growing seeds, one warp walks it: per node the per-seed float32 sums are added in edge order (lanes
// grouped by seed, the first lane of a group adds its group's terms in lane order), the float64 total is a
// tree sum when every partial sum is exactly representable (checked: then every order gives the same bits)
// and an ordered sum otherwise.
#include "common.cuh"
#include "std_sort.cuh"

#include <algorithm>

namespace mcb {

namespace {

constexpr int SEED_THREADS = 1024;
constexpr int SEED_WARPS = SEED_THREADS / 32;
constexpr unsigned FULL = 0xffffffffu;
constexpr int SERIAL_LIMIT = 192;  // ranges this short are selected by one thread
constexpr int CONNECT_UNROLL = 4;

enum { SEEDS_OK = 0, SEEDS_ASSERT = 2, SEEDS_UNLUCKY_DRAW = 3 };

__device__ __forceinline__ uint32_t rng_next(uint32_t& state) {
    state = uint32_t((uint64_t(state) * 48271u) % 2147483647u);
    return state;
}

// ---- the CTA ------------------------------------------------------------------------------------------

struct Control {
    int warp_left[SEED_WARPS];
    int warp_right[SEED_WARPS];
    float staging[32];
    uint32_t pivot;
    int swaps;
    int seed_node;
    int seeds_count;
    int status;
};

// elements are (key << SHIFT | position): compared by key alone, as the reference's comparators
template <typename E, int SHIFT>
struct KeyLess {
    __device__ __forceinline__ bool operator()(E left, E right) const { return (left >> SHIFT) < (right >> SHIFT); }
};

// The warps' exclusive prefix and the total of one count per warp (every warp scans the 32 counts itself).
__device__ __forceinline__ int warp_prefix(const int* per_warp, int warp, int lane, int* total) {
    const int mine = per_warp[lane];
    const int scanned = warp_inclusive_scan(mine, lane);
    *total = __shfl_sync(FULL, scanned, 31);
    return __shfl_sync(FULL, scanned - mine, warp);
}

// Ordered compaction of the i in [0, size) that pass `keep`: emit(rank, i).  Returns the count.
template <typename Keep, typename Emit>
__device__ int cta_compact(int size, Control& C, Keep keep, Emit emit) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int segment = (size + SEED_WARPS - 1) / SEED_WARPS;
    const int begin = min(size, warp * segment), end = min(size, begin + segment);
    int passed = 0;
    for (int base = begin; base < end; base += 32) {
        const int i = base + lane;
        passed += __popc(__ballot_sync(FULL, i < end && keep(i)));
    }
    if (lane == 0) C.warp_left[warp] = passed;
    __syncthreads();
    int total;
    int rank = warp_prefix(C.warp_left, warp, lane, &total);
    for (int base = begin; base < end; base += 32) {
        const int i = base + lane;
        const bool mine = i < end && keep(i);
        const unsigned ballot = __ballot_sync(FULL, mine);
        if (mine) emit(rank + __popc(ballot & ((1u << lane) - 1u)), i);
        rank += __popc(ballot);
    }
    __syncthreads();
    return total;
}

// std::nth_element(a + first, a + nth, a + last) by the whole CTA; `left` / `right` hold a range's stoppers.
template <typename E, int SHIFT>
__device__ void cta_nth_element(E* a, uint16_t* left, uint16_t* right, int first, int nth, int last, Control& C) {
    if (first == last || nth == last) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const KeyLess<E, SHIFT> less;
    int depth_limit = floor_lg(last - first) * 2;
    while (last - first > SERIAL_LIMIT && depth_limit > 0) {
        --depth_limit;
        if (threadIdx.x == 0) {
            median_to_first(a, first, first + 1, first + (last - first) / 2, last - 1, less);
            C.pivot = uint32_t(a[first] >> SHIFT);
        }
        __syncthreads();
        const uint32_t pivot = C.pivot;
        const int low = first + 1;
        const int segment = (last - low + SEED_WARPS - 1) / SEED_WARPS;
        const int begin = min(last, low + warp * segment), end = min(last, begin + segment);
        int lefts = 0, rights = 0;
        for (int base = begin; base < end; base += 32) {
            const int i = base + lane;
            const uint32_t key = i < end ? uint32_t(a[i] >> SHIFT) : 0u;
            lefts += __popc(__ballot_sync(FULL, i < end && key >= pivot));
            rights += __popc(__ballot_sync(FULL, i < end && key <= pivot));
        }
        if (lane == 0) {
            C.warp_left[warp] = lefts;
            C.warp_right[warp] = rights;
        }
        __syncthreads();
        int total_left, total_right;
        int left_rank = warp_prefix(C.warp_left, warp, lane, &total_left);
        int right_rank = warp_prefix(C.warp_right, warp, lane, &total_right);
        const int pairs = min(total_left, total_right);
        if (threadIdx.x == 0) C.swaps = pairs;
        const unsigned below = (1u << lane) - 1u;
        for (int base = begin; base < end; base += 32) {
            const int i = base + lane;
            const uint32_t key = i < end ? uint32_t(a[i] >> SHIFT) : 0u;
            const bool is_left = i < end && key >= pivot, is_right = i < end && key <= pivot;
            const unsigned left_ballot = __ballot_sync(FULL, is_left), right_ballot = __ballot_sync(FULL, is_right);
            if (is_left) left[left_rank + __popc(left_ballot & below)] = uint16_t(i - low);
            if (is_right) right[total_right - 1 - (right_rank + __popc(right_ballot & below))] = uint16_t(i - low);
            left_rank += __popc(left_ballot);
            right_rank += __popc(right_ballot);
        }
        __syncthreads();
        for (int k = threadIdx.x; k < pairs; k += SEED_THREADS) {
            const int l = low + left[k], r = low + right[k];
            if (l < r) {
                const E tmp = a[l];
                a[l] = a[r];
                a[r] = tmp;
            } else if (k == 0 || left[k - 1] < right[k - 1]) {
                C.swaps = k;
            }
        }
        __syncthreads();
        const int swaps = C.swaps;
        const int left_stop = swaps < total_left ? low + left[swaps] : last;
        const int right_stop = swaps >= 1 ? low + right[swaps - 1] : last;
        const int cut = min(left_stop, right_stop);
        if (cut <= nth)
            first = cut;
        else
            last = cut;
        __syncthreads();
    }
    if (threadIdx.x == 0) serial_select(a, first, nth, last, depth_limit, less);
    __syncthreads();
}

// seed sizes, seed weights and the bitmap of touched seeds at the head of dynamic shared memory (16-byte multiple)
__host__ __device__ inline size_t seed_tables_bytes(int seeds_padded) {
    return (size_t(seeds_padded) * 8 + size_t(seeds_padded >> 5) * 4 + 15) & ~size_t(15);
}

struct SeedsParams {
    const float* out_w;
    const int32_t* out_idx;
    const int32_t* out_ptr;
    const float* in_w;
    const int32_t* in_idx;
    const int32_t* in_ptr;
    int nodes;
    uint32_t rng;  // the generator's state after seeding
    int max_seeds;
    float min_quantile, max_quantile;
    int32_t* seed_of;
    int32_t* connected;  // per node: how many of its incoming neighbours have no seed (0 once it has one)
    int32_t* candidates[2];
    int32_t* taken;      // the nodes a seed takes, band capacity + 1
    void* select_global; // selection buffers when they do not fit shared memory, else null
    int select_capacity; // candidates the selection buffers hold
    int band_capacity;   // widest incoming band
    int32_t* result;     // [0] status, [1] seeds count, [2] widest incoming band, [3] largest given seed
};

// One warp: choose_seeds.cpp:164-228 for one node without a seed.  1 connected, 0 not yet, < 0 failure.
__device__ int connect_node(const SeedsParams& P, int node, int seeds_count, int* seed_sizes, float* seed_weights,
                            uint32_t* touched, float* staging, uint32_t& rng, int lane) {
    const int in_begin = P.in_ptr[node], in_count = P.in_ptr[node + 1] - in_begin;
    const int out_begin = P.out_ptr[node], count = in_count + P.out_ptr[node + 1] - out_begin;
    double lane_sum = 0.0;
    int smallest_exponent = 1 << 20;
    bool unusual = false;
    for (int start = 0; start < count; start += 32 * CONNECT_UNROLL) {
        int seed[CONNECT_UNROLL];
        float weight[CONNECT_UNROLL];
#pragma unroll
        for (int j = 0; j < CONNECT_UNROLL; ++j) {
            const int e = start + j * 32 + lane;
            seed[j] = -1;
            weight[j] = 0.0f;
            if (e < count) {
                const bool incoming = e < in_count;
                const int at = incoming ? in_begin + e : out_begin + e - in_count;
                seed[j] = incoming ? P.in_idx[at] : P.out_idx[at];
                weight[j] = incoming ? P.in_w[at] : P.out_w[at];
            }
        }
#pragma unroll
        for (int j = 0; j < CONNECT_UNROLL; ++j)
            if (seed[j] >= 0) seed[j] = P.seed_of[seed[j]];
#pragma unroll
        for (int j = 0; j < CONNECT_UNROLL; ++j) {
            if (start + j * 32 >= count) break;
            const bool valid = seed[j] >= 0;
            const float share = valid ? __fdiv_rn(weight[j], float(seed_sizes[seed[j]])) : 0.0f;
            const unsigned peers = __match_any_sync(FULL, valid ? seed[j] : -1 - lane);
            staging[lane] = share;
            __syncwarp();
            if (valid && lane == __ffs(peers) - 1) {  // the first lane of a seed adds its terms in edge order
                float chain = seed_weights[seed[j]];
                for (unsigned rest = peers; rest; rest &= rest - 1) chain = __fadd_rn(chain, staging[__ffs(rest) - 1]);
                seed_weights[seed[j]] = chain;
                atomicOr(&touched[seed[j] >> 5], 1u << (seed[j] & 31));
            }
            __syncwarp();
            if (valid && share != 0.0f) {
                lane_sum = __dadd_rn(lane_sum, double(share));
                const int exponent = int(__float_as_uint(share) >> 23 & 0xffu);
                unusual |= exponent == 0 || exponent == 255;
                smallest_exponent = min(smallest_exponent, exponent - 127);
            }
        }
    }
#pragma unroll
    for (int offset = 16; offset > 0; offset >>= 1) {
        lane_sum = __dadd_rn(lane_sum, __shfl_xor_sync(FULL, lane_sum, offset));
        smallest_exponent = min(smallest_exponent, __shfl_xor_sync(FULL, smallest_exponent, offset));
    }
    unusual = __any_sync(FULL, unusual);
    double total = lane_sum;
    // every term is a multiple of 2^(smallest - 23); a total below 2^53 of those units means no sum rounded
    if (total != 0.0 && (unusual || !(total <= ldexp(1.0, smallest_exponent + 29)))) {
        total = 0.0;  // the reference's order: float64 += float32, edge by edge
        for (int start = 0; start < count; start += 32) {
            const int e = start + lane;
            float share = 0.0f;
            bool valid = false;
            if (e < count) {
                const bool incoming = e < in_count;
                const int at = incoming ? in_begin + e : out_begin + e - in_count;
                const int seed = P.seed_of[incoming ? P.in_idx[at] : P.out_idx[at]];
                valid = seed >= 0;
                if (valid) share = __fdiv_rn(incoming ? P.in_w[at] : P.out_w[at], float(seed_sizes[seed]));
            }
            for (int source = 0; source < 32; ++source) {
                const float term = __shfl_sync(FULL, share, source);
                if (__shfl_sync(FULL, int(valid), source)) total = __dadd_rn(total, double(term));
            }
        }
    }
    int outcome = 0;
    if (lane == 0) {
        const int words = (seeds_count + 31) >> 5;
        if (total == 0.0) {
            for (int word = 0; word < words; ++word) {  // only zero weights were added, if any
                for (uint32_t bits = touched[word]; bits; bits &= bits - 1) seed_weights[word * 32 + __ffs(bits) - 1] = 0.0f;
                touched[word] = 0u;
            }
        } else {
            // uniform_real_distribution<double>(0, total): generate_canonical<double, 53>, two draws
            const double r = 2147483646.0;
            double sum = __dadd_rn(0.0, __dmul_rn(double(rng_next(rng) - 1u), 1.0));
            sum = __dadd_rn(sum, __dmul_rn(double(rng_next(rng) - 1u), r));
            double canonical = __ddiv_rn(sum, __dmul_rn(r, r));
            if (canonical >= 1.0) canonical = __longlong_as_double(0x3FEFFFFFFFFFFFFFLL);
            double remaining = __dadd_rn(__dmul_rn(canonical, __dsub_rn(total, 0.0)), 0.0);
            int chosen = remaining <= 0.0 ? 0 : -1;  // `weight -= 0` of a seed that shares no edge changes nothing
            if (!(0.0 <= remaining && remaining <= total)) chosen = -2;
            for (int word = 0; word < words; ++word) {
                for (uint32_t bits = touched[word]; bits; bits &= bits - 1) {
                    const int seed = word * 32 + __ffs(bits) - 1;
                    if (chosen == -1) {
                        remaining = __dsub_rn(remaining, double(seed_weights[seed]));
                        if (remaining <= 0.0) chosen = seed;
                    }
                    seed_weights[seed] = 0.0f;
                }
                touched[word] = 0u;
            }
            if (chosen >= 0) {
                P.seed_of[node] = chosen;
                seed_sizes[chosen] += 1;
                outcome = 1;
            } else {
                outcome = chosen == -2 ? -SEEDS_ASSERT : -SEEDS_UNLUCKY_DRAW;
            }
        }
    }
    __syncwarp();
    return __shfl_sync(FULL, outcome, 0);
}

__global__ void seeds_prepare_kernel(const int32_t* __restrict__ in_idx, const int32_t* __restrict__ in_ptr,
                                     const int32_t* __restrict__ seed_of, int nodes, int32_t* __restrict__ connected,
                                     int32_t* __restrict__ result) {
    const int lane = threadIdx.x & 31;
    const int warps = gridDim.x * (blockDim.x >> 5);
    int widest = 0, largest = -1;
    for (int node = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); node < nodes; node += warps) {
        const int begin = in_ptr[node], end = in_ptr[node + 1];
        const int seed = seed_of[node];
        int count = 0;
        if (seed < 0)  // collect_connected_nodes, choose_seeds.cpp:5-26
            for (int at = begin + lane; at < end; at += 32) count += seed_of[in_idx[at]] < 0;
        count = warp_reduce_sum(count);
        if (lane == 0) connected[node] = count;
        widest = max(widest, end - begin);
        largest = max(largest, seed);
    }
    if (lane == 0) {
        atomicMax(&result[2], widest);
        atomicMax(&result[3], largest);
    }
}

__global__ void __launch_bounds__(SEED_THREADS, 1) choose_seeds_kernel(const SeedsParams P) {
    extern __shared__ __align__(16) unsigned char dynamic[];
    __shared__ Control C;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nodes = P.nodes;
    const int seeds_padded = (P.max_seeds + 31) & ~31;
    int* seed_sizes = reinterpret_cast<int*>(dynamic);
    float* seed_weights = reinterpret_cast<float*>(seed_sizes + seeds_padded);
    uint32_t* touched = reinterpret_cast<uint32_t*>(seed_weights + seeds_padded);
    unsigned char* select = P.select_global ? static_cast<unsigned char*>(P.select_global)
                                            : dynamic + seed_tables_bytes(seeds_padded);
    for (int i = threadIdx.x; i < seeds_padded; i += SEED_THREADS) {
        seed_sizes[i] = 0;
        seed_weights[i] = 0.0f;
        if (i < (seeds_padded >> 5)) touched[i] = 0u;
    }
    if (threadIdx.x == 0) C.status = SEEDS_OK;
    __syncthreads();

    const int given = max(P.result[3], -1) + 1;
    int seeds_count = given;
    uint32_t rng = P.rng;  // thread 0's copy is the generator
    int status = SEEDS_OK;

    if (given < P.max_seeds) {
        int current = 0;
        int size = cta_compact(nodes, C, [&](int node) { return P.seed_of[node] < 0; },
                               [&](int rank, int node) { P.candidates[0][rank] = node; });
        int seed_size = 0;
        if (size < P.max_seeds - given)
            status = SEEDS_ASSERT;  // choose_seeds.cpp:318
        else
            seed_size = size / (P.max_seeds - given);  // ceil of an integer quotient, choose_seeds.cpp:319
        uint32_t* order = reinterpret_cast<uint32_t*>(select);
        uint16_t* order_left = reinterpret_cast<uint16_t*>(order + P.select_capacity);
        uint16_t* order_right = order_left + P.select_capacity;
        uint64_t* band = reinterpret_cast<uint64_t*>(select);
        uint16_t* band_left = reinterpret_cast<uint16_t*>(band + P.band_capacity);
        uint16_t* band_right = band_left + P.band_capacity;
        while (status == SEEDS_OK && seeds_count < P.max_seeds) {
            // keep_large_candidates (choose_seeds.cpp:154-162) + the positions of choose_seed_node keyed by size
            const int32_t* source = P.candidates[current];
            int32_t* target = P.candidates[current ^ 1];
            size = cta_compact(size, C, [&](int i) { return P.connected[source[i]] > 0; },
                               [&](int rank, int i) {
                                   const int node = source[i];
                                   target[rank] = node;
                                   order[rank] = uint32_t(P.connected[node]) << 16 | uint32_t(rank);
                               });
            current ^= 1;
            if (size == 0) break;
            // choose_seed_node, choose_seeds.cpp:28-74: size_t times float32 is a float32 product
            const int min_rank = int(floorf(__fmul_rn(float(size - 1), P.min_quantile)));
            const int max_rank = int(ceilf(__fmul_rn(float(size - 1), P.max_quantile)));
            if (!(0 <= min_rank && min_rank <= max_rank && max_rank <= size - 1)) {
                status = SEEDS_ASSERT;
                break;
            }
            cta_nth_element<uint32_t, 16>(order, order_left, order_right, 0, min_rank, size, C);
            cta_nth_element<uint32_t, 16>(order, order_left, order_right, min_rank, max_rank, size, C);
            if (threadIdx.x == 0) {
                const uint32_t selected = rng_next(rng) % uint32_t(max_rank + 1 - min_rank);
                C.seed_node = target[order[min_rank + selected] & 0xffffu];
            }
            __syncthreads();
            const int seed_node = C.seed_node;
            // store_seed_node, choose_seeds.cpp:88-151: the seed takes its heaviest incoming neighbours without a seed
            const int band_begin = P.in_ptr[seed_node], degree = P.in_ptr[seed_node + 1] - band_begin;
            int kept = cta_compact(degree, C, [&](int i) { return P.seed_of[P.in_idx[band_begin + i]] < 0; },
                                   [&](int rank, int i) {
                                       const uint32_t key = ~float_to_ordered(canonical_zero(P.in_w[band_begin + i]));
                                       band[rank] = uint64_t(key) << 32 | uint32_t(i);
                                   });
            if (seed_size < kept) {
                cta_nth_element<uint64_t, 32>(band, band_left, band_right, 0, seed_size, kept, C);
                kept = seed_size;
            }
            for (int i = threadIdx.x; i < kept; i += SEED_THREADS) {
                const int node = P.in_idx[band_begin + int(uint32_t(band[i]))];
                P.seed_of[node] = seeds_count;
                P.connected[node] = 0;
                P.taken[i] = node;
            }
            if (threadIdx.x == 0) {
                P.seed_of[seed_node] = seeds_count;
                P.connected[seed_node] = 0;
                P.taken[kept] = seed_node;
            }
            __syncthreads();
            for (int i = warp; i <= kept; i += SEED_WARPS) {  // the nodes taken leave their neighbours' lists
                const int node = P.taken[i];
                for (int at = P.out_ptr[node] + lane, end = P.out_ptr[node + 1]; at < end; at += 32) {
                    const int other = P.out_idx[at];
                    if (P.seed_of[other] < 0) atomicSub(&P.connected[other], 1);
                }
            }
            __syncthreads();
            ++seeds_count;
        }
        if (status == SEEDS_OK && seeds_count < P.max_seeds) {
            // choose_seeds.cpp:337-364: the best connected of the remaining nodes become seeds of their own
            int rest = cta_compact(nodes, C, [&](int node) { return P.seed_of[node] < 0; },
                                   [&](int rank, int node) { P.candidates[0][rank] = node; });
            if (threadIdx.x == 0) {
                const int32_t* in_ptr = P.in_ptr;
                const int32_t* out_ptr = P.out_ptr;
                serial_sort(P.candidates[0], rest, [=](int32_t left, int32_t right) {
                    const uint64_t left_edges = uint64_t(in_ptr[left + 1] - in_ptr[left] + 1) * uint64_t(out_ptr[left + 1] - out_ptr[left] + 1);
                    const uint64_t right_edges = uint64_t(in_ptr[right + 1] - in_ptr[right] + 1) * uint64_t(out_ptr[right + 1] - out_ptr[right] + 1);
                    return left_edges > right_edges;
                });
                while (rest > 0 && seeds_count < P.max_seeds) P.seed_of[P.candidates[0][--rest]] = seeds_count++;
                C.seeds_count = seeds_count;
            }
            __syncthreads();
            seeds_count = C.seeds_count;
        }
    }

    // do_complete_seeds, choose_seeds.cpp:230-262
    if (status == SEEDS_OK) {
        for (int node = threadIdx.x; node < nodes; node += SEED_THREADS) {
            const int seed = P.seed_of[node];
            if (seed >= 0) atomicAdd(&seed_sizes[seed], 1);
        }
    }
    __syncthreads();
    if (status == SEEDS_OK && warp == 0) {
        int32_t* fresh = P.candidates[0];
        int32_t* stale = P.candidates[1];
        int fresh_count = 0;
        for (int base = 0; base < nodes && status == SEEDS_OK; base += 32) {
            const int node = base + lane;
            unsigned waiting = __ballot_sync(FULL, node < nodes && P.seed_of[node] < 0);
            for (; waiting && status == SEEDS_OK; waiting &= waiting - 1) {
                const int next = base + __ffs(waiting) - 1;
                const int outcome = connect_node(P, next, seeds_count, seed_sizes, seed_weights, touched, C.staging, rng, lane);
                if (outcome < 0) status = -outcome;
                if (outcome == 0) {
                    if (lane == 0) fresh[fresh_count] = next;
                    ++fresh_count;
                }
            }
        }
        while (fresh_count > 0 && status == SEEDS_OK) {
            __syncwarp();
            int32_t* swap = stale;
            stale = fresh;
            fresh = swap;
            const int stale_count = fresh_count;
            fresh_count = 0;
            for (int i = 0; i < stale_count && status == SEEDS_OK; ++i) {
                const int next = stale[i];
                const int outcome = connect_node(P, next, seeds_count, seed_sizes, seed_weights, touched, C.staging, rng, lane);
                if (outcome < 0) status = -outcome;
                if (outcome == 0) {
                    if (lane == 0) fresh[fresh_count] = next;
                    ++fresh_count;
                }
            }
            if (status == SEEDS_OK && !(fresh_count < stale_count)) status = SEEDS_ASSERT;  // choose_seeds.cpp:260
        }
        if (lane == 0) C.status = status;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        P.result[0] = status != SEEDS_OK ? status : C.status;
        P.result[1] = seeds_count;
    }
}

}  // namespace

// Device arrays in, seed_of_nodes filled in place, the seeds count to the host.
int seeds_launch(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w, const int32_t* in_idx,
                 const int32_t* in_ptr, int64_t nodes, uint64_t random_seed, int64_t max_seeds_count,
                 float min_seed_size_quantile, float max_seed_size_quantile, int32_t* seed_of_nodes, int64_t* seeds_count,
                 cudaStream_t stream) {
    MCB_REQUIRE(nodes > 0, "failed assert: nodes_count > 0");  // choose_seeds.cpp:279
    MCB_REQUIRE(nodes <= 65535, "choose_seeds: %lld nodes; at most 65535 are supported", (long long)nodes);
    MCB_REQUIRE(out_ptr && in_ptr && seed_of_nodes && seeds_count, "choose_seeds: null argument");
    MCB_REQUIRE(0 <= min_seed_size_quantile && min_seed_size_quantile <= max_seed_size_quantile &&
                    max_seed_size_quantile <= 1.0f,
                "failed assert: 0 <= min_seed_size_quantile <= max_seed_size_quantile <= 1");  // choose_seeds.cpp:296-298
    MCB_REQUIRE(max_seeds_count >= 0 && max_seeds_count <= 16384, "choose_seeds: %lld seeds; at most 16384 are supported",
                (long long)max_seeds_count);
    const int n = int(nodes);
    Scratch connected, candidates, taken, result, select_global;
    MCB_TRY(connected.alloc(size_t(n) * 4, stream));
    MCB_TRY(candidates.alloc(size_t(n) * 8, stream));
    MCB_TRY(taken.alloc((size_t(n) + 1) * 4, stream));
    MCB_TRY(result.alloc(4 * sizeof(int32_t), stream));
    const int32_t result_init[4] = {0, 0, 0, -1};
    int32_t result_host[4];
    MCB_TRY(upload_small(result.ptr, result_init, sizeof(result_init), stream));
    MCB_PROFILED("seeds_prepare", stream, seeds_prepare_kernel<<<int(std::min<int64_t>(ceil_div(n, 8), int64_t(sm_count()) * 8)), 256, 0, stream>>>(
                     in_idx, in_ptr, seed_of_nodes, n, connected.as<int32_t>(), result.as<int32_t>()));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(result_host, result.ptr, sizeof(result_host), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));

    SeedsParams params;
    params.out_w = out_w, params.out_idx = out_idx, params.out_ptr = out_ptr;
    params.in_w = in_w, params.in_idx = in_idx, params.in_ptr = in_ptr;
    params.nodes = n;
    params.rng = uint32_t(random_seed % 2147483647ull);
    if (params.rng == 0) params.rng = 1;
    MCB_REQUIRE(int64_t(result_host[3]) + 1 <= max_seeds_count, "failed assert: seeds_count <= max_seeds_count");  // :366
    params.max_seeds = int(max_seeds_count);
    params.min_quantile = min_seed_size_quantile, params.max_quantile = max_seed_size_quantile;
    params.seed_of = seed_of_nodes;
    params.connected = connected.as<int32_t>();
    params.candidates[0] = candidates.as<int32_t>();
    params.candidates[1] = candidates.as<int32_t>() + n;
    params.taken = taken.as<int32_t>();
    params.select_capacity = (n + 7) & ~7;
    params.band_capacity = (std::max(result_host[2], 1) + 7) & ~7;
    params.result = result.as<int32_t>();
    const int seeds_padded = (params.max_seeds + 31) & ~31;
    const size_t fixed_bytes = seed_tables_bytes(seeds_padded);
    const size_t select_bytes = std::max(size_t(params.select_capacity) * 8, size_t(params.band_capacity) * 12);
    int device = 0, shared_limit = 0;
    MCB_CUDA(cudaGetDevice(&device));
    MCB_CUDA(cudaDeviceGetAttribute(&shared_limit, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    const size_t budget = size_t(shared_limit) - 2048;  // the static control block
    size_t shared_bytes = fixed_bytes + select_bytes;
    params.select_global = nullptr;
    if (shared_bytes > budget) {
        MCB_TRY(select_global.alloc(select_bytes, stream));
        params.select_global = select_global.ptr;
        shared_bytes = fixed_bytes;
    }
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(choose_seeds_kernel), budget));
    MCB_PROFILED("choose_seeds", stream, choose_seeds_kernel<<<1, SEED_THREADS, shared_bytes, stream>>>(params));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(result_host, result.ptr, sizeof(result_host), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    MCB_REQUIRE(result_host[0] != SEEDS_ASSERT, "choose_seeds: a FastAssert of the reference fails on this input");
    MCB_REQUIRE(result_host[0] != SEEDS_UNLUCKY_DRAW,
                "choose_seeds: a draw fell beyond the rounded seed weights (the reference asserts false, choose_seeds.cpp:227)");
    *seeds_count = result_host[1];
    return MCB_OK;
}

}  // namespace mcb
