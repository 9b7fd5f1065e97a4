// Per-row ordering kernels: rank_matrix (rank.cpp:38-87), rank_rows (rank.cpp:5-36) and
// collect_top (top_per.cpp:5-92).
//
// Tie rule (documented in oracle/mc_oracle.c): the reference orders with unstable std::sort /
// std::nth_element on value-only comparators, so the order of exactly equal values is an artefact
// of libstdc++'s introsort.  Here equal values are ordered by their position (lower first).
//
//   rank_matrix   one CTA (1024 threads) per row; stable LSD radix sort of (key, position) in
//                 shared memory: 4 passes of 8 bits; every warp owns a contiguous segment of the
//                 current order and ranks its elements with __match_any_sync, so a pass needs a
//                 32 x 256 table of per-warp digit counts and no atomics.  Keys stay put in shared
//                 memory; only 16-bit positions ping-pong, so a 24k-wide row (the largest pile,
//                 parameters.py:115) fits one SM.  The last pass writes the ranks straight to the
//                 row in global memory: algorithmic traffic is one read and one write of the matrix.
//   rank_rows     one CTA per row; 4-pass radix select of the rank-th smallest value.
//   collect_top   one CTA per row; radix select of the degree-th largest value, then an ordered
//                 compaction (indices come out ascending, as top_per.cpp:35 sorts them).
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include <cstdlib>

#include "select.cuh"

namespace mcb {

namespace {

constexpr int RM_THREADS = 1024;
constexpr int RM_WARPS = RM_THREADS / 32;

struct RankMatrixParams {
    float* matrix;
    int64_t rows, columns, row_stride;
    int ascending;
    int segment;          // elements per warp, a multiple of 32
    int smem_capacity;    // elements that fit the shared-memory layout (0: use the workspace)
    uint8_t* workspace;   // per CTA: keys u32[capacity] + 2 x positions u16[capacity]
    int64_t workspace_stride;
};

__global__ void __launch_bounds__(RM_THREADS, 1) rank_matrix_kernel(RankMatrixParams p) {
    extern __shared__ __align__(16) uint8_t rm_smem[];
    __shared__ int scan_scratch[34];

    uint32_t* hist = reinterpret_cast<uint32_t*>(rm_smem);  // [RM_WARPS][256]
    const int n = (int)p.columns;
    const int capacity = p.segment * RM_WARPS;  // >= n
    uint8_t* buffers = p.smem_capacity ? rm_smem + RM_WARPS * 256 * sizeof(uint32_t)
                                       : p.workspace + (int64_t)blockIdx.x * p.workspace_stride;
    uint32_t* keys = reinterpret_cast<uint32_t*>(buffers);
    uint16_t* order_a = reinterpret_cast<uint16_t*>(buffers + (size_t)capacity * 4);
    uint16_t* order_b = order_a + capacity;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lower_lanes = lanemask_lt();
    uint32_t* my_hist = hist + warp * 256;
    const int seg_begin = warp * p.segment, seg_end = min(seg_begin + p.segment, n);

    for (int64_t row = blockIdx.x; row < p.rows; row += gridDim.x) {
        float* values = p.matrix + row * p.row_stride;
        uint16_t* source = order_a;
        uint16_t* target = order_b;
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 8 * pass;
            for (int i = tid; i < RM_WARPS * 256; i += RM_THREADS) hist[i] = 0;
            __syncthreads();
            // (1) per-warp digit counts over this warp's segment of the current order
            for (int base = seg_begin; base < seg_end; base += 32) {
                const int i = base + lane;
                uint32_t digit = 0;
                if (i < seg_end) {
                    uint32_t key;
                    if (pass == 0) {
                        key = sort_key(values[i], p.ascending != 0);
                        keys[i] = key;
                    } else {
                        key = keys[source[i]];
                    }
                    digit = (key >> shift) & 255u;
                }
                const uint32_t peers = warp_peers_8bit(digit, i < seg_end);
                if (i < seg_end && (peers & lower_lanes) == 0)
                    my_hist[digit] += __popc(peers);
                __syncwarp();
            }
            __syncthreads();
            // (2) exclusive scan in (digit, warp) order
            {
                int total = 0;
                if (tid < 256)
                    for (int w = 0; w < RM_WARPS; ++w) total += (int)hist[w * 256 + tid];
                int block_total;
                const int start = block_exclusive_scan(tid < 256 ? total : 0, scan_scratch, &block_total);
                if (tid < 256) {
                    uint32_t running = (uint32_t)start;
                    for (int w = 0; w < RM_WARPS; ++w) {
                        const uint32_t count = hist[w * 256 + tid];
                        hist[w * 256 + tid] = running;
                        running += count;
                    }
                }
            }
            __syncthreads();
            // (3) stable scatter
            for (int base = seg_begin; base < seg_end; base += 32) {
                const int i = base + lane;
                uint32_t digit = 0;
                uint32_t position = 0;
                if (i < seg_end) {
                    position = pass == 0 ? (uint32_t)i : (uint32_t)source[i];
                    digit = (keys[position] >> shift) & 255u;
                }
                const uint32_t peers = warp_peers_8bit(digit, i < seg_end);
                uint32_t slot = 0;
                if (i < seg_end)
                    slot = my_hist[digit] + __popc(peers & lower_lanes);
                __syncwarp();
                if (i < seg_end) {
                    if ((peers & lower_lanes) == 0)
                        my_hist[digit] += __popc(peers);
                    if (pass == 3)
                        values[position] = (float)(slot + 1);  // rank.cpp:73-75
                    else
                        target[slot] = (uint16_t)position;
                }
                __syncwarp();
            }
            __syncthreads();
            uint16_t* swap = source;
            source = target;
            target = swap;
        }
    }
}

// ---- rows of up to 24 576 elements (every default pile): a merge sort of (key, position) pairs ---------------------------------
//
// The pair (32-bit sort key << 16 | 16-bit position) is unique per element, so any correct sort gives the
// position-ordered tie rule.  Every thread takes ITEMS elements of the row (strided: coalesced loads), sorts
// them in registers with a bitonic network, and log2(1024) merge levels follow: at each level the threads
// park their sorted run in shared memory and every thread produces ITEMS consecutive outputs of the merge of
// two runs (merge path: a binary search for its split, then a sequential merge).  About a quarter of the
// instructions of the four-pass radix sort above.  The sorted positions are turned into ranks through a
// 16-bit table in shared memory so that the row is written back coalesced.
constexpr int MS_THREADS = 1024;

__device__ __forceinline__ int ms_slot(int element) { return element + (element >> 4); }  // bank-conflict padding

// Batcher's odd-even merge sort as a compare-exchange network over registers.  The iterative construction is
// valid for any ITEMS (63 exchanges for 16, 132 for 24); the list of exchanges is built at compile time so that
// every register index is a constant after unrolling.
template <int ITEMS>
struct SortNetwork {
    int count = 0;
    int low[ITEMS * ITEMS] = {};
    int high[ITEMS * ITEMS] = {};
    constexpr SortNetwork() {
        for (int p = 1; p < ITEMS; p <<= 1)
            for (int k = p; k >= 1; k >>= 1)
                for (int j = k % p; j + k < ITEMS; j += 2 * k)
                    for (int i = 0; i < k; ++i)
                        if (i + j + k < ITEMS && (i + j) / (2 * p) == (i + j + k) / (2 * p)) {
                            low[count] = i + j;
                            high[count] = i + j + k;
                            ++count;
                        }
    }
};

template <int ITEMS>
__device__ __forceinline__ void ms_sort_registers(unsigned long long (&a)[ITEMS]) {
    constexpr SortNetwork<ITEMS> network{};
#pragma unroll
    for (int c = 0; c < network.count; ++c) {
        const unsigned long long x = a[network.low[c]], y = a[network.high[c]];
        const bool swap = x > y;
        a[network.low[c]] = swap ? y : x;
        a[network.high[c]] = swap ? x : y;
    }
}

template <int ITEMS>
__global__ void __launch_bounds__(MS_THREADS, 1) rank_matrix_merge_kernel(float* __restrict__ matrix, int64_t rows, int columns,
                                                                          int64_t row_stride, int ascending,
                                                                          const int* __restrict__ row_list,
                                                                          const int* __restrict__ row_count) {
    extern __shared__ __align__(16) uint8_t rm_smem[];
    unsigned long long* runs = reinterpret_cast<unsigned long long*>(rm_smem);  // ms_slot(MS_THREADS * ITEMS) entries
    uint16_t* rank_of = reinterpret_cast<uint16_t*>(rm_smem);                   // reuses the buffer at the end
    constexpr int CAPACITY = MS_THREADS * ITEMS;
    const int tid = threadIdx.x;
    if (row_list != nullptr) rows = *row_count;  // only the listed rows (the ones the bucket kernel passed on)
    for (int64_t entry = blockIdx.x; entry < rows; entry += gridDim.x) {
        const int64_t row = row_list != nullptr ? row_list[entry] : entry;
        float* values = matrix + row * row_stride;
        unsigned long long a[ITEMS];
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const int i = r * MS_THREADS + tid;
            a[r] = i < columns ? ((unsigned long long)sort_key(values[i], ascending != 0) << 16) | (unsigned)i : ~0ull;
        }
        ms_sort_registers<ITEMS>(a);
        for (int length = ITEMS; length < CAPACITY; length <<= 1) {
            __syncthreads();
#pragma unroll
            for (int r = 0; r < ITEMS; ++r) runs[ms_slot(tid * ITEMS + r)] = a[r];
            __syncthreads();
            const int first = tid * ITEMS;                  // this thread's outputs within the whole array
            const int pair = first / (2 * length) * (2 * length);
            const int diagonal = first - pair;              // outputs of this pair that precede this thread's
            const int left = pair, right = pair + length;   // the two runs
            int low = max(0, diagonal - length), high = min(diagonal, length);
            while (low < high) {
                const int middle = (low + high) >> 1;
                if (runs[ms_slot(left + middle)] < runs[ms_slot(right + diagonal - 1 - middle)])
                    low = middle + 1;
                else
                    high = middle;
            }
            // Sequential merge; one unconditional load per output (index CAPACITY is a padded, never-taken slot).
            int i = left + low, j = right + diagonal - low;
            const int i_end = left + length, j_end = right + length;
            unsigned long long x = runs[ms_slot(i)], y = runs[ms_slot(j)];
#pragma unroll
            for (int r = 0; r < ITEMS; ++r) {
                const bool take_left = j >= j_end || (i < i_end && x < y);
                a[r] = take_left ? x : y;
                i += take_left ? 1 : 0;
                j += take_left ? 0 : 1;
                const unsigned long long next = runs[ms_slot(take_left ? i : j)];
                x = take_left ? next : x;
                y = take_left ? y : next;
            }
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const int sorted_at = tid * ITEMS + r;
            if (sorted_at < columns) rank_of[(unsigned)a[r] & 0xffffu] = (uint16_t)(sorted_at + 1);  // rank.cpp:73-75
        }
        __syncthreads();
        for (int i = tid; i < columns; i += MS_THREADS) values[i] = (float)rank_of[i];
    }
}

// row_list / row_count (device): when given, only those rows are ranked (their count is not known on the host, so
// the grid covers the SMs a few times over and CTAs without a row return at once).
template <int ITEMS>
int rank_matrix_merge_launch(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, cudaStream_t stream,
                             const int* row_list = nullptr, const int* row_count = nullptr) {
    const size_t smem_bytes = (size_t)(MS_THREADS * ITEMS + MS_THREADS * ITEMS / 16 + 1) * 8;
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(rank_matrix_merge_kernel<ITEMS>), (size_t)smem_bytes));
    const int64_t grid = row_list != nullptr ? std::min<int64_t>(rows, (int64_t)sm_count() * 4) : std::min<int64_t>(rows, 1 << 20);
    MCB_PROFILED(row_list != nullptr ? "rank_matrix_ties" : "rank_matrix", stream,
                 rank_matrix_merge_kernel<ITEMS><<<(unsigned)grid, MS_THREADS, smem_bytes, stream>>>(
                     matrix, rows, (int)columns, row_stride, ascending, row_list, row_count));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

// ---- the same rows, faster when the values of a row are spread out: a one-pass bucket sort -------------------------------------
//
// A monotone bin of the value with about one element per bin turns sorting into counting.  Values are not spread
// evenly over their range (most similarities of a cell are small, a few are large), so the bin is the row's own
// distribution function, estimated by a coarse histogram.  Per row, with the elements in registers (strided: coalesced):
//   1. the range: largest value and second smallest (the smallest is the diagonal the caller preset to min - 1,
//      knn_graph.py:288-292, which would otherwise double the range); a histogram of 1024 linear bins over it and its
//      scan: (elements before the bin, elements inside) is a piecewise-linear estimate of every element's rank;
//   2. count: the estimated rank is the fine bin; a shared-memory atomicAdd per element, whose return value is the
//      element's slot inside its bucket;
//   3. an exclusive scan of the counts gives every bucket its place in the sorted order;
//   4. scatter of (key, position) to place + slot;
//   5. every element counts the members of its bucket that sort before it (key, then position: the tie rule of
//      the file) and writes its rank, coalesced, from the registers it was loaded into.
// About 80 instructions per element against 200 for the merge sort.  The cost of step 5 is quadratic in the bucket
// size, so a row with a bucket wider than BK_LIMIT (heavy ties, a constant row) is passed on to the merge sort
// through a list in device memory.  Ranks do not depend on which kernel produced them.
constexpr int BK_LIMIT = 32;
constexpr int BK_COARSE = MS_THREADS;  // one coarse bin per thread
constexpr float BK_COARSE_TOP = 1023.99993896484375f;  // the largest float below BK_COARSE
static_assert(BK_COARSE == 1024, "BK_COARSE_TOP is the float before 1024");

template <int ITEMS>
struct BucketPlan {
    static constexpr int CAPACITY = MS_THREADS * ITEMS;
    static constexpr int BINS = ITEMS <= 16 ? CAPACITY : 16384;
    static constexpr int PER_THREAD = BINS / MS_THREADS;  // bins scanned by one thread: 4, 8 or 16
    static constexpr size_t BASE_BYTES = (size_t)(BINS + 4) * 4;
    static constexpr size_t KEYS_BYTES = (size_t)(CAPACITY + 4) * 4;  // padded: the counting loop reads in fours
    static constexpr size_t POSITIONS_BYTES = (size_t)CAPACITY * 2;
    static constexpr size_t SCRATCH_BYTES = 1024;
    static constexpr size_t COARSE_BYTES = (size_t)BK_COARSE * 8;
    static constexpr size_t SMEM = BASE_BYTES + KEYS_BYTES + POSITIONS_BYTES + SCRATCH_BYTES + COARSE_BYTES;
};

template <int ITEMS>
__global__ void __launch_bounds__(MS_THREADS, 1) rank_matrix_bucket_kernel(float* __restrict__ matrix, int64_t rows, int columns,
                                                                           int64_t row_stride, int ascending,
                                                                           int* __restrict__ passed_rows,
                                                                           int* __restrict__ passed_count) {
    using Plan = BucketPlan<ITEMS>;
    extern __shared__ __align__(16) uint8_t rm_smem[];
    int* base = reinterpret_cast<int*>(rm_smem);                                                       // BINS + 4
    uint32_t* keys = reinterpret_cast<uint32_t*>(rm_smem + Plan::BASE_BYTES);                          // CAPACITY
    uint16_t* positions = reinterpret_cast<uint16_t*>(rm_smem + Plan::BASE_BYTES + Plan::KEYS_BYTES);  // CAPACITY
    int* scratch = reinterpret_cast<int*>(rm_smem + Plan::BASE_BYTES + Plan::KEYS_BYTES + Plan::POSITIONS_BYTES);
    int2* coarse = reinterpret_cast<int2*>(rm_smem + Plan::BASE_BYTES + Plan::KEYS_BYTES + Plan::POSITIONS_BYTES + Plan::SCRATCH_BYTES);
    float* extremes = reinterpret_cast<float*>(scratch + 40);  // three values per warp
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float v[ITEMS];  // the row being ranked (the grid has a CTA per row up to 2^20 rows: the loop is for more)
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        float* values = matrix + row * row_stride;
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const int i = r * MS_THREADS + tid;
            v[r] = i < columns ? values[i] : 0.0f;
        }
        // 1. largest, smallest and second smallest value of the row
        float largest = -INFINITY, least = INFINITY, second = INFINITY;
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const int i = r * MS_THREADS + tid;
            v[r] = canonical_zero(v[r]);
            if (i < columns) {
                largest = fmaxf(largest, v[r]);
                second = fminf(second, fmaxf(least, v[r]));
                least = fminf(least, v[r]);
            }
        }
#pragma unroll
        for (int offset = 16; offset > 0; offset >>= 1) {
            const float other_largest = __shfl_xor_sync(0xffffffffu, largest, offset);
            const float other_least = __shfl_xor_sync(0xffffffffu, least, offset);
            const float other_second = __shfl_xor_sync(0xffffffffu, second, offset);
            largest = fmaxf(largest, other_largest);
            second = fminf(fminf(second, other_second), fmaxf(least, other_least));
            least = fminf(least, other_least);
        }
        __syncthreads();  // the previous row is done with the scratch and the buffers
        if (lane == 0) {
            extremes[warp] = largest;
            extremes[32 + warp] = least;
            extremes[64 + warp] = second;
        }
        // clear the counts
        {
            int4* base4 = reinterpret_cast<int4*>(base);
            for (int i = tid; i < (Plan::BINS + 4) / 4; i += MS_THREADS) base4[i] = make_int4(0, 0, 0, 0);
            coarse[tid] = make_int2(0, 0);
        }
        __syncthreads();
        {
            largest = extremes[lane];
            least = extremes[32 + lane];
            second = extremes[64 + lane];
#pragma unroll
            for (int offset = 16; offset > 0; offset >>= 1) {
                const float other_largest = __shfl_xor_sync(0xffffffffu, largest, offset);
                const float other_least = __shfl_xor_sync(0xffffffffu, least, offset);
                const float other_second = __shfl_xor_sync(0xffffffffu, second, offset);
                largest = fmaxf(largest, other_largest);
                second = fminf(fminf(second, other_second), fmaxf(least, other_least));
                least = fminf(least, other_least);
            }
        }
        // Coarse bins are linear in the value and monotone in the sort order: ascending from the second smallest
        // value up, descending from the largest value down; whatever falls outside is clamped into the first / last.
        const float low = isfinite(second) ? second : least;
        const float range = largest - low;
        const float scale = (range > 0.0f && isfinite(range)) ? (float)BK_COARSE / range : 0.0f;
        const float origin = ascending ? low : largest;
        const float direction = ascending ? scale : -scale;
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const int i = r * MS_THREADS + tid;
            if (i < columns) {
                const float t = fminf(fmaxf((v[r] - origin) * direction, 0.0f), BK_COARSE_TOP);
                atomicAdd(&coarse[__float2int_rz(t)].y, 1);
            }
        }
        __syncthreads();
        {
            // coarse[c] = (elements of earlier coarse bins, elements of this one): a piecewise-linear estimate of the
            // row's distribution function, whose value at an element is roughly its rank
            int total;
            const int count = coarse[tid].y;
            const int before = block_exclusive_scan(count, scratch, &total);
            coarse[tid].x = before;
        }
        __syncthreads();
        // 2. fine bins: the estimated rank, scaled to the number of bins; monotone in the sort order as well
        const float ratio = (float)Plan::BINS / (float)columns;
        uint32_t placed[ITEMS];  // bin | slot inside the bucket << 16 (saturated: such a row is passed on anyway)
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const int i = r * MS_THREADS + tid;
            placed[r] = 0;
            if (i < columns) {
                const float t = fminf(fmaxf((v[r] - origin) * direction, 0.0f), BK_COARSE_TOP);
                const int c = __float2int_rz(t);
                const int2 estimate = coarse[c];
                const float rank_estimate = (float)estimate.x + (t - (float)c) * (float)estimate.y;
                const int bin = min(__float2int_rz(rank_estimate * ratio), Plan::BINS - 1);
                const int slot = atomicAdd(&base[bin], 1);
                placed[r] = (uint32_t)bin | ((uint32_t)min(slot, 0xffff) << 16);
            }
        }
        __syncthreads();
        // 3. exclusive scan of the counts; the widest bucket decides whether the row stays here
        int counts[Plan::PER_THREAD], sum = 0, widest = 0;
        {
            const int4* base4 = reinterpret_cast<const int4*>(base) + tid * (Plan::PER_THREAD / 4);
#pragma unroll
            for (int q = 0; q < Plan::PER_THREAD / 4; ++q) {
                const int4 four = base4[q];
                counts[4 * q] = four.x, counts[4 * q + 1] = four.y, counts[4 * q + 2] = four.z, counts[4 * q + 3] = four.w;
            }
#pragma unroll
            for (int b = 0; b < Plan::PER_THREAD; ++b) {
                sum += counts[b];
                widest = max(widest, counts[b]);
            }
        }
        int total;
        int before = block_exclusive_scan(sum, scratch, &total);
        const int too_wide = __syncthreads_or(widest > BK_LIMIT);
        if (too_wide) {
            if (tid == 0) passed_rows[atomicAdd(passed_count, 1)] = (int)row;
            continue;
        }
        {
            int4* base4 = reinterpret_cast<int4*>(base) + tid * (Plan::PER_THREAD / 4);
#pragma unroll
            for (int q = 0; q < Plan::PER_THREAD / 4; ++q) {
                int4 four;  // the bucket's place in the sorted order, its size above it (the row has at most 24 576 elements)
                four.x = before | (counts[4 * q] << 16), before += counts[4 * q];
                four.y = before | (counts[4 * q + 1] << 16), before += counts[4 * q + 1];
                four.z = before | (counts[4 * q + 2] << 16), before += counts[4 * q + 2];
                four.w = before | (counts[4 * q + 3] << 16), before += counts[4 * q + 3];
                base4[q] = four;
            }
        }
        __syncthreads();
        // 4. scatter
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const int i = r * MS_THREADS + tid;
            if (i < columns) {
                const int at = (base[placed[r] & 0xffffu] & 0xffff) + (int)(placed[r] >> 16);
                keys[at] = sort_key(v[r], ascending != 0);
                positions[at] = (uint16_t)i;
            }
        }
        __syncthreads();
        // 5. rank = elements of earlier buckets + members of the own bucket that sort before
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const int i = r * MS_THREADS + tid;
            if (i < columns) {
                const int packed = base[placed[r] & 0xffffu];
                const int first = packed & 0xffff, last = first + (packed >> 16);
                const uint32_t key = sort_key(v[r], ascending != 0);
                // four members at a time, the ones past the end of the bucket masked (the buffer is padded): most
                // buckets take one round, and the four loads are independent
                int smaller = 0, not_larger = 0;
#pragma unroll 1
                for (int at = first; at < last; at += 4) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const uint32_t other = keys[at + u];
                        const bool inside = at + u < last;
                        smaller += (inside && other < key) ? 1 : 0;
                        not_larger += (inside && other <= key) ? 1 : 0;
                    }
                }
                int earlier = first + smaller;
                if (not_larger - smaller > 1) {  // equal keys besides the element itself: earlier positions first
                    for (int at = first; at < last; ++at) earlier += (keys[at] == key && (int)positions[at] < i) ? 1 : 0;
                }
                values[i] = (float)(earlier + 1);  // rank.cpp:73-75
            }
        }
    }
}

template <int ITEMS>
int rank_matrix_bucket_launch(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, cudaStream_t stream) {
    using Plan = BucketPlan<ITEMS>;
    Scratch passed;  // [0]: count, [1..]: the rows passed on to the merge sort
    MCB_TRY(passed.alloc((size_t)(rows + 1) * sizeof(int), stream));
    MCB_CUDA(cudaMemsetAsync(passed.ptr, 0, sizeof(int), stream));
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(rank_matrix_bucket_kernel<ITEMS>), (size_t)Plan::SMEM));
    MCB_PROFILED("rank_matrix", stream, rank_matrix_bucket_kernel<ITEMS><<<(unsigned)std::min<int64_t>(rows, 1 << 20), MS_THREADS, Plan::SMEM, stream>>>(
                     matrix, rows, (int)columns, row_stride, ascending, passed.as<int>() + 1, passed.as<int>()));
    MCB_LAUNCH_CHECK();
    return rank_matrix_merge_launch<ITEMS>(matrix, rows, columns, row_stride, ascending, stream, passed.as<int>() + 1, passed.as<int>());
}

// MCB_RANK_MERGE=1 in the environment keeps every row on the merge sort (for comparisons).
bool merge_sort_only() {
    static const bool only = [] {
        const char* text = std::getenv("MCB_RANK_MERGE");
        return text != nullptr && text[0] == '1';
    }();
    return only;
}

template <int ITEMS>
int rank_matrix_sorted_launch(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, cudaStream_t stream) {
    if (merge_sort_only()) return rank_matrix_merge_launch<ITEMS>(matrix, rows, columns, row_stride, ascending, stream);
    return rank_matrix_bucket_launch<ITEMS>(matrix, rows, columns, row_stride, ascending, stream);
}

}  // namespace

// In place: every row of matrix[rows, columns] (row stride in elements) becomes its 1-based ranks.
int rank_matrix_launch(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, cudaStream_t stream) {
    if (rows == 0 || columns == 0)
        return MCB_OK;
    MCB_REQUIRE(columns <= 65535, "rank_matrix: rows wider than 65535 elements are not supported (got %lld)", (long long)columns);
    if (columns <= MS_THREADS * 4) return rank_matrix_sorted_launch<4>(matrix, rows, columns, row_stride, ascending, stream);
    if (columns <= MS_THREADS * 8) return rank_matrix_sorted_launch<8>(matrix, rows, columns, row_stride, ascending, stream);
    if (columns <= MS_THREADS * 16) return rank_matrix_sorted_launch<16>(matrix, rows, columns, row_stride, ascending, stream);
    if (columns <= MS_THREADS * 24) return rank_matrix_sorted_launch<24>(matrix, rows, columns, row_stride, ascending, stream);
    RankMatrixParams p{};
    p.matrix = matrix;
    p.rows = rows;
    p.columns = columns;
    p.row_stride = row_stride;
    p.ascending = ascending;
    p.segment = (int)(ceil_div(ceil_div(columns, RM_WARPS), 32) * 32);
    const int64_t capacity = (int64_t)p.segment * RM_WARPS;
    const size_t hist_bytes = RM_WARPS * 256 * sizeof(uint32_t);
    const size_t buffer_bytes = (size_t)capacity * 8;
    const size_t smem_limit = 227 * 1024 - 2048;  // static shared memory of the kernel comes on top
    int64_t grid = rows;
    size_t smem_bytes = hist_bytes;
    Scratch workspace;
    if (hist_bytes + buffer_bytes <= smem_limit) {
        p.smem_capacity = (int)capacity;
        smem_bytes += buffer_bytes;
    } else {
        grid = std::min<int64_t>(rows, sm_count());
        p.workspace_stride = (int64_t)((buffer_bytes + 255) / 256 * 256);
        MCB_TRY(workspace.alloc((size_t)grid * p.workspace_stride, stream));
        p.workspace = workspace.as<uint8_t>();
    }
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(rank_matrix_kernel), (size_t)smem_bytes));
    MCB_PROFILED("rank_matrix", stream, rank_matrix_kernel<<<(unsigned)grid, RM_THREADS, smem_bytes, stream>>>(p));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

namespace {

constexpr int SEL_THREADS = 512;

// rank.cpp:5-36: output[row] = the value of 0-based rank `rank` in ascending order.
__global__ void __launch_bounds__(SEL_THREADS) rank_rows_kernel(const float* __restrict__ matrix, int64_t rows, int64_t columns,
                                                               int64_t row_stride, float* __restrict__ output, int rank,
                                                               int cache_capacity) {
    extern __shared__ __align__(16) uint8_t sel_smem[];
    __shared__ SelectState state;
    uint32_t* cache = reinterpret_cast<uint32_t*>(sel_smem);
    const int n = (int)columns;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const float* values = matrix + row * row_stride;
        int residual;
        uint32_t key;
        if (n <= cache_capacity) {
            for (int i = threadIdx.x; i < n; i += blockDim.x) cache[i] = sort_key(values[i], true);
            __syncthreads();
            key = block_radix_select([&](int i) { return cache[i]; }, n, rank, &state, &residual);
        } else {
            key = block_radix_select([&](int i) { return sort_key(values[i], true); }, n, rank, &state, &residual);
        }
        if (threadIdx.x == 0)
            output[row] = ordered_to_float(key);
    }
}

struct CollectTopParams {
    const float* similarity;
    int64_t rows, columns, row_stride;
    int degree;
    int32_t* out_indices;
    float* out_data;
    int ranks;
    int cache_capacity;
};

// top_per.cpp:5-64.  Shared memory: keys cache u32[cache_capacity] + kept keys u32[degree].
__global__ void __launch_bounds__(SEL_THREADS) collect_top_kernel(CollectTopParams p) {
    extern __shared__ __align__(16) uint8_t sel_smem[];
    __shared__ SelectState state;
    __shared__ int scan_scratch[34];
    __shared__ int carry[2];
    uint32_t* cache = reinterpret_cast<uint32_t*>(sel_smem);
    uint32_t* kept_keys = cache + p.cache_capacity;
    const int n = (int)p.columns, tid = threadIdx.x;
    for (int64_t row = blockIdx.x; row < p.rows; row += gridDim.x) {
        const float* values = p.similarity + row * p.row_stride;
        int32_t* row_indices = p.out_indices + row * p.degree;
        float* row_data = p.out_data + row * p.degree;
        const bool cached = n <= p.cache_capacity;
        if (cached) {
            for (int i = tid; i < n; i += blockDim.x) cache[i] = sort_key(values[i], false);
            __syncthreads();
        }
        auto key_at = [&](int i) { return cached ? cache[i] : sort_key(values[i], false); };
        int residual;
        const uint32_t threshold = block_radix_select(key_at, n, p.degree - 1, &state, &residual);
        const int ties_wanted = residual + 1;  // equal-to-threshold elements to keep, lowest positions first
        if (tid == 0) {
            carry[0] = 0;  // kept so far
            carry[1] = 0;  // ties seen so far
        }
        __syncthreads();
        for (int base = 0; base < n; base += blockDim.x) {
            const int i = base + tid;
            const uint32_t key = i < n ? key_at(i) : 0xffffffffu;
            const int is_tie = (i < n && key == threshold) ? 1 : 0;
            int ties_total, kept_total;
            const int ties_before = carry[1] + block_exclusive_scan(is_tie, scan_scratch, &ties_total);
            const int keep = (i < n && (key < threshold || (is_tie && ties_before < ties_wanted))) ? 1 : 0;
            const int slot = carry[0] + block_exclusive_scan(keep, scan_scratch, &kept_total);
            if (keep) {
                row_indices[slot] = i;
                if (p.ranks)
                    kept_keys[slot] = key;
                else
                    row_data[slot] = values[i];
            }
            __syncthreads();
            if (tid == 0) {
                carry[0] += kept_total;
                carry[1] += ties_total;
            }
            __syncthreads();
        }
        if (p.ranks) {
            // top_per.cpp:44-63: 1-based rank in ASCENDING value order among the kept values (keys are
            // descending, so a larger key is a smaller value); equal values by position.
            for (int e = tid; e < p.degree; e += blockDim.x) {
                const uint32_t mine = kept_keys[e];
                int before = 0;
                for (int f = 0; f < p.degree; ++f) {
                    const uint32_t other = kept_keys[f];
                    before += (other > mine || (other == mine && f < e)) ? 1 : 0;
                }
                row_data[e] = (float)(before + 1);
            }
        }
        __syncthreads();
    }
}

// top_per.cpp:5-64 for rows whose keys fit shared memory (columns <= CT_MAX_CACHED).  Three sweeps over the cached
// keys and a handful of block barriers per row:
//   1. load + histogram of the top 12 key bits (shared-memory atomics);
//   2. the bin holding the degree-th key; its members (few: 2^20 values share a bin) become candidates, and an
//      8/8/4-bit radix select over the candidates alone yields the threshold key and how many of its ties to keep;
//   3. every warp owns a contiguous segment: count kept elements per segment, scan the 16 counts, write.  An
//      element's slot = (# smaller keys before it) + min(# ties before it, ties kept): indices come out ascending
//      and ties are kept lowest position first, without a scan per 512 elements.
constexpr int CT_MAX_CACHED = 32768;
constexpr int CT_BINS = 4096;

struct CollectTopState {
    int bin, rank_in_bin, candidates;
    uint32_t prefix;
    int rank;
    int kept_before[SEL_THREADS / 32 + 1], ties_before[SEL_THREADS / 32 + 1];
    int scan[34];
};

__global__ void __launch_bounds__(SEL_THREADS) collect_top_cached_kernel(CollectTopParams p) {
    extern __shared__ __align__(16) uint8_t sel_smem[];
    __shared__ CollectTopState state;
    const int n = (int)p.columns, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int WARPS = SEL_THREADS / 32;
    uint32_t* keys = reinterpret_cast<uint32_t*>(sel_smem);          // [n]
    uint32_t* candidates = keys + p.cache_capacity;                   // [n]
    int* hist = reinterpret_cast<int*>(candidates + p.cache_capacity);  // [CT_BINS]
    uint32_t* kept_keys = reinterpret_cast<uint32_t*>(hist + CT_BINS);  // [degree] (ranks mode)
    const uint32_t lower_lanes = lanemask_lt();
    const int segment = ((n + WARPS - 1) / WARPS + 31) / 32 * 32;  // elements per warp in sweep 3
    for (int64_t row = blockIdx.x; row < p.rows; row += gridDim.x) {
        const float* values = p.similarity + row * p.row_stride;
        int32_t* row_indices = p.out_indices + row * p.degree;
        float* row_data = p.out_data + row * p.degree;
        for (int i = tid; i < CT_BINS; i += SEL_THREADS) hist[i] = 0;
        if (tid == 0) state.candidates = 0;
        __syncthreads();
        // sweep 1: the row into shared memory (several loads in flight per thread), then the histogram
        for (int base = 0; base < n; base += 4 * SEL_THREADS) {
            float loaded[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = base + u * SEL_THREADS + tid;
                loaded[u] = i < n ? values[i] : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = base + u * SEL_THREADS + tid;
                if (i < n) keys[i] = sort_key(loaded[u], false);
            }
        }
        __syncthreads();
        for (int base = 0; base < n; base += SEL_THREADS) {
            const int i = base + tid;
            const uint32_t key = i < n ? keys[i] : 0;
            // one shared-memory atomic per element: lanes with equal digits serialise in the memory pipe, which is
            // cheaper than building their mask with a dozen ballots (ncu: ballots were 40 % of the instructions)
            if (i < n) atomicAdd(&hist[key >> 20], 1);
        }
        __syncthreads();
        {   // the bin of the (degree - 1)-th key in ascending key order: thread t owns bins 8t .. 8t+7
            int counts[CT_BINS / SEL_THREADS], sum = 0;
#pragma unroll
            for (int b = 0; b < CT_BINS / SEL_THREADS; ++b) {
                counts[b] = hist[tid * (CT_BINS / SEL_THREADS) + b];
                sum += counts[b];
            }
            int total;
            int before = block_exclusive_scan(sum, state.scan, &total);
            const int wanted = p.degree - 1;
            if (wanted >= before && wanted < before + sum) {
#pragma unroll
                for (int b = 0; b < CT_BINS / SEL_THREADS; ++b) {
                    if (wanted >= before && wanted < before + counts[b]) {
                        state.bin = tid * (CT_BINS / SEL_THREADS) + b;
                        state.rank_in_bin = wanted - before;
                    }
                    before += counts[b];
                }
            }
        }
        __syncthreads();
        // sweep 2: the members of that bin
        const uint32_t bin = (uint32_t)state.bin;
        for (int base = 0; base < n; base += SEL_THREADS) {
            const int i = base + tid;
            const uint32_t key = i < n ? keys[i] : 0;
            const bool member = i < n && (key >> 20) == bin;
            const uint32_t mask = __ballot_sync(0xffffffffu, member);
            if (mask != 0) {
                int start = 0;
                if (lane == 0) start = atomicAdd(&state.candidates, __popc(mask));
                start = __shfl_sync(0xffffffffu, start, 0);
                if (member) candidates[start + __popc(mask & lower_lanes)] = key;
            }
        }
        __syncthreads();
        // radix select of the low 20 bits among the candidates: digits of 8, 8 and 4 bits
        const int members = state.candidates;
        if (tid == 0) {
            state.prefix = bin << 20;
            state.rank = state.rank_in_bin;
        }
        uint32_t known_mask = 0xfff00000u;
#pragma unroll 1
        for (int round = 0; round < 3; ++round) {
            const int shift = round == 0 ? 12 : (round == 1 ? 4 : 0);
            const uint32_t digit_mask = round == 2 ? 15u : 255u;
            for (int i = tid; i < 256; i += SEL_THREADS) hist[i] = 0;
            __syncthreads();
            const uint32_t prefix = state.prefix;
            for (int base = 0; base < members; base += SEL_THREADS) {
                const int i = base + tid;
                uint32_t digit = 0;
                bool valid = false;
                if (i < members) {
                    const uint32_t key = candidates[i];
                    valid = (key & known_mask) == prefix;
                    digit = (key >> shift) & digit_mask;
                }
                if (valid) atomicAdd(&hist[digit], 1);
            }
            __syncthreads();
            if (tid < 32) {
                int counts[8], sum = 0;
#pragma unroll
                for (int b = 0; b < 8; ++b) {
                    counts[b] = hist[lane * 8 + b];
                    sum += counts[b];
                }
                const int inclusive = warp_inclusive_scan(sum, lane);
                int before = inclusive - sum;
                const int wanted = state.rank;
                __syncwarp();
                if (wanted >= before && wanted < inclusive) {
#pragma unroll
                    for (int b = 0; b < 8; ++b) {
                        if (wanted >= before && wanted < before + counts[b]) {
                            state.prefix = prefix | ((uint32_t)(lane * 8 + b) << shift);
                            state.rank = wanted - before;
                        }
                        before += counts[b];
                    }
                }
            }
            known_mask |= digit_mask << shift;
            __syncthreads();
        }
        const uint32_t threshold = state.prefix;
        const int ties_wanted = state.rank + 1;  // equal-to-threshold elements to keep, lowest positions first
        // sweep 3a: smaller keys and ties per warp segment
        const int seg_begin = min(warp * segment, n), seg_end = min(seg_begin + segment, n);
        int smaller = 0, ties = 0;
        for (int base = seg_begin; base < seg_end; base += 32) {
            const int i = base + lane;
            const uint32_t key = i < seg_end ? keys[i] : 0xffffffffu;
            smaller += __popc(__ballot_sync(0xffffffffu, i < seg_end && key < threshold));
            ties += __popc(__ballot_sync(0xffffffffu, i < seg_end && key == threshold));
        }
        if (lane == 0) {
            state.kept_before[warp] = smaller;
            state.ties_before[warp] = ties;
        }
        __syncthreads();
        if (tid == 0) {
            int running_smaller = 0, running_ties = 0;
            for (int w = 0; w < WARPS; ++w) {
                const int s_count = state.kept_before[w], t_count = state.ties_before[w];
                state.kept_before[w] = running_smaller;
                state.ties_before[w] = running_ties;
                running_smaller += s_count;
                running_ties += t_count;
            }
        }
        __syncthreads();
        // sweep 3b: write, ascending positions
        smaller = state.kept_before[warp];
        ties = state.ties_before[warp];
        for (int base = seg_begin; base < seg_end; base += 32) {
            const int i = base + lane;
            const uint32_t key = i < seg_end ? keys[i] : 0xffffffffu;
            const bool is_smaller = i < seg_end && key < threshold, is_tie = i < seg_end && key == threshold;
            const uint32_t smaller_mask = __ballot_sync(0xffffffffu, is_smaller), tie_mask = __ballot_sync(0xffffffffu, is_tie);
            const int smaller_before = smaller + __popc(smaller_mask & lower_lanes);
            const int ties_before = ties + __popc(tie_mask & lower_lanes);
            if (is_smaller || (is_tie && ties_before < ties_wanted)) {
                const int slot = smaller_before + min(ties_before, ties_wanted);
                row_indices[slot] = i;
                if (p.ranks)
                    kept_keys[slot] = key;
                else
                    row_data[slot] = values[i];
            }
            smaller += __popc(smaller_mask);
            ties += __popc(tie_mask);
        }
        if (p.ranks) {
            __syncthreads();
            // top_per.cpp:44-63: 1-based rank in ASCENDING value order among the kept values; equal values by position.
            for (int e = tid; e < p.degree; e += SEL_THREADS) {
                const uint32_t mine = kept_keys[e];
                int before = 0;
                for (int f = 0; f < p.degree; ++f) {
                    const uint32_t other = kept_keys[f];
                    before += (other > mine || (other == mine && f < e)) ? 1 : 0;
                }
                row_data[e] = (float)(before + 1);
            }
        }
        __syncthreads();
    }
}

}  // namespace

int rank_rows_launch(const float* matrix, int64_t rows, int64_t columns, int64_t row_stride, float* output, int64_t rank,
                     cudaStream_t stream) {
    if (rows == 0)
        return MCB_OK;
    MCB_REQUIRE(rank >= 0 && rank < columns, "rank_rows: rank %lld is out of range for %lld columns", (long long)rank, (long long)columns);
    MCB_REQUIRE(columns < (1ll << 31), "rank_rows: too many columns");
    int cache_capacity = columns <= 16384 ? (int)columns : (columns <= 48 * 1024 ? (int)columns : 0);
    const size_t smem_bytes = (size_t)cache_capacity * 4;
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(rank_rows_kernel), (size_t)smem_bytes));
    const int64_t grid = std::min<int64_t>(rows, (int64_t)sm_count() * 8);
    MCB_PROFILED("rank_rows", stream, rank_rows_kernel<<<(unsigned)grid, SEL_THREADS, smem_bytes, stream>>>(matrix, rows, columns, row_stride, output, (int)rank,
                                                                        cache_capacity));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

int collect_top_launch(int64_t degree, const float* similarity, int64_t rows, int64_t columns, int64_t row_stride,
                       int32_t* out_indices, float* out_data, int ranks, cudaStream_t stream) {
    if (rows == 0)
        return MCB_OK;
    MCB_REQUIRE(degree > 0 && degree < columns, "collect_top: degree %lld must be in 1..columns-1 (%lld columns)",
                (long long)degree, (long long)columns);  // top_per.cpp:81-82
    MCB_REQUIRE(columns < (1ll << 31), "collect_top: too many columns");
    CollectTopParams p{};
    p.similarity = similarity;
    p.rows = rows;
    p.columns = columns;
    p.row_stride = row_stride;
    p.degree = (int)degree;
    p.out_indices = out_indices;
    p.out_data = out_data;
    p.ranks = ranks;
    const size_t kept_bytes = ranks ? (size_t)degree * 4 : 0;
    MCB_REQUIRE(kept_bytes <= 96 * 1024, "collect_top: degree %lld is too large for the ranks mode", (long long)degree);
    if (columns <= CT_MAX_CACHED && (size_t)columns * 8 + CT_BINS * 4 + kept_bytes <= 220 * 1024) {
        p.cache_capacity = (int)columns;
        const size_t cached_bytes = (size_t)columns * 8 + CT_BINS * 4 + kept_bytes;
        MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(collect_top_cached_kernel), (size_t)cached_bytes));
        const int64_t cached_grid = std::min<int64_t>(rows, (int64_t)sm_count() * 8);
        MCB_PROFILED("collect_top", stream, collect_top_cached_kernel<<<(unsigned)cached_grid, SEL_THREADS, cached_bytes, stream>>>(p));
        MCB_LAUNCH_CHECK();
        return MCB_OK;
    }
    p.cache_capacity = ((size_t)columns * 4 + kept_bytes <= 200 * 1024) ? (int)columns : 0;
    const size_t smem_bytes = (size_t)p.cache_capacity * 4 + kept_bytes;
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(collect_top_kernel), (size_t)smem_bytes));
    const int64_t grid = std::min<int64_t>(rows, (int64_t)sm_count() * 8);
    MCB_PROFILED("collect_top", stream, collect_top_kernel<<<(unsigned)grid, SEL_THREADS, smem_bytes, stream>>>(p));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

}  // namespace mcb

// ---- rank_matrix without any allocation ---------------------------------------------------------------------------------
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same two kernels as
// rank_matrix_launch (bucket sort; rows with wide groups of ties passed on to the merge sort), with the list of passed
// rows in a file-scope device array instead of the stream-ordered pool.  Every GPU that loads the library holds the
// array (257 KiB) until exit; every call shares it, so calls that overlap on the GPU race.

#include <mutex>

#include "../../include/mcb200.h"

namespace mcb {
namespace {

constexpr int64_t NOALLOC_MAX_ROWS = MCB_RANK_MATRIX_NOALLOC_MAX_ROWS;
constexpr int64_t NOALLOC_MAX_COLUMNS = MCB_RANK_MATRIX_NOALLOC_MAX_COLUMNS;
static_assert(NOALLOC_MAX_COLUMNS == MS_THREADS * 24, "the widest row the shared-memory rank kernels take");

// [0]: the count of passed rows (cleared by every call), [64 ...]: the rows.  In host code this name is a stand-in, not
// a device address: the address comes from cudaGetSymbolAddress, once per device.
static __device__ __align__(256) int rank_noalloc_scratch[64 + NOALLOC_MAX_ROWS];

int rank_noalloc_scratch_address(int** address) {
    constexpr int MAX_DEVICES = 64;
    static std::mutex mutex;
    static int* of_device[MAX_DEVICES] = {};
    int device = 0;
    MCB_CUDA(cudaGetDevice(&device));
    MCB_REQUIRE(device >= 0 && device < MAX_DEVICES, "rank_matrix_noalloc: device %d is beyond the %d this build knows", device, MAX_DEVICES);
    std::lock_guard<std::mutex> lock(mutex);
    if (of_device[device] == nullptr) {
        void* pointer = nullptr;
        MCB_CUDA(cudaGetSymbolAddress(&pointer, rank_noalloc_scratch));  // with `device` current: this device's copy
        of_device[device] = static_cast<int*>(pointer);
    }
    *address = of_device[device];
    return MCB_OK;
}

template <int ITEMS>
int rank_matrix_noalloc_launch(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, int* scratch,
                               cudaStream_t stream) {
    using Plan = BucketPlan<ITEMS>;
    int* passed_count = scratch;
    int* passed_rows = scratch + 64;
    MCB_CUDA(cudaMemsetAsync(passed_count, 0, sizeof(int), stream));  // the array keeps the last call's contents
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(rank_matrix_bucket_kernel<ITEMS>), (size_t)Plan::SMEM));
    rank_matrix_bucket_kernel<ITEMS><<<(unsigned)rows, MS_THREADS, Plan::SMEM, stream>>>(matrix, rows, (int)columns, row_stride,
                                                                                        ascending, passed_rows, passed_count);
    MCB_LAUNCH_CHECK();
    const size_t smem_bytes = (size_t)(MS_THREADS * ITEMS + MS_THREADS * ITEMS / 16 + 1) * 8;
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(rank_matrix_merge_kernel<ITEMS>), smem_bytes));
    const int64_t grid = std::min<int64_t>(rows, (int64_t)sm_count() * 4);
    rank_matrix_merge_kernel<ITEMS><<<(unsigned)grid, MS_THREADS, smem_bytes, stream>>>(matrix, rows, (int)columns, row_stride,
                                                                                       ascending, passed_rows, passed_count);
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

}  // namespace
}  // namespace mcb

extern "C" int mcb_dev_rank_matrix_noalloc_f32(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending,
                                               void* stream) {
    using namespace mcb;
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "rank_matrix_noalloc: bad shape");
    MCB_REQUIRE(rows <= NOALLOC_MAX_ROWS, "rank_matrix_noalloc: %lld rows; the static scratch holds at most %lld",
                (long long)rows, (long long)NOALLOC_MAX_ROWS);
    MCB_REQUIRE(columns <= NOALLOC_MAX_COLUMNS, "rank_matrix_noalloc: %lld columns; at most %lld (wider rows need a workspace)",
                (long long)columns, (long long)NOALLOC_MAX_COLUMNS);
    if (rows == 0 || columns == 0)
        return MCB_OK;
    int* scratch = nullptr;
    MCB_TRY(rank_noalloc_scratch_address(&scratch));
    cudaStream_t s = (cudaStream_t)stream;
    if (columns <= MS_THREADS * 4) return rank_matrix_noalloc_launch<4>(matrix, rows, columns, row_stride, ascending, scratch, s);
    if (columns <= MS_THREADS * 8) return rank_matrix_noalloc_launch<8>(matrix, rows, columns, row_stride, ascending, scratch, s);
    if (columns <= MS_THREADS * 16) return rank_matrix_noalloc_launch<16>(matrix, rows, columns, row_stride, ascending, scratch, s);
    return rank_matrix_noalloc_launch<24>(matrix, rows, columns, row_stride, ascending, scratch, s);
}
