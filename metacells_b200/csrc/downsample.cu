// UMI downsampling (multinomial sampling without replacement), bit-exact with the reference's
// downsample.cpp:81-113 (downsample_slice) for the same random_seed.
//
// What the reference does per row: build a binary sum tree over the row's counts and, `samples`
// times, draw r = minstd() % remaining_total, walk the tree to the leaf holding the r-th
// remaining UMI (gene order), and remove it (downsample.cpp:46-79, 107-112).  The walk is only a
// way to find "the r-th remaining UMI", so any order-statistic structure gives identical output.
//
// B200 formulation (one CTA per row, BATCH draws decoded in parallel):
//   * x_d = seed * 48271^(d+1) mod (2^31-1) is a closed form, so every draw's raw value is known
//     up front, and so is its modulus T - d: r_d = x_d % (T - d).
//   * r_d is a rank among the UMIs left after draws 0..d-1.  Inside a batch that starts at d0, the
//     rank relative to the UMIs left at the START of the batch is obtained by replaying the earlier
//     draws of the batch backwards:  q = r_d;  for j = d-1 .. d0:  q += (r_j <= q).
//     (Draw j removed rank r_j in its own coordinate system; everything at or above it shifts.)
//   * q is then located in the inclusive prefix sum of the remaining counts (binary search in
//     shared memory), the leaf count is decremented atomically, and the prefix sum is rebuilt
//     once per batch with a block scan.
// The output is original - remaining, so no separate output accumulator is needed.
#include <algorithm>

#include "common.cuh"

namespace mcb {

namespace {

constexpr int DS_THREADS = 256;  // == draws decoded per batch
constexpr uint32_t MINSTD_A = 48271u;
constexpr uint32_t MINSTD_M = 2147483647u;

__device__ __forceinline__ uint32_t mulmod(uint32_t a, uint32_t b) {
    return (uint32_t)(((uint64_t)a * (uint64_t)b) % MINSTD_M);
}

struct DownsampleParams {
    const float* input;
    float* output;
    const int32_t* indptr;   // nullptr: dense rows of `columns` elements
    const int32_t* samples;  // per row; nullptr: use `samples_all`
    int32_t samples_all;
    int64_t columns;
    int64_t rows;            // one past the last row
    int64_t first_row;       // rows [first_row, rows) are processed
    uint64_t seed;
    int use_row_seed;        // 1: seed + 997 * row (matrix forms); 0: seed as given (array form)
    const int32_t* seed_rows;  // nullable: the row index the seed is made of (a row's index in the matrix it is sampled as part of)
    uint32_t* workspace;     // 2 * capacity uint32 per CTA when rows do not fit shared memory
    int64_t capacity;        // elements per array in `workspace` (0: shared memory only)
    int smem_capacity;       // elements per array that fit shared memory
    int* error_flag;
    const int32_t* indices;  // with column_totals: the column of every element of a compressed matrix
    float* column_totals;    // nullable: += every output value (integers far below 2^24: exact in any order)
    // nullable, with column_totals: per column += value^2 << 16 | (value >= top_threshold), one 64-bit integer add per
    // non-zero output: the exact sum of squares (below 2^48) and the count of elements at or above the threshold (below
    // 2^16 rows) that the feature-gene statistics of pipeline/select.py start from
    unsigned long long* column_stats;
    float top_threshold;
};

__device__ __forceinline__ void add_column_statistics(const DownsampleParams& p, int32_t column, float value) {
    atomicAdd(&p.column_totals[column], value);
    if (p.column_stats) {
        const unsigned long long count = (unsigned long long)value;
        atomicAdd(&p.column_stats[column], (count * count) << 16 | (value >= p.top_threshold ? 1ull : 0ull));
    }
}

__global__ void __launch_bounds__(DS_THREADS) downsample_kernel(DownsampleParams p) {
    extern __shared__ uint32_t smem[];
    __shared__ uint32_t draws[DS_THREADS];
    __shared__ uint32_t pow_a[DS_THREADS];  // 48271^(t+1) mod m
    __shared__ int scan_scratch[34];
    __shared__ unsigned long long total_shared;

    const int tid = threadIdx.x;
    {
        // a^(tid+1) by square-and-multiply (8 bits of exponent)
        uint32_t result = 1, base = MINSTD_A;
        for (int e = tid + 1; e; e >>= 1) {
            if (e & 1)
                result = mulmod(result, base);
            base = mulmod(base, base);
        }
        pow_a[tid] = result;
    }
    __syncthreads();
    const uint32_t pow_batch = pow_a[DS_THREADS - 1];  // a^BATCH

    for (int64_t row = p.first_row + blockIdx.x; row < p.rows; row += gridDim.x) {
        const int64_t lo = p.indptr ? (int64_t)p.indptr[row] : row * p.columns;
        const int64_t hi = p.indptr ? (int64_t)p.indptr[row + 1] : lo + p.columns;
        const int n = (int)(hi - lo);
        const int32_t samples = p.samples ? p.samples[row] : p.samples_all;
        const float* in = p.input + lo;
        float* out = p.output + lo;

        if (samples < 0 || n == 0)
            continue;  // downsample.cpp:86-88: output left untouched
        if (n == 1) {  // downsample.cpp:90-93
            if (tid == 0)
                out[0] = (double)samples < (double)in[0] ? (float)samples : in[0];
            continue;
        }

        uint32_t* count;
        uint32_t* cumulative;
        if (n <= p.smem_capacity) {
            count = smem;
            cumulative = smem + p.smem_capacity;
        } else if ((int64_t)n <= p.capacity) {
            count = p.workspace + (int64_t)blockIdx.x * 2 * p.capacity;
            cumulative = count + p.capacity;
        } else {
            if (tid == 0)
                atomicExch(p.error_flag, 1);
            continue;
        }

        // Items per thread for the prefix sums: odd, so stride-`items` accesses are conflict free.
        const int items = ((n + DS_THREADS - 1) / DS_THREADS) | 1;
        const int begin = min(tid * items, n), end = min(begin + items, n);

        if (tid == 0)
            total_shared = 0;
        __syncthreads();
        {
            unsigned long long partial = 0;
            for (int i = tid; i < n; i += DS_THREADS) {
                uint32_t c = (uint32_t)in[i];  // float -> size_t truncation (downsample.cpp:24)
                count[i] = c;
                partial += c;
            }
            for (int offset = 16; offset > 0; offset >>= 1) partial += __shfl_xor_sync(0xffffffffu, partial, offset);
            if ((tid & 31) == 0)
                atomicAdd(&total_shared, partial);
        }
        __syncthreads();
        const unsigned long long total64 = total_shared;
        __syncthreads();
        if (total64 <= (unsigned long long)samples) {  // downsample.cpp:100-105
            if (out != in)
                for (int i = tid; i < n; i += DS_THREADS) out[i] = in[i];
            continue;
        }
        if (total64 >= 0xffffffffull) {
            if (tid == 0)
                atomicExch(p.error_flag, 2);
            continue;
        }
        const uint32_t total = (uint32_t)total64;

        uint64_t row_seed = p.seed;
        if (p.use_row_seed)
            row_seed = p.seed == 0 ? 0 : p.seed + (uint64_t)(p.seed_rows ? (int64_t)p.seed_rows[row] : row) * 997ull;
        uint32_t state = (uint32_t)(row_seed % MINSTD_M);
        if (state == 0)
            state = 1;  // std::linear_congruential_engine::seed

        for (int first = 0; first < samples; first += DS_THREADS) {
            // (1) inclusive prefix sums of the remaining counts
            {
                int sum = 0;
                for (int i = begin; i < end; ++i) sum += (int)count[i];
                int block_total;
                int running = block_exclusive_scan(sum, scan_scratch, &block_total);
                for (int i = begin; i < end; ++i) {
                    running += (int)count[i];
                    cumulative[i] = (uint32_t)running;
                }
            }
            // (2) raw draws of this batch
            const int batch = min(DS_THREADS, samples - first);
            if (tid < batch) {
                uint32_t x = mulmod(state, pow_a[tid]);
                draws[tid] = x % (total - (uint32_t)(first + tid));  // downsample.cpp:111
            }
            state = mulmod(state, pow_batch);
            __syncthreads();
            // (3) ranks relative to the start of the batch, (4) locate and remove
            if (tid < batch) {
                uint32_t q = draws[tid];
                const int top = (tid & ~31) + 31;
                for (int j = min(top, batch - 1); j >= 0; --j) {
                    uint32_t earlier = draws[j];
                    q += (j < tid && earlier <= q) ? 1u : 0u;
                }
                int left = 0, right = n - 1;  // smallest i with cumulative[i] > q
                while (left < right) {
                    int middle = (left + right) >> 1;
                    if (cumulative[middle] > q)
                        right = middle;
                    else
                        left = middle + 1;
                }
                atomicSub(&count[left], 1u);
            }
            __syncthreads();
        }
        for (int i = tid; i < n; i += DS_THREADS) out[i] = (float)((uint32_t)in[i] - count[i]);
        __syncthreads();
    }
}

// ---- warp-per-row kernel -------------------------------------------------------------------------
//
// Same sampling process, different bookkeeping: the row's T UMIs are one bit each (gene order), a set
// bit = drawn.  "The r-th remaining UMI" is the r-th clear bit.  Draws are decoded 32 at a time (one
// per lane): ranks are rebased to the start of the batch by the backward replay described above
// (31 steps over shuffles instead of 255 over shared memory), located through the per-word prefix
// sums of clear bits (binary search + find-n-th-set-bit), marked, and the prefix sums (T / 32 words)
// are rebuilt by one warp scan.  The result is the population count of every gene's bit range.
// One warp owns a row end to end, no block barriers; rows are handed out through an atomic counter.

constexpr int DW_WARPS = 8;

// Largest row total of the rows that will actually be sampled (sizes the per-warp bitmap).
__global__ void __launch_bounds__(256) max_row_total_kernel(DownsampleParams p, unsigned long long* __restrict__ result) {
    const int lane = threadIdx.x & 31;
    const int64_t row = p.first_row + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= p.rows)
        return;
    const int64_t lo = p.indptr ? (int64_t)p.indptr[row] : row * p.columns;
    const int64_t hi = p.indptr ? (int64_t)p.indptr[row + 1] : lo + p.columns;
    const int32_t samples = p.samples ? p.samples[row] : p.samples_all;
    unsigned long long total = 0;
    for (int64_t i = lo + lane; i < hi; i += 32) total += (uint32_t)p.input[i];
    for (int offset = 16; offset > 0; offset >>= 1) total += __shfl_xor_sync(0xffffffffu, total, offset);
    if (lane == 0 && samples >= 0 && total > (unsigned long long)samples)
        atomicMax(result, total);
}

struct WarpDownsampleParams {
    DownsampleParams base;
    int words_capacity;           // bitmap words per warp that fit shared memory
    unsigned long long* next_row; // dynamic row counter
};

// x * y mod (2^31 - 1) for x, y < 2^31 - 1, by Mersenne folding.
__device__ __forceinline__ uint32_t mulmod_mersenne(uint32_t x, uint32_t y) {
    const uint64_t product = (uint64_t)x * (uint64_t)y;
    uint32_t folded = (uint32_t)(product & MINSTD_M) + (uint32_t)(product >> 31);
    folded = (folded & MINSTD_M) + (folded >> 31);
    return folded >= MINSTD_M ? folded - MINSTD_M : folded;
}

// Position of the (n + 1)-th set bit of `mask` (n < popc(mask)).
__device__ __forceinline__ int nth_set_bit(uint32_t mask, int n) {
    int position = 0;
#pragma unroll
    for (int width = 16; width > 0; width >>= 1) {
        const int low = __popc(mask & ((1u << width) - 1u));
        if (n >= low) {
            n -= low;
            mask >>= width;
            position += width;
        }
    }
    return position;
}

// PREFIX_ARRAY: the clear bits before every word are kept in shared memory and a draw is located by a binary search
// over the words (rows of any size).  Otherwise (rows of up to 64 words per lane) only the clear bits before every
// LANE's words exist, in a register: a draw finds its lane with five shuffles and walks that lane's few words; half the
// shared memory per warp, so twice the warps per SM for a kernel that waits on its own dependent chains.
template <bool PREFIX_ARRAY>
__global__ void __launch_bounds__(DW_WARPS * 32) downsample_warp_kernel(WarpDownsampleParams wp) {
    extern __shared__ uint32_t smem[];
    const DownsampleParams& p = wp.base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* bitmap = smem + (size_t)warp * (PREFIX_ARRAY ? 2 : 1) * wp.words_capacity;  // bit set = UMI drawn (or padding)
    uint32_t* free_before = bitmap + wp.words_capacity;  // PREFIX_ARRAY only: clear bits in words [0, w)

    // a^(lane+1) and a^32
    uint32_t pow_lane = 1, pow_warp = 1;
    {
        uint32_t base = MINSTD_A;
        for (int e = lane + 1; e; e >>= 1) {
            if (e & 1)
                pow_lane = mulmod_mersenne(pow_lane, base);
            base = mulmod_mersenne(base, base);
        }
        pow_warp = __shfl_sync(0xffffffffu, pow_lane, 31);
    }

    for (;;) {
        unsigned long long claimed = 0;
        if (lane == 0)
            claimed = atomicAdd(wp.next_row, 1ull);
        const int64_t row = p.first_row + (int64_t)__shfl_sync(0xffffffffu, claimed, 0);
        if (row >= p.rows)
            break;
        const int64_t lo = p.indptr ? (int64_t)p.indptr[row] : row * p.columns;
        const int64_t hi = p.indptr ? (int64_t)p.indptr[row + 1] : lo + p.columns;
        const int n = (int)(hi - lo);
        const int32_t samples = p.samples ? p.samples[row] : p.samples_all;
        const float* in = p.input + lo;
        float* out = p.output + lo;
        if (samples < 0 || n == 0)
            continue;  // downsample.cpp:86-88
        const int32_t* columns_of_row = p.column_totals ? p.indices + lo : nullptr;
        if (n == 1) {  // downsample.cpp:90-93
            if (lane == 0) {
                const float kept = (double)samples < (double)in[0] ? (float)samples : in[0];
                out[0] = kept;
                if (columns_of_row && kept != 0.0f)
                    add_column_statistics(p, columns_of_row[0], kept);
            }
            continue;
        }
        unsigned long long total64 = 0;
        {
            // four independent loads in flight per lane: the row is streamed, not walked one load at a time
            int i = lane;
            for (; i + 96 < n; i += 128) {
                const float a = in[i], b = in[i + 32], c = in[i + 64], d = in[i + 96];
                total64 += (unsigned long long)(uint32_t)a + (uint32_t)b;
                total64 += (unsigned long long)(uint32_t)c + (uint32_t)d;
            }
            for (; i < n; i += 32) total64 += (uint32_t)in[i];
        }
        for (int offset = 16; offset > 0; offset >>= 1) total64 += __shfl_xor_sync(0xffffffffu, total64, offset);
        if (total64 <= (unsigned long long)samples) {  // downsample.cpp:100-105
            if (out != in)
                for (int i = lane; i < n; i += 32) out[i] = in[i];
            if (columns_of_row)
                for (int i = lane; i < n; i += 32)
                    if (in[i] != 0.0f)
                        add_column_statistics(p, columns_of_row[i], in[i]);
            continue;
        }
        const int words = (int)((total64 + 31) >> 5);
        if (total64 >= 0xffffffffull || words > wp.words_capacity) {
            if (lane == 0)
                atomicExch(p.error_flag, total64 >= 0xffffffffull ? 2 : 3);
            continue;
        }
        const uint32_t total = (uint32_t)total64;
        const int per_lane = (words + 31) >> 5;  // consecutive words owned by a lane in the scans

        for (int w = lane; w < words; w += 32) bitmap[w] = 0;
        __syncwarp();
        if (lane == 0 && (total & 31u))
            bitmap[words - 1] = ~0u << (total & 31u);  // padding bits count as drawn
        __syncwarp();

        uint64_t row_seed = p.seed;
        if (p.use_row_seed)
            row_seed = p.seed == 0 ? 0 : p.seed + (uint64_t)(p.seed_rows ? (int64_t)p.seed_rows[row] : row) * 997ull;
        uint32_t state = (uint32_t)(row_seed % MINSTD_M);
        if (state == 0)
            state = 1;

        for (int first = 0; first < samples; first += 32) {
            // prefix sums of clear bits: per word (PREFIX_ARRAY) or per lane
            uint32_t lane_before;
            {
                const int begin = min(lane * per_lane, words), end = min(begin + per_lane, words);
                int sum = 0;
                for (int w = begin; w < end; ++w) sum += __popc(~bitmap[w]);
                int running = warp_inclusive_scan(sum, lane) - sum;
                lane_before = (uint32_t)running;
                if (PREFIX_ARRAY)
                    for (int w = begin; w < end; ++w) {
                        free_before[w] = (uint32_t)running;
                        running += __popc(~bitmap[w]);
                    }
            }
            __syncwarp();
            const int draw = first + lane;
            const bool active = draw < samples;
            uint32_t r = 0;
            if (active)
                r = mulmod_mersenne(state, pow_lane) % (total - (uint32_t)draw);  // downsample.cpp:111
            state = mulmod_mersenne(state, pow_warp);
            // rank among the UMIs left at the start of the batch
            // replay of the earlier draws of the batch, latest first: lane - 1, lane - 2, ... 0.  shfl.up hands lane l the
            // value of lane l - s together with the predicate "that lane exists", which gates the comparison: three
            // instructions per step (shuffle, compare-and, predicated add)
            uint32_t q = r;
#pragma unroll
            for (int s = 1; s <= 31; ++s) {
                asm volatile(
                    "{\n\t"
                    ".reg .pred exists, below;\n\t"
                    ".reg .b32 earlier;\n\t"
                    "shfl.sync.up.b32 earlier|exists, %1, %2, 0, 0xffffffff;\n\t"
                    "setp.le.and.u32 below, earlier, %0, exists;\n\t"
                    "@below add.u32 %0, %0, 1;\n\t"
                    "}"
                    : "+r"(q)
                    : "r"(r), "r"(s));
            }
            int word = 0;
            uint32_t mark = 0;
            if (PREFIX_ARRAY) {
                if (active) {
                    int left = 0, right = words - 1;  // last word with free_before[w] <= q
                    while (left < right) {
                        const int middle = (left + right + 1) >> 1;
                        if (free_before[middle] <= q)
                            left = middle;
                        else
                            right = middle - 1;
                    }
                    word = left;
                    mark = 1u << nth_set_bit(~bitmap[left], (int)(q - free_before[left]));
                }
            } else {
                // the last lane whose words start at or before rank q (lanes without words start at the row's total)
                int owner = 0;
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const uint32_t before = __shfl_sync(0xffffffffu, lane_before, owner + step);
                    if (before <= q)
                        owner += step;
                }
                uint32_t inside = q - __shfl_sync(0xffffffffu, lane_before, owner);
                if (active) {
                    word = owner * per_lane;
                    uint32_t clear = ~bitmap[word];
                    int count = __popc(clear);
                    while (inside >= (uint32_t)count) {
                        inside -= (uint32_t)count;
                        clear = ~bitmap[++word];
                        count = __popc(clear);
                    }
                    mark = 1u << nth_set_bit(clear, (int)inside);
                }
            }
            __syncwarp();  // every lane has located its draw in the bitmap of the start of the batch
            if (active)
                atomicOr(&bitmap[word], mark);
            __syncwarp();
        }
        if (lane == 0 && (total & 31u))
            bitmap[words - 1] &= ~(~0u << (total & 31u));  // drop the padding again
        __syncwarp();

        // output: drawn UMIs of every gene = set bits in its range
        uint32_t running = 0;
        float ahead = lane < n ? in[lane] : 0.0f;  // the next 32 counts are loaded while these are turned into bit ranges
        for (int base = 0; base < n; base += 32) {
            const int i = base + lane;
            const uint32_t count = (uint32_t)ahead;
            ahead = i + 32 < n ? in[i + 32] : 0.0f;
            const uint32_t inclusive = (uint32_t)warp_inclusive_scan((int)count, lane);
            uint32_t begin = running + inclusive - count, end = begin + count;
            running += __shfl_sync(0xffffffffu, inclusive, 31);
            uint32_t drawn = 0;
            while (begin < end) {
                const uint32_t w = begin >> 5, offset = begin & 31u;
                const uint32_t take = min(32u - offset, end - begin);
                const uint32_t mask = (take == 32u ? ~0u : ((1u << take) - 1u)) << offset;
                drawn += __popc(bitmap[w] & mask);
                begin += take;
            }
            if (i < n) {
                out[i] = (float)drawn;
                if (columns_of_row && drawn)
                    add_column_statistics(p, columns_of_row[i], (float)drawn);
            }
        }
        __syncwarp();
    }
}

// The fallback kernels do not add to the column totals themselves.
__global__ void __launch_bounds__(256) add_column_totals_kernel(DownsampleParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t row = p.first_row + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= p.rows)
        return;
    for (int64_t e = p.indptr[row] + lane; e < p.indptr[row + 1]; e += 32)
        if (p.output[e] != 0.0f)
            add_column_statistics(p, p.indices[e], p.output[e]);
}

}  // namespace

// Rows [first_row, last_row).  known_max_total > 0: the caller knows the largest row total (the pipeline has the cells'
// totals on the host, tools/downsample.py:84-90), which spares a kernel and a stream synchronisation.
int downsample_rows_launch(const float* input, float* output, const int32_t* indptr, const int32_t* samples,
                           int32_t samples_all, int64_t first_row, int64_t last_row, int64_t columns, int64_t max_row_size,
                           uint64_t seed, int use_row_seed, unsigned long long known_max_total, int* deferred_flag,
                           cudaStream_t stream, const int32_t* indices = nullptr, float* column_totals = nullptr,
                           unsigned long long* column_stats = nullptr, float top_threshold = 0.0f, const int32_t* seed_rows = nullptr);

int downsample_launch(const float* input, float* output, const int32_t* indptr, const int32_t* samples,
                      int32_t samples_all, int64_t rows, int64_t columns, int64_t max_row_size, uint64_t seed,
                      int use_row_seed, cudaStream_t stream) {
    return downsample_rows_launch(input, output, indptr, samples, samples_all, 0, rows, columns, max_row_size, seed, use_row_seed,
                                  0, nullptr, stream);
}

int downsample_rows_launch(const float* input, float* output, const int32_t* indptr, const int32_t* samples,
                           int32_t samples_all, int64_t first_row, int64_t last_row, int64_t columns, int64_t max_row_size,
                           uint64_t seed, int use_row_seed, unsigned long long known_max_total, int* deferred_flag,
                           cudaStream_t stream, const int32_t* indices, float* column_totals, unsigned long long* column_stats,
                           float top_threshold, const int32_t* seed_rows) {
    const int64_t rows = last_row - first_row;
    if (rows <= 0)
        return MCB_OK;
    DownsampleParams p{};
    p.input = input;
    p.output = output;
    p.indptr = indptr;
    p.samples = samples;
    p.samples_all = samples_all;
    p.columns = columns;
    p.rows = last_row;
    p.first_row = first_row;
    p.seed = seed;
    p.use_row_seed = use_row_seed;
    p.seed_rows = seed_rows;
    // column totals of the output (compressed input only): added by the kernel that writes it
    MCB_REQUIRE(column_totals == nullptr || (indices != nullptr && indptr != nullptr),
                "downsample: column totals need a compressed matrix");
    p.indices = indices;
    p.column_totals = column_totals;
    MCB_REQUIRE(column_stats == nullptr || column_totals != nullptr, "downsample: column statistics come with the column totals");
    p.column_stats = column_stats;
    p.top_threshold = top_threshold;

    Scratch workspace, flag, counters;
    MCB_TRY(flag.alloc(sizeof(int), stream));
    MCB_CUDA(cudaMemsetAsync(flag.ptr, 0, sizeof(int), stream));
    // deferred_flag (device, sticky): the caller reads it at its next synchronisation instead of this call waiting
    p.error_flag = deferred_flag ? deferred_flag : flag.as<int>();

    // Preferred: one warp per row over a bitmap of the row's UMIs, when the largest row fits shared memory.
    MCB_TRY(counters.alloc(2 * sizeof(unsigned long long), stream));
    MCB_CUDA(cudaMemsetAsync(counters.ptr, 0, 2 * sizeof(unsigned long long), stream));
    unsigned long long max_total = known_max_total;
    if (max_total == 0) {
        MCB_PROFILED("downsample_totals", stream,
                     max_row_total_kernel<<<(unsigned)ceil_div(rows, 8), 256, 0, stream>>>(p, counters.as<unsigned long long>()));
        MCB_LAUNCH_CHECK();
        MCB_TRY(copy_to_host(&max_total, counters.ptr, sizeof(max_total), stream));
        MCB_CUDA(cudaStreamSynchronize(stream));
    }
    if (max_total == 0) {
        // nothing to sample: every row is copied (or left untouched); the warp kernel does that too
        max_total = 1;
    }
    const unsigned long long max_words = (max_total + 31) / 32;
    if (max_words <= 28 * 1024) {
        const bool prefix_array = max_words > 2048;  // more than 64 words per lane: search words, not lanes
        const int capacity = (int)((std::max<unsigned long long>(max_words, 64) + 31) / 32 * 32);
        const int arrays = prefix_array ? 2 : 1;
        int warps = DW_WARPS;
        while (warps > 1 && (size_t)warps * arrays * capacity * sizeof(uint32_t) > 200 * 1024) warps /= 2;
        const size_t smem_bytes = (size_t)warps * arrays * capacity * sizeof(uint32_t);
        WarpDownsampleParams wp{};
        wp.base = p;
        wp.words_capacity = capacity;
        wp.next_row = counters.as<unsigned long long>() + 1;
        int per_sm = (int)std::min<size_t>(16, (220 * 1024) / (smem_bytes + 1024));
        per_sm = std::max(1, std::min(per_sm, 2048 / (warps * 32)));
        const int64_t grid = std::min<int64_t>(ceil_div(rows, warps), (int64_t)sm_count() * per_sm);
        if (prefix_array) {
            MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(downsample_warp_kernel<true>), (size_t)smem_bytes));
            MCB_PROFILED("downsample", stream, downsample_warp_kernel<true><<<(unsigned)grid, warps * 32, smem_bytes, stream>>>(wp));
        } else {
            MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(downsample_warp_kernel<false>), (size_t)smem_bytes));
            MCB_PROFILED("downsample", stream, downsample_warp_kernel<false><<<(unsigned)grid, warps * 32, smem_bytes, stream>>>(wp));
        }
        MCB_LAUNCH_CHECK();
        if (known_max_total > 0 && deferred_flag == nullptr) {
            // the caller's hint sized the bitmap and nobody else reads the flag: a row above the hint must not pass silently
            int host_flag = 0;
            MCB_TRY(copy_to_host(&host_flag, flag.ptr, sizeof(int), stream));
            MCB_CUDA(cudaStreamSynchronize(stream));
            MCB_REQUIRE(host_flag != 3, "downsample: a row totals more than the stated maximum");
            if (host_flag == 2)
                return fail(MCB_ERR_UNSUPPORTED, "downsample: a row totals 2^32 UMIs or more");
        }
        return MCB_OK;
    }

    // Fallback for rows of more than ~900k UMIs: one CTA per row over per-gene counts.
    // Shared-memory capacity classes: 8 B per element; smaller classes allow more CTAs per SM.
    int smem_capacity = 2048;
    while (smem_capacity < max_row_size && smem_capacity < 24576) smem_capacity *= 2;
    if (smem_capacity > 24576)
        smem_capacity = 24576;
    p.smem_capacity = smem_capacity;
    const size_t smem_bytes = (size_t)smem_capacity * 8;

    int64_t grid = rows;
    if (max_row_size > smem_capacity) {
        grid = rows < (int64_t)sm_count() * 2 ? rows : (int64_t)sm_count() * 2;
        p.capacity = max_row_size;
        MCB_TRY(workspace.alloc((size_t)grid * 2 * (size_t)max_row_size * sizeof(uint32_t), stream));
        p.workspace = workspace.as<uint32_t>();
    }
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(downsample_kernel), (size_t)smem_bytes));
    MCB_PROFILED("downsample", stream, downsample_kernel<<<(unsigned)grid, DS_THREADS, smem_bytes, stream>>>(p));
    MCB_LAUNCH_CHECK();
    if (column_totals) {
        add_column_totals_kernel<<<(unsigned)ceil_div(rows, 8), 256, 0, stream>>>(p);
        MCB_LAUNCH_CHECK();
    }
    int host_flag = 0;
    MCB_TRY(copy_to_host(&host_flag, flag.ptr, sizeof(int), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    if (host_flag == 1)
        return fail(MCB_ERR_ARG, "downsample: a row is wider than the stated max_row_size");
    if (host_flag == 2)
        return fail(MCB_ERR_UNSUPPORTED, "downsample: a row totals 2^32 UMIs or more");
    return MCB_OK;
}

}  // namespace mcb
