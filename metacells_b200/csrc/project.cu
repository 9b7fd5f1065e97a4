// Query x atlas projection front-end (BASELINE.json config 5): the Pearson correlation of every
// query row with every atlas row (tools/project.py:249-270 -> ut.cross_corrcoef_rows, computation.py:478-524 ->
// correlate.cpp:433-494) followed, per query row, by the first `k` entries of the descending order of its correlations
// (tools/project.py:310, `np.argsort(-correlations)`, of which the candidates loop :317-331 consumes a prefix).
//
// The reference materialises the whole [query, atlas] matrix in host memory and fully sorts every row.  Here the
// matrix exists only as one block of query rows at a time in HBM: the tcgen05 contraction of correlate_umma.cu writes
// the block, and one CTA per row selects the ordered top k from it in two streaming reads:
//   1. histogram of a monotone 12-bit linear bin of the value (correlations spread evenly over [-1, 1], unlike the
//      bits of their float representation) with shared-memory atomics; the bin of the k-th largest follows;
//   2. every element of that bin or a higher one (a few hundred) goes to shared memory as (key, ~index);
//   3. a bitonic sort of those candidates, descending: larger value first, equal values by ascending index (what a
//      stable argsort of the negated row gives).
// A row whose threshold bin holds more candidates than fit (massive ties) takes an exact path: radix select of the
// k-th key inside the bin, then the equal keys in index order.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "../../include/mcb200.h"

namespace mcb {

int prepare_rows_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, int64_t padded, float* hi,
                     float* lo, double* f64, cudaStream_t stream);
int correlate_umma_launch(const float* a_hi, const float* a_lo, int64_t rows_a, const float* b_hi, const float* b_lo,
                          int64_t rows_b, int64_t padded, float* output, int64_t out_stride, int symmetric, int absolute,
                          unsigned int* min_key, cudaStream_t stream);

namespace {

constexpr int PT_THREADS = 1024;
constexpr int PT_BINS = 4096;
constexpr int PT_CAPACITY = 8192;  // candidates kept in shared memory (64 KiB of 64-bit composites)

struct OrderedTopParams {
    const float* values;  // [rows, row_stride]
    int64_t rows, columns, row_stride;
    int k;
    int32_t* out_indices;  // [rows, k]
    float* out_values;     // [rows, k]
};

struct OrderedTopShared {
    int hist[PT_BINS];
    int warp_sums[33];
    int threshold_bin;  // the bin of the k-th largest value
    int above;          // elements in higher bins (< k)
    int inside;         // elements in the threshold bin
    int cursor;
    uint32_t prefix;  // exact path: the known high bits of the k-th key
    int need;         // exact path: rank (1-based, from the top) still to resolve inside the current group
};

// Monotone (non-strictly) in the value: a larger value never lands in a lower bin.
__device__ __forceinline__ int linear_bin(float value) {
    const int bin = __float2int_rz((value + 1.0f) * (PT_BINS / 2));
    return min(max(bin, 0), PT_BINS - 1);
}

__device__ __forceinline__ uint64_t composite_of(uint32_t key, int64_t index) {
    return ((uint64_t)key << 32) | (uint32_t)~(uint32_t)index;  // descending composite: larger key, then smaller index
}

// Calls visit(value, column) for every element of the row; float4 loads when the row allows them, two in flight.
template <typename Visit>
__device__ __forceinline__ void stream_row(const float* __restrict__ row, int64_t columns, bool vector, Visit visit) {
    const int tid = threadIdx.x;
    int64_t done = 0;
    if (vector) {
        const float4* row4 = reinterpret_cast<const float4*>(row);
        const int64_t quads = columns >> 2;
        int64_t q = tid;
        for (; q + PT_THREADS < quads; q += 2 * PT_THREADS) {
            const float4 a = __ldg(row4 + q);
            const float4 b = __ldg(row4 + q + PT_THREADS);
            visit(a.x, 4 * q), visit(a.y, 4 * q + 1), visit(a.z, 4 * q + 2), visit(a.w, 4 * q + 3);
            const int64_t r = q + PT_THREADS;
            visit(b.x, 4 * r), visit(b.y, 4 * r + 1), visit(b.z, 4 * r + 2), visit(b.w, 4 * r + 3);
        }
        if (q < quads) {
            const float4 a = __ldg(row4 + q);
            visit(a.x, 4 * q), visit(a.y, 4 * q + 1), visit(a.z, 4 * q + 2), visit(a.w, 4 * q + 3);
        }
        done = quads << 2;
    }
    for (int64_t i = done + tid; i < columns; i += PT_THREADS) visit(__ldg(row + i), i);
}

// The digit (from the top) that holds the element of rank `need` (1-based, from the largest digit down) in a
// histogram of `bins` <= PT_BINS entries.  One thread owns 4 consecutive bins, highest bins first.  Every thread
// receives the digit, the count in higher digits and the count in the digit itself.
__device__ void find_from_top(OrderedTopShared* shared, int bins, int need, int* digit, int* above, int* inside) {
    const int tid = threadIdx.x;
    int counts[4], sum = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        const int bin = bins - 1 - (tid * 4 + b);
        counts[b] = bin >= 0 ? shared->hist[bin] : 0;
        sum += counts[b];
    }
    int total;
    int before = block_exclusive_scan(sum, shared->warp_sums, &total);
    if (need > before && need <= before + sum) {
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            if (need > before && need <= before + counts[b]) {
                shared->threshold_bin = bins - 1 - (tid * 4 + b);
                shared->above = before;
                shared->inside = counts[b];
            }
            before += counts[b];
        }
    }
    __syncthreads();
    *digit = shared->threshold_bin;
    *above = shared->above;
    *inside = shared->inside;
    __syncthreads();
}

__global__ void __launch_bounds__(PT_THREADS, 2) ordered_top_kernel(OrderedTopParams p) {
    extern __shared__ __align__(16) unsigned char raw_smem[];
    OrderedTopShared* shared = reinterpret_cast<OrderedTopShared*>(raw_smem);
    uint64_t* candidates = reinterpret_cast<uint64_t*>(raw_smem + ((sizeof(OrderedTopShared) + 15) & ~size_t(15)));
    const int tid = threadIdx.x;

    for (int64_t row_index = blockIdx.x; row_index < p.rows; row_index += gridDim.x) {
        const float* row = p.values + row_index * p.row_stride;
        const bool vector = (reinterpret_cast<uintptr_t>(row) & 15) == 0;

        // 1. histogram of the linear bins
        for (int i = tid; i < PT_BINS; i += PT_THREADS) shared->hist[i] = 0;
        if (tid == 0) shared->cursor = 0;
        __syncthreads();
        stream_row(row, p.columns, vector, [&](float value, int64_t) { atomicAdd(&shared->hist[linear_bin(value)], 1); });
        __syncthreads();
        int threshold_bin, above, inside;
        find_from_top(shared, PT_BINS, p.k, &threshold_bin, &above, &inside);

        int count;
        if (above + inside <= PT_CAPACITY) {
            // 2. the candidates: everything in the threshold bin or above it
            stream_row(row, p.columns, vector, [&](float value, int64_t column) {
                if (linear_bin(value) >= threshold_bin)
                    candidates[atomicAdd(&shared->cursor, 1)] = composite_of(float_to_ordered(canonical_zero(value)), column);
            });
            count = above + inside;
        } else {
            // exact path: radix select (12 + 12 + 8 bits from the top) of the k-th key among the keys of the threshold
            // bin; bins are monotone in the key, so every element of a higher bin has a larger key
            if (tid == 0) {
                shared->prefix = 0;
                shared->need = p.k - above;
            }
            __syncthreads();
            uint32_t known = 0;
            for (int level = 0; level < 3; ++level) {
                const int shift = level == 0 ? 20 : level == 1 ? 8 : 0, bins = level == 2 ? 256 : 4096;
                for (int i = tid; i < bins; i += PT_THREADS) shared->hist[i] = 0;
                __syncthreads();
                const uint32_t prefix = shared->prefix;
                const int need = shared->need;
                stream_row(row, p.columns, vector, [&](float value, int64_t) {
                    const uint32_t key = float_to_ordered(canonical_zero(value));
                    if (linear_bin(value) == threshold_bin && (key & known) == prefix)
                        atomicAdd(&shared->hist[(key >> shift) & (uint32_t)(bins - 1)], 1);
                });
                __syncthreads();
                int digit, higher, same;
                find_from_top(shared, bins, need, &digit, &higher, &same);
                if (tid == 0) {
                    shared->prefix = prefix | ((uint32_t)digit << shift);
                    shared->need = need - higher;
                }
                known |= (uint32_t)(bins - 1) << shift;
                __syncthreads();
            }
            const uint32_t kth_key = shared->prefix;
            const int equal_needed = shared->need;  // of the elements equal to the k-th key, the first this many
            const int larger = p.k - equal_needed;
            stream_row(row, p.columns, vector, [&](float value, int64_t column) {
                const uint32_t key = float_to_ordered(canonical_zero(value));
                if (key > kth_key) candidates[atomicAdd(&shared->cursor, 1)] = composite_of(key, column);
            });
            // the equal keys in index order: a block scan per chunk of PT_THREADS columns
            int taken = 0;
            for (int64_t start = 0; start < p.columns && taken < equal_needed; start += PT_THREADS) {
                const int64_t column = start + tid;
                const bool equal = column < p.columns && float_to_ordered(canonical_zero(row[column])) == kth_key;
                int total;
                const int before = block_exclusive_scan(equal ? 1 : 0, shared->warp_sums, &total);
                if (equal && taken + before < equal_needed) candidates[larger + taken + before] = composite_of(kth_key, column);
                taken += total;
            }
            count = p.k;
        }
        __syncthreads();

        // 3. bitonic sort of the candidates, descending
        int padded = 1;
        while (padded < count) padded <<= 1;
        for (int i = count + tid; i < padded; i += PT_THREADS) candidates[i] = 0;
        __syncthreads();
        for (int size = 2; size <= padded; size <<= 1) {
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                for (int t = tid; t < (padded >> 1); t += PT_THREADS) {
                    const int low = ((t & ~(stride - 1)) << 1) | (t & (stride - 1));
                    const int high = low | stride;
                    const bool descending = (low & size) == 0;
                    const uint64_t a = candidates[low], b = candidates[high];
                    if ((a < b) == descending) {
                        candidates[low] = b;
                        candidates[high] = a;
                    }
                }
                __syncthreads();
            }
        }
        for (int i = tid; i < p.k; i += PT_THREADS) {
            const uint64_t composite = candidates[i];
            p.out_indices[row_index * p.k + i] = (int32_t)~(uint32_t)composite;
            p.out_values[row_index * p.k + i] = ordered_to_float((uint32_t)(composite >> 32));
        }
        __syncthreads();
    }
}

size_t ordered_top_smem() { return ((sizeof(OrderedTopShared) + 15) & ~size_t(15)) + (size_t)PT_CAPACITY * sizeof(uint64_t); }

}  // namespace

// values: device [rows, row_stride]; out_*: device [rows, k].
int ordered_top_launch(const float* values, int64_t rows, int64_t columns, int64_t row_stride, int64_t k, int32_t* out_indices,
                       float* out_values, cudaStream_t stream) {
    if (rows == 0)
        return MCB_OK;
    MCB_REQUIRE(k > 0 && k <= columns, "ordered_top: k %lld must be in 1..%lld (the columns)", (long long)k, (long long)columns);
    MCB_REQUIRE(k <= PT_CAPACITY / 2, "ordered_top: k %lld is larger than %d", (long long)k, PT_CAPACITY / 2);
    MCB_REQUIRE(columns < (1ll << 31), "ordered_top: too many columns");
    OrderedTopParams p{};
    p.values = values;
    p.rows = rows;
    p.columns = columns;
    p.row_stride = row_stride;
    p.k = (int)k;
    p.out_indices = out_indices;
    p.out_values = out_values;
    const size_t smem = ordered_top_smem();
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(ordered_top_kernel), (size_t)smem));
    const int64_t grid = std::min<int64_t>(rows, (int64_t)sm_count() * 16);
    MCB_PROFILED("ordered_top", stream, ordered_top_kernel<<<(unsigned)grid, PT_THREADS, smem, stream>>>(p));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

}  // namespace mcb

using namespace mcb;

namespace {

// Bytes of similarities held at once (one block of query rows against the whole atlas).
int64_t block_budget() {
    const char* text = std::getenv("MCB_PROJECT_BLOCK_BYTES");
    const long long wanted = text ? std::atoll(text) : 0;
    return wanted > 0 ? wanted : (4ll << 30);
}

int project_top_device(const float* query, int64_t query_rows, int64_t query_stride, const float* atlas, int64_t atlas_rows,
                       int64_t atlas_stride, int64_t columns, int64_t k, int32_t* out_indices, float* out_correlations,
                       cudaStream_t stream) {
    if (query_rows == 0)
        return MCB_OK;
    MCB_REQUIRE(k > 0 && k <= atlas_rows, "project_top: k %lld must be in 1..%lld (the atlas rows)", (long long)k, (long long)atlas_rows);
    const int64_t padded = ceil_div(columns > 0 ? columns : 1, 32) * 32;
    Scratch q_hi, q_lo, a_hi, a_lo, block;
    MCB_TRY(q_hi.alloc((size_t)query_rows * padded * sizeof(float), stream));
    MCB_TRY(q_lo.alloc((size_t)query_rows * padded * sizeof(float), stream));
    MCB_TRY(a_hi.alloc((size_t)atlas_rows * padded * sizeof(float), stream));
    MCB_TRY(a_lo.alloc((size_t)atlas_rows * padded * sizeof(float), stream));
    MCB_TRY(prepare_rows_f32(query, query_rows, columns, query_stride, padded, q_hi.as<float>(), q_lo.as<float>(), nullptr, stream));
    MCB_TRY(prepare_rows_f32(atlas, atlas_rows, columns, atlas_stride, padded, a_hi.as<float>(), a_lo.as<float>(), nullptr, stream));
    const int64_t stride = ceil_div(atlas_rows, 4) * 4;  // rows of the block start on 16 bytes
    int64_t block_rows = std::max<int64_t>(128, block_budget() / (stride * 4) / 128 * 128);
    block_rows = std::min(block_rows, query_rows);
    MCB_TRY(block.alloc((size_t)block_rows * stride * sizeof(float), stream));
    for (int64_t first = 0; first < query_rows; first += block_rows) {
        const int64_t rows = std::min(block_rows, query_rows - first);
        MCB_TRY(correlate_umma_launch(q_hi.as<float>() + first * padded, q_lo.as<float>() + first * padded, rows, a_hi.as<float>(),
                                      a_lo.as<float>(), atlas_rows, padded, block.as<float>(), stride, 0, 0, nullptr, stream));
        MCB_TRY(ordered_top_launch(block.as<float>(), rows, atlas_rows, stride, k, out_indices + first * k,
                                   out_correlations + first * k, stream));
    }
    return MCB_OK;
}

}  // namespace

extern "C" {

int mcb_dev_ordered_top_f32(const float* values, int64_t rows, int64_t columns, int64_t row_stride, int64_t k,
                            int32_t* out_indices, float* out_values, void* stream) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "ordered_top: bad shape");
    return ordered_top_launch(values, rows, columns, row_stride, k, out_indices, out_values, (cudaStream_t)stream);
}

int mcb_dev_project_top_f32(const float* query, int64_t query_rows, const float* atlas, int64_t atlas_rows, int64_t columns,
                            int64_t k, int32_t* out_indices, float* out_correlations, void* stream) {
    MCB_REQUIRE(query_rows >= 0 && atlas_rows >= 0 && columns >= 0, "project_top: bad shape");
    return project_top_device(query, query_rows, columns, atlas, atlas_rows, columns, columns, k, out_indices, out_correlations,
                              (cudaStream_t)stream);
}

int mcb_host_project_top_f32(const float* query, int64_t query_rows, int64_t query_stride, const float* atlas, int64_t atlas_rows,
                             int64_t atlas_stride, int64_t columns, int64_t k, int32_t* out_indices, float* out_correlations) {
    MCB_REQUIRE(query_rows >= 0 && atlas_rows >= 0 && columns >= 0 && query_stride >= columns && atlas_stride >= columns,
                "project_top: bad shape");
    if (query_rows == 0)
        return MCB_OK;
    MCB_REQUIRE(k > 0 && k <= atlas_rows, "project_top: k %lld must be in 1..%lld (the atlas rows)", (long long)k, (long long)atlas_rows);
    cudaStream_t stream = nullptr;
    Scratch d_query, d_atlas, d_indices, d_values;
    MCB_TRY(d_query.alloc((size_t)query_rows * columns * sizeof(float), stream));
    MCB_TRY(d_atlas.alloc((size_t)atlas_rows * columns * sizeof(float), stream));
    if (columns > 0) {
        MCB_CUDA(cudaMemcpy2DAsync(d_query.ptr, columns * sizeof(float), query, query_stride * sizeof(float), columns * sizeof(float),
                                   query_rows, cudaMemcpyHostToDevice, stream));
        MCB_CUDA(cudaMemcpy2DAsync(d_atlas.ptr, columns * sizeof(float), atlas, atlas_stride * sizeof(float), columns * sizeof(float),
                                   atlas_rows, cudaMemcpyHostToDevice, stream));
    }
    MCB_TRY(d_indices.alloc((size_t)query_rows * k * sizeof(int32_t), stream));
    MCB_TRY(d_values.alloc((size_t)query_rows * k * sizeof(float), stream));
    MCB_TRY(project_top_device(d_query.as<float>(), query_rows, columns, d_atlas.as<float>(), atlas_rows, columns, columns, k,
                               d_indices.as<int32_t>(), d_values.as<float>(), stream));
    MCB_TRY(copy_to_host(out_indices, d_indices.ptr, (size_t)query_rows * k * sizeof(int32_t), stream));
    MCB_TRY(copy_to_host(out_correlations, d_values.ptr, (size_t)query_rows * k * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

}  // extern "C"
