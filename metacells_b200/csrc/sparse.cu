// Compressed (CSR/CSC "band") kernels of the k-NN graph path:
//   collect_pruned            prune_per.cpp:5-105   keep the `degree` largest values of every band
//   collect_compressed        relayout.cpp:73-108   CSR <-> CSC transpose scatter
//   sort_compressed_indices   relayout.cpp:111-167  sort every band by index
//
// Bands are short (a few hundred to a few thousand elements), so every kernel assigns one CTA to
// a band, stages it in shared memory and streams it back coalesced: algorithmic traffic is one
// read and one write of the (data, indices) pairs.
//
// Tie rule of collect_pruned: among exactly equal values the lower position in the band is kept
// (the reference's std::nth_element leaves this to libstdc++'s introselect; see oracle/mc_oracle.c).
#include <algorithm>

#include "common.cuh"
#include "select.cuh"

namespace mcb {

namespace {

constexpr int SP_THREADS = 256;

// ---- exclusive scan of int32 (single CTA; the arrays here have one entry per band) -------------

constexpr int SCAN_ITEMS = 16;  // consecutive elements per thread: 16 384 per round of the one CTA (a pile's rows: one round)

__global__ void __launch_bounds__(1024) exclusive_scan_kernel(const int32_t* __restrict__ input, int32_t* __restrict__ output,
                                                             int64_t count, int32_t* __restrict__ total_out) {
    // output[0] = 0, output[i + 1] = input[0] + ... + input[i]; output has count + 1 entries.
    __shared__ int scratch[34];
    int carry = 0;  // the same in every thread
    if (threadIdx.x == 0)
        output[0] = 0;
    for (int64_t base = 0; base < count; base += 1024 * SCAN_ITEMS) {
        const int64_t first = base + (int64_t)threadIdx.x * SCAN_ITEMS;
        int values[SCAN_ITEMS], sum = 0;
#pragma unroll
        for (int r = 0; r < SCAN_ITEMS; ++r) values[r] = first + r < count ? input[first + r] : 0;
#pragma unroll
        for (int r = 0; r < SCAN_ITEMS; ++r) sum += values[r];
        int total;
        int running = carry + block_exclusive_scan(sum, scratch, &total);
#pragma unroll
        for (int r = 0; r < SCAN_ITEMS; ++r) {
            running += values[r];
            if (first + r < count)
                output[first + r + 1] = running;
        }
        carry += total;
    }
    if (threadIdx.x == 0 && total_out)
        *total_out = carry;
}

// ---- collect_pruned ------------------------------------------------------------------------------

__global__ void pruned_sizes_kernel(const int32_t* __restrict__ indptr, int64_t bands, int32_t degree, int32_t* __restrict__ sizes) {
    const int64_t band = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (band < bands)
        sizes[band] = min(indptr[band + 1] - indptr[band], degree);  // prune_per.cpp:85-95
}

struct PruneParams {
    const float* data;
    const int32_t* indices;
    const int32_t* indptr;
    int64_t bands;
    int degree;
    float* out_data;
    int32_t* out_indices;
    const int32_t* out_indptr;
    int cache_capacity;
    int warp_limit;  // bands of up to this many entries belong to prune_bands_warp_kernel, wider ones to prune_bands_kernel
};

// ---- bands that fit a warp's registers (every band of a pile's graph: about a thousand entries) ----------------------
//
// One warp per band, no block barriers: the keys sit in registers (entry j * 32 + lane: coalesced), the degree-th
// smallest key is found by a binary radix select over the bits in which the band's keys differ at all (balanced ranks
// are multiples of 1/8 below a few thousand: about 15 bits), one warp reduction per bit, and the kept entries are
// written in their original order with two ballots per 32 entries (equal keys: the first ones, prune_per.cpp:33 modulo
// the tie rule of the oracle).
constexpr int PW_MAX_ROUNDS = 64;

template <int ROUNDS>
__device__ __forceinline__ void prune_band_in_warp(const PruneParams& p, int lo, int n, int out) {
    const int lane = threadIdx.x & 31;
    const uint32_t lower_lanes = lanemask_lt();
    const float* values = p.data + lo;
    const int32_t* indices = p.indices + lo;
    uint32_t keys[ROUNDS];
    uint32_t differ = 0;
    const uint32_t first = sort_key(values[0], false);
#pragma unroll
    for (int j = 0; j < ROUNDS; ++j) {
        const int i = j * 32 + lane;
        keys[j] = i < n ? sort_key(values[i], false) : first;
        differ |= keys[j] ^ first;
    }
    differ = __reduce_or_sync(0xffffffffu, differ);
    uint32_t prefix = first & ~differ, known = ~differ;
    int wanted = p.degree - 1;  // 0-based rank, ascending keys = descending values
    for (int bit = 31 - __clz((int)(differ | 1u)); bit >= 0; --bit) {
        if (((differ >> bit) & 1u) == 0)
            continue;
        int zeros = 0;
#pragma unroll
        for (int j = 0; j < ROUNDS; ++j)
            zeros += (j * 32 + lane < n && (keys[j] & known) == prefix && ((keys[j] >> bit) & 1u) == 0) ? 1 : 0;
        zeros = __reduce_add_sync(0xffffffffu, zeros);
        if (wanted >= zeros) {
            wanted -= zeros;
            prefix |= 1u << bit;
        }
        known |= 1u << bit;
    }
    const uint32_t threshold = prefix;
    const int ties_wanted = wanted + 1;
    int kept = 0, ties_seen = 0;
#pragma unroll
    for (int j = 0; j < ROUNDS; ++j) {
        if (j * 32 >= n)
            break;
        const int i = j * 32 + lane;
        const bool valid = i < n;
        const bool tie = valid && keys[j] == threshold;
        const uint32_t tie_mask = __ballot_sync(0xffffffffu, tie);
        const bool keep = valid && (keys[j] < threshold || (tie && ties_seen + __popc(tie_mask & lower_lanes) < ties_wanted));
        const uint32_t keep_mask = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int slot = out + kept + __popc(keep_mask & lower_lanes);
            p.out_data[slot] = values[i];
            p.out_indices[slot] = indices[i];
        }
        kept += __popc(keep_mask);
        ties_seen += __popc(tie_mask);
    }
}

__global__ void __launch_bounds__(256) prune_bands_warp_kernel(PruneParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t band = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; band < p.bands; band += warps) {
        const int lo = p.indptr[band], n = p.indptr[band + 1] - lo;
        const int out = p.out_indptr[band];
        if (n <= p.degree) {  // prune_per.cpp:24-28
            for (int i = lane; i < n; i += 32) {
                p.out_data[out + i] = p.data[lo + i];
                p.out_indices[out + i] = p.indices[lo + i];
            }
        } else if (n <= 32 * 32) {
            prune_band_in_warp<32>(p, lo, n, out);
        } else if (n <= p.warp_limit) {
            prune_band_in_warp<PW_MAX_ROUNDS>(p, lo, n, out);
        }
    }
}

__global__ void __launch_bounds__(SP_THREADS) prune_bands_kernel(PruneParams p) {
    extern __shared__ __align__(16) uint8_t sp_smem[];
    __shared__ SelectState state;
    __shared__ int scan_scratch[34];
    __shared__ int carry[2];
    uint32_t* cache = reinterpret_cast<uint32_t*>(sp_smem);
    const int tid = threadIdx.x;
    for (int64_t band = blockIdx.x; band < p.bands; band += gridDim.x) {
        const int lo = p.indptr[band], n = p.indptr[band + 1] - lo;
        const int out = p.out_indptr[band];
        const float* values = p.data + lo;
        const int32_t* indices = p.indices + lo;
        if (n <= p.degree || n <= p.warp_limit)
            continue;  // prune_bands_warp_kernel's
        const bool cached = n <= p.cache_capacity;
        if (cached) {
            for (int i = tid; i < n; i += SP_THREADS) cache[i] = sort_key(values[i], false);
            __syncthreads();
        }
        auto key_at = [&](int i) { return cached ? cache[i] : sort_key(values[i], false); };
        int residual;
        const uint32_t threshold = block_radix_select(key_at, n, p.degree - 1, &state, &residual);
        const int ties_wanted = residual + 1;
        if (tid == 0) {
            carry[0] = 0;
            carry[1] = 0;
        }
        __syncthreads();
        for (int base = 0; base < n; base += SP_THREADS) {
            const int i = base + tid;
            const uint32_t key = i < n ? key_at(i) : 0xffffffffu;
            const int is_tie = (i < n && key == threshold) ? 1 : 0;
            int ties_total, kept_total;
            const int ties_before = carry[1] + block_exclusive_scan(is_tie, scan_scratch, &ties_total);
            const int keep = (i < n && (key < threshold || (is_tie && ties_before < ties_wanted))) ? 1 : 0;
            const int slot = carry[0] + block_exclusive_scan(keep, scan_scratch, &kept_total);
            if (keep) {
                p.out_data[out + slot] = values[i];
                p.out_indices[out + slot] = indices[i];
            }
            __syncthreads();
            if (tid == 0) {
                carry[0] += kept_total;
                carry[1] += ties_total;
            }
            __syncthreads();
        }
    }
}

// ---- collect_compressed / sort_compressed_indices -------------------------------------------------

__global__ void count_indices_kernel(const int32_t* __restrict__ indices, int64_t nnz, int32_t* __restrict__ counts) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&counts[indices[e]], 1);
}

// One warp per input band; `cursor[b]` holds the next free offset of output band b and ends up at
// the end of that band, which is the in/out contract of relayout.cpp:73-108 (the caller passes
// output_indptr[1:] holding the band starts, computation.py:268-292).
__global__ void __launch_bounds__(256) scatter_bands_kernel(const float* __restrict__ data, const int32_t* __restrict__ indices,
                                                           const int32_t* __restrict__ indptr, int64_t bands,
                                                           float* __restrict__ out_data, int32_t* __restrict__ out_indices,
                                                           int32_t* __restrict__ cursor) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t band = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; band < bands; band += warps) {
        const int lo = indptr[band], hi = indptr[band + 1];
        for (int e = lo + lane; e < hi; e += 32) {
            const int at = atomicAdd(&cursor[indices[e]], 1);
            out_indices[at] = (int32_t)band;
            out_data[at] = data[e];
        }
    }
}

// Sort every band by index: bitonic sort of (index << 32 | data bits) in shared memory (or, for a
// band wider than the shared-memory capacity, in a per-CTA global workspace).
__global__ void __launch_bounds__(SP_THREADS) sort_bands_kernel(float* __restrict__ data, int32_t* __restrict__ indices,
                                                               const int32_t* __restrict__ indptr, int64_t bands,
                                                               int smem_capacity, unsigned long long* __restrict__ workspace,
                                                               int64_t workspace_capacity) {
    extern __shared__ __align__(16) uint8_t sp_smem[];
    const int tid = threadIdx.x;
    for (int64_t band = blockIdx.x; band < bands; band += gridDim.x) {
        const int lo = indptr[band], n = indptr[band + 1] - lo;
        if (n < 2)
            continue;
        // already sorted?  (the common case after a deterministic scatter)
        int unsorted = 0;
        for (int i = tid; i + 1 < n; i += SP_THREADS) unsorted |= indices[lo + i] > indices[lo + i + 1];
        if (!__syncthreads_or(unsorted))
            continue;
        int padded = 1;
        while (padded < n) padded <<= 1;
        unsigned long long* keys = padded <= smem_capacity ? reinterpret_cast<unsigned long long*>(sp_smem)
                                                            : workspace + (int64_t)blockIdx.x * workspace_capacity;
        for (int i = tid; i < padded; i += SP_THREADS)
            keys[i] = i < n ? (((unsigned long long)(uint32_t)indices[lo + i]) << 32) | __float_as_uint(data[lo + i])
                            : 0xffffffffffffffffull;
        __syncthreads();
        for (int size = 2; size <= padded; size <<= 1) {
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                for (int t = tid; t < (padded >> 1); t += SP_THREADS) {
                    const int i = ((t & ~(stride - 1)) << 1) | (t & (stride - 1));
                    const int j = i | stride;
                    const bool up = (i & size) == 0;
                    const unsigned long long a = keys[i], b = keys[j];
                    if ((a > b) == up) {
                        keys[i] = b;
                        keys[j] = a;
                    }
                }
                __syncthreads();
            }
        }
        for (int i = tid; i < n; i += SP_THREADS) {
            indices[lo + i] = (int32_t)(keys[i] >> 32);
            data[lo + i] = __uint_as_float((uint32_t)keys[i]);
        }
        __syncthreads();
    }
}

__global__ void max_band_kernel(const int32_t* __restrict__ indptr, int64_t bands, int32_t* __restrict__ widest) {
    int local = 0;
    for (int64_t band = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; band < bands; band += (int64_t)gridDim.x * blockDim.x)
        local = max(local, indptr[band + 1] - indptr[band]);
    for (int offset = 16; offset > 0; offset >>= 1) local = max(local, __shfl_xor_sync(0xffffffffu, local, offset));
    if ((threadIdx.x & 31) == 0)
        atomicMax(widest, local);
}


}  // namespace

int exclusive_scan_i32(const int32_t* input, int32_t* output, int64_t count, int32_t* total_out, cudaStream_t stream) {
    MCB_PROFILED("exclusive_scan", stream, exclusive_scan_kernel<<<1, 1024, 0, stream>>>(input, output, count, total_out));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

// out_indptr[bands + 1] is written; out_data / out_indices need room for sum(min(size, degree)).
int collect_pruned_launch(int64_t degree, const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands,
                          float* out_data, int32_t* out_indices, int32_t* out_indptr, cudaStream_t stream) {
    MCB_REQUIRE(degree >= 0 && degree < (1ll << 31), "collect_pruned: bad degree %lld", (long long)degree);
    Scratch sizes;
    MCB_TRY(sizes.alloc((size_t)(bands ? bands : 1) * sizeof(int32_t), stream));
    if (bands > 0) {
        pruned_sizes_kernel<<<(unsigned)ceil_div(bands, 256), 256, 0, stream>>>(indptr, bands, (int32_t)degree, sizes.as<int32_t>());
        MCB_LAUNCH_CHECK();
    }
    MCB_TRY(exclusive_scan_i32(sizes.as<int32_t>(), out_indptr, bands, nullptr, stream));
    if (bands == 0)
        return MCB_OK;
    PruneParams p{};
    p.data = data;
    p.indices = indices;
    p.indptr = indptr;
    p.bands = bands;
    p.degree = (int)degree;
    p.out_data = out_data;
    p.out_indices = out_indices;
    p.out_indptr = out_indptr;
    p.cache_capacity = 8192;
    p.warp_limit = PW_MAX_ROUNDS * 32;
    const size_t smem_bytes = (size_t)p.cache_capacity * 4;
    const int64_t warp_grid = std::min<int64_t>(ceil_div(bands * 32, 256), (int64_t)sm_count() * 8);
    MCB_PROFILED("prune_bands", stream, prune_bands_warp_kernel<<<(unsigned)warp_grid, 256, 0, stream>>>(p));
    MCB_LAUNCH_CHECK();
    const int64_t grid = std::min<int64_t>(bands, (int64_t)sm_count() * 6);
    MCB_PROFILED("prune_wide_bands", stream, prune_bands_kernel<<<(unsigned)grid, SP_THREADS, smem_bytes, stream>>>(p));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

// In/out `cursor` (= output_indptr[1:] of the caller): band starts on entry, band ends on return.
int collect_compressed_launch(const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands, float* out_data,
                              int32_t* out_indices, int32_t* cursor, cudaStream_t stream) {
    if (bands == 0)
        return MCB_OK;
    const int64_t grid = std::min<int64_t>(ceil_div(bands * 32, 256), (int64_t)sm_count() * 16);
    MCB_PROFILED("relayout_scatter", stream, scatter_bands_kernel<<<(unsigned)grid, 256, 0, stream>>>(data, indices, indptr, bands, out_data, out_indices, cursor));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

// widest_band: an upper bound of the band sizes, or < 0 to have it computed (one extra sync).
int sort_compressed_indices_launch(float* data, int32_t* indices, const int32_t* indptr, int64_t bands, int64_t widest_band,
                                   cudaStream_t stream) {
    if (bands == 0)
        return MCB_OK;
    if (widest_band < 0) {
        Scratch widest;
        MCB_TRY(widest.alloc(sizeof(int32_t), stream));
        MCB_CUDA(cudaMemsetAsync(widest.ptr, 0, sizeof(int32_t), stream));
        max_band_kernel<<<(unsigned)std::min<int64_t>(ceil_div(bands, 256), 1024), 256, 0, stream>>>(indptr, bands, widest.as<int32_t>());
        MCB_LAUNCH_CHECK();
        int32_t host = 0;
        MCB_TRY(copy_to_host(&host, widest.ptr, sizeof(int32_t), stream));
        MCB_CUDA(cudaStreamSynchronize(stream));
        widest_band = host;
    }
    int64_t padded = 1;
    while (padded < widest_band) padded <<= 1;
    // 32 KiB of shared memory keeps several CTAs per SM; the rare wider band sorts in global memory.
    const int smem_capacity = (int)std::min<int64_t>(padded, 4096);
    const size_t smem_bytes = (size_t)smem_capacity * 8;
    const int64_t grid = std::min<int64_t>(bands, (int64_t)sm_count() * 4);
    Scratch workspace;
    if (padded > smem_capacity)
        MCB_TRY(workspace.alloc((size_t)grid * padded * 8, stream));
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(sort_bands_kernel), (size_t)smem_bytes));
    MCB_PROFILED("sort_bands", stream, sort_bands_kernel<<<(unsigned)grid, SP_THREADS, smem_bytes, stream>>>(data, indices, indptr, bands, smem_capacity,
                                                                          workspace.as<unsigned long long>(), padded));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

namespace {

// Bands whose indices are DISTINCT (every transpose or merge of a proper compressed matrix): a counting sort
// through a bitmap of the band's indices -- position = number of smaller indices present = popcount prefix.
// Linear in band size + index_range / 32, against the log^2 stages of the bitonic sort above.
__global__ void __launch_bounds__(SP_THREADS) sort_distinct_bands_kernel(float* __restrict__ data, int32_t* __restrict__ indices,
                                                                        const int32_t* __restrict__ indptr, int64_t bands,
                                                                        int index_range) {
    extern __shared__ __align__(16) uint8_t sd_smem[];
    __shared__ int scan_scratch[34];
    const int words = (index_range + 31) >> 5;
    uint32_t* bitmap = reinterpret_cast<uint32_t*>(sd_smem);
    uint32_t* before = bitmap + words;
    unsigned long long* staged = reinterpret_cast<unsigned long long*>(sd_smem + (((size_t)words * 8 + 15) & ~(size_t)15));
    const int tid = threadIdx.x;
    for (int64_t band = blockIdx.x; band < bands; band += gridDim.x) {
        const int lo = indptr[band], count = indptr[band + 1] - lo;
        if (count <= 1)
            continue;
        for (int w = tid; w < words; w += SP_THREADS) bitmap[w] = 0;
        __syncthreads();
        for (int e = tid; e < count; e += SP_THREADS) {
            const uint32_t index = (uint32_t)indices[lo + e];
            staged[e] = ((unsigned long long)index << 32) | __float_as_uint(data[lo + e]);
            atomicOr(&bitmap[index >> 5], 1u << (index & 31));
        }
        __syncthreads();
        int running = 0;
        for (int base = 0; base < words; base += SP_THREADS) {
            const int w = base + tid;
            const int present = w < words ? __popc(bitmap[w]) : 0;
            int total;
            const int exclusive = block_exclusive_scan(present, scan_scratch, &total);
            if (w < words)
                before[w] = (uint32_t)(running + exclusive);
            running += total;
        }
        __syncthreads();
        for (int e = tid; e < count; e += SP_THREADS) {
            const unsigned long long entry = staged[e];
            const uint32_t index = (uint32_t)(entry >> 32);
            const int at = lo + (int)before[index >> 5] + __popc(bitmap[index >> 5] & ((1u << (index & 31)) - 1u));
            indices[at] = (int32_t)index;
            data[at] = __uint_as_float((uint32_t)entry);
        }
        __syncthreads();
    }
}

}  // namespace

// As sort_compressed_indices_launch for bands of distinct indices below index_range (see the kernel); falls back to
// the general sort when a band or the index range is too large for the shared-memory plan.
int sort_distinct_indices_launch(float* data, int32_t* indices, const int32_t* indptr, int64_t bands, int64_t widest_band,
                                 int64_t index_range, cudaStream_t stream) {
    if (bands == 0)
        return MCB_OK;
    if (index_range > (1 << 17))
        return sort_compressed_indices_launch(data, indices, indptr, bands, widest_band, stream);
    if (widest_band < 0 || widest_band > 2048) {  // a loose bound would waste shared memory: measure
        Scratch widest;
        MCB_TRY(widest.alloc(sizeof(int32_t), stream));
        MCB_CUDA(cudaMemsetAsync(widest.ptr, 0, sizeof(int32_t), stream));
        max_band_kernel<<<(unsigned)std::min<int64_t>(ceil_div(bands, 256), 1024), 256, 0, stream>>>(indptr, bands, widest.as<int32_t>());
        MCB_LAUNCH_CHECK();
        int32_t host = 0;
        MCB_TRY(copy_to_host(&host, widest.ptr, sizeof(int32_t), stream));
        MCB_CUDA(cudaStreamSynchronize(stream));
        widest_band = host;
    }
    if (widest_band > 8192)
        return sort_compressed_indices_launch(data, indices, indptr, bands, widest_band, stream);
    const size_t words = (size_t)((index_range + 31) >> 5);
    const size_t smem_bytes = ((words * 8 + 15) & ~(size_t)15) + (size_t)widest_band * 8;
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(sort_distinct_bands_kernel), (size_t)smem_bytes));
    const int64_t grid = std::min<int64_t>(bands, (int64_t)sm_count() * 8);
    MCB_PROFILED("sort_bands", stream, sort_distinct_bands_kernel<<<(unsigned)grid, SP_THREADS, smem_bytes, stream>>>(
                     data, indices, indptr, bands, (int)index_range));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

// Full transpose: counts -> out_indptr[bands_out + 1], scatter, sort.  The relayout of
// computation.py:251-303 (bincount + cumsum + collect_compressed + sort_compressed_indices).
int relayout_launch(const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands_in, int64_t bands_out,
                    int64_t nnz, float* out_data, int32_t* out_indices, int32_t* out_indptr, cudaStream_t stream) {
    Scratch counts, cursor;
    MCB_TRY(counts.alloc((size_t)(bands_out + 1) * sizeof(int32_t), stream));
    MCB_TRY(cursor.alloc((size_t)(bands_out + 1) * sizeof(int32_t), stream));
    MCB_CUDA(cudaMemsetAsync(counts.ptr, 0, (size_t)(bands_out + 1) * sizeof(int32_t), stream));
    if (nnz > 0) {
        MCB_PROFILED("relayout_count", stream, count_indices_kernel<<<(unsigned)std::min<int64_t>(ceil_div(nnz, 256), (int64_t)sm_count() * 16), 256, 0, stream>>>(
            indices, nnz, counts.as<int32_t>()));
        MCB_LAUNCH_CHECK();
    }
    MCB_TRY(exclusive_scan_i32(counts.as<int32_t>(), out_indptr, bands_out, nullptr, stream));
    MCB_CUDA(cudaMemcpyAsync(cursor.ptr, out_indptr, (size_t)(bands_out ? bands_out : 1) * sizeof(int32_t), cudaMemcpyDeviceToDevice, stream));
    MCB_TRY(collect_compressed_launch(data, indices, indptr, bands_in, out_data, out_indices, cursor.as<int32_t>(), stream));
    return sort_distinct_indices_launch(out_data, out_indices, out_indptr, bands_out, bands_in, bands_in, stream);
}

}  // namespace mcb
