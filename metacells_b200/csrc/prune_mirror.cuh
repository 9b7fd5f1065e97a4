// _prune_ranks (knn_graph.py:343-392) without the transposition.
//
// The reference prunes the balanced ranks twice: per "column" first, on the CSR arrays read as CSC
// (`to_layout(..., symmetric=True)`, knn_graph.py:371-374: band i keeps its best `incoming_degree` entries, the set
// S_i), then it really transposes (:377) and every band r of the result keeps the best `outgoing_degree` of the entries
// (i, value B'[i][r]) with r in S_i (:380).  B' is symmetric except for the preserved entry of every row, so the
// entries band r can receive are the columns of its own CSR row plus the rows whose preserved column is r:
//
//   row_cuts_kernel      per row i: the cut of its first prune, (key of the last value kept, column of the last
//                        kept among equal values): "r in S_i" is a comparison with the cut of row i;
//   mirror_prune_kernel  per row r: every entry (r, i) looks up the cut of row i, takes the value the transposed
//                        entry holds (B'[i][r]: the preserved entry of row i is raised to 1, knn_graph.py:331), and the
//                        best `outgoing_degree` survivors are written in column order, equal values by lower column
//                        as in the transposed band; rows whose preserved column is r but that are not among r's own
//                        columns compete too (their entries come back with merge_preserved).
//
// Bands of up to 2048 entries sit in a warp's registers, wider ones (a few hub cells of a pile) are read again in
// every pass.  A row with so many foreign preserved entries that a lane would hold two makes the caller fall back to
// prune -> relayout -> prune.  Included by knn_graph.cu only.
#pragma once

#include "common.cuh"
#include "select.cuh"

namespace mcb {

namespace {

struct RowCut {
    uint32_t key;       // sort_key(value, descending) of the last value the row keeps; 0xffffffff: the row keeps everything
    int32_t column;     // among values equal to it, the last column kept
    int32_t preserved;  // the row's preserved column
    int32_t unused;
};

struct MirrorParams {
    const float* data;  // balanced CSR (sorted columns)
    const int32_t* indices;
    const int32_t* indptr;
    int64_t n;
    int incoming_degree, outgoing_degree;
    const int32_t* best_index;  // preserved column of every row
    const float* best_value;    // its value, raised to 1
    const float* best_raw;      // its value as computed
    const int32_t* reverse_ptr;   // rows whose preserved column is this row
    const int32_t* reverse_list;
    RowCut* cuts;
    float* padded_data;  // [n, outgoing_degree]
    int32_t* padded_indices;
    int32_t* sizes;
    int* overflow;  // set when a lane would have to hold two foreign preserved entries of one row
    int32_t* class_rows;  // [3, n]: the rows of every band-width class (up to 1024 / up to 2048 / wider), in any order
    int* class_counts;    // [3], filled where the rows' sizes are computed (balance_rows_kernel)
};

constexpr int MP_MAX_ROUNDS = 64;  // bands of up to 2048 entries

template <int ROUNDS>
__device__ __forceinline__ void row_cut_in_warp(const MirrorParams& p, int64_t row, int lo, int n) {
    const int lane = threadIdx.x & 31;
    const uint32_t lower_lanes = lanemask_lt();
    const float* values = p.data + lo;
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
    int wanted = p.incoming_degree - 1;
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
    // the (wanted + 1)-th entry, in column order, among the entries equal to the threshold is the last one kept
    int ties_seen = 0;
#pragma unroll
    for (int j = 0; j < ROUNDS; ++j) {
        if (j * 32 >= n)
            break;
        const int i = j * 32 + lane;
        const bool tie = i < n && keys[j] == prefix;
        const uint32_t tie_mask = __ballot_sync(0xffffffffu, tie);
        if (tie && ties_seen + __popc(tie_mask & lower_lanes) == wanted) {
            RowCut cut;
            cut.key = prefix;
            cut.column = p.indices[lo + i];
            cut.preserved = p.best_index[row];
            cut.unused = 0;
            p.cuts[row] = cut;
        }
        ties_seen += __popc(tie_mask);
    }
}

// A band wider than a warp's registers hold (rare: a few hub cells): the keys are read again in every pass.
__device__ void row_cut_streamed(const MirrorParams& p, int64_t row, int lo, int n) {
    const int lane = threadIdx.x & 31;
    const uint32_t lower_lanes = lanemask_lt();
    const float* values = p.data + lo;
    const uint32_t first = sort_key(values[0], false);
    uint32_t differ = 0;
    for (int i = lane; i < n; i += 32) differ |= sort_key(values[i], false) ^ first;
    differ = __reduce_or_sync(0xffffffffu, differ);
    uint32_t prefix = first & ~differ, known = ~differ;
    int wanted = p.incoming_degree - 1;
    for (int bit = 31 - __clz((int)(differ | 1u)); bit >= 0; --bit) {
        if (((differ >> bit) & 1u) == 0)
            continue;
        int zeros = 0;
        for (int i = lane; i < n; i += 32) {
            const uint32_t key = sort_key(values[i], false);
            zeros += ((key & known) == prefix && ((key >> bit) & 1u) == 0) ? 1 : 0;
        }
        zeros = __reduce_add_sync(0xffffffffu, zeros);
        if (wanted >= zeros) {
            wanted -= zeros;
            prefix |= 1u << bit;
        }
        known |= 1u << bit;
    }
    int ties_seen = 0;
    for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        const bool tie = i < n && sort_key(values[i], false) == prefix;
        const uint32_t tie_mask = __ballot_sync(0xffffffffu, tie);
        if (tie && ties_seen + __popc(tie_mask & lower_lanes) == wanted) {
            RowCut cut;
            cut.key = prefix;
            cut.column = p.indices[lo + i];
            cut.preserved = p.best_index[row];
            cut.unused = 0;
            p.cuts[row] = cut;
        }
        ties_seen += __popc(tie_mask);
    }
}

// Three launches by band width, as mirror_prune_kernel below: ROUNDS = 32 (up to 1024 entries), MP_MAX_ROUNDS (up to 2048),
// 0 (wider: streamed), each over the list of the rows of its class.
template <int ROUNDS>
__global__ void __launch_bounds__(256) row_cuts_kernel(MirrorParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    constexpr int BAND_CLASS = ROUNDS == 32 ? 0 : (ROUNDS == MP_MAX_ROUNDS ? 1 : 2);
    const int32_t* rows = p.class_rows + BAND_CLASS * p.n;
    const int count = p.class_counts[BAND_CLASS];
    for (int64_t at = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; at < count; at += warps) {
        const int64_t row = rows[at];
        const int lo = p.indptr[row], n = p.indptr[row + 1] - lo;
        if (n <= p.incoming_degree) {  // prune_per.cpp:24-28: the band is kept as it is
            if (lane == 0) {
                RowCut cut;
                cut.key = 0xffffffffu;
                cut.column = 0x7fffffff;
                cut.preserved = p.best_index[row];
                cut.unused = 0;
                p.cuts[row] = cut;
            }
        } else if (ROUNDS == 32) {
            row_cut_in_warp<32>(p, row, lo, n);
        } else if (ROUNDS == MP_MAX_ROUNDS) {
            row_cut_in_warp<MP_MAX_ROUNDS>(p, row, lo, n);
        } else {
            row_cut_streamed(p, row, lo, n);
        }
    }
}

// Is (transposed value `key`, band `row`) among the entries row `other` keeps in its first prune?
__device__ __forceinline__ bool kept_by(const RowCut& cut, uint32_t key, int64_t row) {
    return key < cut.key || (key == cut.key && row <= (int64_t)cut.column);
}

// An entry (row, other) seen from the other side: the key of the value the transposed entry holds, and whether that
// entry exists and survives the first prune of its own row (`cut`: the cut of row `other`).
__device__ __forceinline__ bool transposed_entry(const MirrorParams& p, int64_t row, int32_t other, float value, const RowCut& cut,
                                                 int32_t own_preserved, float own_raw, uint32_t* key_out) {
    const bool mine = other == own_preserved;  // the entry is this row's preserved one: stored raised
    const float as_computed = mine ? own_raw : value;
    const bool passes = mine ? (as_computed > 0.0f && other != row) : true;  // knn_graph.py:333-335
    const bool theirs = cut.preserved == row;  // the transposed entry is row `other`'s preserved one
    const float transposed = theirs ? p.best_value[other] : as_computed;
    const uint32_t key = sort_key(transposed, false);
    *key_out = key;
    return (passes || theirs) && (key < cut.key || (key == cut.key && row <= (int64_t)cut.column));
}

__device__ __forceinline__ RowCut load_cut(const MirrorParams& p, int32_t other) {
    const int4 raw = reinterpret_cast<const int4*>(p.cuts)[other];  // one 128-bit load
    RowCut cut;
    cut.key = (uint32_t)raw.x, cut.column = raw.y, cut.preserved = raw.z, cut.unused = raw.w;
    return cut;
}

__device__ __forceinline__ bool transposed_entry(const MirrorParams& p, int64_t row, const float* values, const int32_t* indices,
                                                 int i, int32_t own_preserved, float own_raw, uint32_t* key_out) {
    const int32_t other = indices[i];
    return transposed_entry(p, row, other, values[i], load_cut(p, other), own_preserved, own_raw, key_out);
}

// The rows whose preserved column is `row` without being one of its columns (their preserved entry failed the
// threshold): they compete in the second prune of band `row`.  A lane holds at most one; more set the overflow flag.
__device__ __forceinline__ void foreign_preserved(const MirrorParams& p, int64_t row, const int32_t* indices, int n,
                                                  uint32_t* extra_key, int32_t* extra_row, uint32_t* all_or, uint32_t* all_and,
                                                  int* count) {
    const int lane = threadIdx.x & 31;
    const int begin = p.reverse_ptr[row], end = p.reverse_ptr[row + 1];
    for (int at = begin + lane; at < end; at += 32) {
        const int32_t other = p.reverse_list[at];
        int left = 0, right = n;
        while (left < right) {
            const int middle = (left + right) >> 1;
            if (indices[middle] < other)
                left = middle + 1;
            else
                right = middle;
        }
        if (left < n && indices[left] == other)
            continue;
        const uint32_t key = sort_key(p.best_value[other], false);
        if (!kept_by(p.cuts[other], key, row))
            continue;
        if (*extra_row >= 0) {
            atomicExch(p.overflow, 1);
            continue;
        }
        *extra_key = key;
        *extra_row = other;
        *all_or |= key;
        *all_and &= key;
        *count += 1;
    }
}

template <int ROUNDS>
__device__ __forceinline__ void mirror_prune_in_warp(const MirrorParams& p, int64_t row, int lo, int n) {
    const int lane = threadIdx.x & 31;
    const uint32_t lower_lanes = lanemask_lt();
    const float* values = p.data + lo;
    const int32_t* indices = p.indices + lo;
    const int32_t own_preserved = p.best_index[row];
    const float own_raw = p.best_raw[row];
    uint32_t keys[ROUNDS];
#pragma unroll
    for (int j = 0; j < ROUNDS; ++j) keys[j] = 0;
    uint64_t members = 0;  // bit j: entry j * 32 + lane takes part in the second prune
    uint32_t all_or = 0, all_and = 0xffffffffu;
    int count = 0;
    // eight rounds at a time: their columns first, then the eight cuts (independent gathers in flight), then the tests
    constexpr int CHUNK = 8;
#pragma unroll
    for (int first = 0; first < ROUNDS; first += CHUNK) {
        if (first * 32 >= n)
            break;
        int32_t other[CHUNK];
        float value[CHUNK];
        RowCut cut[CHUNK];
#pragma unroll
        for (int u = 0; u < CHUNK; ++u) {
            const int i = (first + u) * 32 + lane;
            other[u] = i < n ? indices[i] : -1;
            value[u] = i < n ? values[i] : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < CHUNK; ++u) cut[u] = load_cut(p, max(other[u], 0));
#pragma unroll
        for (int u = 0; u < CHUNK; ++u) {
            const int j = first + u;
            keys[j] = 0;
            if (other[u] >= 0 && transposed_entry(p, row, other[u], value[u], cut[u], own_preserved, own_raw, &keys[j])) {
                members |= 1ull << j;
                all_or |= keys[j];
                all_and &= keys[j];
                ++count;
            }
        }
    }
    // rows whose preserved column is this row without being one of its columns
    uint32_t extra_key = 0;
    int32_t extra_row = -1;
    foreign_preserved(p, row, indices, n, &extra_key, &extra_row, &all_or, &all_and, &count);
    count = __reduce_add_sync(0xffffffffu, count);
    float* out_data = p.padded_data + row * p.outgoing_degree;
    int32_t* out_indices = p.padded_indices + row * p.outgoing_degree;
    uint32_t threshold = 0xffffffffu;  // keep every member
    int ties_wanted = 0x7fffffff;
    bool extra_ties = false;
    if (count > p.outgoing_degree) {
        all_or = __reduce_or_sync(0xffffffffu, all_or);
        all_and = __reduce_and_sync(0xffffffffu, all_and);
        const uint32_t differ = all_or ^ all_and;
        uint32_t prefix = all_and & ~differ, known = ~differ;
        int wanted = p.outgoing_degree - 1;
        for (int bit = 31 - __clz((int)(differ | 1u)); bit >= 0; --bit) {
            if (((differ >> bit) & 1u) == 0)
                continue;
            int zeros = (extra_row >= 0 && (extra_key & known) == prefix && ((extra_key >> bit) & 1u) == 0) ? 1 : 0;
#pragma unroll
            for (int j = 0; j < ROUNDS; ++j)
                zeros += ((members >> j & 1ull) && (keys[j] & known) == prefix && ((keys[j] >> bit) & 1u) == 0) ? 1 : 0;
            zeros = __reduce_add_sync(0xffffffffu, zeros);
            if (wanted >= zeros) {
                wanted -= zeros;
                prefix |= 1u << bit;
            }
            known |= 1u << bit;
        }
        threshold = prefix;
        ties_wanted = wanted + 1;
        extra_ties = __any_sync(0xffffffffu, extra_row >= 0 && extra_key == threshold);
    }
    int kept = 0, ties_seen = 0;
#pragma unroll
    for (int j = 0; j < ROUNDS; ++j) {
        if (j * 32 >= n)
            break;
        const int i = j * 32 + lane;
        const bool member = members >> j & 1ull;
        const bool tie = member && keys[j] == threshold;
        const uint32_t tie_mask = __ballot_sync(0xffffffffu, tie);
        int earlier = ties_seen + __popc(tie_mask & lower_lanes);
        if (extra_ties) {  // equal values keep the order of the transposed band: by row index
            const int32_t column = i < n ? indices[i] : 0;
            for (uint32_t holders = __ballot_sync(0xffffffffu, extra_row >= 0 && extra_key == threshold); holders;
                 holders &= holders - 1) {
                const int32_t extra = __shfl_sync(0xffffffffu, extra_row, __ffs(holders) - 1);
                earlier += extra < column ? 1 : 0;
            }
        }
        const bool keep = member && (keys[j] < threshold || (tie && earlier < ties_wanted));
        const uint32_t keep_mask = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int slot = kept + __popc(keep_mask & lower_lanes);
            out_data[slot] = ordered_to_float(~keys[j]);
            out_indices[slot] = indices[i];
        }
        kept += __popc(keep_mask);
        ties_seen += __popc(tie_mask);
    }
    if (lane == 0)
        p.sizes[row] = kept;
}

// The same for a band wider than a warp's registers hold.  `cache` (shared memory of the warp, n entries) keeps the key
// of every member (0xffffffff: not a member) after the first pass; without it the entries are looked up again in every
// pass.
__device__ void mirror_prune_streamed(const MirrorParams& p, int64_t row, int lo, int n, uint32_t* cache) {
    const int lane = threadIdx.x & 31;
    const uint32_t lower_lanes = lanemask_lt();
    const float* values = p.data + lo;
    const int32_t* indices = p.indices + lo;
    const int32_t own_preserved = p.best_index[row];
    const float own_raw = p.best_raw[row];
    uint32_t all_or = 0, all_and = 0xffffffffu;
    int count = 0;
    for (int i = lane; i < n; i += 32) {
        uint32_t key;
        const bool member = transposed_entry(p, row, values, indices, i, own_preserved, own_raw, &key);
        if (member) {
            all_or |= key;
            all_and &= key;
            ++count;
        }
        if (cache) cache[i] = member ? key : 0xffffffffu;
    }
    __syncwarp();
    uint32_t extra_key = 0;
    int32_t extra_row = -1;
    foreign_preserved(p, row, indices, n, &extra_key, &extra_row, &all_or, &all_and, &count);
    count = __reduce_add_sync(0xffffffffu, count);
    float* out_data = p.padded_data + row * p.outgoing_degree;
    int32_t* out_indices = p.padded_indices + row * p.outgoing_degree;
    uint32_t threshold = 0xffffffffu;
    int ties_wanted = 0x7fffffff;
    bool extra_ties = false;
    if (count > p.outgoing_degree) {
        all_or = __reduce_or_sync(0xffffffffu, all_or);
        all_and = __reduce_and_sync(0xffffffffu, all_and);
        const uint32_t differ = all_or ^ all_and;
        uint32_t prefix = all_and & ~differ, known = ~differ;
        int wanted = p.outgoing_degree - 1;
        for (int bit = 31 - __clz((int)(differ | 1u)); bit >= 0; --bit) {
            if (((differ >> bit) & 1u) == 0)
                continue;
            int zeros = (extra_row >= 0 && (extra_key & known) == prefix && ((extra_key >> bit) & 1u) == 0) ? 1 : 0;
            for (int i = lane; i < n; i += 32) {
                uint32_t key;
                bool member;
                if (cache) {
                    key = cache[i];
                    member = key != 0xffffffffu;
                } else {
                    member = transposed_entry(p, row, values, indices, i, own_preserved, own_raw, &key);
                }
                zeros += (member && (key & known) == prefix && ((key >> bit) & 1u) == 0) ? 1 : 0;
            }
            zeros = __reduce_add_sync(0xffffffffu, zeros);
            if (wanted >= zeros) {
                wanted -= zeros;
                prefix |= 1u << bit;
            }
            known |= 1u << bit;
        }
        threshold = prefix;
        ties_wanted = wanted + 1;
        extra_ties = __any_sync(0xffffffffu, extra_row >= 0 && extra_key == threshold);
    }
    int kept = 0, ties_seen = 0;
    for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        uint32_t key = 0;
        bool member = false;
        if (i < n) {
            if (cache) {
                key = cache[i];
                member = key != 0xffffffffu;
            } else {
                member = transposed_entry(p, row, values, indices, i, own_preserved, own_raw, &key);
            }
        }
        const bool tie = member && key == threshold;
        const uint32_t tie_mask = __ballot_sync(0xffffffffu, tie);
        int earlier = ties_seen + __popc(tie_mask & lower_lanes);
        if (extra_ties) {
            const int32_t column = i < n ? indices[i] : 0;
            for (uint32_t holders = __ballot_sync(0xffffffffu, extra_row >= 0 && extra_key == threshold); holders;
                 holders &= holders - 1) {
                const int32_t extra = __shfl_sync(0xffffffffu, extra_row, __ffs(holders) - 1);
                earlier += extra < column ? 1 : 0;
            }
        }
        const bool keep = member && (key < threshold || (tie && earlier < ties_wanted));
        const uint32_t keep_mask = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int slot = kept + __popc(keep_mask & lower_lanes);
            out_data[slot] = ordered_to_float(~key);
            out_indices[slot] = indices[i];
        }
        kept += __popc(keep_mask);
        ties_seen += __popc(tie_mask);
    }
    if (lane == 0)
        p.sizes[row] = kept;
}

// Three launches, by band width, so that the narrow bands (most of them) do not pay for the registers of the wide ones:
// ROUNDS = 32: up to 1024 entries; 64: up to 2048; 0: the few wider ones, keys in shared memory (MP_CACHE per warp).
constexpr int MP_CACHE = 8192;

template <int ROUNDS>
__global__ void __launch_bounds__(128, ROUNDS == 32 ? 4 : (ROUNDS == 0 ? 1 : 3)) mirror_prune_kernel(MirrorParams p) {
    extern __shared__ uint32_t mp_cache[];
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    constexpr int BAND_CLASS = ROUNDS == 32 ? 0 : (ROUNDS == MP_MAX_ROUNDS ? 1 : 2);
    const int32_t* rows = p.class_rows + BAND_CLASS * p.n;
    const int count = p.class_counts[BAND_CLASS];
    // the warps of the first CTAs take one row of the class each: every resident warp works, whatever share of the
    // rows the class holds
    for (int64_t at = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; at < count; at += warps) {
        const int64_t row = rows[at];
        const int lo = p.indptr[row], n = p.indptr[row + 1] - lo;
        if (ROUNDS == 32)
            mirror_prune_in_warp<32>(p, row, lo, n);
        else if (ROUNDS == MP_MAX_ROUNDS)
            mirror_prune_in_warp<MP_MAX_ROUNDS>(p, row, lo, n);
        else
            mirror_prune_streamed(p, row, lo, n, n <= MP_CACHE ? mp_cache + (threadIdx.x >> 5) * MP_CACHE : nullptr);
    }
}

__global__ void __launch_bounds__(256) compact_rows_kernel(const float* __restrict__ padded_data,
                                                          const int32_t* __restrict__ padded_indices, int64_t n, int stride,
                                                          const int32_t* __restrict__ indptr, float* __restrict__ data,
                                                          int32_t* __restrict__ indices) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n)
        return;
    const int lo = indptr[row], size = indptr[row + 1] - lo;
    for (int i = lane; i < size; i += 32) {
        data[lo + i] = padded_data[row * stride + i];
        indices[lo + i] = padded_indices[row * stride + i];
    }
}

}  // namespace

}  // namespace mcb
