// Balanced-rank k-NN graph construction on the device: the numpy/scipy of
// tools/knn_graph.py:285-407 between the native calls, fused into a few passes.
//
//   prepare_similarity   knn_graph.py:288-292: diagonal <- min(similarity) - 1 (in place)
//   balance (3 kernels)  knn_graph.py:300-340: b = sqrt(r * r^T) in float32, the 2^21 offset round trip
//                        (which quantises b to 1/8), per-row argmax ("preserved" edge), threshold
//                        at max_rank, dense -> CSR.  r * r^T is symmetric, so 64 x 64 rank tiles are
//                        processed in pairs (I, J) / (J, I): each tile is read once per pass and
//                        serves the rows of I and the rows of J.  Pass 1 counts the surviving
//                        entries per (row, column tile) and takes the argmax, pass 2 scans, pass 3
//                        recomputes and writes the CSR in place: 2 reads of the rank matrix and one
//                        write of the CSR, against >= 8 full read+write passes of the numpy code.
//   merge_preserved      knn_graph.py:357-369, 383-388: element-wise maximum with the preserved edge of
//                        every row and with its transpose.
//   weigh_edges          knn_graph.py:395-407: divide every row by its float32 sum.
#include <cmath>
#include <cstring>
#include <type_traits>
#include <algorithm>

#include "numpy_sum.cuh"
#include "common.cuh"
#include "select.cuh"
#include "prune_mirror.cuh"
#include "knn_graph.cuh"

namespace mcb {

int exclusive_scan_i32(const int32_t* input, int32_t* output, int64_t count, int32_t* total_out, cudaStream_t stream);
int sort_compressed_indices_launch(float* data, int32_t* indices, const int32_t* indptr, int64_t bands, int64_t widest_band,
                                   cudaStream_t stream);
int sort_distinct_indices_launch(float* data, int32_t* indices, const int32_t* indptr, int64_t bands, int64_t widest_band,
                                 int64_t index_range, cudaStream_t stream);

namespace {

constexpr int BT = 64;  // rank tile edge
constexpr int BAL_THREADS = 256;
constexpr float OFFSET_2_21 = 2097152.0f;

// ---- prepare_similarity ------------------------------------------------------------------------

__global__ void min_matrix_kernel(const float* __restrict__ matrix, int64_t rows, int64_t columns, int64_t row_stride,
                                  uint32_t* __restrict__ min_key) {
    uint32_t local = 0xffffffffu;
    const int64_t total = rows * columns;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x)
        local = min(local, float_to_ordered(matrix[(e / columns) * row_stride + e % columns]));
    for (int offset = 16; offset > 0; offset >>= 1) local = min(local, __shfl_xor_sync(0xffffffffu, local, offset));
    if ((threadIdx.x & 31) == 0)
        atomicMin(min_key, local);
}

__global__ void fill_diagonal_below_min_kernel(float* __restrict__ matrix, int64_t rows, int64_t row_stride,
                                               const uint32_t* __restrict__ min_key) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row < rows)
        matrix[row * row_stride + row] = __fadd_rn(ordered_to_float(*min_key), -1.0f);  // knn_graph.py:290-292
}

// ---- balance -------------------------------------------------------------------------------------

struct BalanceParams {
    const float* ranks;  // [n, n] outgoing ranks (row stride n)
    int64_t n;
    int tiles;           // ceil(n / BT)
    float shift;         // float32(max_rank + 1 - 2^21), knn_graph.py:325
    int32_t* counts;     // [n, tiles]: pass 1 counts, after pass 2 exclusive offsets inside the row
    unsigned long long* row_best;  // [n]: (ordered(t) << 32) | ~column  -> max t, lowest column
    const int32_t* indptr;         // pass 3
    float* out_data;
    int32_t* out_indices;
    const int32_t* best_index;     // pass 3: the preserved column of every row
    float keep_product;            // the largest float32 r_ij * r_ji whose balanced rank passes the threshold (0: none)
    // One pass over the ranks instead of two: the count pass also parks what it found -- per (row, column tile) the 64-bit
    // mask of the entries that pass and, packed at the start of the tile's 64 slots, their values -- and
    // balance_scatter_kernel places them once the offsets are known (null: balance_kernel<true> reads the ranks again).
    unsigned long long* masks;     // [n, tiles]
    float* staged;                 // [n, tiles, BT]
    const float* best_value;       // scatter: the value of the preserved entry of every row, raised to 1
};

__device__ __forceinline__ void tile_of_block(int block, int* tile_i, int* tile_j) {
    // block = J * (J + 1) / 2 + I with I <= J
    int j = (int)((sqrtf(8.0f * (float)block + 1.0f) - 1.0f) * 0.5f);
    while ((int64_t)j * (j + 1) / 2 > block) --j;
    while ((int64_t)(j + 1) * (j + 2) / 2 <= block) ++j;
    *tile_j = j;
    *tile_i = block - (int)((int64_t)j * (j + 1) / 2);
}

// t = 2^21 - sqrt(r_ij * r_ji), every step rounded to float32 as numpy does (knn_graph.py:308-317).
__device__ __forceinline__ float balanced_t_of_product(float product) { return __fadd_rn(-__fsqrt_rn(product), OFFSET_2_21); }
__device__ __forceinline__ float balanced_t(float r_ij, float r_ji) { return balanced_t_of_product(__fmul_rn(r_ij, r_ji)); }

// t lies in [2^20, 2^21) (ranks are below 2^20), where float32 steps by 1/8: 23 bits hold it exactly, and with the
// column inside the tile in the low 6 bits (inverted: numpy's first-maximum rule) a row's best entry of a tile is one
// 32-bit warp maximum.
__device__ __forceinline__ uint32_t pack_tile_best(float t, int column_in_tile) {
    return ((uint32_t)__float2int_rz(__fmul_rn(__fadd_rn(t, -1048576.0f), 8.0f)) << 6) | (uint32_t)(BT - 1 - column_in_tile);
}
static_assert(BT == 64, "pack_tile_best keeps the column of a tile in 6 bits");

template <bool FILL>
__global__ void __launch_bounds__(BAL_THREADS) balance_kernel(BalanceParams p) {
    __shared__ float tile_a[BT][BT + 1];  // ranks[I rows][J columns]
    __shared__ float tile_b[BT][BT + 1];  // ranks[J rows][I columns]
    int tile_i, tile_j;
    tile_of_block(blockIdx.x, &tile_i, &tile_j);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t n = p.n;
    const int64_t i0 = (int64_t)tile_i * BT, j0 = (int64_t)tile_j * BT;
    if ((n & 3) == 0 && i0 + BT <= n && j0 + BT <= n) {
        // whole tiles of a matrix whose rows start on 16-byte boundaries: 128-bit loads, 16 lanes per tile row
        const int quad = (threadIdx.x & 15) * 4;
        for (int r = threadIdx.x >> 4; r < BT; r += BAL_THREADS / 16) {
            const float4 a = *reinterpret_cast<const float4*>(p.ranks + (i0 + r) * n + j0 + quad);
            const float4 b = *reinterpret_cast<const float4*>(p.ranks + (j0 + r) * n + i0 + quad);
            tile_a[r][quad] = a.x, tile_a[r][quad + 1] = a.y, tile_a[r][quad + 2] = a.z, tile_a[r][quad + 3] = a.w;
            tile_b[r][quad] = b.x, tile_b[r][quad + 1] = b.y, tile_b[r][quad + 2] = b.z, tile_b[r][quad + 3] = b.w;
        }
    } else {
        for (int r = warp; r < BT; r += BAL_THREADS / 32) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int c = lane + 32 * half;
                tile_a[r][c] = (i0 + r < n && j0 + c < n) ? p.ranks[(i0 + r) * n + j0 + c] : 0.0f;
                tile_b[r][c] = (j0 + r < n && i0 + c < n) ? p.ranks[(j0 + r) * n + i0 + c] : 0.0f;
            }
        }
    }
    __syncthreads();
    const uint32_t lower_lanes = lanemask_lt();
    const int sides = tile_i == tile_j ? 1 : 2;
    // whole tiles off the diagonal (all but a few hundred of the 31 375 of a 16 000-cell pile) need neither the test
    // against the matrix's edge nor the one against the diagonal: the loop is compiled both ways
    const bool careful_tile = tile_i == tile_j || i0 + BT > n || j0 + BT > n;
    auto rows_of_side = [&](int side, auto careful_tag) {
        constexpr bool CAREFUL = decltype(careful_tag)::value;
        // side 0: rows of I, columns of J; side 1: rows of J, columns of I
        const int64_t row0 = side == 0 ? i0 : j0, col0 = side == 0 ? j0 : i0;
        const int column_tile = side == 0 ? tile_j : tile_i;
        const float(*mine_tile)[BT + 1] = side == 0 ? tile_a : tile_b;
        const float(*other_tile)[BT + 1] = side == 0 ? tile_b : tile_a;
        for (int r = warp; r < BT; r += BAL_THREADS / 32) {
            const int64_t row = row0 + r;
            if (CAREFUL && row >= n)
                break;
            // The threshold on the balanced rank (knn_graph.py:333-334) is a threshold on the float32 product: every step
            // from the product to the comparison is monotone, and the host found the largest product that passes with the
            // same float32 steps.  Only the entries that pass (a few percent) pay for the square root.
            float product[2];
            bool valid[2], passes[2], kept[2];
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int c = lane + 32 * half;
                const int64_t column = col0 + c;
                product[half] = __fmul_rn(mine_tile[r][c], other_tile[c][r]);
                valid[half] = !CAREFUL || column < n;
                passes[half] = valid[half] && product[half] <= p.keep_product;
                kept[half] = passes[half] && (!CAREFUL || column != row);
            }
            if (!FILL) {
                // the row's largest t can only be an entry that passes, when there is one (rows without are redone
                // by balance_best_kernel)
                uint32_t best = 0;
                float t[2] = {0.0f, 0.0f};  // the square root is taken once, for the row's best entry and for what is parked
#pragma unroll
                for (int half = 0; half < 2; ++half)
                    if (passes[half]) {
                        t[half] = balanced_t_of_product(product[half]);
                        best = max(best, pack_tile_best(t[half], lane + 32 * half));
                    }
                best = __reduce_max_sync(0xffffffffu, best);
                const uint32_t mask0 = __ballot_sync(0xffffffffu, kept[0]), mask1 = __ballot_sync(0xffffffffu, kept[1]);
                const int count = __popc(mask0) + __popc(mask1);
                if (p.staged != nullptr) {
                    float* slots = p.staged + (row * p.tiles + column_tile) * BT;
#pragma unroll
                    for (int half = 0; half < 2; ++half)
                        if (kept[half])
                            slots[(half ? __popc(mask0) : 0) + __popc((half ? mask1 : mask0) & lower_lanes)] =
                                __fadd_rn(t[half], p.shift);
                    if (lane == 0) p.masks[row * p.tiles + column_tile] = (unsigned long long)mask0 | ((unsigned long long)mask1 << 32);
                }
                if (lane == 0) {
                    p.counts[row * p.tiles + column_tile] = count;
                    if (best != 0) {
                        const float t = __fadd_rn(__fmul_rn((float)(best >> 6), 0.125f), 1048576.0f);
                        const uint32_t column = (uint32_t)(col0 + (BT - 1 - (int)(best & (BT - 1))));
                        atomicMax(&p.row_best[row], ((unsigned long long)float_to_ordered(t) << 32) | (uint32_t)~column);
                    }
                }
            } else {
                const int64_t preserved = p.best_index[row];
#pragma unroll
                for (int half = 0; half < 2; ++half) kept[half] = kept[half] || (valid[half] && col0 + lane + 32 * half == preserved);
                const uint32_t mask0 = __ballot_sync(0xffffffffu, kept[0]), mask1 = __ballot_sync(0xffffffffu, kept[1]);
                const int64_t base = (int64_t)p.indptr[row] + p.counts[row * p.tiles + column_tile];
#pragma unroll
                for (int half = 0; half < 2; ++half)
                    if (kept[half]) {
                        const int64_t column = col0 + lane + 32 * half;
                        float value = __fadd_rn(balanced_t_of_product(product[half]), p.shift);
                        if (column == preserved)
                            value = fmaxf(value, 1.0f);  // knn_graph.py:331
                        const int64_t at = base + (half ? __popc(mask0) : 0) + __popc((half ? mask1 : mask0) & lower_lanes);
                        p.out_data[at] = value;
                        p.out_indices[at] = (int32_t)column;
                    }
            }
        }
    };
    for (int side = 0; side < sides; ++side) {
        if (careful_tile)
            rows_of_side(side, std::true_type());
        else
            rows_of_side(side, std::false_type());
    }
}

// The fill pass without the ranks: one warp per row, one lane per column tile; the entries the count pass parked go to
// their places in column order, the preserved entry raised to 1 where it is among them and inserted where the threshold
// had dropped it (balance_rows_kernel counted it in).
__global__ void __launch_bounds__(256) balance_scatter_kernel(BalanceParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= p.n)
        return;
    const int64_t preserved = p.best_index[row];
    const int preserved_tile = (int)(preserved / BT), preserved_bit = (int)(preserved % BT);
    const float preserved_value = p.best_value[row];
    const int64_t first = p.indptr[row];
    for (int base = 0; base < p.tiles; base += 32) {
        const int tile = base + lane;
        if (tile >= p.tiles)
            continue;
        unsigned long long mask = p.masks[row * p.tiles + tile];
        const float* slots = p.staged + (row * p.tiles + tile) * BT;
        int64_t at = first + p.counts[row * p.tiles + tile];
        const int64_t column0 = (int64_t)tile * BT;
        const bool holds_preserved = tile == preserved_tile;
        bool pending = holds_preserved && !(mask >> preserved_bit & 1ull);  // the threshold dropped the preserved entry
        int slot = 0;
        while (mask != 0) {
            const int bit = __ffsll((long long)mask) - 1;
            mask &= mask - 1;
            if (pending && bit > preserved_bit) {
                p.out_data[at] = preserved_value;
                p.out_indices[at] = (int32_t)preserved;
                ++at;
                pending = false;
            }
            float value = slots[slot++];
            if (holds_preserved && bit == preserved_bit)
                value = fmaxf(value, 1.0f);  // knn_graph.py:331
            p.out_data[at] = value;
            p.out_indices[at] = (int32_t)(column0 + bit);
            ++at;
        }
        if (pending) {
            p.out_data[at] = preserved_value;
            p.out_indices[at] = (int32_t)preserved;
        }
    }
}

// Rows none of whose entries pass the threshold (a k that is tiny against the pile): the largest t over the whole row.
__global__ void __launch_bounds__(256) balance_best_kernel(BalanceParams p) {
    __shared__ unsigned long long warp_best[8];
    const int64_t row = blockIdx.x;
    if (p.row_best[row] != 0)
        return;
    unsigned long long best = 0;
    for (int64_t column = threadIdx.x; column < p.n; column += blockDim.x) {
        const float t = balanced_t(p.ranks[row * p.n + column], p.ranks[column * p.n + row]);
        best = max(best, ((unsigned long long)float_to_ordered(t) << 32) | (uint32_t)~(uint32_t)column);
    }
    for (int offset = 16; offset > 0; offset >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, offset));
    if ((threadIdx.x & 31) == 0) warp_best[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) best = max(best, warp_best[w]);
        p.row_best[row] = best;
    }
}

// Pass 2: one warp per row.  Decodes the argmax, adds the preserved entry to its tile's count when the
// threshold dropped it, turns the counts into exclusive offsets and writes the row's size.
__global__ void __launch_bounds__(256) balance_rows_kernel(BalanceParams p, int32_t* __restrict__ row_sizes,
                                                          int32_t* __restrict__ best_index, float* __restrict__ best_value,
                                                          float* __restrict__ best_raw, int32_t* __restrict__ widest,
                                                          int32_t* __restrict__ class_rows, int* __restrict__ class_counts) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= p.n)
        return;
    const unsigned long long best = p.row_best[row];
    const float t = ordered_to_float((uint32_t)(best >> 32));
    const int64_t column = (int64_t)(uint32_t)(~(uint32_t)best);
    const float value = __fadd_rn(t, p.shift);
    const bool counted = column != row && value > 0.0f;
    int32_t* counts = p.counts + row * p.tiles;
    int running = 0;
    for (int base = 0; base < p.tiles; base += 32) {
        const int tile = base + lane;
        int count = tile < p.tiles ? counts[tile] : 0;
        if (!counted && tile == (int)(column / BT))
            count += 1;
        const int inclusive = warp_inclusive_scan(count, lane);
        if (tile < p.tiles)
            counts[tile] = running + inclusive - count;
        running += __shfl_sync(0xffffffffu, inclusive, 31);
    }
    if (lane == 0) {
        row_sizes[row] = running;
        best_index[row] = (int32_t)column;
        best_value[row] = fmaxf(value, 1.0f);
        best_raw[row] = value;
        atomicMax(widest, running);
        // the row joins the list of its band-width class: the pruning kernels walk lists without gaps, in any order
        const int band_class = running <= 32 * 32 ? 0 : (running <= 32 * MP_MAX_ROUNDS ? 1 : 2);
        class_rows[band_class * p.n + atomicAdd(&class_counts[band_class], 1)] = (int32_t)row;
    }
}

// ---- merge_preserved ---------------------------------------------------------------------------------

__global__ void reverse_count_kernel(const int32_t* __restrict__ best_index, int64_t n, int32_t* __restrict__ counts) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row < n)
        atomicAdd(&counts[best_index[row]], 1);
}

__global__ void reverse_scatter_kernel(const int32_t* __restrict__ best_index, int64_t n, int32_t* __restrict__ cursor,
                                       int32_t* __restrict__ reverse_list) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row < n)
        reverse_list[atomicAdd(&cursor[best_index[row]], 1)] = (int32_t)row;
}

// Position of `column` in the sorted indices[0..n), or -1.
__device__ __forceinline__ int find_column(const int32_t* __restrict__ indices, int n, int32_t column) {
    int left = 0, right = n;
    while (left < right) {
        const int middle = (left + right) >> 1;
        if (indices[middle] < column)
            left = middle + 1;
        else
            right = middle;
    }
    return (left < n && indices[left] == column) ? left : -1;
}

struct MergeParams {
    const float* data;  // pruned CSR (sorted indices)
    const int32_t* indices;
    const int32_t* indptr;
    int64_t n;
    const int32_t* best_index;  // preserved column of every row
    const float* best_value;    // preserved value of every row
    const int32_t* reverse_ptr;   // rows whose preserved column is this row
    const int32_t* reverse_list;
    int32_t* out_sizes;           // count pass
    const int32_t* out_indptr;    // fill pass
    float* out_data;
    int32_t* out_indices;
};

// One warp per row.  FILL = false: size of the merged row.  FILL = true: existing entries first (values
// raised where a preserved edge lands on them), new entries appended; the caller sorts rows afterwards.
template <bool FILL>
__global__ void __launch_bounds__(256) merge_preserved_kernel(MergeParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= p.n)
        return;
    const int lo = p.indptr[row], n = p.indptr[row + 1] - lo;
    const int32_t* indices = p.indices + lo;
    const int32_t own_column = p.best_index[row];
    const int own_at = find_column(indices, n, own_column);
    const int own_new = own_at < 0 ? 1 : 0;
    const int reverse_lo = p.reverse_ptr[row], reverse_n = p.reverse_ptr[row + 1] - reverse_lo;
    float* out_data = nullptr;
    int32_t* out_indices = nullptr;
    if (FILL) {
        out_data = p.out_data + p.out_indptr[row];
        out_indices = p.out_indices + p.out_indptr[row];
        for (int e = lane; e < n; e += 32) {
            out_data[e] = p.data[lo + e];
            out_indices[e] = indices[e];
        }
        __syncwarp();
        if (lane == 0) {
            const float value = p.best_value[row];
            if (own_at >= 0) {
                out_data[own_at] = fmaxf(out_data[own_at], value);
            } else {
                out_data[n] = value;
                out_indices[n] = own_column;
            }
        }
        __syncwarp();
    }
    int appended = 0;
    for (int base = 0; base < reverse_n; base += 32) {
        const int e = base + lane;
        int32_t column = -1;
        int at = -1;
        bool is_new = false;
        if (e < reverse_n) {
            column = p.reverse_list[reverse_lo + e];
            if (column == own_column)
                at = own_at >= 0 ? own_at : n;  // lands on this row's own preserved edge
            else
                at = find_column(indices, n, column);
            is_new = at < 0;
        }
        const uint32_t new_mask = __ballot_sync(0xffffffffu, is_new);
        if (FILL && e < reverse_n) {
            const float value = p.best_value[column];
            if (is_new) {
                const int slot = n + own_new + appended + __popc(new_mask & lanemask_lt());
                out_data[slot] = value;
                out_indices[slot] = column;
            } else {
                // positive floats order like their bit patterns
                atomicMax(reinterpret_cast<int*>(&out_data[at]), __float_as_int(value));
            }
        }
        appended += __popc(new_mask);
    }
    if (!FILL && lane == 0)
        p.out_sizes[row] = n + own_new + appended;
}

// ---- weigh_edges ---------------------------------------------------------------------------------------

// float32 row sums as scipy's csr sum(axis=1) makes them (np.add.reduceat: numpy's pairwise order, numpy_sum.cuh); scale =
// 1 / sum.  A sum that is exact in float32 is the same in every order, and the balanced ranks of a pile's graph (multiples
// of 1/8, totals below 2^21) are: the left-to-right sum is kept whenever the float64 sum says it is exact, numpy's order is
// followed where it is not (graphs with k beyond ~1000: the grouping piles of a small atlas, k = n - 1 = 1864 -- there the
// left-to-right sum is off by up to 4e-6 and the optimiser's "outgoing weights sum to 1" assertion fails).
__global__ void row_scale_kernel(const float* __restrict__ data, const int32_t* __restrict__ indptr, int64_t n,
                                 float* __restrict__ scale, int32_t* __restrict__ min_size) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int size = 0x7fffffff;
    if (row < n) {
        const int lo = indptr[row], hi = indptr[row + 1];
        float sum = 0.0f;
        double wide = 0.0;
        for (int e = lo; e < hi; ++e) {
            const float value = data[e];
            sum = __fadd_rn(sum, value);
            wide = __dadd_rn(wide, (double)value);
        }
        if ((double)sum != wide && hi > lo)
            sum = numpy_add_reduce_f32(data + lo, hi - lo);
        scale[row] = __frcp_rn(sum);  // np.reciprocal
        size = hi - lo;
    }
    for (int offset = 16; offset > 0; offset >>= 1) size = min(size, __shfl_xor_sync(0xffffffffu, size, offset));
    if ((threadIdx.x & 31) == 0 && min_size)
        atomicMin(min_size, size);
}

__global__ void __launch_bounds__(256) scale_rows_kernel(const float* __restrict__ data, const int32_t* __restrict__ indptr,
                                                        int64_t n, const float* __restrict__ scale, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n)
        return;
    const float factor = scale[row];
    for (int e = indptr[row] + lane; e < indptr[row + 1]; e += 32) out[e] = __fmul_rn(data[e], factor);
}


}  // namespace

int rank_matrix_launch(float* matrix, int64_t rows, int64_t columns, int64_t row_stride, int ascending, cudaStream_t stream);
int collect_pruned_launch(int64_t degree, const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands,
                          float* out_data, int32_t* out_indices, int32_t* out_indptr, cudaStream_t stream);
int relayout_launch(const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands_in, int64_t bands_out,
                    int64_t nnz, float* out_data, int32_t* out_indices, int32_t* out_indptr, cudaStream_t stream);

static int read_i32(const int32_t* device, int32_t* host, cudaStream_t stream) {
    return copy_to_host(host, device, sizeof(int32_t), stream);  // waits for the stream
}

// knn_graph.py:285-297 (_rank_outgoing) on a private copy: diagonal below the minimum, descending ranks.
// known_min (optional, device): the minimum of `work` as a float_to_ordered key, when the producer of the
// similarity already has it (the correlate epilogue of the pile path); otherwise one more pass over the matrix.
int rank_outgoing_launch(float* work, int64_t n, cudaStream_t stream, const uint32_t* known_min) {
    if (n == 0)
        return MCB_OK;
    Scratch min_key;
    if (known_min == nullptr) {
        MCB_TRY(min_key.alloc(sizeof(uint32_t), stream));
        MCB_CUDA(cudaMemsetAsync(min_key.ptr, 0xff, sizeof(uint32_t), stream));
        MCB_PROFILED("min_matrix", stream, min_matrix_kernel<<<(unsigned)std::min<int64_t>(ceil_div(n * n, 1024), (int64_t)sm_count() * 8), 256, 0, stream>>>(
            work, n, n, n, min_key.as<uint32_t>()));
        MCB_LAUNCH_CHECK();
        known_min = min_key.as<uint32_t>();
    }
    fill_diagonal_below_min_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, stream>>>(work, n, n, known_min);
    MCB_LAUNCH_CHECK();
    return rank_matrix_launch(work, n, n, n, 0, stream);
}

// The largest float32 product whose balanced rank passes `2^21 - sqrt(product) + shift > 0`, through the same float32
// steps as the kernels (and numpy): the steps are monotone, so a bisection over the bit patterns finds it.  Products are
// at least 1 (ranks start at 1); 0 means that nothing passes.
static float largest_kept_product(float shift) {
    auto passes = [shift](uint32_t bits) {
        float product;
        memcpy(&product, &bits, sizeof(product));
        volatile float root = sqrtf(product);
        volatile float t = OFFSET_2_21 - root;
        volatile float value = t + shift;
        return value > 0.0f;
    };
    uint32_t low = 0x3f800000u, high = 0x7f000000u;  // 1.0f .. 2^127
    if (!passes(low))
        return 0.0f;
    if (passes(high))
        low = high;
    while (low + 1 < high) {
        const uint32_t middle = low + (high - low) / 2;
        if (passes(middle))
            low = middle;
        else
            high = middle;
    }
    float product;
    memcpy(&product, &low, sizeof(product));
    return product;
}

int balance_ranks_launch(const float* ranks, int64_t n, float shift, BalancedRanks* out, cudaStream_t stream) {
    const int tiles = (int)ceil_div(n, BT);
    const int64_t blocks = (int64_t)tiles * (tiles + 1) / 2;
    MCB_REQUIRE(blocks < (1ll << 31), "balance_ranks: too many tiles");
    Scratch counts, row_best, row_sizes;
    MCB_TRY(counts.alloc((size_t)n * tiles * sizeof(int32_t), stream));
    MCB_TRY(row_best.alloc((size_t)n * sizeof(unsigned long long), stream));
    MCB_TRY(row_sizes.alloc((size_t)n * sizeof(int32_t), stream));
    MCB_TRY(out->indptr.alloc((size_t)(n + 1) * sizeof(int32_t), stream));
    MCB_TRY(out->best_index.alloc((size_t)n * sizeof(int32_t), stream));
    MCB_TRY(out->best_value.alloc((size_t)n * sizeof(float), stream));
    MCB_TRY(out->best_raw.alloc((size_t)n * sizeof(float), stream));
    Scratch widest;
    MCB_TRY(widest.alloc(sizeof(int32_t), stream));
    MCB_CUDA(cudaMemsetAsync(widest.ptr, 0, sizeof(int32_t), stream));
    MCB_CUDA(cudaMemsetAsync(row_best.ptr, 0, (size_t)n * sizeof(unsigned long long), stream));
    BalanceParams p{};
    p.ranks = ranks;
    p.n = n;
    p.tiles = tiles;
    p.shift = shift;
    p.keep_product = largest_kept_product(shift);
    MCB_REQUIRE(n < (1 << 20), "balance_ranks: %lld rows are more than the packed row maximum holds", (long long)n);
    p.counts = counts.as<int32_t>();
    p.row_best = row_best.as<unsigned long long>();
    Scratch masks, staged;
    const size_t staged_bytes = (size_t)n * tiles * BT * sizeof(float);
    if (staged_bytes <= ((size_t)8 << 30) && getenv("MCB_BALANCE_TWO_PASSES") == nullptr) {
        MCB_TRY(masks.alloc((size_t)n * tiles * sizeof(unsigned long long), stream));
        MCB_TRY(staged.alloc(staged_bytes, stream));
        p.masks = masks.as<unsigned long long>();
        p.staged = staged.as<float>();
    }
    MCB_PROFILED("balance_count", stream, balance_kernel<false><<<(unsigned)blocks, BAL_THREADS, 0, stream>>>(p));
    MCB_LAUNCH_CHECK();
    MCB_PROFILED("balance_best", stream, balance_best_kernel<<<(unsigned)n, 256, 0, stream>>>(p));
    MCB_LAUNCH_CHECK();
    MCB_TRY(out->class_rows.alloc((size_t)3 * n * sizeof(int32_t), stream));
    MCB_TRY(out->class_counts.alloc(4 * sizeof(int), stream));
    MCB_CUDA(cudaMemsetAsync(out->class_counts.ptr, 0, 4 * sizeof(int), stream));
    MCB_PROFILED("balance_rows", stream, balance_rows_kernel<<<(unsigned)ceil_div(n, 8), 256, 0, stream>>>(p, row_sizes.as<int32_t>(), out->best_index.as<int32_t>(),
                                                                     out->best_value.as<float>(), out->best_raw.as<float>(), widest.as<int32_t>(),
                                                                     out->class_rows.as<int32_t>(), out->class_counts.as<int>()));
    MCB_LAUNCH_CHECK();
    MCB_TRY(exclusive_scan_i32(row_sizes.as<int32_t>(), out->indptr.as<int32_t>(), n, nullptr, stream));
    int32_t nnz = 0, widest_host = 0;
    MCB_TRY(read_i32(out->indptr.as<int32_t>() + n, &nnz, stream));
    MCB_TRY(read_i32(widest.as<int32_t>(), &widest_host, stream));
    out->nnz = nnz;
    out->widest = widest_host;
    MCB_TRY(out->data.alloc((size_t)nnz * sizeof(float), stream));
    MCB_TRY(out->indices.alloc((size_t)nnz * sizeof(int32_t), stream));
    p.indptr = out->indptr.as<int32_t>();
    p.out_data = out->data.as<float>();
    p.out_indices = out->indices.as<int32_t>();
    p.best_index = out->best_index.as<int32_t>();
    p.best_value = out->best_value.as<float>();
    if (p.staged != nullptr)
        MCB_PROFILED("balance_fill", stream, balance_scatter_kernel<<<(unsigned)ceil_div(n, 8), 256, 0, stream>>>(p));
    else
        MCB_PROFILED("balance_fill", stream, balance_kernel<true><<<(unsigned)blocks, BAL_THREADS, 0, stream>>>(p));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

// knn_graph.py:343-392 (_prune_ranks) after the balanced CSR exists.
int prune_ranks_launch(const BalancedRanks& balanced, int64_t n, int64_t incoming_degree, int64_t outgoing_degree,
                       PrunedRanks* out, cudaStream_t stream) {
    // (4) knn_graph.py:383-388: maximum with the preserved edges and their transpose
    Scratch reverse_counts, reverse_ptr, reverse_cursor, reverse_list, sizes;
    MCB_TRY(reverse_counts.alloc((size_t)n * sizeof(int32_t), stream));
    MCB_TRY(reverse_ptr.alloc((size_t)(n + 1) * sizeof(int32_t), stream));
    MCB_TRY(reverse_cursor.alloc((size_t)n * sizeof(int32_t), stream));
    MCB_TRY(reverse_list.alloc((size_t)n * sizeof(int32_t), stream));
    MCB_TRY(sizes.alloc((size_t)n * sizeof(int32_t), stream));
    MCB_CUDA(cudaMemsetAsync(reverse_counts.ptr, 0, (size_t)n * sizeof(int32_t), stream));
    const unsigned row_blocks = (unsigned)ceil_div(n, 256), warp_blocks = (unsigned)ceil_div(n, 8);
    reverse_count_kernel<<<row_blocks, 256, 0, stream>>>(balanced.best_index.as<int32_t>(), n, reverse_counts.as<int32_t>());
    MCB_LAUNCH_CHECK();
    MCB_TRY(exclusive_scan_i32(reverse_counts.as<int32_t>(), reverse_ptr.as<int32_t>(), n, nullptr, stream));
    MCB_CUDA(cudaMemcpyAsync(reverse_cursor.ptr, reverse_ptr.ptr, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToDevice, stream));
    reverse_scatter_kernel<<<row_blocks, 256, 0, stream>>>(balanced.best_index.as<int32_t>(), n, reverse_cursor.as<int32_t>(),
                                                          reverse_list.as<int32_t>());
    MCB_LAUNCH_CHECK();
    Scratch out_data, out_indices, out_indptr;
    // (1)-(3) without the transposition (prune_mirror.cuh) when every band fits a warp's registers
    bool mirrored = outgoing_degree >= 1 && getenv("MCB_PRUNE_RELAYOUT") == nullptr;
    if (mirrored) {
        Scratch cuts, padded_data, padded_indices, kept_sizes;
        MCB_TRY(cuts.alloc((size_t)n * sizeof(RowCut), stream));
        MCB_TRY(padded_data.alloc((size_t)n * outgoing_degree * sizeof(float), stream));
        MCB_TRY(padded_indices.alloc((size_t)n * outgoing_degree * sizeof(int32_t), stream));
        MCB_TRY(kept_sizes.alloc((size_t)n * sizeof(int32_t), stream));

        MirrorParams mp{};
        mp.data = balanced.data.as<float>();
        mp.indices = balanced.indices.as<int32_t>();
        mp.indptr = balanced.indptr.as<int32_t>();
        mp.n = n;
        mp.incoming_degree = (int)incoming_degree;
        mp.outgoing_degree = (int)outgoing_degree;
        mp.best_index = balanced.best_index.as<int32_t>();
        mp.best_value = balanced.best_value.as<float>();
        mp.best_raw = balanced.best_raw.as<float>();
        mp.reverse_ptr = reverse_ptr.as<int32_t>();
        mp.reverse_list = reverse_list.as<int32_t>();
        mp.cuts = cuts.as<RowCut>();
        mp.padded_data = padded_data.as<float>();
        mp.padded_indices = padded_indices.as<int32_t>();
        mp.sizes = kept_sizes.as<int32_t>();
        mp.overflow = balanced.class_counts.as<int>() + 3;
        mp.class_rows = balanced.class_rows.as<int32_t>();
        mp.class_counts = balanced.class_counts.as<int>();
        const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(n * 32, 256), (int64_t)sm_count() * 8);
        const unsigned narrow_grid = (unsigned)std::min<int64_t>(ceil_div(n * 32, 128), (int64_t)sm_count() * 16);
        // The few bands wider than 2048 entries (the hub cells of a pile) take one warp each for a long time: in both
        // steps they start first, on the thread's second stream, and the two classes that hold the rows run beside them.
        cudaStream_t side = nullptr;
        cudaEvent_t fork = nullptr, join = nullptr;
        const bool hubs = balanced.widest > 32 * MP_MAX_ROUNDS;
        if (hubs) {
            MCB_TRY(thread_side_stream(&side, &fork, &join));
            MCB_CUDA(cudaEventRecord(fork, stream));
            MCB_CUDA(cudaStreamWaitEvent(side, fork, 0));
            MCB_PROFILED("row_cuts_widest", side, row_cuts_kernel<0><<<(unsigned)std::min<int64_t>(grid, sm_count()), 256, 0, side>>>(mp));
            MCB_LAUNCH_CHECK();
            MCB_CUDA(cudaEventRecord(join, side));
        }
        MCB_PROFILED("row_cuts", stream, row_cuts_kernel<32><<<grid, 256, 0, stream>>>(mp));
        MCB_LAUNCH_CHECK();
        if (balanced.widest > 32 * 32) {
            MCB_PROFILED("row_cuts_wide", stream, row_cuts_kernel<MP_MAX_ROUNDS><<<grid, 256, 0, stream>>>(mp));
            MCB_LAUNCH_CHECK();
        }
        if (hubs) {  // every cut is there before any row is pruned
            MCB_CUDA(cudaStreamWaitEvent(stream, join, 0));
            MCB_CUDA(cudaEventRecord(fork, stream));
            MCB_CUDA(cudaStreamWaitEvent(side, fork, 0));
            const size_t cache_bytes = (size_t)4 * MP_CACHE * sizeof(uint32_t);
            MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(mirror_prune_kernel<0>), cache_bytes));
            const unsigned hub_grid = (unsigned)std::min<int64_t>(narrow_grid, sm_count());
            MCB_PROFILED("mirror_prune_widest", side, mirror_prune_kernel<0><<<hub_grid, 128, cache_bytes, side>>>(mp));
            MCB_LAUNCH_CHECK();
            MCB_CUDA(cudaEventRecord(join, side));
        }
        MCB_PROFILED("mirror_prune", stream, mirror_prune_kernel<32><<<narrow_grid, 128, 0, stream>>>(mp));
        MCB_LAUNCH_CHECK();
        if (balanced.widest > 32 * 32) {
            MCB_PROFILED("mirror_prune_wide", stream, mirror_prune_kernel<MP_MAX_ROUNDS><<<narrow_grid, 128, 0, stream>>>(mp));
            MCB_LAUNCH_CHECK();
        }
        if (hubs) MCB_CUDA(cudaStreamWaitEvent(stream, join, 0));
        MCB_TRY(out_indptr.alloc((size_t)(n + 1) * sizeof(int32_t), stream));
        MCB_TRY(exclusive_scan_i32(kept_sizes.as<int32_t>(), out_indptr.as<int32_t>(), n, nullptr, stream));
        int32_t kept_nnz = 0, overflowed = 0;
        MCB_TRY(read_i32(out_indptr.as<int32_t>() + n, &kept_nnz, stream));
        MCB_TRY(read_i32(balanced.class_counts.as<int32_t>() + 3, &overflowed, stream));
        if (overflowed) {
            mirrored = false;  // a row with more than 32 foreign preserved entries: the general path
        } else {
            MCB_TRY(out_data.alloc((size_t)kept_nnz * sizeof(float), stream));
            MCB_TRY(out_indices.alloc((size_t)kept_nnz * sizeof(int32_t), stream));
            MCB_PROFILED("compact_rows", stream, compact_rows_kernel<<<warp_blocks, 256, 0, stream>>>(
                             padded_data.as<float>(), padded_indices.as<int32_t>(), n, (int)outgoing_degree, out_indptr.as<int32_t>(),
                             out_data.as<float>(), out_indices.as<int32_t>()));
            MCB_LAUNCH_CHECK();
        }
    }
    if (!mirrored) {
        Scratch fresh_data, fresh_indices, fresh_indptr;
        std::swap(out_data.ptr, fresh_data.ptr), std::swap(out_data.stream, fresh_data.stream);
        std::swap(out_indices.ptr, fresh_indices.ptr), std::swap(out_indices.stream, fresh_indices.stream);
        std::swap(out_indptr.ptr, fresh_indptr.ptr), std::swap(out_indptr.stream, fresh_indptr.stream);
        // (1) knn_graph.py:371-374: the CSR arrays reinterpreted as CSC; keep the best `incoming_degree` per column
        Scratch in_data, in_indices, in_indptr;
        MCB_TRY(in_data.alloc((size_t)balanced.nnz * sizeof(float), stream));
        MCB_TRY(in_indices.alloc((size_t)balanced.nnz * sizeof(int32_t), stream));
        MCB_TRY(in_indptr.alloc((size_t)(n + 1) * sizeof(int32_t), stream));
        MCB_TRY(collect_pruned_launch(incoming_degree, balanced.data.as<float>(), balanced.indices.as<int32_t>(),
                                      balanced.indptr.as<int32_t>(), n, in_data.as<float>(), in_indices.as<int32_t>(),
                                      in_indptr.as<int32_t>(), stream));
        int32_t in_nnz = 0;
        MCB_TRY(read_i32(in_indptr.as<int32_t>() + n, &in_nnz, stream));
        // (2) knn_graph.py:377: CSC -> CSR
        Scratch row_data, row_indices, row_indptr;
        MCB_TRY(row_data.alloc((size_t)in_nnz * sizeof(float), stream));
        MCB_TRY(row_indices.alloc((size_t)in_nnz * sizeof(int32_t), stream));
        MCB_TRY(row_indptr.alloc((size_t)(n + 1) * sizeof(int32_t), stream));
        trace_lap("prune: incoming", stream);
        MCB_TRY(relayout_launch(in_data.as<float>(), in_indices.as<int32_t>(), in_indptr.as<int32_t>(), n, n, in_nnz,
                                row_data.as<float>(), row_indices.as<int32_t>(), row_indptr.as<int32_t>(), stream));
        trace_lap("prune: relayout", stream);
        // (3) knn_graph.py:380: keep the best `outgoing_degree` per row
        MCB_TRY(out_data.alloc((size_t)in_nnz * sizeof(float), stream));
        MCB_TRY(out_indices.alloc((size_t)in_nnz * sizeof(int32_t), stream));
        MCB_TRY(out_indptr.alloc((size_t)(n + 1) * sizeof(int32_t), stream));
        MCB_TRY(collect_pruned_launch(outgoing_degree, row_data.as<float>(), row_indices.as<int32_t>(), row_indptr.as<int32_t>(), n,
                                      out_data.as<float>(), out_indices.as<int32_t>(), out_indptr.as<int32_t>(), stream));
        trace_lap("prune: outgoing", stream);
    }
    MergeParams m{};
    m.data = out_data.as<float>();
    m.indices = out_indices.as<int32_t>();
    m.indptr = out_indptr.as<int32_t>();
    m.n = n;
    m.best_index = balanced.best_index.as<int32_t>();
    m.best_value = balanced.best_value.as<float>();
    m.reverse_ptr = reverse_ptr.as<int32_t>();
    m.reverse_list = reverse_list.as<int32_t>();
    m.out_sizes = sizes.as<int32_t>();
    MCB_PROFILED("merge_count", stream, merge_preserved_kernel<false><<<warp_blocks, 256, 0, stream>>>(m));
    MCB_LAUNCH_CHECK();
    MCB_TRY(out->indptr.alloc((size_t)(n + 1) * sizeof(int32_t), stream));
    MCB_TRY(exclusive_scan_i32(sizes.as<int32_t>(), out->indptr.as<int32_t>(), n, nullptr, stream));
    int32_t nnz = 0;
    MCB_TRY(read_i32(out->indptr.as<int32_t>() + n, &nnz, stream));
    out->nnz = nnz;
    MCB_TRY(out->data.alloc((size_t)nnz * sizeof(float), stream));
    MCB_TRY(out->indices.alloc((size_t)nnz * sizeof(int32_t), stream));
    m.out_indptr = out->indptr.as<int32_t>();
    m.out_data = out->data.as<float>();
    m.out_indices = out->indices.as<int32_t>();
    MCB_PROFILED("merge_fill", stream, merge_preserved_kernel<true><<<warp_blocks, 256, 0, stream>>>(m));
    MCB_LAUNCH_CHECK();
    // a row's own columns and the appended preserved edges are distinct
    return sort_distinct_indices_launch(out->data.as<float>(), out->indices.as<int32_t>(), out->indptr.as<int32_t>(), n, -1, n,
                                        stream);
}

// knn_graph.py:395-407 (_weigh_edges); also reports the smallest row size (knn_graph.py:262-263).
int weigh_edges_launch(const float* data, const int32_t* indptr, int64_t n, float* weights, int32_t* min_degree_host,
                       cudaStream_t stream) {
    Scratch scale, min_size;
    MCB_TRY(scale.alloc((size_t)n * sizeof(float), stream));
    MCB_TRY(min_size.alloc(sizeof(int32_t), stream));
    MCB_CUDA(cudaMemsetAsync(min_size.ptr, 0x7f, sizeof(int32_t), stream));
    MCB_PROFILED("row_scale", stream, row_scale_kernel<<<(unsigned)ceil_div(n, 128), 128, 0, stream>>>(data, indptr, n, scale.as<float>(), min_size.as<int32_t>()));
    MCB_LAUNCH_CHECK();
    if (weights) {
        MCB_PROFILED("scale_rows", stream, scale_rows_kernel<<<(unsigned)ceil_div(n, 8), 256, 0, stream>>>(data, indptr, n, scale.as<float>(), weights));
        MCB_LAUNCH_CHECK();
    }
    if (min_degree_host)
        MCB_TRY(read_i32(min_size.as<int32_t>(), min_degree_host, stream));
    return MCB_OK;
}

}  // namespace mcb
