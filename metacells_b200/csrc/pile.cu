// Device-resident pile: the part of compute_direct_metacells (pipeline/direct.py:26-290) between the
// UMI matrix and the k-NN graph, kept in HBM from the H2D copy of the CSR slice to the D2H copy of
// the graph.
//
//   column_totals       per-gene totals of the downsampled CSR (the stand-in statistic the feature
//                       gene selection of pipeline/select.py starts from); counts are integers far
//                       below 2^24, so float32 atomics are exact and order independent.
//   log_fraction_dense  direct.py:242-249: dense slice of the feature genes, + 1/7, log2.  The
//                       inputs are small non-negative integers, so log2(c + 1/7) is a table made by
//                       calling numpy's own np.log2 on the host: bit-exact by construction.  One warp
//                       per cell; the dense row is first filled with table[0] and then the non-zeros
//                       are scattered, so HBM sees one coalesced write of the row and one read of the
//                       CSR row.
#include <algorithm>

#include "common.cuh"

namespace mcb {

namespace {

__global__ void column_totals_kernel(const float* __restrict__ data, const int32_t* __restrict__ indices, int64_t nnz,
                                     float* __restrict__ totals) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x) {
        const float value = data[e];
        if (value != 0.0f)
            atomicAdd(&totals[indices[e]], value);
    }
}

__global__ void __launch_bounds__(256) log_fraction_dense_kernel(const float* __restrict__ data, const int32_t* __restrict__ indices,
                                                                const int32_t* __restrict__ indptr, int64_t rows,
                                                                const int32_t* __restrict__ slot_of_gene, int64_t features,
                                                                const float* __restrict__ table, int table_size,
                                                                float regularization, float* __restrict__ dense,
                                                                int* __restrict__ error_flag) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows)
        return;
    float* out = dense + row * features;
    const float zero_value = table[0];
    for (int64_t c = lane; c < features; c += 32) out[c] = zero_value;
    __syncwarp();
    // four independent chains per lane: the gene's slot and the count are loaded for all of them before any is used
    constexpr int UNROLL = 4;
    const int end = indptr[row + 1];
    for (int base = indptr[row] + lane; base < end; base += 32 * UNROLL) {
        int slot[UNROLL];
        float value[UNROLL];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const int e = base + 32 * j;
            slot[j] = e < end ? indices[e] : -1;
            value[j] = e < end ? data[e] : 0.0f;
        }
#pragma unroll
        for (int j = 0; j < UNROLL; ++j)
            if (slot[j] >= 0) slot[j] = slot_of_gene[slot[j]];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            if (slot[j] < 0)
                continue;
            const int count = (int)value[j];
            float result;
            if (value[j] >= 0.0f && (float)count == value[j] && count < table_size) {
                result = table[count];
            } else {
                result = log2f(__fadd_rn(value[j], regularization));  // not covered by the numpy-made table
                if (error_flag)
                    atomicExch(error_flag, 1);
            }
            out[slot[j]] = result;
        }
    }
}

}  // namespace

int column_totals_launch(const float* data, const int32_t* indices, int64_t nnz, int64_t columns, float* totals,
                         cudaStream_t stream) {
    MCB_CUDA(cudaMemsetAsync(totals, 0, (size_t)columns * sizeof(float), stream));
    if (nnz == 0)
        return MCB_OK;
    const int64_t grid = std::min<int64_t>(ceil_div(nnz, 256), (int64_t)sm_count() * 16);
    MCB_PROFILED("column_totals", stream, column_totals_kernel<<<(unsigned)grid, 256, 0, stream>>>(data, indices, nnz, totals));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

// dense[rows, features] = table[count] for the genes with slot_of_gene >= 0.  *outside_table (device
// int, may be null) is set when a value was not a non-negative integer below table_size.
int log_fraction_dense_launch(const float* data, const int32_t* indices, const int32_t* indptr, int64_t rows,
                              const int32_t* slot_of_gene, int64_t features, const float* table, int64_t table_size,
                              float regularization, float* dense, int* outside_table, cudaStream_t stream) {
    if (rows == 0 || features == 0)
        return MCB_OK;
    MCB_REQUIRE(table_size >= 1 && table_size < (1ll << 31), "log_fraction: the table needs at least the entry for 0");
    MCB_PROFILED("log_fraction", stream, log_fraction_dense_kernel<<<(unsigned)ceil_div(rows, 8), 256, 0, stream>>>(data, indices, indptr, rows, slot_of_gene, features,
                                                                               table, (int)table_size, regularization, dense,
                                                                               outside_table));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

}  // namespace mcb
