// C-ABI entry points for the Pearson correlation family (reference correlate.cpp:350-527).
#include "common.cuh"
#include "../../include/mcb200.h"

namespace mcb {
int prepare_rows_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, int64_t padded, float* hi,
                     float* lo, double* f64, cudaStream_t stream);
int prepare_rows_f64(const double* input, int64_t rows, int64_t columns, int64_t row_stride, int64_t padded, float* hi,
                     float* lo, double* f64, cudaStream_t stream);
int dot_rows_f64(const double* x, const double* y, int64_t rows_x, int64_t rows_y, int64_t padded, float* output,
                 int64_t out_stride, int symmetric, cudaStream_t stream);
int pairs_rows_f32(const float* first, const float* second, int64_t rows, int64_t columns, float* output, cudaStream_t stream);
int pairs_rows_f64(const double* first, const double* second, int64_t rows, int64_t columns, float* output, cudaStream_t stream);
int correlate_umma_launch(const float* a_hi, const float* a_lo, int64_t rows_a, const float* b_hi, const float* b_lo,
                          int64_t rows_b, int64_t padded, float* output, int64_t out_stride, int symmetric, int absolute,
                          unsigned int* min_key, cudaStream_t stream);
}  // namespace mcb

using namespace mcb;

namespace {

int64_t padded_width(int64_t columns) { return ceil_div(columns > 0 ? columns : 1, 32) * 32; }

// float32 rows -> tcgen05 3xTF32 contraction.
int correlate_f32_device(const float* first, int64_t first_rows, int64_t first_stride, const float* second,
                         int64_t second_rows, int64_t second_stride, int64_t columns, float* output, int64_t out_stride,
                         int symmetric, int absolute, cudaStream_t stream, unsigned int* min_key = nullptr) {
    const int64_t padded = padded_width(columns);
    Scratch a_hi, a_lo, b_hi, b_lo;
    MCB_TRY(a_hi.alloc((size_t)first_rows * padded * sizeof(float), stream));
    MCB_TRY(a_lo.alloc((size_t)first_rows * padded * sizeof(float), stream));
    MCB_TRY(prepare_rows_f32(first, first_rows, columns, first_stride, padded, a_hi.as<float>(), a_lo.as<float>(), nullptr, stream));
    trace_lap("correlate: prepare", stream);
    const float* bh = a_hi.as<float>();
    const float* bl = a_lo.as<float>();
    if (!symmetric) {
        MCB_TRY(b_hi.alloc((size_t)second_rows * padded * sizeof(float), stream));
        MCB_TRY(b_lo.alloc((size_t)second_rows * padded * sizeof(float), stream));
        MCB_TRY(prepare_rows_f32(second, second_rows, columns, second_stride, padded, b_hi.as<float>(), b_lo.as<float>(), nullptr, stream));
        bh = b_hi.as<float>();
        bl = b_lo.as<float>();
    }
    return correlate_umma_launch(a_hi.as<float>(), a_lo.as<float>(), first_rows, bh, bl, symmetric ? first_rows : second_rows,
                                 padded, output, out_stride, symmetric, absolute, min_key, stream);
}

// float64 rows -> float64 SIMT contraction (the reference's float64 instantiation is fp64 throughout).
int correlate_f64_device(const double* first, int64_t first_rows, int64_t first_stride, const double* second,
                         int64_t second_rows, int64_t second_stride, int64_t columns, float* output, int64_t out_stride,
                         int symmetric, cudaStream_t stream) {
    const int64_t padded = ceil_div(columns > 0 ? columns : 1, 16) * 16;
    Scratch a, b;
    MCB_TRY(a.alloc((size_t)first_rows * padded * sizeof(double), stream));
    MCB_TRY(prepare_rows_f64(first, first_rows, columns, first_stride, padded, nullptr, nullptr, a.as<double>(), stream));
    const double* y = a.as<double>();
    if (!symmetric) {
        MCB_TRY(b.alloc((size_t)second_rows * padded * sizeof(double), stream));
        MCB_TRY(prepare_rows_f64(second, second_rows, columns, second_stride, padded, nullptr, nullptr, b.as<double>(), stream));
        y = b.as<double>();
    }
    return dot_rows_f64(a.as<double>(), y, first_rows, symmetric ? first_rows : second_rows, padded, output, out_stride,
                        symmetric, stream);
}

// Upload a strided host matrix as a dense device matrix.
template <typename T>
int upload_matrix(Scratch& device, const T* host, int64_t rows, int64_t columns, int64_t row_stride, cudaStream_t stream) {
    MCB_TRY(device.alloc((size_t)rows * columns * sizeof(T), stream));
    if (rows > 0 && columns > 0)
        MCB_CUDA(cudaMemcpy2DAsync(device.ptr, columns * sizeof(T), host, row_stride * sizeof(T), columns * sizeof(T), rows,
                                   cudaMemcpyHostToDevice, stream));
    return MCB_OK;
}

}  // namespace

namespace {
// np.fill_diagonal(similarity, 1) (computation.py:466) for a block of rows of the full matrix.
__global__ void unit_diagonal_kernel(float* __restrict__ block, int64_t rows, int64_t stride, int64_t first_row, int64_t columns) {
    const int64_t row = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (row < rows && first_row + row < columns) block[row * stride + first_row + row] = 1.0f;
}
}  // namespace

namespace mcb {
// The symmetric similarity of the pile path: as mcb_dev_correlate_dense_f32, and *min_key (device, preset to
// 0xffffffff) receives the matrix minimum as a float_to_ordered key (fused np.min of knn_graph.py:290).
int correlate_dense_with_min(const float* input, int64_t rows, int64_t columns, int64_t row_stride, float* output,
                             int64_t out_stride, int absolute, unsigned int* min_key, cudaStream_t stream) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns && out_stride >= rows, "correlate_dense: bad shape");
    if (rows == 0)
        return MCB_OK;
    return correlate_f32_device(input, rows, row_stride, nullptr, 0, 0, columns, output, out_stride, 1, absolute, stream, min_key);
}
}  // namespace mcb

extern "C" {

int mcb_dev_similarity_block_f32(const float* rows_block, int64_t block_rows, int64_t first_row, const float* all_rows,
                                 int64_t all_count, int64_t columns, float* output, int64_t out_stride, int absolute,
                                 void* stream) {
    MCB_REQUIRE(block_rows >= 0 && all_count >= 0 && columns >= 0 && out_stride >= all_count && first_row >= 0 &&
                    first_row + block_rows <= all_count,
                "similarity_block: bad shape");
    if (block_rows == 0)
        return MCB_OK;
    MCB_TRY(correlate_f32_device(rows_block, block_rows, columns, all_rows, all_count, columns, columns, output, out_stride, 0,
                                 absolute, (cudaStream_t)stream));
    unit_diagonal_kernel<<<(unsigned)ceil_div(block_rows, 256), 256, 0, (cudaStream_t)stream>>>(output, block_rows, out_stride,
                                                                                               first_row, all_count);
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

int mcb_dev_correlate_dense_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, float* output,
                                int64_t out_stride, int absolute, void* stream) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns && out_stride >= rows, "correlate_dense: bad shape");
    return correlate_f32_device(input, rows, row_stride, nullptr, 0, 0, columns, output, out_stride, 1, absolute,
                                (cudaStream_t)stream);
}

int mcb_dev_cross_correlate_dense_f32(const float* first, int64_t first_rows, const float* second, int64_t second_rows,
                                      int64_t columns, float* output, int64_t out_stride, void* stream) {
    MCB_REQUIRE(first_rows >= 0 && second_rows >= 0 && columns >= 0 && out_stride >= second_rows, "cross_correlate_dense: bad shape");
    return correlate_f32_device(first, first_rows, columns, second, second_rows, columns, columns, output, out_stride, 0, 0,
                                (cudaStream_t)stream);
}

int mcb_dev_pairs_correlate_dense_f32(const float* first, const float* second, int64_t rows, int64_t columns, float* output,
                                      void* stream) {
    MCB_REQUIRE(rows >= 0 && columns >= 0, "pairs_correlate_dense: bad shape");
    return pairs_rows_f32(first, second, rows, columns, output, (cudaStream_t)stream);
}

int mcb_host_correlate_dense_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, float* output) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "correlate_dense: bad shape");
    if (rows == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch d_in, d_out;
    MCB_TRY(upload_matrix(d_in, input, rows, columns, row_stride, stream));
    MCB_TRY(d_out.alloc((size_t)rows * rows * sizeof(float), stream));
    MCB_TRY(correlate_f32_device(d_in.as<float>(), rows, columns, nullptr, 0, 0, columns, d_out.as<float>(), rows, 1, 0, stream));
    MCB_TRY(copy_to_host(output, d_out.ptr, (size_t)rows * rows * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_correlate_dense_f64(const double* input, int64_t rows, int64_t columns, int64_t row_stride, float* output) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && row_stride >= columns, "correlate_dense: bad shape");
    if (rows == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch d_in, d_out;
    MCB_TRY(upload_matrix(d_in, input, rows, columns, row_stride, stream));
    MCB_TRY(d_out.alloc((size_t)rows * rows * sizeof(float), stream));
    MCB_TRY(correlate_f64_device(d_in.as<double>(), rows, columns, nullptr, 0, 0, columns, d_out.as<float>(), rows, 1, stream));
    MCB_TRY(copy_to_host(output, d_out.ptr, (size_t)rows * rows * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_cross_correlate_dense_f32(const float* first, int64_t first_rows, int64_t first_stride, const float* second,
                                       int64_t second_rows, int64_t second_stride, int64_t columns, float* output) {
    MCB_REQUIRE(first_rows >= 0 && second_rows >= 0 && columns >= 0 && first_stride >= columns && second_stride >= columns,
                "cross_correlate_dense: bad shape");
    if (first_rows == 0 || second_rows == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch d_a, d_b, d_out;
    MCB_TRY(upload_matrix(d_a, first, first_rows, columns, first_stride, stream));
    MCB_TRY(upload_matrix(d_b, second, second_rows, columns, second_stride, stream));
    MCB_TRY(d_out.alloc((size_t)first_rows * second_rows * sizeof(float), stream));
    MCB_TRY(correlate_f32_device(d_a.as<float>(), first_rows, columns, d_b.as<float>(), second_rows, columns, columns,
                                 d_out.as<float>(), second_rows, 0, 0, stream));
    MCB_TRY(copy_to_host(output, d_out.ptr, (size_t)first_rows * second_rows * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_cross_correlate_dense_f64(const double* first, int64_t first_rows, int64_t first_stride, const double* second,
                                       int64_t second_rows, int64_t second_stride, int64_t columns, float* output) {
    MCB_REQUIRE(first_rows >= 0 && second_rows >= 0 && columns >= 0 && first_stride >= columns && second_stride >= columns,
                "cross_correlate_dense: bad shape");
    if (first_rows == 0 || second_rows == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch d_a, d_b, d_out;
    MCB_TRY(upload_matrix(d_a, first, first_rows, columns, first_stride, stream));
    MCB_TRY(upload_matrix(d_b, second, second_rows, columns, second_stride, stream));
    MCB_TRY(d_out.alloc((size_t)first_rows * second_rows * sizeof(float), stream));
    MCB_TRY(correlate_f64_device(d_a.as<double>(), first_rows, columns, d_b.as<double>(), second_rows, columns, columns,
                                 d_out.as<float>(), second_rows, 0, stream));
    MCB_TRY(copy_to_host(output, d_out.ptr, (size_t)first_rows * second_rows * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_pairs_correlate_dense_f32(const float* first, int64_t first_stride, const float* second, int64_t second_stride,
                                       int64_t rows, int64_t columns, float* output) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && first_stride >= columns && second_stride >= columns, "pairs_correlate_dense: bad shape");
    if (rows == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch d_a, d_b, d_out;
    MCB_TRY(upload_matrix(d_a, first, rows, columns, first_stride, stream));
    MCB_TRY(upload_matrix(d_b, second, rows, columns, second_stride, stream));
    MCB_TRY(d_out.alloc((size_t)rows * sizeof(float), stream));
    MCB_TRY(pairs_rows_f32(d_a.as<float>(), d_b.as<float>(), rows, columns, d_out.as<float>(), stream));
    MCB_TRY(copy_to_host(output, d_out.ptr, (size_t)rows * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_pairs_correlate_dense_f64(const double* first, int64_t first_stride, const double* second, int64_t second_stride,
                                       int64_t rows, int64_t columns, float* output) {
    MCB_REQUIRE(rows >= 0 && columns >= 0 && first_stride >= columns && second_stride >= columns, "pairs_correlate_dense: bad shape");
    if (rows == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch d_a, d_b, d_out;
    MCB_TRY(upload_matrix(d_a, first, rows, columns, first_stride, stream));
    MCB_TRY(upload_matrix(d_b, second, rows, columns, second_stride, stream));
    MCB_TRY(d_out.alloc((size_t)rows * sizeof(float), stream));
    MCB_TRY(pairs_rows_f64(d_a.as<double>(), d_b.as<double>(), rows, columns, d_out.as<float>(), stream));
    MCB_TRY(copy_to_host(output, d_out.ptr, (size_t)rows * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

}  // extern "C"
