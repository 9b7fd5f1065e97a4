// C-ABI entry points for downsample (host-buffer and device-pointer forms).
#include "common.cuh"
#include "../../include/mcb200.h"

namespace mcb {
int downsample_launch(const float* input, float* output, const int32_t* indptr, const int32_t* samples,
                      int32_t samples_all, int64_t rows, int64_t columns, int64_t max_row_size, uint64_t seed,
                      int use_row_seed, cudaStream_t stream);
}

using namespace mcb;

extern "C" {

int mcb_dev_downsample_compressed_f32_i32_f32(const float* data, const int32_t* indptr, float* out,
                                              const int32_t* samples, int64_t bands, int64_t max_band_size,
                                              uint64_t random_seed, void* stream) {
    MCB_REQUIRE(bands >= 0 && max_band_size >= 0, "downsample_compressed: negative size");
    return downsample_launch(data, out, indptr, samples, 0, bands, 0, max_band_size, random_seed, 1,
                             (cudaStream_t)stream);
}

int mcb_host_downsample_compressed_f32_i32_f32(const float* data, const int32_t* indptr, float* out,
                                               const int32_t* samples, int64_t bands, uint64_t random_seed) {
    MCB_REQUIRE(bands >= 0, "downsample_compressed: negative bands count");
    if (bands == 0)
        return MCB_OK;
    const int64_t nnz = indptr[bands];
    int64_t widest = 0;
    for (int64_t b = 0; b < bands; ++b) {
        MCB_REQUIRE(indptr[b + 1] >= indptr[b], "downsample_compressed: indptr is not monotone at band %lld", (long long)b);
        widest = indptr[b + 1] - indptr[b] > widest ? indptr[b + 1] - indptr[b] : widest;
    }
    cudaStream_t stream = nullptr;
    Scratch d_data, d_out, d_indptr, d_samples;
    MCB_TRY(d_data.alloc(nnz * sizeof(float), stream));
    MCB_TRY(d_out.alloc(nnz * sizeof(float), stream));
    MCB_TRY(d_indptr.alloc((bands + 1) * sizeof(int32_t), stream));
    MCB_TRY(d_samples.alloc(bands * sizeof(int32_t), stream));
    MCB_CUDA(cudaMemcpyAsync(d_data.ptr, data, nnz * sizeof(float), cudaMemcpyHostToDevice, stream));
    // Rows the reference leaves untouched must keep the caller's bytes: start from `out` itself.
    MCB_CUDA(cudaMemcpyAsync(d_out.ptr, out, nnz * sizeof(float), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_indptr.ptr, indptr, (bands + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_samples.ptr, samples, bands * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    MCB_TRY(downsample_launch(d_data.as<float>(), d_out.as<float>(), d_indptr.as<int32_t>(), d_samples.as<int32_t>(), 0,
                              bands, 0, widest, random_seed, 1, stream));
    MCB_TRY(copy_to_host(out, d_out.ptr, nnz * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_dev_downsample_dense_f32_f32(const float* input, float* output, const int32_t* samples, int64_t rows,
                                     int64_t columns, uint64_t random_seed, void* stream) {
    MCB_REQUIRE(rows >= 0 && columns >= 0, "downsample_dense: negative size");
    return downsample_launch(input, output, nullptr, samples, 0, rows, columns, columns, random_seed, 1,
                             (cudaStream_t)stream);
}

int mcb_host_downsample_dense_f32_f32(const float* input, float* output, const int32_t* samples, int64_t rows,
                                      int64_t columns, uint64_t random_seed) {
    MCB_REQUIRE(rows >= 0 && columns >= 0, "downsample_dense: negative size");
    if (rows == 0 || columns == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    const size_t bytes = (size_t)rows * columns * sizeof(float);
    Scratch d_in, d_out, d_samples;
    MCB_TRY(d_in.alloc(bytes, stream));
    MCB_TRY(d_out.alloc(bytes, stream));
    MCB_TRY(d_samples.alloc(rows * sizeof(int32_t), stream));
    MCB_CUDA(cudaMemcpyAsync(d_in.ptr, input, bytes, cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_out.ptr, output, bytes, cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_samples.ptr, samples, rows * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    MCB_TRY(downsample_launch(d_in.as<float>(), d_out.as<float>(), nullptr, d_samples.as<int32_t>(), 0, rows, columns,
                              columns, random_seed, 1, stream));
    MCB_TRY(copy_to_host(output, d_out.ptr, bytes, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_downsample_array_f32_f32(const float* input, float* output, int64_t size, int32_t samples,
                                      uint64_t random_seed) {
    MCB_REQUIRE(size >= 0, "downsample_array: negative size");
    if (size == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    Scratch d_in, d_out;
    MCB_TRY(d_in.alloc(size * sizeof(float), stream));
    MCB_TRY(d_out.alloc(size * sizeof(float), stream));
    MCB_CUDA(cudaMemcpyAsync(d_in.ptr, input, size * sizeof(float), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_out.ptr, output, size * sizeof(float), cudaMemcpyHostToDevice, stream));
    MCB_TRY(downsample_launch(d_in.as<float>(), d_out.as<float>(), nullptr, nullptr, samples, 1, size, size, random_seed,
                              0, stream));
    MCB_TRY(copy_to_host(output, d_out.ptr, size * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

}  // extern "C"
