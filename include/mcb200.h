/*
 * mcb200.h -- C ABI of libmcb200.so: the B200 (sm_100a) implementation of the metacells per-pile
 * k-NN similarity-graph hot path.  Plain pointers and sizes only.
 *
 * Every entry point replaces one function of the reference's pybind11 module
 * `metacells.extensions` (reference file:line cited per function; paths relative to
 * tanaylab/metacells v0.9.5).  Two forms exist for each operation:
 *
 *   mcb_host_<op>(...)          pointers are HOST memory (numpy buffers).  The call copies to the
 *                               GPU, runs the kernels, copies back and returns when done: the exact
 *                               contract of the reference function it replaces.
 *   mcb_dev_<op>(..., stream)   pointers are DEVICE memory; work is enqueued on `stream`
 *                               (a cudaStream_t passed as void*).  Used by the device-resident
 *                               pipeline so a pile never leaves HBM between stages.
 *
 * All functions return 0 on success; otherwise a non-zero code, with a message available from
 * mcb_last_error() (thread local).  Where the reference calls std::exit(1) on a failed assertion
 * (extensions.h:59-72) this library returns MCB_ERR_ARG instead.
 *
 * Type suffixes follow the reference's symbol names: f32 = float32_t, f64 = float64_t,
 * i32 = int32_t.  Only the instantiations the hot path uses are provided.
 */
#ifndef MCB200_H
#define MCB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------------------- */

const char* mcb_last_error(void);
int mcb_version(void);
int mcb_device_count(void);
/* Select the CUDA device of the calling thread; fails unless it is compute capability 10.x. */
int mcb_set_device(int device);
int mcb_synchronize(void* stream);

/* ---- downsample.cpp ------------------------------------------------------------------------- */

/* downsample_compressed<float32,int32,float32>  (downsample.cpp:171-192).
 * data/out: nnz values; indptr: bands+1; samples: bands.  Rows with samples < 0 or no elements
 * leave `out` untouched, as the reference does (:86-88). */
int mcb_host_downsample_compressed_f32_i32_f32(const float* data, const int32_t* indptr, float* out,
                                               const int32_t* samples, int64_t bands, uint64_t random_seed);
/* max_band_size: an upper bound of indptr[b+1]-indptr[b] (the host form computes it itself). */
int mcb_dev_downsample_compressed_f32_i32_f32(const float* data, const int32_t* indptr, float* out,
                                              const int32_t* samples, int64_t bands, int64_t max_band_size,
                                              uint64_t random_seed, void* stream);

/* downsample_dense<float32,float32>  (downsample.cpp:131-151): rows x columns, C contiguous. */
int mcb_host_downsample_dense_f32_f32(const float* input, float* output, const int32_t* samples, int64_t rows,
                                      int64_t columns, uint64_t random_seed);
int mcb_dev_downsample_dense_f32_f32(const float* input, float* output, const int32_t* samples, int64_t rows,
                                     int64_t columns, uint64_t random_seed, void* stream);

/* downsample_array<float32,float32>  (downsample.cpp:116-128): the seed is used as given. */
int mcb_host_downsample_array_f32_f32(const float* input, float* output, int64_t size, int32_t samples,
                                      uint64_t random_seed);

/* ---- correlate.cpp --------------------------------------------------------------------------- */

/* correlate_dense<F>  (correlate.cpp:350-431): all-pairs Pearson r between the rows of
 * input[rows, columns] (row stride in elements >= columns).  output: float32 [rows, rows],
 * symmetric, unit diagonal, 0 for constant rows, clamped to [-1, 1]; NaN results are stored as 0
 * (the reference's caller does that at computation.py:465).
 * float32 input runs on the tensor cores (tcgen05, 3xTF32 split); float64 input in float64 FMA. */
int mcb_host_correlate_dense_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, float* output);
int mcb_host_correlate_dense_f64(const double* input, int64_t rows, int64_t columns, int64_t row_stride, float* output);
/* Device form; absolute != 0 stores |r| (fuses tools/similarity.py:214-215 "abs_pearson"). */
int mcb_dev_correlate_dense_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, float* output,
                                int64_t out_stride, int absolute, void* stream);

/* cross_correlate_dense<F>  (correlate.cpp:433-494): output float32 [first_rows, second_rows]. */
int mcb_host_cross_correlate_dense_f32(const float* first, int64_t first_rows, int64_t first_stride, const float* second,
                                       int64_t second_rows, int64_t second_stride, int64_t columns, float* output);
int mcb_host_cross_correlate_dense_f64(const double* first, int64_t first_rows, int64_t first_stride, const double* second,
                                       int64_t second_rows, int64_t second_stride, int64_t columns, float* output);
int mcb_dev_cross_correlate_dense_f32(const float* first, int64_t first_rows, const float* second, int64_t second_rows,
                                      int64_t columns, float* output, int64_t out_stride, void* stream);

/* pairs_correlate_dense<F>  (correlate.cpp:496-527): output[i] = r(first[i, :], second[i, :]). */
int mcb_host_pairs_correlate_dense_f32(const float* first, int64_t first_stride, const float* second, int64_t second_stride,
                                       int64_t rows, int64_t columns, float* output);
int mcb_host_pairs_correlate_dense_f64(const double* first, int64_t first_stride, const double* second, int64_t second_stride,
                                       int64_t rows, int64_t columns, float* output);
int mcb_dev_pairs_correlate_dense_f32(const float* first, const float* second, int64_t rows, int64_t columns, float* output,
                                      void* stream);

/* @@APPEND@@ */

#ifdef __cplusplus
}
#endif

#endif /* MCB200_H */
