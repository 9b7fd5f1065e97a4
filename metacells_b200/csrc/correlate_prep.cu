// Row preparation for Pearson correlation, and the float64 SIMT contraction.
//
// The reference computes r = (n Sxy - Sx Sy) / sqrt((n Sxx - Sx^2)(n Syy - Sy^2)) with Sx, Sxx in
// float64 (correlate.cpp:14-27, 188-203).  For log-fraction data the mean dwarfs the spread, so
// that form cancels badly.  Here every row is centred by its float64 mean and scaled to unit L2
// norm first; then r is simply the dot product of two prepared rows, every partial sum is <= 1
// in magnitude, and a constant row becomes all zeros, which yields the reference's "0 for a
// constant row" without a special case (correlate.cpp:196-201).
//
//   prepare_rows<T>   : one warp per row; float64 mean and norm; writes either
//                         (a) the TF32 hi/lo split for the tcgen05 kernel (correlate_umma.cu), or
//                         (b) float64 normalised rows for the SIMT kernel below.
//   dot_rows_f64      : C = X Y^T in float64 FMA, 64x64 tiles; serves correlate_dense_float64_t
//                       and is the independent on-device check of the tensor-core path.
//   pairs_kernel      : r between row i of A and row i of B (pairs_correlate_dense).
#include "common.cuh"

namespace mcb {

namespace {

__device__ __forceinline__ double warp_sum(double value) {
#pragma unroll
    for (int offset = 16; offset > 0; offset >>= 1) value += __shfl_xor_sync(0xffffffffu, value, offset);
    return value;
}

__device__ __forceinline__ float round_to_tf32(float value) {
    uint32_t bits;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(bits) : "f"(value));
    return __uint_as_float(bits);
}

// One warp per row.  out_hi/out_lo: [rows, padded] float (zero padded columns), or out_f64.
template <typename T>
__global__ void __launch_bounds__(256) prepare_rows_kernel(const T* __restrict__ input, int64_t rows, int64_t columns,
                                                          int64_t row_stride, int64_t padded, float* __restrict__ out_hi,
                                                          float* __restrict__ out_lo, double* __restrict__ out_f64) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows)
        return;
    const T* in = input + row * row_stride;
    double sum = 0;
    for (int64_t c = lane; c < columns; c += 32) sum += (double)in[c];
    sum = warp_sum(sum);
    const double mean = sum / (double)columns;
    double squares = 0;
    for (int64_t c = lane; c < columns; c += 32) {
        double centred = (double)in[c] - mean;
        squares = fma(centred, centred, squares);
    }
    squares = warp_sum(squares);
    const double scale = squares > 0 ? 1.0 / sqrt(squares) : 0.0;
    if (out_f64) {
        double* out = out_f64 + row * padded;
        for (int64_t c = lane; c < padded; c += 32) out[c] = c < columns ? ((double)in[c] - mean) * scale : 0.0;
    } else {
        float* hi = out_hi + row * padded;
        float* lo = out_lo + row * padded;
        for (int64_t c = lane; c < padded; c += 32) {
            float high = 0.0f, low = 0.0f;
            if (c < columns) {
                double value = ((double)in[c] - mean) * scale;
                high = round_to_tf32((float)value);
                low = round_to_tf32((float)(value - (double)high));
            }
            hi[c] = high;
            lo[c] = low;
        }
    }
}

constexpr int DT = 64;  // output tile
constexpr int DK = 16;  // k slab

// C[i, j] = clamp(sum_k X[i, k] Y[j, k]); symmetric: only tiles with bj <= bi, mirrored on store.
__global__ void __launch_bounds__(256) dot_rows_f64_kernel(const double* __restrict__ x, const double* __restrict__ y,
                                                          int64_t rows_x, int64_t rows_y, int64_t padded,
                                                          float* __restrict__ output, int64_t out_stride, int symmetric) {
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (symmetric && bj > bi)
        return;
    __shared__ double tile_x[DK][DT + 1];
    __shared__ double tile_y[DK][DT + 1];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    double acc[4][4] = {};
    for (int64_t k0 = 0; k0 < padded; k0 += DK) {
        for (int e = threadIdx.x; e < DT * DK; e += 256) {
            int r = e / DK, k = e % DK;
            int64_t gi = (int64_t)bi * DT + r, gj = (int64_t)bj * DT + r;
            tile_x[k][r] = gi < rows_x ? x[gi * padded + k0 + k] : 0.0;
            tile_y[k][r] = gj < rows_y ? y[gj * padded + k0 + k] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < DK; ++k) {
            double a[4], b[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                a[u] = tile_x[k][ty * 4 + u];
                b[u] = tile_y[k][tx * 4 + u];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) acc[u][v] = fma(a[u], b[v], acc[u][v]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            int64_t gi = (int64_t)bi * DT + ty * 4 + u, gj = (int64_t)bj * DT + tx * 4 + v;
            if (gi >= rows_x || gj >= rows_y)
                continue;
            float r = (float)acc[u][v];
            r = fminf(fmaxf(r, -1.0f), 1.0f);
            if (symmetric) {
                if (gi == gj)
                    r = 1.0f;  // correlate.cpp:374-376
                output[gi * out_stride + gj] = r;
                output[gj * out_stride + gi] = r;
            } else {
                output[gi * out_stride + gj] = r;
            }
        }
}

// One warp per pair of rows; everything in float64 (correlate.cpp:496-527 via :29-62).
template <typename T>
__global__ void __launch_bounds__(256) pairs_kernel(const T* __restrict__ first, const T* __restrict__ second,
                                                   int64_t rows, int64_t columns, float* __restrict__ output) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows)
        return;
    const T* a = first + row * columns;
    const T* b = second + row * columns;
    double sa = 0, sb = 0;
    for (int64_t c = lane; c < columns; c += 32) {
        sa += (double)a[c];
        sb += (double)b[c];
    }
    sa = warp_sum(sa) / (double)columns;
    sb = warp_sum(sb) / (double)columns;
    double aa = 0, bb = 0, ab = 0;
    for (int64_t c = lane; c < columns; c += 32) {
        double ca = (double)a[c] - sa, cb = (double)b[c] - sb;
        aa = fma(ca, ca, aa);
        bb = fma(cb, cb, bb);
        ab = fma(ca, cb, ab);
    }
    aa = warp_sum(aa);
    bb = warp_sum(bb);
    ab = warp_sum(ab);
    if (lane == 0) {
        double factors = sqrt(aa * bb);
        float r = 0.0f;
        if (factors != 0) {
            r = (float)(ab / factors);
            r = fminf(fmaxf(r, -1.0f), 1.0f);
        }
        output[row] = r;
    }
}

}  // namespace

int prepare_rows_f32(const float* input, int64_t rows, int64_t columns, int64_t row_stride, int64_t padded, float* hi,
                     float* lo, double* f64, cudaStream_t stream) {
    if (rows == 0)
        return MCB_OK;
    MCB_PROFILED("correlate_prepare", stream, prepare_rows_kernel<float><<<(unsigned)ceil_div(rows, 8), 256, 0, stream>>>(input, rows, columns, row_stride, padded, hi, lo, f64));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

int prepare_rows_f64(const double* input, int64_t rows, int64_t columns, int64_t row_stride, int64_t padded, float* hi,
                     float* lo, double* f64, cudaStream_t stream) {
    if (rows == 0)
        return MCB_OK;
    MCB_PROFILED("correlate_prepare", stream, prepare_rows_kernel<double><<<(unsigned)ceil_div(rows, 8), 256, 0, stream>>>(input, rows, columns, row_stride, padded, hi, lo, f64));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

int dot_rows_f64(const double* x, const double* y, int64_t rows_x, int64_t rows_y, int64_t padded, float* output,
                 int64_t out_stride, int symmetric, cudaStream_t stream) {
    if (rows_x == 0 || rows_y == 0)
        return MCB_OK;
    dim3 grid((unsigned)ceil_div(rows_y, DT), (unsigned)ceil_div(rows_x, DT));
    MCB_PROFILED("correlate_f64", stream, dot_rows_f64_kernel<<<grid, 256, 0, stream>>>(x, y, rows_x, rows_y, padded, output, out_stride, symmetric));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

int pairs_rows_f32(const float* first, const float* second, int64_t rows, int64_t columns, float* output, cudaStream_t stream) {
    if (rows == 0)
        return MCB_OK;
    pairs_kernel<float><<<(unsigned)ceil_div(rows, 8), 256, 0, stream>>>(first, second, rows, columns, output);
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

int pairs_rows_f64(const double* first, const double* second, int64_t rows, int64_t columns, float* output, cudaStream_t stream) {
    if (rows == 0)
        return MCB_OK;
    pairs_kernel<double><<<(unsigned)ceil_div(rows, 8), 256, 0, stream>>>(first, second, rows, columns, output);
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

}  // namespace mcb
