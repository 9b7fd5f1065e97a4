// Pearson correlation as a tcgen05 contraction:  C = Xn Yn^T  with Xn, Yn the centred, unit-norm
// rows produced by prepare_rows (correlate_prep.cu), in FP32-emulated precision:
//
//     x = hi + lo (both TF32)      x.y ~= hi.hi + hi.lo + lo.hi        (3 x kind::tf32 MMA)
//
// Replaces correlate_dense / cross_correlate_dense (reference correlate.cpp:350-494).
//
// Kernel shape (one 128 x 128 output tile per CTA, 192 threads):
//   warp 0      TMA producer: per k-block of 32 floats loads four 128 x 32 tiles (A_hi, A_lo, B_hi,
//               B_lo; 128-byte swizzle) into a 3-stage ring, signalling `full[stage]`.
//   warp 1      allocates 128 TMEM columns; one lane issues 12 tcgen05.mma per k-block
//               (4 k-steps of 8 x 3 split terms) and tcgen05.commit's the stage back to `empty`.
//   warps 2..5  epilogue: tcgen05.ld the fp32 accumulator (lane = tile row), clamp to [-1, 1],
//               NaN -> 0, optional |r|, diagonal = 1, store the tile and (symmetric case, off
//               diagonal tiles) its mirror image, so only the lower triangle of tiles is computed,
//               as the reference computes only the lower triangle of pairs (correlate.cpp:381-403).
#include <cuda.h>

#include "common.cuh"

namespace mcb {

namespace {

constexpr int BM = 128;
constexpr int BN = 128;
constexpr int BK = 32;  // floats: 128 bytes = one swizzle row
constexpr int STAGES = 3;
constexpr int TILE_BYTES = BM * BK * 4;          // 16 KiB
constexpr int STAGE_BYTES = 4 * TILE_BYTES;      // A_hi, A_lo, B_hi, B_lo
constexpr int UMMA_THREADS = 192;
constexpr int TMEM_COLUMNS = 128;
constexpr size_t UMMA_SMEM = (size_t)STAGES * STAGE_BYTES + 1024 /* alignment slack */ + 256 /* barriers */;

// kind::tf32, D = F32, A/B = TF32, both K-major, M = 128, N = 128 (cute::UMMA::InstrDescriptor).
constexpr uint32_t INSTR_DESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ uint32_t smem_u32(const void* pointer) { return (uint32_t)__cvta_generic_to_shared(pointer); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
// K-major, 128-byte swizzle: 8-row groups are 1024 bytes apart (SBO), version 1 (sm_100).
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_address) {
    return (uint64_t)((smem_address & 0x3ffffu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(INSTR_DESC), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

struct UmmaParams {
    float* output;
    int64_t out_stride;
    int64_t rows_a, rows_b;
    int k_blocks;
    int tiles_b;    // tiles along B (cross case)
    int symmetric;  // lower-triangle tiles, mirrored store, unit diagonal
    int absolute;   // store |r|
};

__device__ __forceinline__ float finish(float r, int absolute) {
    r = fminf(fmaxf(r, -1.0f), 1.0f);  // also maps NaN to -1 then ... handled below
    return absolute ? fabsf(r) : r;
}

}  // namespace

__global__ void __launch_bounds__(UMMA_THREADS, 1)
    correlate_umma_kernel(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                          const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo,
                          UmmaParams p) {
    extern __shared__ uint8_t raw_smem[];
    const uint32_t base = (smem_u32(raw_smem) + 1023u) & ~1023u;
    const uint32_t bar_base = base + STAGES * STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    const uint32_t tmem_full_bar = bar_base + 8u * (2 * STAGES);
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    int tile_i, tile_j;
    if (p.symmetric) {
        const int t = blockIdx.x;
        int i = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
        while ((int64_t)i * (i + 1) / 2 > t) --i;
        while ((int64_t)(i + 1) * (i + 2) / 2 <= t) ++i;
        tile_i = i;
        tile_j = t - (int)((int64_t)i * (i + 1) / 2);
    } else {
        tile_i = blockIdx.x / p.tiles_b;
        tile_j = blockIdx.x % p.tiles_b;
    }

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        mbar_init(tmem_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a_hi) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a_lo) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b_hi) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b_lo) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLUMNS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < p.k_blocks; ++kb) {
                const int s = kb % STAGES;
                if (kb >= STAGES)
                    mbar_wait(empty_bar(s), ((kb / STAGES) - 1) & 1);
                const uint32_t stage = base + s * STAGE_BYTES;
                mbar_expect_tx(full_bar(s), STAGE_BYTES);
                tma_load_2d(stage + 0 * TILE_BYTES, &map_a_hi, full_bar(s), kb * BK, tile_i * BM);
                tma_load_2d(stage + 1 * TILE_BYTES, &map_a_lo, full_bar(s), kb * BK, tile_i * BM);
                tma_load_2d(stage + 2 * TILE_BYTES, &map_b_hi, full_bar(s), kb * BK, tile_j * BN);
                tma_load_2d(stage + 3 * TILE_BYTES, &map_b_lo, full_bar(s), kb * BK, tile_j * BN);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            for (int kb = 0; kb < p.k_blocks; ++kb) {
                const int s = kb % STAGES;
                mbar_wait(full_bar(s), (kb / STAGES) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t stage = base + s * STAGE_BYTES;
                const uint64_t a_hi = umma_desc(stage + 0 * TILE_BYTES), a_lo = umma_desc(stage + 1 * TILE_BYTES);
                const uint64_t b_hi = umma_desc(stage + 2 * TILE_BYTES), b_lo = umma_desc(stage + 3 * TILE_BYTES);
#pragma unroll
                for (int k = 0; k < BK / 8; ++k) {
                    const uint64_t advance = (uint64_t)((k * 8 * 4) >> 4);  // bytes >> 4 inside the swizzle row
                    // small cross terms first, the dominant hi.hi term last
                    umma_tf32(tmem_base, a_lo + advance, b_hi + advance, (kb | k) ? 1u : 0u);
                    umma_tf32(tmem_base, a_hi + advance, b_lo + advance, 1u);
                    umma_tf32(tmem_base, a_hi + advance, b_hi + advance, 1u);
                }
                umma_commit(empty_bar(s));
            }
            umma_commit(tmem_full_bar);
        }
    } else {
        const int quarter = warp & 3;  // TMEM lanes 32 * quarter .. + 31 are visible to this warp
        mbar_wait(tmem_full_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int row_in_tile = quarter * 32 + lane;
        const int64_t global_row = (int64_t)tile_i * BM + row_in_tile;
        const bool mirror = p.symmetric && tile_i != tile_j;
        for (int chunk = 0; chunk < BN / 16; ++chunk) {
            uint32_t v[16];
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(chunk * 16);
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                  "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                : "r"(taddr)
                : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const int64_t global_col = (int64_t)tile_j * BN + chunk * 16 + c;
                float r = __uint_as_float(v[c]);
                r = (r != r) ? 0.0f : r;  // computation.py:465 maps NaN to 0
                r = finish(r, p.absolute);
                if (p.symmetric && global_row == global_col)
                    r = 1.0f;
                if (global_row < p.rows_a && global_col < p.rows_b) {
                    p.output[global_row * p.out_stride + global_col] = r;
                    if (mirror)
                        p.output[global_col * p.out_stride + global_row] = r;
                }
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLUMNS) : "memory");
    }
}

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// libcuda is reached through the runtime so that the library loads on machines without a driver.
int encode_fn(EncodeTiledFn* fn) {
    static EncodeTiledFn cached = nullptr;
    if (!cached) {
        void* entry = nullptr;
        cudaDriverEntryPointQueryResult query;
        MCB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &entry, cudaEnableDefault, &query));
        if (!entry || query != cudaDriverEntryPointSuccess)
            return fail(MCB_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
        cached = (EncodeTiledFn)entry;
    }
    *fn = cached;
    return MCB_OK;
}

// [rows, padded] float32, row major; box = 32 columns x 128 rows; 128-byte swizzle; OOB rows read 0.
int make_map(CUtensorMap* map, const float* data, int64_t rows, int64_t padded) {
    EncodeTiledFn encode;
    MCB_TRY(encode_fn(&encode));
    cuuint64_t dims[2] = {(cuuint64_t)padded, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)padded * sizeof(float)};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)BM};
    cuuint32_t element_strides[2] = {1, 1};
    CUresult result = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)data, dims, strides, box, element_strides,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (result != CUDA_SUCCESS)
        return fail(MCB_ERR_CUDA, "cuTensorMapEncodeTiled failed with code %d", (int)result);
    return MCB_OK;
}

}  // namespace

// a_* / b_*: prepared rows [rows, padded] (padded a multiple of 32).  For the symmetric case pass the
// same arrays for a and b.  output: [rows_a, out_stride >= rows_b].
int correlate_umma_launch(const float* a_hi, const float* a_lo, int64_t rows_a, const float* b_hi, const float* b_lo,
                          int64_t rows_b, int64_t padded, float* output, int64_t out_stride, int symmetric, int absolute,
                          cudaStream_t stream) {
    if (rows_a == 0 || rows_b == 0)
        return MCB_OK;
    MCB_REQUIRE(padded > 0 && padded % BK == 0, "correlate: padded width %lld is not a positive multiple of %d", (long long)padded, BK);
    CUtensorMap map_a_hi, map_a_lo, map_b_hi, map_b_lo;
    MCB_TRY(make_map(&map_a_hi, a_hi, rows_a, padded));
    MCB_TRY(make_map(&map_a_lo, a_lo, rows_a, padded));
    MCB_TRY(make_map(&map_b_hi, b_hi, rows_b, padded));
    MCB_TRY(make_map(&map_b_lo, b_lo, rows_b, padded));
    UmmaParams p{};
    p.output = output;
    p.out_stride = out_stride;
    p.rows_a = rows_a;
    p.rows_b = rows_b;
    p.k_blocks = (int)(padded / BK);
    p.symmetric = symmetric;
    p.absolute = absolute;
    const int64_t tiles_a = ceil_div(rows_a, BM), tiles_b = ceil_div(rows_b, BN);
    p.tiles_b = (int)tiles_b;
    const int64_t grid = symmetric ? tiles_a * (tiles_a + 1) / 2 : tiles_a * tiles_b;
    MCB_REQUIRE(grid < (1ll << 31), "correlate: too many tiles");
    MCB_CUDA(cudaFuncSetAttribute(correlate_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)UMMA_SMEM));
    correlate_umma_kernel<<<(unsigned)grid, UMMA_THREADS, UMMA_SMEM, stream>>>(map_a_hi, map_a_lo, map_b_hi, map_b_lo, p);
    MCB_CUDA(cudaGetLastError());
    return MCB_OK;
}

}  // namespace mcb
