// Pearson correlation as a tcgen05 contraction:  C = Xn Yn^T  with Xn, Yn the centred, unit-norm
// rows produced by prepare_rows (correlate_prep.cu), in FP32-emulated precision:
//
//     x = hi + lo (both TF32)      x.y ~= hi.hi + hi.lo + lo.hi        (3 x kind::tf32 MMA)
//
// Replaces correlate_dense / cross_correlate_dense (reference correlate.cpp:350-494).
//
// Persistent kernel, one CTA per SM, 192 threads, 128 x 128 output tiles taken from a host-built
// tile list (super-blocks of 8 x 8 tiles so that a wave of CTAs re-uses row panels out of L2;
// symmetric case: lower-triangle tiles only, as the reference computes only the lower triangle
// of pairs, correlate.cpp:381-403):
//   warp 0      TMA producer: per k-block of 32 floats loads four 128 x 32 tiles (A_hi, A_lo, B_hi,
//               B_lo; 128-byte swizzle) into a 3-stage ring, signalling `full[stage]`.
//   warp 1      allocates 256 TMEM columns (two 128-column accumulators); one lane issues 12
//               tcgen05.mma per k-block (4 k-steps of 8 x 3 split terms), tcgen05.commit's the
//               stage back to `empty`, and after every CHUNK k-blocks commits `acc_full[buffer]`
//               and switches to the other accumulator.
//   warps 2..5  drain + epilogue: the tensor core adds into its FP32 accumulator with truncation,
//               which over the ~10^3 MMA steps of one dot product costs ~1e-5 (measured).  So each
//               chunk's partial sum is moved TMEM -> registers (tcgen05.ld, lane = tile row) and
//               summed there in round-to-nearest FP32, while the MMA warp already fills the other
//               accumulator.  At the end of a tile: clamp to [-1, 1], NaN -> 0, optional |r|,
//               diagonal = 1, store the tile and (symmetric case, off-diagonal tiles) its mirror.
#include <cuda.h>

#include <algorithm>
#include <vector>

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
constexpr int TMEM_COLUMNS = 256;  // two accumulators of BN columns
constexpr int CHUNK = 4;           // k-blocks accumulated in TMEM between two drains (K = 128)
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
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

struct UmmaParams {
    float* output;
    int64_t out_stride;
    int64_t rows_a, rows_b;
    const int2* tiles;  // (tile_i, tile_j) in processing order
    int tiles_count;
    int k_blocks;
    int symmetric;  // lower-triangle tiles, mirrored store, unit diagonal
    int absolute;   // store |r|
    unsigned int* min_key;  // optional: receives min over the stored values, as float_to_ordered keys (atomicMin)
};

__device__ __forceinline__ float finish(float r, int absolute) {
    r = (r != r) ? 0.0f : r;           // computation.py:465 maps NaN to 0
    r = fminf(fmaxf(r, -1.0f), 1.0f);  // correlate.cpp:199-200
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
    auto acc_full_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + b); };
    auto acc_empty_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + 2 + b); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunks_per_tile = (p.k_blocks + CHUNK - 1) / CHUNK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(acc_full_bar(b), 1);
            mbar_init(acc_empty_bar(b), 4);  // one arrival per drain warp
        }
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
            uint32_t it = 0;  // k-blocks issued so far by this CTA
            for (int t = blockIdx.x; t < p.tiles_count; t += gridDim.x) {
                const int2 tile = p.tiles[t];
                for (int kb = 0; kb < p.k_blocks; ++kb, ++it) {
                    const uint32_t s = it % STAGES;
                    if (it >= STAGES)
                        mbar_wait(empty_bar(s), ((it / STAGES) - 1) & 1);
                    const uint32_t stage = base + s * STAGE_BYTES;
                    mbar_expect_tx(full_bar(s), STAGE_BYTES);
                    tma_load_2d(stage + 0 * TILE_BYTES, &map_a_hi, full_bar(s), kb * BK, tile.x * BM);
                    tma_load_2d(stage + 1 * TILE_BYTES, &map_a_lo, full_bar(s), kb * BK, tile.x * BM);
                    tma_load_2d(stage + 2 * TILE_BYTES, &map_b_hi, full_bar(s), kb * BK, tile.y * BN);
                    tma_load_2d(stage + 3 * TILE_BYTES, &map_b_lo, full_bar(s), kb * BK, tile.y * BN);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            uint32_t it = 0, chunk_index = 0;
            for (int t = blockIdx.x; t < p.tiles_count; t += gridDim.x) {
                for (int kb0 = 0; kb0 < p.k_blocks; kb0 += CHUNK, ++chunk_index) {
                    const uint32_t buffer = chunk_index & 1;
                    if (chunk_index >= 2)
                        mbar_wait(acc_empty_bar(buffer), ((chunk_index >> 1) - 1) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t tmem_d = tmem_base + buffer * BN;
                    const int kb1 = min(kb0 + CHUNK, p.k_blocks);
                    for (int kb = kb0; kb < kb1; ++kb, ++it) {
                        const uint32_t s = it % STAGES;
                        mbar_wait(full_bar(s), (it / STAGES) & 1);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint32_t stage = base + s * STAGE_BYTES;
                        const uint64_t a_hi = umma_desc(stage + 0 * TILE_BYTES), a_lo = umma_desc(stage + 1 * TILE_BYTES);
                        const uint64_t b_hi = umma_desc(stage + 2 * TILE_BYTES), b_lo = umma_desc(stage + 3 * TILE_BYTES);
#pragma unroll
                        for (int k = 0; k < BK / 8; ++k) {
                            const uint64_t advance = (uint64_t)((k * 8 * 4) >> 4);  // bytes >> 4 inside the swizzle row
                            umma_tf32(tmem_d, a_lo + advance, b_hi + advance, (kb > kb0 || k > 0) ? 1u : 0u);
                            umma_tf32(tmem_d, a_hi + advance, b_lo + advance, 1u);
                            umma_tf32(tmem_d, a_hi + advance, b_hi + advance, 1u);
                        }
                        umma_commit(empty_bar(s));
                    }
                    umma_commit(acc_full_bar(buffer));
                }
            }
        }
    } else {
        const int quarter = warp & 3;  // TMEM lanes 32 * quarter .. + 31 are visible to this warp
        const int row_in_tile = quarter * 32 + lane;
        uint32_t chunk_index = 0;
        float least = __uint_as_float(0x7f800000u);  // min over the values this thread stores (np.min, knn_graph.py:290)
        for (int t = blockIdx.x; t < p.tiles_count; t += gridDim.x) {
            const int2 tile = p.tiles[t];
            float acc[BN];
#pragma unroll
            for (int c = 0; c < BN; ++c) acc[c] = 0.0f;
            for (int chunk = 0; chunk < chunks_per_tile; ++chunk, ++chunk_index) {
                const uint32_t buffer = chunk_index & 1;
                mbar_wait(acc_full_bar(buffer), (chunk_index >> 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int part = 0; part < BN / 16; ++part) {
                    uint32_t v[16];
                    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + buffer * BN + (uint32_t)(part * 16);
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                        : "r"(taddr)
                        : "memory");
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int c = 0; c < 16; ++c) acc[part * 16 + c] += __uint_as_float(v[c]);
                }
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0)
                    mbar_arrive(acc_empty_bar(buffer));
            }
            // tile epilogue
            const int64_t global_row = (int64_t)tile.x * BM + row_in_tile;
            const int64_t col0 = (int64_t)tile.y * BN;
            const bool mirror = p.symmetric && tile.x != tile.y;
            const bool row_ok = global_row < p.rows_a;
#pragma unroll
            for (int c = 0; c < BN; ++c) {
                float r = finish(acc[c], p.absolute);
                if (p.symmetric && global_row == col0 + c)
                    r = 1.0f;  // correlate.cpp:374-376
                acc[c] = r;
            }
            const bool diagonal_tile = p.symmetric && tile.x == tile.y;
            if (p.min_key != nullptr && row_ok) {
#pragma unroll
                for (int c = 0; c < BN; ++c)
                    if (col0 + c < p.rows_b && (!diagonal_tile || c <= row_in_tile))
                        least = fminf(least, acc[c]);
            }
            if (row_ok) {
                float* out_row = p.output + global_row * p.out_stride + col0;
                if (diagonal_tile) {
                    // (i, j) and (j, i) of a diagonal tile are accumulated in different term orders; keep
                    // the lower triangle only and mirror it, so the result is exactly symmetric.
                    float* mirror_at = p.output + col0 * p.out_stride + global_row;
#pragma unroll
                    for (int c = 0; c < BN; ++c) {
                        if (c <= row_in_tile)
                            out_row[c] = acc[c];
                        if (c < row_in_tile)
                            *mirror_at = acc[c];
                        mirror_at += p.out_stride;
                    }
                } else if (col0 + BN <= p.rows_b && ((reinterpret_cast<uintptr_t>(out_row) & 15) == 0)) {
#pragma unroll
                    for (int c = 0; c < BN; c += 4)
                        *reinterpret_cast<float4*>(out_row + c) = make_float4(acc[c], acc[c + 1], acc[c + 2], acc[c + 3]);
                } else {
#pragma unroll
                    for (int c = 0; c < BN; ++c)
                        if (col0 + c < p.rows_b)
                            out_row[c] = acc[c];
                }
            }
            if (mirror) {
                // lanes hold consecutive rows: for a fixed column the warp writes 128 contiguous bytes
                float* mirror_at = p.output + col0 * p.out_stride + global_row;
                const int valid = row_ok ? (int)min((int64_t)BN, p.rows_b - col0) : 0;
#pragma unroll
                for (int c = 0; c < BN; ++c) {
                    if (c < valid)
                        *mirror_at = acc[c];
                    mirror_at += p.out_stride;
                }
            }
        }
        if (p.min_key != nullptr) {
            uint32_t key = float_to_ordered(least);
            for (int offset = 16; offset > 0; offset >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, offset));
            if (lane == 0)
                atomicMin(p.min_key, key);
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

// Processing order of the output tiles: super-blocks of SUPER x SUPER tiles, so that the CTAs
// running at the same time share row panels through L2 instead of streaming the whole operand.
static void tile_order(int64_t tiles_a, int64_t tiles_b, int symmetric, std::vector<int2>& order) {
    const int64_t SUPER = 8;
    for (int64_t si = 0; si < tiles_a; si += SUPER)
        for (int64_t sj = 0; sj < (symmetric ? si + 1 : tiles_b); sj += SUPER)
            for (int64_t i = si; i < std::min(si + SUPER, tiles_a); ++i)
                for (int64_t j = sj; j < std::min(sj + SUPER, tiles_b); ++j)
                    if (!symmetric || j <= i)
                        order.push_back(make_int2((int)i, (int)j));
}

// a_* / b_*: prepared rows [rows, padded] (padded a multiple of 32).  For the symmetric case pass the
// same arrays for a and b.  output: [rows_a, out_stride >= rows_b].
int correlate_umma_launch(const float* a_hi, const float* a_lo, int64_t rows_a, const float* b_hi, const float* b_lo,
                          int64_t rows_b, int64_t padded, float* output, int64_t out_stride, int symmetric, int absolute,
                          unsigned int* min_key, cudaStream_t stream) {
    if (rows_a == 0 || rows_b == 0)
        return MCB_OK;
    MCB_REQUIRE(padded > 0 && padded % BK == 0, "correlate: padded width %lld is not a positive multiple of %d", (long long)padded, BK);
    CUtensorMap map_a_hi, map_a_lo, map_b_hi, map_b_lo;
    MCB_TRY(make_map(&map_a_hi, a_hi, rows_a, padded));
    MCB_TRY(make_map(&map_a_lo, a_lo, rows_a, padded));
    MCB_TRY(make_map(&map_b_hi, b_hi, rows_b, padded));
    MCB_TRY(make_map(&map_b_lo, b_lo, rows_b, padded));
    const int64_t tiles_a = ceil_div(rows_a, BM), tiles_b = ceil_div(rows_b, BN);
    MCB_REQUIRE(tiles_a * tiles_b < (1ll << 30), "correlate: too many tiles");
    std::vector<int2> order;
    tile_order(tiles_a, tiles_b, symmetric, order);
    Scratch tiles;
    MCB_TRY(tiles.alloc(order.size() * sizeof(int2), stream));
    // staged before the call returns (page-locked ring or the driver's own staging), so `order` may go out of scope
    MCB_TRY(upload_small(tiles.ptr, order.data(), order.size() * sizeof(int2), stream));
    UmmaParams p{};
    p.output = output;
    p.out_stride = out_stride;
    p.rows_a = rows_a;
    p.rows_b = rows_b;
    p.tiles = tiles.as<int2>();
    p.tiles_count = (int)order.size();
    p.k_blocks = (int)(padded / BK);
    p.symmetric = symmetric;
    p.absolute = absolute;
    p.min_key = min_key;
    const int grid = (int)std::min<int64_t>((int64_t)order.size(), sm_count());
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(correlate_umma_kernel), (size_t)UMMA_SMEM));
    MCB_PROFILED("correlate_umma", stream, correlate_umma_kernel<<<(unsigned)grid, UMMA_THREADS, UMMA_SMEM, stream>>>(map_a_hi, map_a_lo, map_b_hi, map_b_lo, p));
    MCB_LAUNCH_CHECK();
    return MCB_OK;
}

}  // namespace mcb
