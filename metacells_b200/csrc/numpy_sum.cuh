// numpy_sum.cuh -- numpy's float32 `np.add.reduceat` of one segment, add for add.
//
// scipy's `csr_matrix.sum(axis=1)` (the `ut.sum_per` of tools/knn_graph.py:399) is `np.add.reduceat(data, indptr[rows])`:
// the segment's first value plus numpy's pairwise sum of the rest (numpy/_core/src/umath/loops_utils.h.src,
// @TYPE@_pairwise_sum: fewer than 8 values one after the other from 0, up to 128 values through 8 accumulators combined as
// ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) and the rest one after the other, more than that split at n/2 rounded down to a
// multiple of 8).  The order only matters when the sum is not exact in float32: row totals of balanced ranks (multiples of
// 1/8) are exact below 2^21, which graphs with k > ~1000 pass.  Host and device share this code so that a CPU test can pin it
// against numpy itself (mcb_host_numpy_add_reduce_f32, tests/test_numpy_sum.py).
#pragma once

#if defined(__CUDA_ARCH__)
#define MCB_NUMPY_ADD(a, b) __fadd_rn((a), (b))
#else
#define MCB_NUMPY_ADD(a, b) ((a) + (b))
#endif

#if defined(__CUDACC__)
#define MCB_NUMPY_HD __host__ __device__
#else
#define MCB_NUMPY_HD
#endif

namespace mcb {

// up to 128 values: numpy's unrolled block
MCB_NUMPY_HD inline float numpy_block_sum_f32(const float* a, int n) {
    if (n < 8) {
        float sum = 0.0f;
        for (int i = 0; i < n; ++i) sum = MCB_NUMPY_ADD(sum, a[i]);
        return sum;
    }
    float r[8];
    for (int j = 0; j < 8; ++j) r[j] = a[j];
    int i = 8;
    for (; i < n - (n % 8); i += 8)
        for (int j = 0; j < 8; ++j) r[j] = MCB_NUMPY_ADD(r[j], a[i + j]);
    float sum = MCB_NUMPY_ADD(MCB_NUMPY_ADD(MCB_NUMPY_ADD(r[0], r[1]), MCB_NUMPY_ADD(r[2], r[3])),
                              MCB_NUMPY_ADD(MCB_NUMPY_ADD(r[4], r[5]), MCB_NUMPY_ADD(r[6], r[7])));
    for (; i < n; ++i) sum = MCB_NUMPY_ADD(sum, a[i]);
    return sum;
}

// numpy's recursion (left half, right half, their sum) with a stack of its own: device code keeps a known frame size
MCB_NUMPY_HD inline float numpy_pairwise_sum_f32(const float* a, int n) {
    struct Frame {
        int start, count, phase;
        float left;
    };
    Frame stack[32];
    int top = 0;
    float value = 0.0f;  // what the frame that just finished returned
    stack[top++] = Frame{0, n, 0, 0.0f};
    while (top > 0) {
        Frame& frame = stack[top - 1];
        int half = frame.count / 2;
        half -= half % 8;
        if (frame.phase == 0) {
            if (frame.count <= 128) {
                value = numpy_block_sum_f32(a + frame.start, frame.count);
                --top;
            } else {
                frame.phase = 1;
                stack[top++] = Frame{frame.start, half, 0, 0.0f};
            }
        } else if (frame.phase == 1) {
            frame.left = value;
            frame.phase = 2;
            stack[top++] = Frame{frame.start + half, frame.count - half, 0, 0.0f};
        } else {
            value = MCB_NUMPY_ADD(frame.left, value);
            --top;
        }
    }
    return value;
}

// np.add.reduceat over one segment of n >= 1 values
MCB_NUMPY_HD inline float numpy_add_reduce_f32(const float* a, int n) {
    return n == 1 ? a[0] : MCB_NUMPY_ADD(a[0], numpy_pairwise_sum_f32(a + 1, n - 1));
}

}  // namespace mcb
