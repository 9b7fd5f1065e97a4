// Shared helpers for the sm_100a kernels of libmcb200.so (error plumbing, warp/block primitives,
// monotone float keys).  Nothing in here is a public symbol; the C-ABI lives in include/mcb200.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>

#define MCB_OK 0
#define MCB_ERR_CUDA 1
#define MCB_ERR_ARG 2
#define MCB_ERR_UNSUPPORTED 3

namespace mcb {

// Thread-local last error message, read through mcb_last_error().
char* error_buffer();
int fail(int code, const char* format, ...);

#define MCB_CUDA(expr)                                                                              \
    do {                                                                                            \
        cudaError_t mcb_err__ = (expr);                                                             \
        if (mcb_err__ != cudaSuccess)                                                               \
            return ::mcb::fail(MCB_ERR_CUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr,          \
                               cudaGetErrorString(mcb_err__));                                      \
    } while (0)

// After every kernel launch: counts the launch (mcb_launch_count) and surfaces launch errors.
#define MCB_LAUNCH_CHECK()                                                                           \
    do {                                                                                            \
        ::mcb::count_launch();                                                                      \
        MCB_CUDA(cudaGetLastError());                                                               \
    } while (0)

// Runs a kernel launch statement inside a ProfileScope (timed only when mcb_profile_enable(1)).
#define MCB_PROFILED(name, stream, ...)                    \
    do {                                                   \
        ::mcb::ProfileScope mcb_scope__((name), (stream)); \
        __VA_ARGS__;                                       \
    } while (0)

#define MCB_TRY(expr)                  \
    do {                               \
        int mcb_code__ = (expr);       \
        if (mcb_code__ != MCB_OK)      \
            return mcb_code__;         \
    } while (0)

#define MCB_REQUIRE(cond, ...)                           \
    do {                                                 \
        if (!(cond))                                     \
            return ::mcb::fail(MCB_ERR_ARG, __VA_ARGS__); \
    } while (0)

// Once per device: keep freed scratch in the stream-ordered pool instead of returning it to the
// driver at every synchronisation (the default release threshold is 0).
int prepare_memory_pool();

// Device scratch that frees itself on the stream it was allocated on (stream-ordered pool).
struct Scratch {
    void* ptr = nullptr;
    cudaStream_t stream = nullptr;
    Scratch() = default;
    Scratch(const Scratch&) = delete;
    Scratch& operator=(const Scratch&) = delete;
    int alloc(size_t bytes, cudaStream_t s) {
        stream = s;
        MCB_TRY(::mcb::prepare_memory_pool());
        MCB_CUDA(cudaMallocAsync(&ptr, bytes ? bytes : 16, s));
        return MCB_OK;
    }
    template <typename T>
    T* as() const {
        return reinterpret_cast<T*>(ptr);
    }
    ~Scratch() {
        if (ptr)
            cudaFreeAsync(ptr, stream);
    }
};

// Device -> host copy of a result the caller waits for.  Waits for the stream FIRST, whatever the host memory is: a copy
// into pageable memory waits for the stream's earlier work inside the driver call, and a copy into page-locked memory
// that is queued behind a long kernel sits at the head of a copy-engine queue the uploads of other host threads' streams
// line up in; either way calls that should run side by side end up one after the other (measured: 16 optimisations in
// 11 s instead of 3.8 s).  Small results then go through page-locked staging (no second staging inside the driver).
int pinned_staging(void** buffer);  // 256 page-locked bytes of the calling host thread, or nullptr
static inline int copy_to_host(void* host, const void* device, size_t bytes, cudaStream_t stream) {
    MCB_CUDA(cudaStreamSynchronize(stream));
    void* staging = nullptr;
    if (bytes > 0 && bytes <= 256)
        MCB_TRY(pinned_staging(&staging));
    if (staging != nullptr) {
        MCB_CUDA(cudaMemcpyAsync(staging, device, bytes, cudaMemcpyDeviceToHost, stream));
        MCB_CUDA(cudaStreamSynchronize(stream));
        memcpy(host, staging, bytes);
        return MCB_OK;
    }
    if (bytes)
        MCB_CUDA(cudaMemcpyAsync(host, device, bytes, cudaMemcpyDeviceToHost, stream));
    return MCB_OK;
}

int upload_small(void* device, const void* host, size_t bytes, cudaStream_t stream);  // host -> device, not through the copy engine when small
int sm_count();
int allow_dynamic_shared(const void* kernel, size_t bytes);  // opt in to `bytes` of dynamic shared memory (only ever raised)
int thread_stream(cudaStream_t* stream);  // the calling host thread's own stream on the current device
int thread_side_stream(cudaStream_t* side, cudaEvent_t* fork, cudaEvent_t* join);  // that thread's second stream, with a fork and a join event
// Developer aid: with MCB_TRACE=1 in the environment, synchronises `stream` and prints the host time
// since the previous call on stderr.
void trace_lap(const char* name, cudaStream_t stream);
void count_launch();

// Optional per-kernel timing (mcb_profile_enable): CUDA events around a launch on its own stream.
struct ProfileScope {
    int slot = -1;
    cudaStream_t stream = nullptr;
    ProfileScope(const char* name, cudaStream_t stream);
    ~ProfileScope();
};

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

#ifdef __CUDACC__

// Monotone map float -> uint32: a < b (as floats, no NaN)  <=>  key(a) < key(b); -0.0 < +0.0.
__device__ __forceinline__ uint32_t float_to_ordered(float value) {
    uint32_t bits = __float_as_uint(value);
    return bits ^ ((bits & 0x80000000u) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t key) {
    uint32_t bits = key ^ ((key & 0x80000000u) ? 0x80000000u : 0xffffffffu);
    return __uint_as_float(bits);
}
// As std::less on floats treats -0.0 == +0.0, canonicalise zeros before building keys.
__device__ __forceinline__ float canonical_zero(float value) { return value == 0.0f ? 0.0f : value; }

__device__ __forceinline__ int warp_inclusive_scan(int value, int lane) {
#pragma unroll
    for (int offset = 1; offset < 32; offset <<= 1) {
        int other = __shfl_up_sync(0xffffffffu, value, offset);
        if (lane >= offset)
            value += other;
    }
    return value;
}

__device__ __forceinline__ int warp_reduce_sum(int value) {
#pragma unroll
    for (int offset = 16; offset > 0; offset >>= 1) value += __shfl_xor_sync(0xffffffffu, value, offset);
    return value;
}

// Exclusive block scan of one int per thread; `warp_sums` is shared scratch of >= 33 ints.
// Returns the exclusive prefix of this thread; *total receives the block total (all threads).
__device__ __forceinline__ int block_exclusive_scan(int value, int* warp_sums, int* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = (blockDim.x + 31) >> 5;
    int inclusive = warp_inclusive_scan(value, lane);
    if (lane == 31)
        warp_sums[warp] = inclusive;
    __syncthreads();
    if (warp == 0) {
        int sum = lane < warps ? warp_sums[lane] : 0;
        int scanned = warp_inclusive_scan(sum, lane);
        warp_sums[lane] = scanned - sum;
        if (lane == 31)
            warp_sums[32] = scanned;
    }
    __syncthreads();
    int result = warp_sums[warp] + inclusive - value;
    *total = warp_sums[32];
    __syncthreads();
    return result;
}

#endif  // __CUDACC__

}  // namespace mcb
