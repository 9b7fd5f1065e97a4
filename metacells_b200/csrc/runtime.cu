// Library runtime: error reporting, device selection and cached device properties.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "../../include/mcb200.h"

namespace mcb {

char* error_buffer() {
    static thread_local char buffer[1024] = {0};
    return buffer;
}

int fail(int code, const char* format, ...) {
    va_list args;
    va_start(args, format);
    vsnprintf(error_buffer(), 1024, format, args);
    va_end(args);
    return code;
}

int sm_count() {
    static int cached = 0;
    if (cached == 0) {
        int device = 0;
        if (cudaGetDevice(&device) != cudaSuccess ||
            cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || cached <= 0)
            cached = 148;
    }
    return cached;
}

namespace {
std::atomic<long long> g_launches{0};
std::atomic<int> g_profile{0};
struct ProfileRecord {
    std::string name;
    cudaEvent_t start, stop;
};
std::mutex g_profile_mutex;
std::vector<ProfileRecord> g_records;
}  // namespace

int prepare_memory_pool() {
    static std::atomic<unsigned long long> prepared{0};  // bit per device
    int device = 0;
    MCB_CUDA(cudaGetDevice(&device));
    const unsigned long long bit = 1ull << (device & 63);
    if (prepared.load(std::memory_order_relaxed) & bit)
        return MCB_OK;
    cudaMemPool_t pool;
    MCB_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    unsigned long long threshold = ~0ull;
    MCB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold));
    // Piles of different host threads run on streams of their own: never make one stream wait for another just to take
    // over a block it freed (the pool grows a little instead).
    int no = 0;
    MCB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolReuseAllowInternalDependencies, &no));
    prepared.fetch_or(bit);
    return MCB_OK;
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per kernel and device, and again only to raise it: setting
// the attribute while instances of the kernel run waits for them, which queues up calls that would otherwise run side
// by side (one CTA each, on their own streams).
int allow_dynamic_shared(const void* kernel, size_t bytes) {
    static std::mutex mutex;
    static std::vector<std::pair<std::pair<int, const void*>, size_t>> allowed;
    int device = 0;
    MCB_CUDA(cudaGetDevice(&device));
    std::lock_guard<std::mutex> lock(mutex);
    for (auto& entry : allowed)
        if (entry.first.first == device && entry.first.second == kernel) {
            if (entry.second >= bytes)
                return MCB_OK;
            MCB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes)));
            entry.second = bytes;
            return MCB_OK;
        }
    MCB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes)));
    allowed.push_back({{device, kernel}, bytes});
    return MCB_OK;
}

// 256 page-locked bytes per host thread for the small read-backs (counts, flags) every stage ends with.  One block for
// the whole process, allocated at the first call (page-locking memory waits for running kernels, so worker threads must
// not each do it while other threads' optimisations run); a thread takes the next slot, and a process that has used all
// of them gets nullptr: its read-backs go through pageable memory.
int pinned_staging(void** buffer) {
    constexpr size_t SLOT = 256, SLOTS = 16384;
    static std::mutex mutex;
    static unsigned char* block = nullptr;
    static bool tried = false;
    static std::atomic<size_t> next{0};
    static thread_local void* mine = nullptr;
    static thread_local bool asked = false;
    if (!asked) {
        asked = true;
        static const bool disabled = getenv("MCB_NO_STAGING") != nullptr;  // developer aid: read back through pageable memory
        if (disabled) {
            *buffer = nullptr;
            return MCB_OK;
        }
        {
            std::lock_guard<std::mutex> lock(mutex);
            if (!tried) {
                tried = true;
                void* allocated = nullptr;
                if (cudaHostAlloc(&allocated, SLOT * SLOTS, cudaHostAllocPortable) == cudaSuccess)
                    block = static_cast<unsigned char*>(allocated);
                else
                    cudaGetLastError();  // not fatal: the pageable path remains
            }
        }
        const size_t slot = next.fetch_add(1);
        if (block != nullptr && slot < SLOTS)
            mine = block + slot * SLOT;
    }
    *buffer = mine;
    return MCB_OK;
}

// Host -> device for the small arrays a stage takes from the host (gene lists, orders, tile lists, flags) WITHOUT the copy
// engine.  The engine serves copies in the order they were submitted: once the next pile's CSR (hundreds of MB, a stream of
// its own) is in its queue, a 4 KB cudaMemcpyAsync of the current pile's stream waits behind all of it, and the current
// pile's kernels with it -- upload and compute end up one after the other (measured: 16 ms per pile instead of 9).  So the
// bytes go into a page-locked slice of the calling host thread and a kernel of the stream reads them over the link.
// The slice is a ring; before it wraps the thread waits for the last kernel that read from it (long done by then).
namespace {
__global__ void upload_small_kernel(uint32_t* __restrict__ device, const uint32_t* __restrict__ host, size_t words) {
    for (size_t at = blockIdx.x * size_t(blockDim.x) + threadIdx.x; at < words; at += size_t(gridDim.x) * blockDim.x)
        device[at] = host[at];
}
}  // namespace

int upload_small(void* device, const void* host, size_t bytes, cudaStream_t stream) {
    constexpr size_t SLICE = size_t(1) << 20, SLICES = 64, LIMIT = SLICE / 2;
    struct Ring {
        unsigned char* slice = nullptr;
        size_t cursor = 0;
        cudaEvent_t last = nullptr;
        cudaStream_t stream = nullptr;
        int device = -1;  // the device `last` belongs to
        bool pending = false, asked = false;
    };
    static std::mutex mutex;
    static unsigned char* block = nullptr;
    static bool tried = false;
    static std::atomic<size_t> next{0};
    static thread_local Ring ring;
    if (bytes == 0)
        return MCB_OK;
    static const bool disabled = getenv("MCB_NO_SMALL_UPLOADS") != nullptr;  // developer aid: everything through the copy engine
    const bool fits = !disabled && bytes <= LIMIT && bytes % 4 == 0 && reinterpret_cast<uintptr_t>(device) % 4 == 0;
    if (fits && !ring.asked) {
        ring.asked = true;
        {
            std::lock_guard<std::mutex> lock(mutex);
            if (!tried) {
                tried = true;
                void* allocated = nullptr;
                if (cudaHostAlloc(&allocated, SLICE * SLICES, cudaHostAllocPortable | cudaHostAllocMapped) == cudaSuccess)
                    block = static_cast<unsigned char*>(allocated);
                else
                    cudaGetLastError();  // not fatal: the copy engine remains
            }
        }
        const size_t slot = next.fetch_add(1);
        if (block != nullptr && slot < SLICES)
            ring.slice = block + slot * SLICE;
    }
    if (fits && ring.slice != nullptr) {  // the event is the current device's: a thread that moved to another GPU gets a new one
        int device = 0;
        MCB_CUDA(cudaGetDevice(&device));
        if (device != ring.device) {
            if (ring.pending)
                MCB_CUDA(cudaEventSynchronize(ring.last));
            if (ring.last != nullptr)
                cudaEventDestroy(ring.last);
            ring.last = nullptr;
            ring.pending = false;
            ring.cursor = 0;
            MCB_CUDA(cudaEventCreateWithFlags(&ring.last, cudaEventDisableTiming));
            ring.device = device;
        }
    }
    if (!fits || ring.slice == nullptr) {
        MCB_CUDA(cudaMemcpyAsync(device, host, bytes, cudaMemcpyHostToDevice, stream));
        return MCB_OK;
    }
    const size_t rounded = (bytes + 255) / 256 * 256;
    if (ring.pending && (ring.stream != stream || ring.cursor + rounded > SLICE)) {
        MCB_CUDA(cudaEventSynchronize(ring.last));  // every kernel that read the ring was launched on ring.stream before it
        ring.pending = false;
        ring.cursor = 0;
    }
    if (ring.cursor + rounded > SLICE)
        ring.cursor = 0;
    unsigned char* staged = ring.slice + ring.cursor;
    memcpy(staged, host, bytes);
    const size_t words = bytes / 4;
    const int blocks = int(std::min<size_t>((words + 255) / 256, 64));
    MCB_PROFILED("upload_small", stream, upload_small_kernel<<<blocks, 256, 0, stream>>>(
                     static_cast<uint32_t*>(device), reinterpret_cast<const uint32_t*>(staged), words));
    MCB_LAUNCH_CHECK();
    MCB_CUDA(cudaEventRecord(ring.last, stream));
    ring.cursor += rounded;
    ring.stream = stream;
    ring.pending = true;
    return MCB_OK;
}

// A stream of the calling host thread's own (per device): host entry points that may be called from several
// threads at once (one pile each) run side by side instead of queueing on the legacy default stream.  The stream
// outlives its thread on purpose: piles and graphs created by a worker thread keep using it (their scratch is freed on
// it) after the thread has gone; keep worker threads in a pool rather than spawning one per call.
int thread_stream(cudaStream_t* stream) {
    static thread_local cudaStream_t streams[64] = {};
    int device = 0;
    MCB_CUDA(cudaGetDevice(&device));
    cudaStream_t& mine = streams[device & 63];
    if (mine == nullptr)
        MCB_CUDA(cudaStreamCreateWithFlags(&mine, cudaStreamNonBlocking));
    *stream = mine;
    return MCB_OK;
}

// A second stream of the calling host thread with a fork and a join event: a short kernel that would otherwise be a
// latency tail of its own runs beside the kernels that follow it on the main stream.
int thread_side_stream(cudaStream_t* side, cudaEvent_t* fork, cudaEvent_t* join) {
    struct Side {
        cudaStream_t stream = nullptr;
        cudaEvent_t fork = nullptr, join = nullptr;
    };
    static thread_local Side sides[64];
    int device = 0;
    MCB_CUDA(cudaGetDevice(&device));
    Side& mine = sides[device & 63];
    if (mine.stream == nullptr) {
        MCB_CUDA(cudaStreamCreateWithFlags(&mine.stream, cudaStreamNonBlocking));
        MCB_CUDA(cudaEventCreateWithFlags(&mine.fork, cudaEventDisableTiming));
        MCB_CUDA(cudaEventCreateWithFlags(&mine.join, cudaEventDisableTiming));
    }
    *side = mine.stream;
    *fork = mine.fork;
    *join = mine.join;
    return MCB_OK;
}

void trace_lap(const char* name, cudaStream_t stream) {
    static const char* mode = getenv("MCB_TRACE");  // "host": the host's own time between laps, no synchronisation
    if (mode == nullptr)
        return;
    static thread_local std::chrono::steady_clock::time_point last = std::chrono::steady_clock::now();
    if (strcmp(mode, "host") != 0)
        cudaStreamSynchronize(stream);
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[mcb trace] %-28s %8.3f ms\n", name, std::chrono::duration<double, std::milli>(now - last).count());
    last = now;
}

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

ProfileScope::ProfileScope(const char* name, cudaStream_t s) : stream(s) {
    if (!g_profile.load(std::memory_order_relaxed))
        return;
    ProfileRecord record;
    record.name = name;
    if (cudaEventCreate(&record.start) != cudaSuccess || cudaEventCreate(&record.stop) != cudaSuccess)
        return;
    cudaEventRecord(record.start, stream);
    std::lock_guard<std::mutex> lock(g_profile_mutex);
    g_records.push_back(record);
    slot = (int)g_records.size() - 1;
}

ProfileScope::~ProfileScope() {
    if (slot < 0)
        return;
    std::lock_guard<std::mutex> lock(g_profile_mutex);
    if (slot < (int)g_records.size())
        cudaEventRecord(g_records[slot].stop, stream);
}

}  // namespace mcb

extern "C" {

long long mcb_launch_count(void) { return mcb::g_launches.load(); }

// Grows the device's stream-ordered pool to hold `bytes` of scratch now (one allocation, freed again: the pool keeps
// it), so that calls running side by side later do not each stop to take memory from the driver while kernels run.
int mcb_reserve_scratch(long long bytes) {
    MCB_REQUIRE(bytes >= 0, "mcb_reserve_scratch: negative size");
    if (bytes == 0)
        return MCB_OK;
    cudaStream_t stream = nullptr;
    MCB_TRY(mcb::thread_stream(&stream));
    {
        mcb::Scratch block;
        MCB_TRY(block.alloc(size_t(bytes), stream));
    }
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

// The CUDA device of the calling thread (a new host thread starts on device 0: workers of a process that drives
// another GPU select it with mcb_set_device first).
int mcb_current_device(int* device) {
    MCB_REQUIRE(device != nullptr, "mcb_current_device: null argument");
    MCB_CUDA(cudaGetDevice(device));
    return MCB_OK;
}

int mcb_device_sm_count(int* count) {
    MCB_REQUIRE(count != nullptr, "mcb_device_sm_count: null argument");
    *count = mcb::sm_count();
    return *count > 0 ? MCB_OK : mcb::fail(MCB_ERR_CUDA, "mcb_device_sm_count: no device");
}

int mcb_profile_enable(int enable) {
    mcb::g_profile.store(enable ? 1 : 0);
    return MCB_OK;
}

// Sum of the recorded durations of `name` (milliseconds) and their count; synchronises the device.
int mcb_profile_read(const char* name, double* total_ms, long long* count) {
    MCB_CUDA(cudaDeviceSynchronize());
    std::lock_guard<std::mutex> lock(mcb::g_profile_mutex);
    double total = 0;
    long long n = 0;
    for (auto& record : mcb::g_records)
        if (record.name == name) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, record.start, record.stop) == cudaSuccess) {
                total += ms;
                ++n;
            }
        }
    if (total_ms) *total_ms = total;
    if (count) *count = n;
    return MCB_OK;
}

// Names seen so far, separated by newlines, into buffer (truncated to size - 1 characters).
int mcb_profile_names(char* buffer, long long size) {
    std::lock_guard<std::mutex> lock(mcb::g_profile_mutex);
    std::string names;
    for (auto& record : mcb::g_records)
        if (("\n" + names).find("\n" + record.name + "\n") == std::string::npos)
            names += record.name + "\n";
    if (buffer && size > 0) {
        const size_t n = std::min<size_t>(names.size(), (size_t)size - 1);
        memcpy(buffer, names.data(), n);
        buffer[n] = 0;
    }
    return MCB_OK;
}

int mcb_profile_reset(void) {
    MCB_CUDA(cudaDeviceSynchronize());
    std::lock_guard<std::mutex> lock(mcb::g_profile_mutex);
    for (auto& record : mcb::g_records) {
        cudaEventDestroy(record.start);
        cudaEventDestroy(record.stop);
    }
    mcb::g_records.clear();
    return MCB_OK;
}


const char* mcb_last_error(void) { return mcb::error_buffer(); }

int mcb_version(void) { return 1; }

int mcb_device_count(void) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess)
        return 0;
    return count;
}

int mcb_set_device(int device) {
    MCB_CUDA(cudaSetDevice(device));
    int major = 0, minor = 0;
    MCB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    MCB_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
    if (major != 10)
        return mcb::fail(MCB_ERR_UNSUPPORTED, "libmcb200 is built for sm_100a only; device %d is sm_%d%d", device, major,
                         minor);
    return MCB_OK;
}

int mcb_synchronize(void* stream) {
    MCB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return MCB_OK;
}

}  // extern "C"
