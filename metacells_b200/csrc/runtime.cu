// Library runtime: error reporting, device selection and cached device properties.
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

}  // namespace mcb

extern "C" {

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
