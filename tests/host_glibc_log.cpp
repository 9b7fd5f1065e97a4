// CPU harness of tests/test_glibc_log.py: metacells_b200/csrc/glibc_log.cuh compiled as plain C++
// (g++ -O2 -mfma -ffp-contract=off), so the restated operation sequence can be pinned against libm
// without a GPU.  Test infrastructure only.
#include "../metacells_b200/csrc/glibc_log.cuh"

extern "C" void host_glibc_log(const double* x, double* y, long long n) {
    for (long long i = 0; i < n; ++i) y[i] = glibc_log::log(x[i]);
}
