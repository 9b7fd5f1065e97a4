// Test helper: libstdc++'s own std::nth_element / std::sort over positions compared by key, for pinning the
// restatements in oracle/mc_seeds.c (and through it the device code).  Compiled by tests/test_oracle_seeds.py.
#include <algorithm>
#include <cstdint>

extern "C" void host_nth_element(int64_t* positions, const int64_t* keys, int64_t first, int64_t nth, int64_t last) {
    std::nth_element(positions + first, positions + nth, positions + last,
                     [&](int64_t left, int64_t right) { return keys[left] < keys[right]; });
}

extern "C" void host_sort(int64_t* positions, const int64_t* keys, int64_t count) {
    std::sort(positions, positions + count, [&](int64_t left, int64_t right) { return keys[left] < keys[right]; });
}
