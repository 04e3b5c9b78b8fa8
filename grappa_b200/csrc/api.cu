// C-ABI plumbing: error string, launch counter, ABI version, host-side CSR builder.
#include "common.cuh"
#include <vector>

namespace grappa {
thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};
}  // namespace grappa

using namespace grappa;

extern "C" const char* grappa_last_error(void) { return g_err; }
extern "C" int grappa_abi_version(void) { return 2; }
extern "C" int64_t grappa_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

// Stable counting sort by key (host).  See include/grappa_b200.h.
extern "C" int grappa_csr_build(int64_t n, int64_t n_keys, const int64_t* keys, int32_t* ptr, int32_t* order) {
    GRAPPA_CHECK_ARG(n >= 0 && n_keys >= 0 && ptr && (n == 0 || (keys && order)), "bad arguments");
    GRAPPA_CHECK_ARG(n < 2147483647LL, "too many items for int32 offsets");
    for (int64_t k = 0; k <= n_keys; ++k) ptr[k] = 0;
    for (int64_t i = 0; i < n; ++i) {
        if (keys[i] < 0 || keys[i] >= n_keys) return fail("grappa_csr_build: %s key %lld out of range [0,%lld)", "", (long long)keys[i], (long long)n_keys);
        ptr[keys[i] + 1]++;
    }
    for (int64_t k = 0; k < n_keys; ++k) ptr[k + 1] += ptr[k];
    std::vector<int32_t> fill(ptr, ptr + n_keys);
    for (int64_t i = 0; i < n; ++i) order[fill[keys[i]]++] = (int32_t)i;
    return 0;
}
