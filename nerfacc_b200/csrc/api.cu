// api.cu — version, error text and launch accounting of libnerfacc_b200.
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace nfb {
static thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}
}  // namespace nfb

extern "C" {
NFB_API int nfb_version(void) { return 100; /* 0.1.0 */ }
NFB_API int nfb_compiled_arch(void) { return 100; }
NFB_API const char* nfb_last_error(void) { return nfb::g_err; }
NFB_API long long nfb_launch_count(void) { return nfb::g_launches.load(std::memory_order_relaxed); }
}
