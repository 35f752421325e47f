#include "host.h"
#include <cstdlib>
#include <atomic>
#include <stdlib.h>
#include <string.h>

namespace b200 {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
    static int cached = 0;
    if (cached == 0) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
            cached = n;
        else
            cached = 148;
        // B200SD_SM_LIMIT=<k>: size every persistent grid for k SMs instead of all of them, leaving the rest to kernels
        // of other streams (the NCCL all-reduce of the data-parallel gradient exchange, tools/overlap_probe.py)
        if (const char* e = getenv("B200SD_SM_LIMIT")) {
            const int k = atoi(e) & ~1;
            if (k >= 8 && k < cached) cached = k;
        }
    }
    return cached;
}

bool pdl_enabled() {
    static int cached = -1;
    if (cached < 0) {
        const char* e = getenv("B200SD_PDL");
        cached = (e && e[0] == '1') ? 1 : 0;   // measured on B200: no gain inside CUDA graphs, so opt-in
    }
    return cached == 1;
}

}  // namespace b200

extern "C" {
const char* b200sd_last_error(void) { return b200::g_err; }
int b200sd_version(void) { return 100; }
int64_t b200sd_launch_count(void) { return b200::g_launches.load(std::memory_order_relaxed); }
}
