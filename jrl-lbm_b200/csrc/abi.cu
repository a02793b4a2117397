// libjrl_lbm.so: version, per-thread error message, device queries.
#include <cstring>

#include "common.cuh"

namespace jrl {

static thread_local char g_error[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int sm_count()
{
    static int n[64] = {};  // per device: a process may drive several GPUs (jrl_native.device_of)
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
        return 148;
    if (n[dev] == 0 && cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
        n[dev] = 148;
    return n[dev];
}

}  // namespace jrl

extern "C" {

int jrl_version(void) { return JRL_LBM_VERSION; }

const char *jrl_last_error(void) { return jrl::g_error; }

}  // extern "C"
