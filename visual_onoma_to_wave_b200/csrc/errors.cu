// error message storage + trivial entry points
#include "common.cuh"

static thread_local char g_err[512] = "";

void vo2w_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* vo2w_last_error(void) { return g_err; }
extern "C" int vo2w_version(void) { return 100; }

int vo2w_sm_count() {
    static int n_sms[VO2W_MAX_DEV];
    const int dev = vo2w_cur_dev();
    if (!n_sms[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        n_sms[dev] = n;
    }
    return n_sms[dev];
}

extern "C" int vo2w_device_check(void) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        vo2w_set_error("no CUDA device");
        return VO2W_E_NODEVICE;
    }
    static int checked[64];                       // cudaGetDeviceProperties costs milliseconds: ask once per device
    if (dev >= 0 && dev < 64 && checked[dev]) return VO2W_OK;
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess || major != 10) {
        vo2w_set_error("device is not sm_100 (compute capability 10.x)");
        return VO2W_E_NODEVICE;
    }
    if (dev >= 0 && dev < 64) checked[dev] = 1;
    return VO2W_OK;
}
