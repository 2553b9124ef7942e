// ABI bookkeeping for libweclip_sm100.so.
#include "common.cuh"

namespace wc {
thread_local char g_err[512] = "";
}

extern "C" int wc_abi_version(void) { return WC_ABI_VERSION; }
extern "C" const char* wc_last_error(void) { return wc::g_err; }
extern "C" int wc_device_ok(void) {
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
    return major == 10;
}
