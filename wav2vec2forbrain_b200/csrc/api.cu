// error plumbing + version + device check of libb2t_sm100.so
#include "common.cuh"
#include "../../include/b2t_sm100.h"
#include <stdarg.h>

static thread_local char g_err[512] = "";

void b2t_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* b2t_last_error(void) { return g_err; }
extern "C" int b2t_abi_version(void) { return 1; }

extern "C" int b2t_check_device(int dev) {
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, dev);
    if (e != cudaSuccess) {
        b2t_set_error("b2t_check_device: %s", cudaGetErrorString(e));
        return B2T_EARG;
    }
    if (prop.major != 10) {
        b2t_set_error("b2t_check_device: device %d is sm_%d%d, this library is sm_100a only (no fallback)", dev, prop.major,
                      prop.minor);
        return B2T_EARG;
    }
    return 0;
}
