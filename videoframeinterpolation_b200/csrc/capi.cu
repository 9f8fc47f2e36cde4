// Library-wide entry points: version, error string, launch counter.
#include <stdarg.h>

#include "vfi_common.cuh"

namespace vfi {

static thread_local char g_err[512] = "ok";
static thread_local int64_t g_launches = 0;

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void note_launch() { ++g_launches; }

}  // namespace vfi

extern "C" {

int vfi_version(void) { return VFI_B200_VERSION; }

const char* vfi_last_error_string(void) { return vfi::g_err; }

int vfi_device_supported(void) {
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
    return major == 10 ? 1 : 0;
}

int64_t vfi_launch_count(void) { return vfi::g_launches; }
void vfi_launch_count_reset(void) { vfi::g_launches = 0; }

}  // extern "C"
