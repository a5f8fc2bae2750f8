// Error buffer, device properties and ABI version for libb200vit.
#include "common.cuh"

#include <stdarg.h>
#include <string.h>

namespace b200vit {

static thread_local char g_err[512] = "";

char* last_error_buf() { return g_err; }

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

// SMs left to something else (a communication kernel running beside the compute stream). The persistent kernels of this
// library size their grids to sm_count() CTAs with a STATIC split of the work: a CTA that finds its SM occupied starts
// only when another CTA has finished its whole share, which doubles the kernel's run time -- so a concurrent kernel
// must get SMs of its own, not borrow ours. See b200vit_set_sm_reserve().
static int g_sm_reserve = 0;

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    int n = cached[dev] - g_sm_reserve;
    n &= ~1;                    // CTA pairs
    return n < 2 ? 2 : n;
}

}  // namespace b200vit

extern "C" {

int b200vit_abi_version(void) { return 1; }

int b200vit_set_sm_reserve(int sms) {
    if (sms < 0 || sms > 64) return b200vit::set_error(-1, "set_sm_reserve: %d is outside [0, 64]", sms);
    const int prev = b200vit::g_sm_reserve;
    b200vit::g_sm_reserve = sms;
    return prev >= 0 ? 0 : 0;
}

const char* b200vit_last_error(void) { return b200vit::last_error_buf(); }

}
