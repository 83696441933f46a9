// Error plumbing and ABI version of libdvf_b200.
#include "common.cuh"

namespace dvf {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what, int line) {
    set_error("CUDA error '%s' (%d) in %s at line %d", cudaGetErrorString(e), (int)e, what, line);
    return DVF_ERR_CUDA;
}
}  // namespace dvf

extern "C" int dvf_abi_version(void) { return 1; }
extern "C" const char* dvf_last_error(void) { return dvf::g_err; }
