// Version / error plumbing of the C ABI (include/quber_b200.h).
#include "common.cuh"

namespace quber {
static thread_local char g_last_cuda_error[512] = "";
void set_last_cuda_error(cudaError_t e, const char *where) {
    snprintf(g_last_cuda_error, sizeof(g_last_cuda_error), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
}
}  // namespace quber

extern "C" int quber_version(void) { return QUBER_ABI_VERSION; }

extern "C" const char *quber_error_string(int code) {
    switch (code) {
        case QUBER_OK: return "ok";
        case QUBER_E_INVALID: return "invalid argument";
        case QUBER_E_WORKSPACE: return "workspace missing, misaligned or too small";
        case QUBER_E_CUDA: return "CUDA runtime / launch failure";
        case QUBER_E_ALIGN: return "pointer or row pitch not aligned as documented";
        case QUBER_E_UNSUPPORTED: return "shape not supported by the kernels";
        default: return "unknown error code";
    }
}

extern "C" const char *quber_last_cuda_error(void) { return quber::g_last_cuda_error; }
