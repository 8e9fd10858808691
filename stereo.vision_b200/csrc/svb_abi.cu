// ABI bookkeeping: version and error strings.
#include "svb_common.cuh"

extern "C" int svb_version(void) { return 1; }

extern "C" const char *svb_error_string(int code) {
    switch (code) {
        case SVB_OK: return "ok";
        case SVB_E_NULL: return "a required pointer is NULL";
        case SVB_E_SHAPE: return "invalid shape (F >= 1, 2 <= H,W <= 2048, step >= 1, T,S >= 1)";
        case SVB_E_ARG: return "invalid argument";
        case SVB_E_SCRATCH: return "scratch buffer too small";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "unknown error";
}
