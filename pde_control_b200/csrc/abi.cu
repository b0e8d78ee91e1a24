// abi.cu -- library-level C-ABI entry points (version, error strings, device probe).
#include "pdec_common.cuh"

extern "C" int pdec_version(void) { return PDEC_VERSION; }

extern "C" const char* pdec_error_string(int code) {
    switch (code) {
        case PDEC_OK: return "ok";
        case PDEC_ERR_INVALID_ARGUMENT: return "invalid argument";
        case PDEC_ERR_UNSUPPORTED_SIZE: return "unsupported size (row or tile does not fit the kernel's shared-memory plan)";
        case PDEC_ERR_CUDA_LAUNCH: return "CUDA launch failed";
        case PDEC_ERR_WORKSPACE_TOO_SMALL: return "workspace too small";
        case PDEC_ERR_NO_DEVICE: return "no CUDA device";
        case PDEC_ERR_NOT_CONVERGED: return "solver did not converge";
        case PDEC_ERR_STATIC_SCRATCH_TOO_SMALL: return "B * T exceeds PDEC_STATIC_MAX_PAIRS, the capacity of the library's static scratch array";
        default: return "unknown error";
    }
}

extern "C" int pdec_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();  // clear the sticky "no device" error
        return 0;
    }
    return n;
}
