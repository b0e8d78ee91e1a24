// pdec_burgers_pair_loss_static: the fused Burgers force loss + gradient for callers that cannot allocate
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has no counterpart
// in the reference.
//
// The loss partials and the "CTAs done" counter of the existing kernels live in a file-scope device array of
// PDEC_STATIC_SCRATCH_BYTES instead of a caller-owned workspace.  Every GPU that loads libpdec.so holds that array
// until the process exits.  All calls on a device share it: calls that overlap on the GPU (two streams) race.
#include <cuda_runtime.h>

#include <mutex>

#include "../../include/pdec.h"

// 1 MiB of loss partials (B * T <= 131072 frame pairs of float2: BASELINE config 2, s[33, 4096, 128], exactly) + the counter block
static __device__ __align__(256) unsigned char g_static_scratch[PDEC_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_mutex;
void* g_scratch_ptr[kMaxDevices] = {};       // device address of g_static_scratch on each device (0: not looked up yet)

// In host code `g_static_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
int scratch_address(void** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return PDEC_ERR_NO_DEVICE;
    if (dev < 0 || dev >= kMaxDevices) return PDEC_ERR_UNSUPPORTED_SIZE;
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_ptr[dev]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (e != cudaSuccess || !ptr) {
            (void)cudaGetLastError();
            return PDEC_ERR_CUDA_LAUNCH;
        }
        g_scratch_ptr[dev] = ptr;
    }
    *out = g_scratch_ptr[dev];
    return PDEC_OK;
}

}  // namespace

extern "C" int pdec_burgers_pair_loss_static(const float* s, float* grad_s, float* losses, int T, int B, int n, int mode, double dt,
                                             double viscosity, float box_lower, float box_size, int substeps, int resubstep,
                                             float gl2, float gl1, void* stream) {
    if (B < 1 || T < 1 || !s || !losses) return PDEC_ERR_INVALID_ARGUMENT;
    // the limit, checked before anything is enqueued: B * T frame pairs must fit the static array
    if ((unsigned long long)B * (unsigned long long)T > (unsigned long long)PDEC_STATIC_MAX_PAIRS) return PDEC_ERR_STATIC_SCRATCH_TOO_SMALL;
    const size_t need = pdec_burgers_pair_loss_workspace_bytes(B, n, T);
    if (need > (size_t)PDEC_STATIC_SCRATCH_BYTES) return PDEC_ERR_STATIC_SCRATCH_TOO_SMALL;
    void* scratch = nullptr;
    const int rc = scratch_address(&scratch);
    if (rc != PDEC_OK) return rc;
    // The array keeps the previous call's contents.  What must start at zero is the "CTAs done" counter behind the partials:
    // the existing entry clears it on `stream` (cudaMemsetAsync: no allocation) before the launch; every partial the
    // reduction reads is written by this call's kernel first.
    return pdec_burgers_pair_loss(s, grad_s, losses, T, B, n, mode, dt, viscosity, box_lower, box_size, substeps, resubstep, gl2, gl1,
                                  scratch, (size_t)PDEC_STATIC_SCRATCH_BYTES, stream);
}
