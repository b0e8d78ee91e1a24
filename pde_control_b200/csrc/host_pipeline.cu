// host_pipeline.cu -- the force loss fed from HOST arrays (what the reference's session.run(feed_dict=...) does every
// optimisation step, src/control/control_training.py:113-142 -> PF/tf/session.py): the batch is cut into row chunks and the
// three legs of every chunk -- H2D of the frames, the fused pair-loss kernel, D2H of the gradient -- run on three streams,
// so PCIe is busy in both directions at once and the kernel hides behind the copies.  The frame-major host layout
// [T+1, B, n] makes a chunk of rows a pitched 2-D region; cudaMemcpy2DAsync moves it with the copy engines, no packing pass.
//
// When the fused unit-box kernel applies (pdec_burgers_pair_loss_window_supported) and the pipeline was created with
// frames_per_slice > 0, the batch is cut along TIME instead: slice k = pairs [k m, (k+1) m) of ALL rows.  Its frames are one
// contiguous block on the host and on the device, so both PCIe legs are plain large 1-D copies (measured: pitched D2H pieces
// reach ~35 GB/s, contiguous ones 49 GB/s with both directions busy, scripts/diag/pcie_probe.py); the kernel of slice k needs
// only frames <= (k+1) m and completes the gradient of frames [k m, (k+1) m).
//
// A pipeline object owns its streams, events and device staging (allocated once at create time); the call itself allocates
// nothing, never synchronises the device and is ordered on the caller's stream like every other entry of the library: work
// enqueued on `stream` before the call happens-before the first copy, and `stream` waits for the last copy.
#include <cstdlib>
#include <new>
#include "pdec_common.cuh"

namespace {

constexpr int SLOTS = 3;

struct Slot {
    float* s = nullptr;          // [T+1, rows, n] staging of the frames
    float* g = nullptr;          // [T+1, rows, n] staging of the gradient
    cudaEvent_t h2d = nullptr, kernel = nullptr, d2h = nullptr;
};

}  // namespace

struct pdec_host_pipeline {
    int rows, n, T;
    int max_chunks;
    Slot slot[SLOTS];
    cudaStream_t in = nullptr, compute = nullptr, out = nullptr;
    cudaEvent_t begin = nullptr, end = nullptr;
    void* ws = nullptr;
    size_t ws_bytes = 0;
    float* chunk_losses = nullptr;   // [max(max_chunks, T)][2] device
    float* losses = nullptr;         // [2] device
    // frame-sliced mode
    int slice = 0;                   // pairs per slice (0: mode not available)
    long long max_batch = 0;
    float* full_s = nullptr;         // [T+1, max_batch, n]
    float* full_g = nullptr;
    cudaEvent_t* ev_h2d = nullptr;   // [T]
    cudaEvent_t* ev_kernel = nullptr;
};

namespace pdec {

__host__ __device__ inline float chunk_scale(int c, int nchunks, int rows, int B) {      // rows of chunk c / B
    return (float)((double)((c + 1 < nchunks) ? rows : B - c * rows) / (double)B);
}

// L = sum_c scale_c * L_c in ascending chunk order (one thread: the order is the definition of the result)
__global__ void combine_chunk_losses_kernel(const float* __restrict__ chunk_losses, int nchunks, int rows, int B,
                                            float* __restrict__ losses) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float l2 = 0.0f, l1 = 0.0f;
        for (int c = 0; c < nchunks; ++c) {
            const float sc = rows > 0 ? chunk_scale(c, nchunks, rows, B) : 1.0f;      // rows == 0: time slices, already per B
            l2 += chunk_losses[2 * c] * sc;
            l1 += chunk_losses[2 * c + 1] * sc;
        }
        losses[0] = l2;
        losses[1] = l1;
    }
}

}  // namespace pdec

#define HP_TRY(expr)                                                   \
    do {                                                               \
        if ((expr) != cudaSuccess) { cudaGetLastError(); goto fail; }  \
    } while (0)

extern "C" int pdec_host_pipeline_destroy(pdec_host_pipeline* h) {
    if (!h) return PDEC_OK;
    for (int i = 0; i < SLOTS; ++i) {
        if (h->slot[i].s) cudaFree(h->slot[i].s);
        if (h->slot[i].g) cudaFree(h->slot[i].g);
        if (h->slot[i].h2d) cudaEventDestroy(h->slot[i].h2d);
        if (h->slot[i].kernel) cudaEventDestroy(h->slot[i].kernel);
        if (h->slot[i].d2h) cudaEventDestroy(h->slot[i].d2h);
    }
    if (h->full_s) cudaFree(h->full_s);
    if (h->full_g) cudaFree(h->full_g);
    for (int i = 0; i < h->T; ++i) {
        if (h->ev_h2d && h->ev_h2d[i]) cudaEventDestroy(h->ev_h2d[i]);
        if (h->ev_kernel && h->ev_kernel[i]) cudaEventDestroy(h->ev_kernel[i]);
    }
    delete[] h->ev_h2d;
    delete[] h->ev_kernel;
    if (h->ws) cudaFree(h->ws);
    if (h->chunk_losses) cudaFree(h->chunk_losses);
    if (h->losses) cudaFree(h->losses);
    if (h->begin) cudaEventDestroy(h->begin);
    if (h->end) cudaEventDestroy(h->end);
    if (h->in) cudaStreamDestroy(h->in);
    if (h->compute) cudaStreamDestroy(h->compute);
    if (h->out) cudaStreamDestroy(h->out);
    delete h;
    return PDEC_OK;
}

extern "C" int pdec_host_pipeline_create(pdec_host_pipeline** out, int chunk_rows, int n, int T, int max_chunks, int frames_per_slice) {
    if (!out || chunk_rows < 1 || n < 1 || T < 1 || max_chunks < 1 || frames_per_slice < 0) return PDEC_ERR_INVALID_ARGUMENT;
    pdec_host_pipeline* h = new (std::nothrow) pdec_host_pipeline();
    if (!h) return PDEC_ERR_INVALID_ARGUMENT;
    h->rows = chunk_rows; h->n = n; h->T = T; h->max_chunks = max_chunks;
    {
        const size_t bytes = (size_t)(T + 1) * chunk_rows * n * sizeof(float);
        for (int i = 0; i < SLOTS; ++i) {
            HP_TRY(cudaMalloc(&h->slot[i].s, bytes));
            HP_TRY(cudaMalloc(&h->slot[i].g, bytes));
            HP_TRY(cudaEventCreateWithFlags(&h->slot[i].h2d, cudaEventDisableTiming));
            HP_TRY(cudaEventCreateWithFlags(&h->slot[i].kernel, cudaEventDisableTiming));
            HP_TRY(cudaEventCreateWithFlags(&h->slot[i].d2h, cudaEventDisableTiming));
        }
        h->max_batch = (long long)chunk_rows * max_chunks;
        if (frames_per_slice > 0) {
            if (h->max_batch > (1LL << 24)) { pdec_host_pipeline_destroy(h); return PDEC_ERR_INVALID_ARGUMENT; }
            h->slice = frames_per_slice;
            const size_t full = (size_t)(T + 1) * (size_t)h->max_batch * n * sizeof(float);
            HP_TRY(cudaMalloc(&h->full_s, full));
            HP_TRY(cudaMalloc(&h->full_g, full));
            h->ev_h2d = new (std::nothrow) cudaEvent_t[T]();
            h->ev_kernel = new (std::nothrow) cudaEvent_t[T]();
            if (!h->ev_h2d || !h->ev_kernel) goto fail;
            for (int i = 0; i < T; ++i) {
                HP_TRY(cudaEventCreateWithFlags(&h->ev_h2d[i], cudaEventDisableTiming));
                HP_TRY(cudaEventCreateWithFlags(&h->ev_kernel[i], cudaEventDisableTiming));
            }
        }
        h->ws_bytes = pdec_burgers_pair_loss_workspace_bytes(frames_per_slice > 0 ? (int)h->max_batch : chunk_rows, n, T);
        HP_TRY(cudaMalloc(&h->ws, h->ws_bytes));
        HP_TRY(cudaMalloc(&h->chunk_losses, sizeof(float) * 2 * (max_chunks > T ? max_chunks : T)));
        HP_TRY(cudaMalloc(&h->losses, sizeof(float) * 2));
        HP_TRY(cudaStreamCreateWithFlags(&h->in, cudaStreamNonBlocking));
        HP_TRY(cudaStreamCreateWithFlags(&h->compute, cudaStreamNonBlocking));
        HP_TRY(cudaStreamCreateWithFlags(&h->out, cudaStreamNonBlocking));
        HP_TRY(cudaEventCreateWithFlags(&h->begin, cudaEventDisableTiming));
        HP_TRY(cudaEventCreateWithFlags(&h->end, cudaEventDisableTiming));
    }
    *out = h;
    return PDEC_OK;
fail:
    pdec_host_pipeline_destroy(h);
    return PDEC_ERR_CUDA_LAUNCH;
}

extern "C" int pdec_burgers_pair_loss_host(pdec_host_pipeline* h, const float* s_host, float* grad_host, float* losses_host, int T,
                                           int B, int n, int mode, double dt, double viscosity, float box_lower, float box_size,
                                           int substeps, int resubstep, float gl2, float gl1, void* stream) {
    if (!h || !s_host || !losses_host || B < 1) return PDEC_ERR_INVALID_ARGUMENT;
    if (T != h->T || n != h->n) return PDEC_ERR_INVALID_ARGUMENT;
    const int rows = h->rows;
    const int nchunks = (B + rows - 1) / rows;
    if (nchunks > h->max_chunks) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    cudaStream_t user = (cudaStream_t)stream;
    const size_t host_pitch = (size_t)B * n * sizeof(float);
    int rc = PDEC_OK;
    float* grad_direct = nullptr;
    HP_TRY(cudaEventRecord(h->begin, user));
    HP_TRY(cudaStreamWaitEvent(h->in, h->begin, 0));
    HP_TRY(cudaStreamWaitEvent(h->compute, h->begin, 0));
    HP_TRY(cudaStreamWaitEvent(h->out, h->begin, 0));
    if (h->slice > 0 && B <= h->max_batch && pdec_burgers_pair_loss_window_supported(n, mode, box_lower, box_size, substeps)) {
        // frame-sliced mode: contiguous frame blocks in both directions
        const size_t frame = (size_t)B * n;                         // floats per frame
        const int nsl = (T + h->slice - 1) / h->slice;
        for (int c = 0; c < nsl; ++c) {
            const int t0 = c * h->slice, t1 = (t0 + h->slice < T) ? t0 + h->slice : T;
            const int first = c == 0 ? 0 : t0 + 1;                  // frame t0 came with the previous slice
            HP_TRY(cudaMemcpyAsync(h->full_s + first * frame, s_host + first * frame, (size_t)(t1 - first + 1) * frame * sizeof(float),
                                   cudaMemcpyHostToDevice, h->in));
            HP_TRY(cudaEventRecord(h->ev_h2d[c], h->in));
            HP_TRY(cudaStreamWaitEvent(h->compute, h->ev_h2d[c], 0));
            rc = pdec_burgers_pair_loss_window(h->full_s, grad_host ? h->full_g : nullptr, h->chunk_losses + 2 * c, T, B, n, t0, t1, mode, dt,
                                               viscosity, box_lower, box_size, substeps, resubstep, gl2, gl1, h->ws, h->ws_bytes, h->compute);
            if (rc != PDEC_OK) return rc;
            HP_TRY(cudaEventRecord(h->ev_kernel[c], h->compute));
            if (grad_host) {
                const int last = t1 == T ? T : t1 - 1;              // frames whose gradient is complete after this slice
                HP_TRY(cudaStreamWaitEvent(h->out, h->ev_kernel[c], 0));
                HP_TRY(cudaMemcpyAsync(grad_host + t0 * frame, h->full_g + t0 * frame, (size_t)(last - t0 + 1) * frame * sizeof(float),
                                       cudaMemcpyDeviceToHost, h->out));
            }
        }
        HP_TRY(cudaStreamWaitEvent(h->out, h->ev_kernel[nsl - 1], 0));
        pdec::combine_chunk_losses_kernel<<<1, 32, 0, h->out>>>(h->chunk_losses, nsl, 0, B, h->losses);
        HP_TRY(cudaPeekAtLastError());
        HP_TRY(cudaMemcpyAsync(losses_host, h->losses, 2 * sizeof(float), cudaMemcpyDeviceToHost, h->out));
        HP_TRY(cudaEventRecord(h->end, h->out));
        HP_TRY(cudaStreamWaitEvent(user, h->end, 0));
        return PDEC_OK;
    }
    // direct gradient: can the device address grad_host?  Then the fused kernel stores there itself (no staging, no D2H leg)
    if (grad_host && getenv("PDEC_PIPE_STAGED_D2H") == nullptr && pdec_burgers_pair_loss_window_supported(n, mode, box_lower, box_size, substeps)) {
        cudaPointerAttributes attr;
        if (cudaPointerGetAttributes(&attr, grad_host) == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer != nullptr)
            grad_direct = static_cast<float*>(attr.devicePointer);
        else
            cudaGetLastError();
    }
    for (int c = 0; c < nchunks; ++c) {
        Slot& sl = h->slot[c % SLOTS];
        const int b0 = c * rows, bc = (c + 1 < nchunks) ? rows : B - b0;
        const size_t width = (size_t)bc * n * sizeof(float);
        // leg 1: frames of rows [b0, b0 + bc) -> staging (the slot's previous kernel must have read it)
        if (c >= SLOTS) HP_TRY(cudaStreamWaitEvent(h->in, sl.kernel, 0));
        HP_TRY(cudaMemcpy2DAsync(sl.s, width, s_host + (size_t)b0 * n, host_pitch, width, (size_t)T + 1, cudaMemcpyHostToDevice, h->in));
        HP_TRY(cudaEventRecord(sl.h2d, h->in));
        // leg 2: fused force loss + gradient of the chunk (the slot's previous gradient must have left)
        HP_TRY(cudaStreamWaitEvent(h->compute, sl.h2d, 0));
        const float scale = pdec::chunk_scale(c, nchunks, rows, B);      // the kernel normalises by its own row count; rescale to 1/B
        if (grad_direct) {
            rc = pdec_burgers_pair_loss_window_strided(sl.s, grad_direct + (size_t)b0 * n, h->chunk_losses + 2 * c, T, bc, n, 0, T, mode, dt,
                                                       viscosity, box_lower, box_size, substeps, resubstep, gl2 * scale, gl1 * scale,
                                                       (long long)B * n, h->ws, h->ws_bytes, h->compute);
            if (rc != PDEC_OK) return rc;
            HP_TRY(cudaEventRecord(sl.kernel, h->compute));
            continue;
        }
        if (c >= SLOTS && grad_host) HP_TRY(cudaStreamWaitEvent(h->compute, sl.d2h, 0));
        rc = pdec_burgers_pair_loss(sl.s, grad_host ? sl.g : nullptr, h->chunk_losses + 2 * c, T, bc, n, mode, dt, viscosity, box_lower,
                                    box_size, substeps, resubstep, gl2 * scale, gl1 * scale, h->ws, h->ws_bytes, h->compute);
        if (rc != PDEC_OK) return rc;
        HP_TRY(cudaEventRecord(sl.kernel, h->compute));
        // leg 3: gradient -> host
        if (grad_host) {
            HP_TRY(cudaStreamWaitEvent(h->out, sl.kernel, 0));
            HP_TRY(cudaMemcpy2DAsync(grad_host + (size_t)b0 * n, host_pitch, sl.g, width, width, (size_t)T + 1, cudaMemcpyDeviceToHost, h->out));
            HP_TRY(cudaEventRecord(sl.d2h, h->out));
        }
    }
    // the two loss scalars: chunk losses combined in chunk order, then to the host
    HP_TRY(cudaStreamWaitEvent(h->out, h->slot[(nchunks - 1) % SLOTS].kernel, 0));
    pdec::combine_chunk_losses_kernel<<<1, 32, 0, h->out>>>(h->chunk_losses, nchunks, rows, B, h->losses);
    HP_TRY(cudaPeekAtLastError());
    HP_TRY(cudaMemcpyAsync(losses_host, h->losses, 2 * sizeof(float), cudaMemcpyDeviceToHost, h->out));
    HP_TRY(cudaEventRecord(h->end, h->out));
    HP_TRY(cudaStreamWaitEvent(user, h->end, 0));
    return PDEC_OK;
fail:
    return PDEC_ERR_CUDA_LAUNCH;
}
