// halo.cu -- halo rows of row-slab decomposed fields exchanged over NVLink peer memory (config 5, 2048 x 2048 over 2/4/8 GPUs).
//
// The reference has no distributed code: this replaces the halo exchange a slab-decomposed IncompressibleFlow.step
// (PF/physics/fluid.py:90-114) needs around advect.semi_lagrangian (advect.py:24-40: rows within max|v| dt / dx + 1 of a slab seam)
// and StaggeredGrid.gradient (staggered_grid.py:219-231: one row of the pressure).
//
// Transport: the sender's kernel STORES its boundary rows straight into the receiver's inbox (a peer-mapped allocation,
// pdec_comm_alloc / pdec_comm_open); every float travels in an 8-byte word together with the exchange's sequence number and the
// receiver's kernel polls each word until the tag matches (bounded) -- no fence, no atomic, no host round trip, no NCCL call, the same
// tagged-word protocol as the slab CG (cg_tiled.cu; a flag hop over NVLink behind fence.sys costs 4.4 us, a tagged word 1.15 us:
// scripts/diag/p2p_latency.cu).  Two slots by sequence parity: a sender is never more than one exchange ahead of its receiver, because
// its next push is stream-ordered behind its own pull of the current exchange.
//
// Inbox layout (uint64 words): [slot 0..1][side 0..1][B * max_rows * row_len]; side 0 = rows arriving from the rank BELOW (they become
// this rank's lower halo), side 1 = from the rank ABOVE.
#include "pdec_common.cuh"

namespace pdec {

__device__ __forceinline__ unsigned long long halo_ld(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void halo_st(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void __launch_bounds__(256) halo_push_kernel(const float* __restrict__ field, unsigned long long* __restrict__ inbox, int B, long long elem_stride,
                                                        int row_len, int row_first, int nrows, unsigned seq) {
    const long long per_elem = (long long)nrows * row_len, total = per_elem * B;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const long long b = q / per_elem, r = q - b * per_elem;
        const float v = field[b * elem_stride + (long long)row_first * row_len + r];
        halo_st(inbox + q, ((unsigned long long)seq << 32) | (unsigned long long)__float_as_uint(v));
    }
}

__global__ void __launch_bounds__(256) halo_pull_kernel(float* __restrict__ field, const unsigned long long* __restrict__ inbox, int B, long long elem_stride,
                                                        int row_len, int row_first, int nrows, unsigned seq, int accumulate, int* __restrict__ status) {
    const long long per_elem = (long long)nrows * row_len, total = per_elem * B;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        unsigned long long w = halo_ld(inbox + q);
        // bounded poll (about 10 s): a peer that never arrives must not hang the GPU; the caller reads `status`
        for (long long spin = 0; (unsigned)(w >> 32) != seq && spin < (1LL << 27); ++spin) {
            __nanosleep(64);
            w = halo_ld(inbox + q);
        }
        if ((unsigned)(w >> 32) != seq) {
            if (status != nullptr) atomicExch(status, -1);
            continue;
        }
        const long long b = q / per_elem, r = q - b * per_elem;
        float* dst = field + b * elem_stride + (long long)row_first * row_len + r;
        const float v = __uint_as_float((unsigned)(w & 0xffffffffu));
        *dst = accumulate ? *dst + v : v;
    }
}

}  // namespace pdec

using namespace pdec;

static size_t halo_side_words(int B, int row_len, int max_rows) { return (size_t)B * max_rows * row_len; }

extern "C" size_t pdec_halo_inbox_bytes(int B, int row_len, int max_rows) {
    if (B < 1 || row_len < 1 || max_rows < 1) return 0;
    return 4 * halo_side_words(B, row_len, max_rows) * sizeof(unsigned long long);
}

extern "C" int pdec_halo_inbox_init(void* inbox, int B, int row_len, int max_rows, void* stream) {
    if (!inbox || pdec_halo_inbox_bytes(B, row_len, max_rows) == 0) return PDEC_ERR_INVALID_ARGUMENT;
    // tag 0 is never sent: a zeroed inbox holds no valid word
    if (cudaMemsetAsync(inbox, 0, pdec_halo_inbox_bytes(B, row_len, max_rows), (cudaStream_t)stream) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    return PDEC_OK;
}

static int halo_grid(long long total) {
    long long g = (total + 255) / 256;
    return (int)(g < 1 ? 1 : (g > 148 * 4 ? 148 * 4 : g));
}

extern "C" int pdec_halo_push_rows(const float* field, int B, int rows_total, int row_len, int row_first, int nrows, void* peer_inbox, int side,
                                   unsigned seq, int max_rows, void* stream) {
    if (B < 1 || row_len < 1 || nrows < 0 || nrows > max_rows || row_first < 0 || row_first + nrows > rows_total || side < 0 || side > 1 || seq == 0)
        return PDEC_ERR_INVALID_ARGUMENT;
    if (nrows == 0) return PDEC_OK;
    if (!field || !peer_inbox) return PDEC_ERR_INVALID_ARGUMENT;
    unsigned long long* box = reinterpret_cast<unsigned long long*>(peer_inbox) + ((size_t)(seq & 1u) * 2 + side) * halo_side_words(B, row_len, max_rows);
    const long long total = (long long)B * nrows * row_len;
    halo_push_kernel<<<halo_grid(total), 256, 0, (cudaStream_t)stream>>>(field, box, B, (long long)rows_total * row_len, row_len, row_first, nrows, seq);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_halo_pull_rows(float* field, int B, int rows_total, int row_len, int row_first, int nrows, const void* my_inbox, int side,
                                   unsigned seq, int accumulate, int max_rows, int* status, void* stream) {
    if (B < 1 || row_len < 1 || nrows < 0 || nrows > max_rows || row_first < 0 || row_first + nrows > rows_total || side < 0 || side > 1 || seq == 0)
        return PDEC_ERR_INVALID_ARGUMENT;
    if (nrows == 0) return PDEC_OK;
    if (!field || !my_inbox) return PDEC_ERR_INVALID_ARGUMENT;
    const unsigned long long* box = reinterpret_cast<const unsigned long long*>(my_inbox) + ((size_t)(seq & 1u) * 2 + side) * halo_side_words(B, row_len, max_rows);
    const long long total = (long long)B * nrows * row_len;
    halo_pull_kernel<<<halo_grid(total), 256, 0, (cudaStream_t)stream>>>(field, box, B, (long long)rows_total * row_len, row_len, row_first, nrows, seq, accumulate, status);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}
