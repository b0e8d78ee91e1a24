// umma_probe.cu -- one 128 x 128 x 128 TF32 GEMM on tcgen05 (cta_group::1, both operands K-major in shared memory, no swizzle), checked
// against the CPU on tf32-truncated inputs.  Pins, before the product kernel relies on them: the shared-memory matrix descriptor
// (start / LBO / SBO in 16-byte units, version 1), the instruction descriptor for kind::tf32, the canonical K-major "interleave" layout
// (8 rows x 16 bytes core matrices; LBO = step along K, SBO = step along M/N), a PADDED LBO (144 bytes) for operand A, the TMEM
// accumulator layout (lane = row, column = n) read back with tcgen05.ld 32x32b.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/diag/umma_probe.bin scripts/diag/umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

#define A_LBO 144u
#define A_SBO (32u * A_LBO)
#define B_LBO 128u
#define B_SBO (32u * B_LBO)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fffu);
    d |= (uint64_t)((lbo >> 4) & 0x3fffu) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fffu) << 32;
    d |= (uint64_t)1 << 46;                                  // descriptor version 1 (Blackwell)
    return d;                                                // base offset 0, LBO mode 0, layout type 0 = no swizzle
}
// element (r, k) of a K-major operand: r = row (M or N index), k = reduction index
__device__ __forceinline__ uint32_t canon(uint32_t r, uint32_t k, uint32_t lbo, uint32_t sbo) {
    return (r >> 3) * sbo + (k >> 2) * lbo + (r & 7u) * 16u + (k & 3u) * 4u;
}

__global__ void __launch_bounds__(128, 1) probe(const float* __restrict__ A, const float* __restrict__ Bnk, float* __restrict__ D, int* status) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA = smem;
    unsigned char* sB = smem + 16 * A_SBO;
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int q = tid; q < 128 * 128; q += 128) {
        const int r = q >> 7, k = q & 127;
        *reinterpret_cast<float*>(sA + canon(r, k, A_LBO, A_SBO)) = A[q];
        *reinterpret_cast<float*>(sB + canon(r, k, B_LBO, B_SBO)) = Bnk[q];
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes of the operands -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_base;
    if (tid == 0) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (16u << 17) | (8u << 24);     // D f32, A/B tf32 K-major, N = 128, M = 128
        for (int ks = 0; ks < 16; ++ks) {                             // K = 8 per instruction = two 16-byte units
            const uint64_t da = make_desc(smem_u32(sA) + ks * 2 * A_LBO, A_LBO, A_SBO);
            const uint64_t db = make_desc(smem_u32(sB) + ks * 2 * B_LBO, B_LBO, B_SBO);
            const uint32_t acc = ks > 0 ? 1u : 0u;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                         ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    // bounded wait on the commit
    {
        uint32_t done = 0;
        for (long long spin = 0; !done && spin < (1LL << 24); ++spin)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        if (!done) { if (tid == 0) *status = -1; }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t v[32];
        const uint32_t taddr = tm + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, "
                     "%21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                       "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
                       "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 32; ++j) D[(size_t)tid * 128 + c0 + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tm) : "memory");
}

static float tf32_trunc(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xffffe000u; memcpy(&x, &u, 4); return x; }

int main() {
    const int n = 128;
    float *hA = (float*)malloc(n * n * 4), *hB = (float*)malloc(n * n * 4), *hD = (float*)malloc(n * n * 4);
    srand(1);
    for (int i = 0; i < n * n; ++i) { hA[i] = (float)rand() / RAND_MAX - 0.5f; hB[i] = (float)rand() / RAND_MAX - 0.5f; }
    float *dA, *dB, *dD; int* dS;
    cudaMalloc(&dA, n * n * 4); cudaMalloc(&dB, n * n * 4); cudaMalloc(&dD, n * n * 4); cudaMalloc(&dS, 4);
    cudaMemcpy(dA, hA, n * n * 4, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, n * n * 4, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, n * n * 4); cudaMemset(dS, 0, 4);
    const size_t smem = 16 * A_SBO + 16 * B_SBO + 1024;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    probe<<<1, 128, smem>>>(dA, dB, dD, dS);
    cudaError_t e = cudaDeviceSynchronize();
    int st = 0; cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost);
    printf("launch: %s, status %d\n", cudaGetErrorString(e), st);
    if (e != cudaSuccess) return 1;
    cudaMemcpy(hD, dD, n * n * 4, cudaMemcpyDeviceToHost);
    double worst_t = 0, worst_f = 0, scale = 0;
    for (int m = 0; m < n; ++m)
        for (int j = 0; j < n; ++j) {
            double rt = 0, rf = 0;
            for (int k = 0; k < n; ++k) {
                rt += (double)tf32_trunc(hA[m * n + k]) * (double)tf32_trunc(hB[j * n + k]);
                rf += (double)hA[m * n + k] * (double)hB[j * n + k];
            }
            worst_t = fmax(worst_t, fabs(hD[m * n + j] - rt)); worst_f = fmax(worst_f, fabs(hD[m * n + j] - rf)); scale = fmax(scale, fabs(rf));
        }
    printf("max |D - truncated-tf32 reference| = %.3e, max |D - exact| = %.3e, max |exact| = %.3e\n", worst_t, worst_f, scale);
    printf("D[0][0..3] = %g %g %g %g ; D[5][7] = %g\n", hD[0], hD[1], hD[2], hD[3], hD[5 * n + 7]);
    return 0;
}
