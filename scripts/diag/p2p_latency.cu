// Diagnostic: GPU<->GPU signalling latency over peer memory (two GPUs, one process).  Bounded spins.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/diag/p2p_latency scripts/diag/p2p_latency.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

// mode 0: plain volatile flag store; 1: data store + fence.sys + flag store; 2: atomicAdd_system on flag; 3: like 1 but 32 data stores
__global__ void pingpong(volatile u64* mine, volatile u64* theirs, float* their_data, int me, int rounds, int mode, long long budget, long long* cycles) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    long long t0 = clock64();
    for (int i = 1; i <= rounds; ++i) {
        if ((i & 1) == me) {          // my turn to signal
            if (mode == 1 || mode == 3) {
                const int n = mode == 3 ? 32 : 1;
                for (int j = 0; j < n; ++j) their_data[j * 32] = (float)i;
                __threadfence_system();
            }
            if (mode == 2) atomicAdd_system((u64*)theirs, 1ull);
            else *theirs = (u64)i;
        } else {
            long long s = clock64();
            while (*mine < (u64)i) { if (clock64() - s > budget) { cycles[0] = -1; return; } }
            if (mode == 1 || mode == 3) __threadfence_system();
        }
    }
    cycles[0] = clock64() - t0;
}

int main() {
    int n = 0; CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("need 2 GPUs\n"); return 0; }
    int can01 = 0, can10 = 0, perf = -1, atom = -1;
    cudaDeviceCanAccessPeer(&can01, 0, 1); cudaDeviceCanAccessPeer(&can10, 1, 0);
    cudaDeviceGetP2PAttribute(&perf, cudaDevP2PAttrPerformanceRank, 0, 1);
    cudaDeviceGetP2PAttribute(&atom, cudaDevP2PAttrNativeAtomicSupported, 0, 1);
    printf("canAccessPeer 0->1 %d 1->0 %d perfRank %d nativeAtomics %d\n", can01, can10, perf, atom);
    u64* flag[2]; float* data[2]; long long* cyc[2]; cudaStream_t st[2]; int khz[2];
    for (int d = 0; d < 2; ++d) {
        CK(cudaSetDevice(d)); CK(cudaDeviceEnablePeerAccess(1 - d, 0));
        CK(cudaMalloc(&flag[d], 256)); CK(cudaMalloc(&data[d], 4096 * 4)); CK(cudaMalloc(&cyc[d], 8)); CK(cudaStreamCreate(&st[d]));
        cudaDeviceGetAttribute(&khz[d], cudaDevAttrClockRate, d);
    }
    const int rounds = 2000;
    for (int mode = 0; mode < 4; ++mode) {
        for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaMemset(flag[d], 0, 256)); CK(cudaDeviceSynchronize()); }
        for (int d = 0; d < 2; ++d) {
            CK(cudaSetDevice(d));
            pingpong<<<1, 32, 0, st[d]>>>(flag[d], flag[1 - d], data[1 - d], d, rounds, mode, 4000000000LL, cyc[d]);
        }
        long long c[2];
        for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); CK(cudaMemcpy(&c[d], cyc[d], 8, cudaMemcpyDeviceToHost)); }
        printf("mode %d: %lld cycles / %d one-way hops = %.0f cycles = %.2f us per hop (clock %d kHz)\n", mode, c[0], rounds, (double)c[0] / rounds,
               (double)c[0] / rounds / khz[0] * 1e3, khz[0]);
    }
    // local reference: fence cost with only local stores outstanding
    return 0;
}
