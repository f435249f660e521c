// imma_probe.cu -- throughput of the legacy warp-level mma.sync (m16n8k32, u8 x s8 -> s32) on sm_100a:
// cycles per instruction and SM sub-partition with 1..8 warps per sub-partition, alone and next to
// FFMA work.  (development tool)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/imma_probe tools/labs/imma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ void imma(int (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2], const int (&c)[4]) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "r"(c[0]), "r"(c[1]), "r"(c[2]), "r"(c[3]));
}

template <int FFMA_PER_IMMA>
__global__ void k_imma(int reps, unsigned long long *out, int *sink) {
    uint32_t a[4], b[6][2];
    int c[4] = {1, 2, 3, 4}, d[6][4];
    for (int i = 0; i < 4; i++) a[i] = threadIdx.x * 2654435761u + i;
    for (int j = 0; j < 6; j++) { b[j][0] = threadIdx.x + 17 * j; b[j][1] = threadIdx.x * 3 + j; }
    float f[8];
    for (int i = 0; i < 8; i++) f[i] = threadIdx.x * 0.001f + i;
    __syncthreads();
    const unsigned long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
#pragma unroll
        for (int j = 0; j < 6; j++) {
            imma(d[j], a, b[j], c);
#pragma unroll
            for (int k = 0; k < FFMA_PER_IMMA; k++) f[k & 7] = fmaf(f[k & 7], 1.0001f, 0.5f);
        }
        a[0] ^= (uint32_t)d[0][0]; a[1] ^= (uint32_t)d[3][1];    // dependency across iterations
    }
    __syncthreads();
    const unsigned long long t1 = clock64();
    int acc = 0;
    for (int j = 0; j < 6; j++) for (int i = 0; i < 4; i++) acc ^= d[j][i];
    float fs = 0; for (int i = 0; i < 8; i++) fs += f[i];
    if (acc == 0x7fffffff && fs == 1.234f) sink[0] = acc;
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int F>
static void run(int warps_per_smsp) {
    const int reps = 2000, threads = warps_per_smsp * 4 * 32;
    unsigned long long *out; int *sink;
    CK(cudaMalloc(&out, 148 * 8)); CK(cudaMalloc(&sink, 4));
    k_imma<F><<<148, threads>>>(reps, out, sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); exit(2); }
    unsigned long long h[148]; CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
    double s = 0; for (int i = 0; i < 148; i++) s += (double)h[i];
    s /= 148;
    const double immas_per_smsp = (double)reps * 6 * warps_per_smsp;
    printf("%d warps per sub-partition, %2d FFMA per IMMA: %7.2f cycles per IMMA.16832 per sub-partition (%6.0f MAC/clk/SM)%s\n",
           warps_per_smsp, F, s / immas_per_smsp, 4.0 * 4096.0 * immas_per_smsp / s,
           F ? "" : "");
    cudaFree(out); cudaFree(sink);
}

int main() {
    for (int w : {1, 2, 4, 8}) run<0>(w);
    for (int w : {2, 4}) { run<4>(w); run<8>(w); run<16>(w); run<32>(w); }
    return 0;
}
