// pipe_probe6.cu -- does the integer dot-product instruction (dp4a / SASS IDP.4A: four 8-bit
// multiply-adds per lane) give the 9-tap filter a cheaper route than packed fp32?  The samples
// are bytes; with the taps split into 8-bit limbs a tap costs limbs/4 lane-operations instead of
// one.  What decides it is the issue rate of IDP.4A on a sub-partition and whether it runs on a
// pipe beside the FMA pipe.  Measured here: IDP.4A alone, FFMA2 alone, and the two interleaved
// 1:1 and 2:1, 16 warps per SM, independent accumulators (no dependency stalls).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/pipe_probe6 tools/labs/pipe_probe6.cu
#include <cstdint>
#include <cstdio>

#define FMA2(acc, x, immhex) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(x), "l"(immhex))
#define DP4A(acc, a, b) asm volatile("dp4a.s32.s32 %0, %1, %2, %0;" : "+r"(acc) : "r"(a), "r"(b))
#define DP2A(acc, a, b) asm volatile("dp2a.lo.s32.s32 %0, %1, %2, %0;" : "+r"(acc) : "r"(a), "r"(b))

// NF packed fp32 FMAs and ND dp4a (or dp2a when TWO) per inner step, eight independent chains each
template <int NF, int ND, bool TWO>
__global__ void __launch_bounds__(512, 1) probe(uint32_t *out, long long *cyc, uint32_t seed) {
    unsigned long long acc[8], x[8];
    int iacc[8];
    uint32_t a[8], b[8];
    for (int i = 0; i < 8; i++) {
        float u = 1.0f + seed * (i + 1) + threadIdx.x, v = 0.5f + seed * (i + 2);
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(u), "f"(v));
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(x[i]) : "f"(v * 1e-3f), "f"(u * 1.1e-3f));
        iacc[i] = seed * (i + 3);
        a[i] = seed * 0x01020304u * (i + 5) + threadIdx.x;
        b[i] = seed * 0x11213141u * (i + 9);
    }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < 1024; it++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
#pragma unroll
            for (int f = 0; f < NF; f++) FMA2(acc[(k + f) & 7], x[k], 0x3e6b12343e6b1234ull);
#pragma unroll
            for (int d = 0; d < ND; d++) {
                if (TWO) DP2A(iacc[(k + d) & 7], a[k], b[(k + d) & 7]);
                else DP4A(iacc[(k + d) & 7], a[k], b[(k + d) & 7]);
            }
        }
    }
    const long long t1 = clock64();
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s += (uint32_t)acc[i] ^ (uint32_t)(acc[i] >> 32) ^ (uint32_t)iacc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int NF, int ND, bool TWO>
void run(const char *nm) {
    uint32_t *out; long long *cyc;
    cudaMalloc(&out, sizeof(uint32_t) * 148 * 512);
    cudaMalloc(&cyc, sizeof(long long) * 148);
    for (int rep = 0; rep < 2; rep++) probe<NF, ND, TWO><<<148, 512>>>(out, cyc, 1u);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += (double)h[i]; avg /= 148;
    // 1024 iterations x 8 steps, 4 warps per sub-partition
    printf("  %-44s %5.2f cycles per step per warp on a sub-partition (%d FFMA2 + %d %s)\n", nm,
           avg / (1024.0 * 8 * 4), NF, ND, TWO ? "IDP.2A" : "IDP.4A");
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<1, 0, false>("FFMA2 alone");
    run<0, 1, false>("IDP.4A alone");
    run<0, 1, true>("IDP.2A alone");
    run<1, 1, false>("FFMA2 + IDP.4A");
    run<2, 1, false>("2 FFMA2 + IDP.4A");
    run<1, 2, false>("FFMA2 + 2 IDP.4A");
    run<0, 4, false>("4 IDP.4A");
    return 0;
}
