// pipe_probe4.cu -- issue rates of the packed-fp16 instructions (HFMA2 with an immediate or a
// register multiplier, HADD2, HMUL2, the half2 compare) beside FFMA / FFMA2 / PRMT, alone and
// interleaved, to see whether an fp16 first pass of the filter could issue two values per cycle.
// 16 warps per SM (4 per sub-partition), 8 independent chains per lane.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/pipe_probe4 tools/labs/pipe_probe4.cu
#include <cstdint>
#include <cstdio>
#include <cuda_fp16.h>

enum Op { H_FMA_IMM, H_FMA_REG, H_ADD, H_MUL_IMM, H_SET, F_FMA_IMM, F_FMA_REG, F2_FMA_IMM, P_RMT, NOP_ };

template <int OP>
__device__ __forceinline__ void one(uint32_t &a, uint32_t x, uint32_t c) {
    if (OP == H_FMA_IMM) asm volatile("fma.rn.f16x2 %0, %1, %2, %0;" : "+r"(a) : "r"(x), "r"(0x2c662e66u));
    if (OP == H_FMA_REG) asm volatile("fma.rn.f16x2 %0, %1, %2, %0;" : "+r"(a) : "r"(x), "r"(c));
    if (OP == H_ADD) asm volatile("add.rn.f16x2 %0, %1, %0;" : "+r"(a) : "r"(x));
    if (OP == H_MUL_IMM) asm volatile("mul.rn.f16x2 %0, %0, %1;" : "+r"(a) : "r"(0x3c003bffu));
    if (OP == H_SET) asm volatile("set.lt.u32.f16x2 %0, %0, %1;" : "+r"(a) : "r"(x));
    if (OP == F_FMA_IMM) asm volatile("fma.rn.f32 %0, %1, 0f3c90da64, %0;" : "+r"(a) : "r"(x));
    if (OP == F_FMA_REG) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+r"(a) : "r"(x), "r"(c));
    if (OP == P_RMT) asm volatile("prmt.b32 %0, %0, %1, 0x7250;" : "+r"(a) : "r"(x));
}

template <int OPA, int NA, int OPB, int NB>
__global__ void __launch_bounds__(512, 1) probe(uint32_t *out, long long *cyc, uint32_t seed) {
    uint32_t a[8], b[8], x[8], c = seed * 3 + 0x2c002c00u;
    unsigned long long p[8], px[8];
    for (int i = 0; i < 8; i++) {
        a[i] = 0x3c003c00u + seed * (i + 1) + threadIdx.x;
        b[i] = 0x38003800u + seed * (i + 3) + threadIdx.x;
        x[i] = 0x1c001c00u + seed * (i + 5);
        p[i] = ((unsigned long long)a[i] << 32) | b[i];
        px[i] = ((unsigned long long)x[i] << 32) | x[i];
    }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < 1024; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < (NA > NB ? NA : NB); i++) {
                const int k = (r * 2 + i) & 7;
                if (i < NA) {
                    if (OPA == F2_FMA_IMM)
                        asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p[k]) : "l"(px[k]), "l"(0x3c90da643c90da64ull));
                    else one<OPA>(a[k], x[k], c);
                }
                if (i < NB) one<OPB>(b[k], x[(k + 3) & 7], c);
            }
        }
    }
    const long long t1 = clock64();
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s += a[i] ^ b[i] ^ (uint32_t)p[i] ^ (uint32_t)(p[i] >> 32);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OPA, int NA, int OPB, int NB>
void run(const char *nm) {
    uint32_t *out; long long *cyc;
    cudaMalloc(&out, sizeof(uint32_t) * 148 * 512);
    cudaMalloc(&cyc, sizeof(long long) * 148);
    for (int rep = 0; rep < 2; rep++) probe<OPA, NA, OPB, NB><<<148, 512>>>(out, cyc, 1u);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += (double)h[i]; avg /= 148;
    printf("  %-40s %6.2f cycles per group per SMSP (A x%d, B x%d)\n", nm, avg / (1024.0 * 4 * 4), NA, NB);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<H_FMA_IMM, 1, NOP_, 0>("HFMA2 imm multiplier");
    run<H_FMA_REG, 1, NOP_, 0>("HFMA2 3 regs");
    run<H_ADD, 1, NOP_, 0>("HADD2");
    run<H_MUL_IMM, 1, NOP_, 0>("HMUL2 imm");
    run<H_SET, 1, NOP_, 0>("HSET2 (compare to mask)");
    run<F_FMA_IMM, 1, NOP_, 0>("FFMA imm");
    run<F_FMA_REG, 1, NOP_, 0>("FFMA 3 regs");
    run<F2_FMA_IMM, 1, NOP_, 0>("FFMA2 imm");
    run<P_RMT, 1, NOP_, 0>("PRMT");
    run<H_FMA_IMM, 1, H_ADD, 1>("HFMA2 imm + HADD2");
    run<H_FMA_IMM, 1, P_RMT, 1>("HFMA2 imm + PRMT");
    run<H_ADD, 1, P_RMT, 1>("HADD2 + PRMT");
    run<H_FMA_IMM, 2, P_RMT, 1>("2 HFMA2 imm + PRMT");
    run<F_FMA_IMM, 1, P_RMT, 1>("FFMA imm + PRMT");
    run<F_FMA_IMM, 2, P_RMT, 1>("2 FFMA imm + PRMT");
    run<F2_FMA_IMM, 1, P_RMT, 1>("FFMA2 imm + PRMT");
    run<F2_FMA_IMM, 1, H_FMA_IMM, 1>("FFMA2 imm + HFMA2 imm");
    run<F_FMA_IMM, 1, H_FMA_IMM, 1>("FFMA imm + HFMA2 imm");
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
