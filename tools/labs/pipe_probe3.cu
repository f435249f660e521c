// pipe_probe3.cu -- do packed-fp32 and fp64 instructions overlap when their operands are
// distinct registers (as in the demod body), or only in the 2-operand form of microbench2?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/pipe_probe3 tools/labs/pipe_probe3.cu
#include <cstdint>
#include <cstdio>

enum { F2_IMM = 1, F2_REG = 2, D_FMA = 4, D_ADD = 8, D_FMA_IMM = 16, F_ADD2 = 32 };

template <int MIX, int NF, int ND>
__global__ void __launch_bounds__(512, 1) probe(float *out, long long *cyc, float seed) {
    unsigned long long fa[8], fx[8], fc;
    double da[8], dx[8], dc = seed * 0.37;
    for (int i = 0; i < 8; i++) {
        float a = seed + i + threadIdx.x, b = seed * 0.5f + i;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(fa[i]) : "f"(a), "f"(b));
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(fx[i]) : "f"(b * 1e-3f), "f"(a * 1e-3f));
        da[i] = a; dx[i] = b * 1e-3;
    }
    asm volatile("mov.b64 %0, {%1, %1};" : "=l"(fc) : "f"(seed * 0.25f));
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < 1024; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < (NF > ND ? NF : ND); i++) {
                const int k = (r * 2 + i) & 7;
                if (i < NF) {
                    if (MIX & F2_IMM) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(fa[k]) : "l"(fx[k]), "l"(0x3c90da643c90da64ull));
                    if (MIX & F2_REG) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(fa[k]) : "l"(fx[k]), "l"(fc));
                    if (MIX & F_ADD2) asm volatile("add.rn.f32x2 %0, %1, %0;" : "+l"(fa[k]) : "l"(fx[k]));
                }
                if (i < ND) {
                    if (MIX & D_FMA) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(da[k]) : "d"(dx[k]), "d"(dc));
                    if (MIX & D_FMA_IMM) asm volatile("fma.rn.f64 %0, %1, 0d3F921B4C7F0C2A3D, %0;" : "+d"(da[k]) : "d"(dx[k]));
                    if (MIX & D_ADD) asm volatile("add.rn.f64 %0, %1, %0;" : "+d"(da[k]) : "d"(dx[k]));
                }
            }
        }
    }
    const long long t1 = clock64();
    float s = 0; double d = 0;
    for (int i = 0; i < 8; i++) {
        float a, b;
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(fa[i]));
        s += a + b; d += da[i];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)d;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MIX, int NF, int ND>
void run(const char *nm) {
    float *out; long long *cyc;
    cudaMalloc(&out, sizeof(float) * 148 * 512);
    cudaMalloc(&cyc, sizeof(long long) * 148);
    for (int rep = 0; rep < 2; rep++) probe<MIX, NF, ND><<<148, 512>>>(out, cyc, 1.0001f);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += (double)h[i]; avg /= 148;
    printf("  %-34s %6.2f cycles per group per SMSP (packed x%d, fp64 x%d)\n", nm, avg / (1024.0 * 4 * 4), NF, ND);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<F2_IMM, 1, 0>("FFMA2 imm");
    run<F2_REG, 1, 0>("FFMA2 3 regs");
    run<F_ADD2, 1, 0>("FADD2");
    run<D_FMA, 0, 1>("DFMA 3 regs");
    run<D_FMA_IMM, 0, 1>("DFMA const");
    run<D_ADD, 0, 1>("DADD");
    run<F2_IMM | D_FMA, 1, 1>("FFMA2 imm + DFMA 3 regs");
    run<F2_IMM | D_FMA_IMM, 1, 1>("FFMA2 imm + DFMA const");
    run<F2_IMM | D_ADD, 1, 1>("FFMA2 imm + DADD");
    run<F_ADD2 | D_ADD, 1, 1>("FADD2 + DADD");
    run<F2_REG | D_FMA, 1, 1>("FFMA2 3 regs + DFMA 3 regs");
    run<F2_IMM | D_FMA_IMM, 2, 1>("2 FFMA2 imm + DFMA const");
    run<F2_IMM | D_FMA_IMM, 2, 2>("2 FFMA2 imm + 2 DFMA const");
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
