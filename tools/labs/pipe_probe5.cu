// pipe_probe5.cu -- is the register file's read bandwidth what makes "every other instruction
// add" beside packed fp32?  Same FFMA2 count, with the multiplicand either a distinct register
// pair per instruction (direct-form filter) or one pair shared by eight consecutive instructions
// (transposed-form filter: the operand-reuse cache serves it), alone and with 1 or 2 PRMT / SHF
// per FFMA2.  16 warps per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/pipe_probe5 tools/labs/pipe_probe5.cu
#include <cstdint>
#include <cstdio>

#define FMA2(acc, x, immhex) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(x), "l"(immhex))
#define ADD2(acc, x) asm volatile("add.rn.f32x2 %0, %1, %0;" : "+l"(acc) : "l"(x))

template <bool SHARED, bool ADD, int NP>
__global__ void __launch_bounds__(512, 1) probe(uint32_t *out, long long *cyc, uint32_t seed) {
    unsigned long long acc[8], x[8];
    uint32_t p[4], q[4];
    for (int i = 0; i < 8; i++) {
        float a = 1.0f + seed * (i + 1) + threadIdx.x, b = 0.5f + seed * (i + 2);
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(a), "f"(b));
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(x[i]) : "f"(b * 1e-3f), "f"(a * 1.1e-3f));
    }
    for (int i = 0; i < 4; i++) { p[i] = seed * (i + 7) + threadIdx.x; q[i] = seed * (i + 11); }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < 1024; it++) {
#pragma unroll
        for (int r = 0; r < 2; r++) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const unsigned long long xx = SHARED ? x[r] : x[k];
                if (ADD) ADD2(acc[k], xx);
                else switch (k) {
                    case 0: FMA2(acc[k], xx, 0x3c90da643c90da64ull); break;
                    case 1: FMA2(acc[k], xx, 0x3d454f2a3d454f2aull); break;
                    case 2: FMA2(acc[k], xx, 0x3dfab9c03dfab9c0ull); break;
                    case 3: FMA2(acc[k], xx, 0x3e4a25703e4a2570ull); break;
                    case 4: FMA2(acc[k], xx, 0x3e6b12343e6b1234ull); break;
                    case 5: FMA2(acc[k], xx, 0x3e4a25713e4a2571ull); break;
                    case 6: FMA2(acc[k], xx, 0x3dfab9c13dfab9c1ull); break;
                    default: FMA2(acc[k], xx, 0x3d454f2b3d454f2bull); break;
                }
                if (NP >= 1) asm volatile("prmt.b32 %0, %0, %1, 0x7250;" : "+r"(p[k & 3]) : "r"(q[k & 3]));
                if (NP >= 2) asm volatile("shf.l.wrap.b32 %0, %0, %1, 3;" : "+r"(q[(k + 1) & 3]) : "r"(p[(k + 2) & 3]));
            }
        }
    }
    const long long t1 = clock64();
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s += (uint32_t)acc[i] ^ (uint32_t)(acc[i] >> 32);
    for (int i = 0; i < 4; i++) s += p[i] ^ q[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <bool SHARED, bool ADD, int NP>
void run(const char *nm) {
    uint32_t *out; long long *cyc;
    cudaMalloc(&out, sizeof(uint32_t) * 148 * 512);
    cudaMalloc(&cyc, sizeof(long long) * 148);
    for (int rep = 0; rep < 2; rep++) probe<SHARED, ADD, NP><<<148, 512>>>(out, cyc, 1u);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += (double)h[i]; avg /= 148;
    printf("  %-52s %5.2f cycles per packed instruction per SMSP\n", nm, avg / (1024.0 * 16 * 4));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<false, false, 0>("FFMA2 imm, distinct multiplicands");
    run<false, false, 1>("FFMA2 imm, distinct multiplicands, + PRMT");
    run<false, false, 2>("FFMA2 imm, distinct multiplicands, + PRMT + SHF");
    run<true, false, 0>("FFMA2 imm, shared multiplicand");
    run<true, false, 1>("FFMA2 imm, shared multiplicand, + PRMT");
    run<true, false, 2>("FFMA2 imm, shared multiplicand, + PRMT + SHF");
    run<false, true, 0>("FADD2, distinct addends");
    run<false, true, 1>("FADD2, distinct addends, + PRMT");
    run<false, true, 2>("FADD2, distinct addends, + PRMT + SHF");
    run<true, true, 0>("FADD2, shared addend");
    run<true, true, 1>("FADD2, shared addend, + PRMT");
    run<true, true, 2>("FADD2, shared addend, + PRMT + SHF");
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
