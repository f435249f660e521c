// microbench2.cu -- which pipes the candidate instructions of the demod kernel use on
// sm_100a and whether they overlap with FFMA.  Each probe runs N_A independent
// chains of instruction A and N_B of instruction B per loop iteration and prints
// cycles per (A+B) group per SMSP.  If A and B run on different pipes the group
// cost is max(costA, costB); on one pipe (or one dispatch port) it is the sum.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench2 microbench2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITER 2048

enum Op { NONE, FFMA, FADD, FHADD, I2FU8, PRMT, SHFW, LOP3, IADD3, IMAD, FFMA2, FSETP_VOTE, LDS32, HADD2, HFMA2, I2F_B1 };

template <int OP>
__device__ __forceinline__ void step(float &f, uint32_t &u, float c, uint32_t k, const uint32_t *sm) {
    if (OP == FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(f) : "f"(c));
    if (OP == FADD) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f) : "f"(c));
    if (OP == FHADD) {
        unsigned short h = (unsigned short)u;
        asm volatile("add.rn.f32.f16 %0, %1, %0;" : "+f"(f) : "h"(h));
    }
    if (OP == I2FU8) asm volatile("{ .reg .b32 t; bfe.u32 t, %1, 8, 8; cvt.rn.f32.u32 %0, t; }" : "=f"(f) : "r"(u));
    if (OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x5140;" : "+r"(u) : "r"(k));
    if (OP == SHFW) asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(u) : "r"(k));
    if (OP == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(k), "r"(k + 1));
    if (OP == IADD3) asm volatile("add.u32 %0, %0, %1;" : "+r"(u) : "r"(k));
    if (OP == IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(u) : "r"(k));
    if (OP == FFMA2) {
        unsigned long long a = ((unsigned long long)__float_as_uint(f) << 32) | u;
        unsigned long long cc = ((unsigned long long)__float_as_uint(c) << 32) | __float_as_uint(c);
        asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(a) : "l"(cc));
        f = __uint_as_float((uint32_t)(a >> 32));
        u = (uint32_t)a;
    }
    if (OP == FSETP_VOTE) {
        uint32_t b;
        asm volatile("{ .reg .pred p; setp.lt.f32 p, %1, %2; vote.sync.ballot.b32 %0, p, 0xffffffff; }"
                     : "=r"(b) : "f"(f), "f"(c));
        u ^= b;
    }
    if (OP == LDS32) { u = sm[(u + k) & 1023]; }
    if (OP == I2F_B1) { f = (float)((u >> 8) & 0xFFu); asm volatile("" : "+f"(f), "+r"(u)); }
    if (OP == HADD2) asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(u) : "r"(k));
    if (OP == HFMA2) asm volatile("fma.rn.f16x2 %0, %0, %1, %1;" : "+r"(u) : "r"(k));
}

template <int A, int NA, int B, int NB>
__global__ void probe(float *out, long long *cycles, float seed, uint32_t kk) {
    __shared__ uint32_t sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = i * 7u;
    float fa[8], fb[8];
    uint32_t ua[8], ub[8];
    for (int i = 0; i < 8; i++) {
        fa[i] = seed + i + threadIdx.x; fb[i] = seed * 2 + i;
        ua[i] = threadIdx.x * 2654435761u + i; ub[i] = threadIdx.x * 40503u + i;
    }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < (NA > NB ? NA : NB); i++) {
                if (i < NA) step<A>(fa[(r * NA + i) & 7], ua[(r * NA + i) & 7], seed, kk, sm);
                if (i < NB) step<B>(fb[(r * NB + i) & 7], ub[(r * NB + i) & 7], seed, kk, sm);
            }
        }
    }
    long long t1 = clock64();
    float s = 0; uint32_t v = 0;
    for (int i = 0; i < 8; i++) { s += fa[i] + fb[i]; v ^= ua[i] ^ ub[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)v;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int A, int NA, int B, int NB>
void run(const char *name) {
    const int warps = 16;
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float *out; long long *cyc;
    cudaMalloc(&out, sizeof(float) * sms * warps * 32);
    cudaMalloc(&cyc, sizeof(long long) * sms);
    for (int rep = 0; rep < 2; rep++) probe<A, NA, B, NB><<<sms, warps * 32>>>(out, cyc, 1.0001f, 0x01020304u);
    cudaDeviceSynchronize();
    long long h[1024]; cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < sms; i++) avg += h[i]; avg /= sms;
    // groups per SMSP = ITER * 4 * (warps / 4)
    double groups = (double)ITER * 4 * (warps / 4);
    printf("%-28s cycles per group per SMSP = %6.2f   (A x%d, B x%d)\n", name, avg / groups, NA, NB);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<FFMA, 1, NONE, 0>("FFMA");
    run<FADD, 1, NONE, 0>("FADD");
    run<FHADD, 1, NONE, 0>("FHADD (add.f32.f16)");
    run<I2FU8, 1, NONE, 0>("BFE+I2F (byte->f32)");
    run<I2F_B1, 1, NONE, 0>("I2F.U8 .B1");
    run<FFMA, 4, I2F_B1, 1>("4 FFMA + 1 I2F.U8");
    run<PRMT, 1, NONE, 0>("PRMT");
    run<SHFW, 1, NONE, 0>("SHF.L.W");
    run<LOP3, 1, NONE, 0>("LOP3");
    run<IADD3, 1, NONE, 0>("IADD");
    run<IMAD, 1, NONE, 0>("IMAD");
    run<FFMA2, 1, NONE, 0>("FFMA2");
    run<HADD2, 1, NONE, 0>("HADD2");
    run<HFMA2, 1, NONE, 0>("HFMA2");
    run<FSETP_VOTE, 1, NONE, 0>("FSETP+VOTE");
    run<LDS32, 1, NONE, 0>("LDS.32 (dependent addr)");
    run<FFMA, 2, PRMT, 1>("2 FFMA + 1 PRMT");
    run<FFMA, 2, SHFW, 1>("2 FFMA + 1 SHF");
    run<FFMA, 2, LOP3, 2>("2 FFMA + 2 LOP3");
    run<FFMA, 2, IADD3, 2>("2 FFMA + 2 IADD");
    run<FFMA, 2, FHADD, 1>("2 FFMA + 1 FHADD");
    run<FFMA, 2, FHADD, 2>("2 FFMA + 2 FHADD");
    run<FFMA, 4, I2FU8, 1>("4 FFMA + 1 BFE+I2F");
    run<FFMA, 2, HADD2, 2>("2 FFMA + 2 HADD2");
    run<FFMA, 4, PRMT, 1>("4 FFMA + 1 PRMT");
    run<FFMA, 4, PRMT, 2>("4 FFMA + 2 PRMT");
    run<FFMA, 4, SHFW, 2>("4 FFMA + 2 SHF");
    run<FFMA, 4, FSETP_VOTE, 1>("4 FFMA + 1 FSETP+VOTE");
    run<FFMA, 4, LDS32, 1>("4 FFMA + 1 LDS");
    run<FFMA2, 1, PRMT, 1>("1 FFMA2 + 1 PRMT");
    run<FFMA2, 2, FFMA, 2>("2 FFMA2 + 2 FFMA");
    run<PRMT, 1, SHFW, 1>("1 PRMT + 1 SHF");
    run<PRMT, 1, LOP3, 2>("1 PRMT + 2 LOP3");
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
