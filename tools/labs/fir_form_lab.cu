// fir_form_lab.cu -- the demod body in steady state (registers only, no memory): per packed step
// and component 2 PRMT + 1 FADD2-imm widen a pair of bytes, then the 9-tap symmetric filter either
//   FORM 0  direct:      4 FADD2 (pair sums) + 1 FFMA2 (centre, residual) + 4 FFMA2-imm
//   FORM 1  transposed:  9 FFMA2-imm that share one multiplicand (the new sample) and add into
//                        nine rotating accumulators; no symmetric pre-add
// followed by the cross product (FMUL2 + FFMA2) and two SHF per step.  Same instruction count and
// the same FMA-pipe cycles; FORM 1 reads fewer registers if the operand-reuse cache serves the
// shared multiplicand.  Prints cycles per packed step per SMSP (4 warps each).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/fir_form_lab tools/labs/fir_form_lab.cu
#include <cstdint>
#include <cstdio>

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 widen(uint32_t wa, uint32_t wb, uint32_t sel) {
    const uint32_t lo = __byte_perm(wa, 0x4B000000u, sel), hi = __byte_perm(wb, 0x4B000000u, sel);
    const u64 v = ((u64)hi << 32) | lo;
    return add2(v, 0xCB00007FCB00007Full);       // -(2^23 + 127) in both halves
}

#define C0 0x3c90da643c90da64ull
#define C1 0x3d454f2a3d454f2aull
#define C2 0x3dfab9c03dfab9c0ull
#define C3 0x3e4a25703e4a2570ull
#define C4 0x3e6b12343e6b1234ull
#define RES 0x3bfae1483bfae148ull

template <int FORM>
__global__ void __launch_bounds__(512, 1) lab(uint32_t *out, long long *cyc, uint32_t seed, int iters) {
    // component 0 = re, 1 = im
    u64 x[2][9], acc[2][9], prev[2];
    uint32_t bits_a = 0, bits_b = 0;
    uint32_t wa = seed * 0x9E3779B9u + threadIdx.x * 0x85EBCA6Bu, wb = wa ^ 0x5bd1e995u;
    for (int c = 0; c < 2; c++) {
        for (int i = 0; i < 9; i++) { x[c][i] = pk(0.f, 0.f); acc[c][i] = pk(0.f, 0.f); }
        prev[c] = pk(0.f, 0.f);
    }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int n = 0; n < 36; n++) {               // 36 packed steps, register names rotate mod 9
            if ((n & 1) == 0) { wa = wa * 1664525u + 1013904223u; wb ^= wa >> 7; }   // a new word pair every 2 steps
            u64 f[2];
#pragma unroll
            for (int c = 0; c < 2; c++) {
                const u64 xn = widen(wa, wb, 0x7650u + ((n & 1) * 2 + c));   // byte (n&1)*2+c -> low byte of 0x4B0000uu
                if (FORM == 0) {
                    x[c][n % 9] = xn;                                   // newest; x[(n-j)%9] is j steps old
                    const u64 s0 = add2(x[c][n % 9], x[c][(n + 1) % 9]);
                    const u64 s1 = add2(x[c][(n + 8) % 9], x[c][(n + 2) % 9]);
                    const u64 s2 = add2(x[c][(n + 7) % 9], x[c][(n + 3) % 9]);
                    const u64 s3 = add2(x[c][(n + 6) % 9], x[c][(n + 4) % 9]);
                    u64 a = fma2(x[c][(n + 5) % 9], C4, RES);
                    a = fma2(s0, C0, a); a = fma2(s1, C1, a); a = fma2(s2, C2, a); a = fma2(s3, C3, a);
                    f[c] = a;
                } else {
                    // output m lives in slot m % 9; this sample is tap j of output n - j
                    acc[c][(n + 1) % 9] = fma2(xn, C0, acc[c][(n + 1) % 9]);   // j = 8: completes
                    acc[c][(n + 2) % 9] = fma2(xn, C1, acc[c][(n + 2) % 9]);
                    acc[c][(n + 3) % 9] = fma2(xn, C2, acc[c][(n + 3) % 9]);
                    acc[c][(n + 4) % 9] = fma2(xn, C3, acc[c][(n + 4) % 9]);
                    acc[c][(n + 5) % 9] = fma2(xn, C4, acc[c][(n + 5) % 9]);
                    acc[c][(n + 6) % 9] = fma2(xn, C3, acc[c][(n + 6) % 9]);
                    acc[c][(n + 7) % 9] = fma2(xn, C2, acc[c][(n + 7) % 9]);
                    acc[c][(n + 8) % 9] = fma2(xn, C1, acc[c][(n + 8) % 9]);
                    acc[c][(n + 0) % 9] = fma2(xn, C0, RES);                   // j = 0: a new output starts
                    f[c] = acc[c][(n + 1) % 9];
                }
            }
            // cross product of consecutive filter outputs and the two decisions
            const u64 p = mul2(prev[1], f[0]);
            const u64 num = fma2(prev[0], f[1], p);   // (sign of p folded elsewhere in the real kernel)
            prev[0] = f[0]; prev[1] = f[1];
            bits_a = __funnelshift_l((uint32_t)num, bits_a, 1);
            bits_b = __funnelshift_l((uint32_t)(num >> 32), bits_b, 1);
        }
    }
    const long long t1 = clock64();
    uint32_t s = bits_a ^ bits_b;
    for (int c = 0; c < 2; c++) for (int i = 0; i < 9; i++) s += (uint32_t)(x[c][i] >> 3) ^ (uint32_t)(acc[c][i] >> 35);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int FORM>
void run(const char *nm) {
    uint32_t *out; long long *cyc;
    cudaMalloc(&out, sizeof(uint32_t) * 148 * 512);
    cudaMalloc(&cyc, sizeof(long long) * 148);
    const int iters = 512;
    for (int rep = 0; rep < 2; rep++) lab<FORM><<<148, 512>>>(out, cyc, 1u, iters);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += (double)h[i]; avg /= 148;
    const double per_step = avg / (iters * 36.0 * 4);
    printf("  %-28s %6.2f cycles per packed step per SMSP = %5.2f per sample\n", nm, per_step, per_step / 2);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("direct (symmetric pre-add)");
    run<1>("transposed (shared operand)");
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
