// tri_lab.cu -- compute-only probe: how fast does the demod body run when NP runs per lane
// are held as packed f32x2 pairs (FMA pipe) and ND further runs are computed in fp64 (the
// FP64 pipe issues beside the FMA pipe on this part: profiles/r01_demod_lab.txt).
// Windows are read from shared memory filled once; nothing is copied from global memory.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/tri_lab tools/labs/tri_lab.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define C0 0.017682261285
#define C1 0.048171339939
#define C2 0.122424706672
#define C3 0.197408519126
#define C4 0.228626345955

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 cfma(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }

#ifndef CONV
#define CONV 0      // 0: every byte PRMT + centring add; 1: even window indices by I2F.S8 of (u ^ 0x80)
#endif
__host__ __device__ constexpr double residual(int q, int comp) {
    const double c[9] = {C0, C1, C2, C3, C4, C3, C2, C1, C0};
#if CONV == 1
    // even indices are centred at 128 (u - 127.4 = (u - 128) + 0.6), odd ones at 127
    const double sre[4] = {0.6, 0.4, -0.6, -0.4}, sim[4] = {0.6, -0.4, -0.6, 0.4};
#else
    const double sre[4] = {-0.4, 0.4, 0.4, -0.4}, sim[4] = {-0.4, -0.4, 0.4, 0.4};
#endif
    double acc = 0.0;
    for (int m = 0; m < 9; m++) acc += c[m] * (comp ? sim[(q + 2 + m) & 3] : sre[(q + 2 + m) & 3]);
    return acc;
}

// ---- packed fp32 pair of runs ---------------------------------------------------
// byte B of x as a signed 8-bit integer -> float: compiles to one I2F.S8 with a byte selector
template <int B>
__device__ __forceinline__ float s8f(uint32_t x) {
    float a;
    asm("{ .reg .b32 t; .reg .b16 h; bfe.s32 t, %1, %2, 8; cvt.s8.s32 h, t; cvt.rn.f32.s16 %0, h; }"
        : "=f"(a) : "r"(x), "n"(8 * B));
    return a;
}
#if CONV == 2
// fp16 route: one PRMT per sample builds (1024 + I, 1024 + Q) as a half pair, a mixed-precision
// add (float + half, SASS FHADD) centres each component: half the PRMTs, scalar adds
__device__ float g_k1151 = 1151.0f;
template <int H, bool NEG>
__device__ __forceinline__ float centred(uint32_t h2, float k1151) {
    float r;
    const unsigned short h = (unsigned short)(H ? (h2 >> 16) : (h2 & 0xFFFFu));
    if (NEG)
        asm("{ .reg .b16 t; neg.f16 t, %1; add.rn.f32.f16 %0, t, %2; }" : "=f"(r) : "h"(h), "f"(k1151));
    else
        asm("add.rn.f32.f16 %0, %1, %2;" : "=f"(r) : "h"(h), "f"(-k1151));
    return r;
}
#endif
template <int I>
__device__ __forceinline__ void conv_pair(uint32_t wa, uint32_t wb, float2 &re, float2 &im) {
#if CONV == 2
    {
        const float k = g_k1151;
        const uint32_t ha = __byte_perm(wa, 0x64646464u, (I & 1) ? 0x4342 : 0x4140);
        const uint32_t hb = __byte_perm(wb, 0x64646464u, (I & 1) ? 0x4342 : 0x4140);
        constexpr int ph = I & 3;
        if (ph == 0) { re = f2(centred<0, false>(ha, k), centred<0, false>(hb, k)); im = f2(centred<1, false>(ha, k), centred<1, false>(hb, k)); }
        else if (ph == 1) { re = f2(centred<1, true>(ha, k), centred<1, true>(hb, k)); im = f2(centred<0, false>(ha, k), centred<0, false>(hb, k)); }
        else if (ph == 2) { re = f2(centred<0, true>(ha, k), centred<0, true>(hb, k)); im = f2(centred<1, true>(ha, k), centred<1, true>(hb, k)); }
        else { re = f2(centred<1, false>(ha, k), centred<1, false>(hb, k)); im = f2(centred<0, true>(ha, k), centred<0, true>(hb, k)); }
        return;
    }
#endif
#if CONV == 1
    if ((I & 1) == 0) {          // bytes 0, 1 of the word; XU pipe, no PRMT, no centring add
        const uint32_t xa = wa ^ 0x80808080u, xb = wb ^ 0x80808080u;
        const float2 vi = f2(s8f<0>(xa), s8f<0>(xb));
        const float2 vq = f2(s8f<1>(xa), s8f<1>(xb));
        if ((I & 3) == 0) { re = vi; im = vq; }
        else { re = f2(-vi.x, -vi.y); im = f2(-vq.x, -vq.y); }
        return;
    }
#endif
    constexpr uint32_t selI = (I & 1) ? 0x7442u : 0x7440u, selQ = (I & 1) ? 0x7443u : 0x7441u;
    const float2 mI = f2(__uint_as_float(__byte_perm(wa, 0x4B000000u, selI)),
                         __uint_as_float(__byte_perm(wb, 0x4B000000u, selI)));
    const float2 mQ = f2(__uint_as_float(__byte_perm(wa, 0x4B000000u, selQ)),
                         __uint_as_float(__byte_perm(wb, 0x4B000000u, selQ)));
    const float2 C = f2(8388735.0f, 8388735.0f), nC = f2(-8388735.0f, -8388735.0f);
    constexpr int ph = I & 3;
    if (ph == 0) { re = cadd(mI, nC); im = cadd(mQ, nC); }
    else if (ph == 1) { re = cadd(C, f2(-mQ.x, -mQ.y)); im = cadd(mI, nC); }
    else if (ph == 2) { re = cadd(C, f2(-mI.x, -mI.y)); im = cadd(C, f2(-mQ.x, -mQ.y)); }
    else { re = cadd(mQ, nC); im = cadd(C, f2(-mI.x, -mI.y)); }
}
template <int I>
struct ConvPair {
    static __device__ __forceinline__ void run(const uint32_t (&wa)[20], const uint32_t (&wb)[20],
                                               float2 (&re)[39], float2 (&im)[39]) {
        conv_pair<I>(wa[I >> 1], wb[I >> 1], re[I], im[I]);
        ConvPair<I + 1>::run(wa, wb, re, im);
    }
};
template <>
struct ConvPair<39> {
    static __device__ __forceinline__ void run(const uint32_t (&)[20], const uint32_t (&)[20],
                                               float2 (&)[39], float2 (&)[39]) {}
};
__device__ __forceinline__ float2 fir_pair(const float2 (&r)[39], int q, float off) {
    const float c0 = (float)C0, c1 = (float)C1, c2 = (float)C2, c3 = (float)C3, c4 = (float)C4;
    float2 acc = cfma(f2(c4, c4), r[q + 6], f2(off, off));
    acc = cfma(f2(c0, c0), cadd(r[q + 2], r[q + 10]), acc);
    acc = cfma(f2(c1, c1), cadd(r[q + 3], r[q + 9]), acc);
    acc = cfma(f2(c2, c2), cadd(r[q + 4], r[q + 8]), acc);
    acc = cfma(f2(c3, c3), cadd(r[q + 5], r[q + 7]), acc);
    return acc;
}
__device__ __forceinline__ void body_pair(const uint8_t *pa, const uint8_t *pb, uint32_t &oa, uint32_t &ob) {
    uint32_t wa[20], wb[20];
#pragma unroll
    for (int k = 0; k < 10; k++) {
        const uint2 t = reinterpret_cast<const uint2 *>(pa)[k];
        const uint2 u = reinterpret_cast<const uint2 *>(pb)[k];
        wa[2 * k] = t.x; wa[2 * k + 1] = t.y; wb[2 * k] = u.x; wb[2 * k + 1] = u.y;
    }
    float2 re[39], im[39];
    ConvPair<2>::run(wa, wb, re, im);
    uint32_t bits_a = 0, bits_b = 0;
    float2 pre = f2(0.f, 0.f), pim = f2(0.f, 0.f);
#pragma unroll
    for (int q = 0; q < 29; q++) {
        const float2 are = fir_pair(re, q, (float)residual(q & 3, 0));
        const float2 aim = fir_pair(im, q, (float)residual(q & 3, 1));
        if (q > 0) {
            const float2 t = __fmul2_rn(pre, aim);
            const float2 num = cfma(pim, are, f2(-t.x, -t.y));
            bits_a = __funnelshift_l(__float_as_uint(num.x), bits_a, 1);
            bits_b = __funnelshift_l(__float_as_uint(num.y), bits_b, 1);
        }
        pre = are; pim = aim;
    }
    oa = __brev(bits_a) >> 4; ob = __brev(bits_b) >> 4;
}

// ---- one run in fp64 ------------------------------------------------------------
// byte -> double without a conversion instruction: (lo = u, hi = 0x43300000) is 2^52 + u
template <int I>
__device__ __forceinline__ void conv_d(uint32_t w, double &re, double &im) {
    constexpr uint32_t selI = (I & 1) ? 0x4442u : 0x4440u, selQ = (I & 1) ? 0x4443u : 0x4441u;
    const double mI = __hiloint2double(0x43300000, (int)__byte_perm(w, 0u, selI));
    const double mQ = __hiloint2double(0x43300000, (int)__byte_perm(w, 0u, selQ));
    const double C = 4503599627370623.0;                         // 2^52 + 127
    constexpr int ph = I & 3;
    if (ph == 0) { re = mI - C; im = mQ - C; }
    else if (ph == 1) { re = C - mQ; im = mI - C; }
    else if (ph == 2) { re = C - mI; im = C - mQ; }
    else { re = mQ - C; im = C - mI; }
}
template <int I>
struct ConvD {
    static __device__ __forceinline__ void run(const uint32_t (&w)[20], double (&re)[39], double (&im)[39]) {
        conv_d<I>(w[I >> 1], re[I], im[I]);
        ConvD<I + 1>::run(w, re, im);
    }
};
template <>
struct ConvD<39> {
    static __device__ __forceinline__ void run(const uint32_t (&)[20], double (&)[39], double (&)[39]) {}
};
__device__ __forceinline__ double fir_d(const double (&r)[39], int q, double off) {
    double acc = fma(C4, r[q + 6], off);
    acc = fma(C0, r[q + 2] + r[q + 10], acc);
    acc = fma(C1, r[q + 3] + r[q + 9], acc);
    acc = fma(C2, r[q + 4] + r[q + 8], acc);
    acc = fma(C3, r[q + 5] + r[q + 7], acc);
    return acc;
}
__device__ __forceinline__ void body_d(const uint8_t *p, uint32_t &o) {
    uint32_t w[20];
#pragma unroll
    for (int k = 0; k < 10; k++) {
        const uint2 t = reinterpret_cast<const uint2 *>(p)[k];
        w[2 * k] = t.x; w[2 * k + 1] = t.y;
    }
    double re[39], im[39];
    ConvD<2>::run(w, re, im);
    uint32_t bits = 0;
    double pre = 0.0, pim = 0.0;
#pragma unroll
    for (int q = 0; q < 29; q++) {
        const double are = fir_d(re, q, residual(q & 3, 0));
        const double aim = fir_d(im, q, residual(q & 3, 1));
        if (q > 0) {
            const double num = fma(pim, are, -(pre * aim));
            bits = __funnelshift_l((uint32_t)__double2hiint(num), bits, 1);
        }
        pre = are; pim = aim;
    }
    o = __brev(bits) >> 4;
}

// fused: the q loop computes every run of the lane, so packed fp32 and fp64 instructions
// sit side by side in the instruction stream
template <int NP, int ND>
__device__ __forceinline__ uint32_t body_mix(const uint8_t *base, int seg) {
    uint32_t wa[NP ? NP : 1][20], wb[NP ? NP : 1][20], wd[ND ? ND : 1][20];
#pragma unroll
    for (int k = 0; k < 10; k++) {
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const uint2 t = reinterpret_cast<const uint2 *>(base + (2 * p) * seg)[k];
            const uint2 u = reinterpret_cast<const uint2 *>(base + (2 * p + 1) * seg)[k];
            wa[p][2 * k] = t.x; wa[p][2 * k + 1] = t.y; wb[p][2 * k] = u.x; wb[p][2 * k + 1] = u.y;
        }
#pragma unroll
        for (int d = 0; d < ND; d++) {
            const uint2 t = reinterpret_cast<const uint2 *>(base + (2 * NP + d) * seg)[k];
            wd[d][2 * k] = t.x; wd[d][2 * k + 1] = t.y;
        }
    }
    float2 re[NP ? NP : 1][39], im[NP ? NP : 1][39];
    double dre[ND ? ND : 1][39], dim[ND ? ND : 1][39];
#pragma unroll
    for (int p = 0; p < NP; p++) ConvPair<2>::run(wa[p], wb[p], re[p], im[p]);
#pragma unroll
    for (int d = 0; d < ND; d++) ConvD<2>::run(wd[d], dre[d], dim[d]);
    uint32_t bits_a[NP ? NP : 1] = {}, bits_b[NP ? NP : 1] = {}, bits_d[ND ? ND : 1] = {};
    float2 pre[NP ? NP : 1], pim[NP ? NP : 1];
    double dpre[ND ? ND : 1], dpim[ND ? ND : 1];
#pragma unroll
    for (int q = 0; q < 29; q++) {
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const float2 are = fir_pair(re[p], q, (float)residual(q & 3, 0));
            const float2 aim = fir_pair(im[p], q, (float)residual(q & 3, 1));
            if (q > 0) {
                const float2 t = __fmul2_rn(pre[p], aim);
                const float2 num = cfma(pim[p], are, f2(-t.x, -t.y));
                bits_a[p] = __funnelshift_l(__float_as_uint(num.x), bits_a[p], 1);
                bits_b[p] = __funnelshift_l(__float_as_uint(num.y), bits_b[p], 1);
            }
            pre[p] = are; pim[p] = aim;
        }
#pragma unroll
        for (int d = 0; d < ND; d++) {
            const double are = fir_d(dre[d], q, residual(q & 3, 0));
            const double aim = fir_d(dim[d], q, residual(q & 3, 1));
            if (q > 0) {
                const double num = fma(dpim[d], are, -(dpre[d] * aim));
                bits_d[d] = __funnelshift_l((uint32_t)__double2hiint(num), bits_d[d], 1);
            }
            dpre[d] = are; dpim[d] = aim;
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int p = 0; p < NP; p++) acc ^= (__brev(bits_a[p]) >> 4) + 3 * (__brev(bits_b[p]) >> 4);
#pragma unroll
    for (int d = 0; d < ND; d++) acc ^= 5 * (__brev(bits_d[d]) >> 4);
    return acc;
}

template <int NP, int ND, int WARPS, int CTAS>
__global__ void __launch_bounds__(WARPS * 32, CTAS) k_probe(uint32_t *out, int iters, long long *cyc) {
    constexpr int RUNS = 2 * NP + ND;
    constexpr int SEG = WARPS * 32 * 56;                         // bytes per run segment
    extern __shared__ __align__(16) uint8_t buf[];
    for (int i = threadIdx.x; i < (RUNS * SEG + 256) / 4; i += blockDim.x) {
        uint32_t x = (uint32_t)i * 2654435761u + blockIdx.x;
        x ^= x >> 15; x *= 0x85EBCA6Bu; x ^= x >> 13;
        reinterpret_cast<uint32_t *>(buf)[i] = (x & 0x1F1F1F1Fu) + 0x70707070u;
    }
    __syncthreads();
    uint32_t acc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        const uint8_t *base = buf + 56 * threadIdx.x + 8 * (it & 7);
        acc ^= body_mix<NP, ND>(base, SEG);
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int NP, int ND, int WARPS, int CTAS>
void run(const char *name, uint32_t *out, long long *cyc, int iters) {
    constexpr int RUNS = 2 * NP + ND;
    const size_t smem = (size_t)RUNS * WARPS * 32 * 56 + 256;
    auto k = k_probe<NP, ND, WARPS, CTAS>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<148 * CTAS, WARPS * 32, smem>>>(out, 2, cyc);
    cudaEventRecord(e0);
    k<<<148 * CTAS, WARPS * 32, smem>>>(out, iters, cyc);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    // lane-samples per SM sub-partition: each lane decides RUNS*28 samples per iteration
    const double lane_samples = (double)iters * RUNS * 28 * (WARPS * CTAS / 4.0);
    const double cycles = ms * 1e-3 * 1.965e9;
    const double gsps = (double)iters * RUNS * 28 * WARPS * 32 * CTAS * 148 / (ms * 1e-3) / 1e9;
    printf("%-34s regs %3d spill %4zu B  %7.3f ms  %6.2f cyc/sample/lane@1965  %7.1f Gsample/s = %5.1f%% of 6442.6 GB/s  %s\n",
           name, fa.numRegs, (size_t)fa.localSizeBytes, ms, cycles / lane_samples, gsps,
           100.0 * 2 * gsps / 6442.6, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main(int argc, char **argv) {
    const int iters = argc > 1 ? atoi(argv[1]) : 2000;
    uint32_t *out; long long *cyc;
    cudaMalloc(&out, 148 * 4 * 1024 * 4);
    cudaMalloc(&cyc, 148 * 4 * 8);
    run<1, 0, 8, 2>("2 runs f32x2          8w x2", out, cyc, iters);
    run<1, 0, 16, 1>("2 runs f32x2         16w x1", out, cyc, iters);
    run<0, 1, 8, 2>("1 run f64             8w x2", out, cyc, iters);
    run<1, 1, 8, 2>("2 f32x2 + 1 f64       8w x2", out, cyc, iters);
    run<1, 1, 16, 1>("2 f32x2 + 1 f64      16w x1", out, cyc, iters);
    run<1, 1, 8, 1>("2 f32x2 + 1 f64       8w x1", out, cyc, iters);
    run<1, 1, 12, 1>("2 f32x2 + 1 f64      12w x1", out, cyc, iters);
    run<2, 1, 8, 1>("4 f32x2 + 1 f64       8w x1", out, cyc, iters);
    run<2, 1, 8, 2>("4 f32x2 + 1 f64       8w x2", out, cyc, iters);
    run<2, 1, 12, 1>("4 f32x2 + 1 f64      12w x1", out, cyc, iters);
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
