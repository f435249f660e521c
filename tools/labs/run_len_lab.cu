// run_len_lab.cu -- compute-only probe: the demod body with RUN consecutive samples per run
// (the product uses 28): longer runs re-convert and re-filter less halo (9 samples per run).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/run_len_lab tools/labs/run_len_lab.cu
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

__host__ __device__ constexpr double residual(int q, int comp) {
    const double c[9] = {C0, C1, C2, C3, C4, C3, C2, C1, C0};
    const double sre[4] = {-0.4, 0.4, 0.4, -0.4}, sim[4] = {-0.4, -0.4, 0.4, 0.4};
    double acc = 0.0;
    for (int m = 0; m < 9; m++) acc += c[m] * (comp ? sim[(q + 2 + m) & 3] : sre[(q + 2 + m) & 3]);
    return acc;
}

template <int I>
__device__ __forceinline__ void conv_pair(uint32_t wa, uint32_t wb, float2 &re, float2 &im) {
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
template <int I, int N, int NW>
struct ConvPair {
    static __device__ __forceinline__ void run(const uint32_t (&wa)[NW], const uint32_t (&wb)[NW],
                                               float2 (&re)[N], float2 (&im)[N]) {
        conv_pair<I>(wa[I >> 1], wb[I >> 1], re[I], im[I]);
        ConvPair<I + 1, N, NW>::run(wa, wb, re, im);
    }
};
template <int N, int NW>
struct ConvPair<N, N, NW> {
    static __device__ __forceinline__ void run(const uint32_t (&)[NW], const uint32_t (&)[NW],
                                               float2 (&)[N], float2 (&)[N]) {}
};
template <int N>
__device__ __forceinline__ float2 fir_pair(const float2 (&r)[N], int q, float off) {
    const float c0 = (float)C0, c1 = (float)C1, c2 = (float)C2, c3 = (float)C3, c4 = (float)C4;
    float2 acc = cfma(f2(c4, c4), r[q + 6], f2(off, off));
    acc = cfma(f2(c0, c0), cadd(r[q + 2], r[q + 10]), acc);
    acc = cfma(f2(c1, c1), cadd(r[q + 3], r[q + 9]), acc);
    acc = cfma(f2(c2, c2), cadd(r[q + 4], r[q + 8]), acc);
    acc = cfma(f2(c3, c3), cadd(r[q + 5], r[q + 7]), acc);
    return acc;
}

template <int RUN>
__device__ __forceinline__ uint32_t body(const uint8_t *pa, const uint8_t *pb) {
    constexpr int N = RUN + 11;                 // window samples 0 .. RUN + 10
    constexpr int NW = (N + 1) / 2;             // 32-bit words (two samples each)
    constexpr int NL = (NW + 1) / 2;            // 8-byte loads
    uint32_t wa[2 * NL], wb[2 * NL];
#pragma unroll
    for (int k = 0; k < NL; k++) {
        const uint2 t = reinterpret_cast<const uint2 *>(pa)[k];
        const uint2 u = reinterpret_cast<const uint2 *>(pb)[k];
        wa[2 * k] = t.x; wa[2 * k + 1] = t.y; wb[2 * k] = u.x; wb[2 * k + 1] = u.y;
    }
    float2 re[N], im[N];
    ConvPair<2, N, 2 * NL>::run(wa, wb, re, im);
    uint32_t acc = 0, bits_a = 0, bits_b = 0;
    float2 pre = f2(0.f, 0.f), pim = f2(0.f, 0.f);
#pragma unroll
    for (int q = 0; q < RUN + 1; q++) {
        const float2 are = fir_pair<N>(re, q, (float)residual(q & 3, 0));
        const float2 aim = fir_pair<N>(im, q, (float)residual(q & 3, 1));
        if (q > 0) {
            const float2 t = __fmul2_rn(pre, aim);
            const float2 num = cfma(pim, are, f2(-t.x, -t.y));
            bits_a = __funnelshift_l(__float_as_uint(num.x), bits_a, 1);
            bits_b = __funnelshift_l(__float_as_uint(num.y), bits_b, 1);
            if (q % 28 == 0) { acc ^= (__brev(bits_a) >> 4) + 3 * (__brev(bits_b) >> 4); bits_a = bits_b = 0; }
        }
        pre = are; pim = aim;
    }
    return acc;
}

template <int RUN, int WARPS, int CTAS>
__global__ void __launch_bounds__(WARPS * 32, CTAS) k_probe(uint32_t *out, int iters) {
    constexpr int STRIDE = 2 * RUN;                              // bytes per run
    constexpr int SEG = WARPS * 32 * STRIDE;
    extern __shared__ __align__(16) uint8_t buf[];
    for (int i = threadIdx.x; i < (2 * SEG + 512) / 4; i += blockDim.x) {
        uint32_t x = (uint32_t)i * 2654435761u + blockIdx.x;
        x ^= x >> 15; x *= 0x85EBCA6Bu; x ^= x >> 13;
        reinterpret_cast<uint32_t *>(buf)[i] = (x & 0x1F1F1F1Fu) + 0x70707070u;
    }
    __syncthreads();
    uint32_t acc = 0;
    for (int it = 0; it < iters; it++) {
        const uint8_t *base = buf + STRIDE * threadIdx.x + 8 * (it & 7);
        acc ^= body<RUN>(base, base + SEG);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int RUN, int WARPS, int CTAS>
void run(const char *name, uint32_t *out, int iters) {
    const size_t smem = (size_t)2 * WARPS * 32 * 2 * RUN + 512;
    auto k = k_probe<RUN, WARPS, CTAS>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<148 * CTAS, WARPS * 32, smem>>>(out, 2);
    cudaEventRecord(e0);
    k<<<148 * CTAS, WARPS * 32, smem>>>(out, iters);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double lane_samples = (double)iters * 2 * RUN * (WARPS * CTAS / 4.0);
    printf("%-28s regs %3d spill %4zu B  %7.3f ms  %6.2f cyc/sample/lane@1965  %s\n", name, fa.numRegs,
           (size_t)fa.localSizeBytes, ms, ms * 1e-3 * 1.965e9 / lane_samples, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main(int argc, char **argv) {
    const int iters = argc > 1 ? atoi(argv[1]) : 1000;
    uint32_t *out;
    cudaMalloc(&out, 148 * 2 * 1024 * 4);
    run<28, 16, 1>("run 28, 16w x1", out, 2 * iters);
    run<56, 16, 1>("run 56, 16w x1", out, iters);
    run<56, 8, 2>("run 56,  8w x2", out, iters);
    run<84, 16, 1>("run 84, 16w x1", out, iters * 2 / 3);
    run<112, 8, 1>("run 112, 8w x1", out, iters / 2);
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
