// mma_probe2.cu -- what does one tcgen05.mma kind::i8 (M = 128, K = 32) cost when the issue loop
// costs nothing?  (development tool; answers VERDICT r02 "What's weak" 1)
//
// The round-2 probe (mma_probe.cu) rebuilt the 64-bit descriptor, took a run-time modulo and moved
// eight values into uniform registers per MMA: ~50 SASS instructions of one thread between two
// UTCIMMAs, and it measured that thread (a flat 168 cycles whatever N).  Here:
//   * N, the accumulator stage count and the number of MMAs per batch are template parameters;
//     the loop is fully unrolled, the accumulator column of MMA r is a compile-time constant;
//   * the A and B descriptors are built once; MMA r uses A descriptor + a compile-time constant in
//     the start-address field (the CUTLASS idiom: low word += bytes >> 4);
//   * one elected lane of a fully converged warp issues (elect.sync), so descriptors stay in the
//     uniform datapath; no clock64 inside a batch; ONE tcgen05.commit per batch.
// Reported per shape: cycles per MMA as the slope between a batch of R_LO and a batch of R_HI MMAs
// (removes the launch-to-drain latency of the last MMA, which is also printed).
// `mma_probe2 sass` is not a mode: check the SASS with
//   cuobjdump -sass tools/labs/mma_probe2 | grep -c UTCIMMA      (and the lines between two of them)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tools/labs/mma_probe2 tools/labs/mma_probe2.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>
#include "../../rtldavis_b200/csrc/davis_async.cuh"
using namespace davis;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mk_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
           (1ULL << 46) | ((uint64_t)layout << 61);
}
__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}

// KIND 0: kind::i8 (s8 x s8 -> s32), 1: kind::f16 (f16 x f16 -> f32)
template <int N, int KIND>
__device__ __forceinline__ void mma(uint32_t d, uint64_t ad, uint64_t bd) {
    constexpr uint32_t idesc_i8 = (2u << 4) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint32_t idesc_f16 = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    if constexpr (KIND == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(d), "l"(ad), "l"(bd), "r"(idesc_i8), "r"(1u) : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(d), "l"(ad), "l"(bd), "r"(idesc_f16), "r"(1u) : "memory");
}

template <int N, int KIND, int R, int STAGES, int STRIDE, int A_STEP>
__device__ __forceinline__ void batch(uint32_t tmem, uint64_t ad, uint64_t bd, uint32_t bar) {
#pragma unroll
    for (int r = 0; r < R; r++)
        mma<N, KIND>(tmem + (uint32_t)((r % STAGES) * STRIDE), ad + (uint64_t)(((r & 7) * A_STEP) >> 4), bd);
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// LAYOUT 0: A = no-swizzle Hankel view (LBO 16, SBO 128, the FIR's raw-byte operand), B no-swizzle
// LAYOUT 1: A = B = swizzle-128B K-major (what a GEMM library uses)
template <int N, int KIND, int R_LO, int R_HI, int LAYOUT>
__global__ void __launch_bounds__(32) k_probe2(unsigned long long *out) {
    extern __shared__ __align__(1024) uint8_t sm[];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint32_t tmem_base;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(sm)[i] = i * 2654435761u;
    if (threadIdx.x == 0) { mbar_init(smem_addr(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&tmem_base)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base;
    constexpr int STRIDE = N <= 64 ? 64 : N;                      // accumulator columns per stage
    constexpr int STAGES = 512 / STRIDE;
    const uint32_t a0 = (smem_addr(sm) + 1023u) & ~1023u, b0 = a0 + 64 * 1024;
    const uint32_t bar_a = smem_addr(&bar);
    constexpr int A_STEP = LAYOUT == 0 ? 2048 : 8192;
    const uint64_t ad = LAYOUT == 0 ? mk_desc(a0, 16, 128, 0) : mk_desc(a0, 16, 1024, 2);
    const uint64_t bd = LAYOUT == 0 ? mk_desc(b0, N * 16, 128, 0) : mk_desc(b0, 16, 1024, 2);
    unsigned long long best_lo = ~0ull, best_hi = ~0ull;
    uint32_t phase = 0;
    for (int trial = 0; trial < 4; trial++) {
        unsigned long long t0 = clock64();
        if (elect_one()) batch<N, KIND, R_LO, STAGES, STRIDE, A_STEP>(tmem, ad, bd, bar_a);
        __syncwarp();
        mbar_wait(bar_a, phase); phase ^= 1;
        unsigned long long t1 = clock64();
        if (t1 - t0 < best_lo) best_lo = t1 - t0;
        t0 = clock64();
        if (elect_one()) batch<N, KIND, R_HI, STAGES, STRIDE, A_STEP>(tmem, ad, bd, bar_a);
        __syncwarp();
        mbar_wait(bar_a, phase); phase ^= 1;
        t1 = clock64();
        if (t1 - t0 < best_hi) best_hi = t1 - t0;
    }
    if (threadIdx.x == 0) { out[2 * blockIdx.x] = best_lo; out[2 * blockIdx.x + 1] = best_hi; }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

template <int N, int KIND, int LAYOUT>
static void run(const char *name, int ctas) {
    constexpr int R_LO = 32, R_HI = 160;
    unsigned long long *out; CK(cudaMalloc(&out, 148 * 16));
    const size_t smem = 97 * 1024 + 1024;
    auto k = k_probe2<N, KIND, R_LO, R_HI, LAYOUT>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<ctas, 32, smem>>>(out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%-40s %s\n", name, cudaGetErrorString(e)); exit(2); }
    unsigned long long h[296]; CK(cudaMemcpy(h, out, ctas * 16, cudaMemcpyDeviceToHost));
    double lo = 0, hi = 0; for (int i = 0; i < ctas; i++) { lo += (double)h[2 * i]; hi += (double)h[2 * i + 1]; }
    lo /= ctas; hi /= ctas;
    const double per = (hi - lo) / (R_HI - R_LO);
    const double macs = 128.0 * N * (KIND == 0 ? 32 : 16);
    printf("%-34s N=%3d  %7.1f cycles/MMA (slope %d..%d)  fixed %6.0f  => %6.0f MAC/clk/SM  [batch of %d: %.1f/MMA]\n",
           name, N, per, R_LO, R_HI, lo - per * R_LO, macs / per, R_HI, hi / R_HI);
    cudaFree(out);
}

int main(int argc, char **argv) {
    const int ctas = argc > 1 ? atoi(argv[1]) : 148;
    printf("tcgen05.mma M=128, one elected lane, fully unrolled issue, one commit per batch; %d CTAs (1 per SM)\n", ctas);
    run<16, 0, 0>("i8 K=32, A Hankel no-swizzle", ctas);
    run<32, 0, 0>("i8 K=32, A Hankel no-swizzle", ctas);
    run<48, 0, 0>("i8 K=32, A Hankel no-swizzle", ctas);
    run<64, 0, 0>("i8 K=32, A Hankel no-swizzle", ctas);
    run<128, 0, 0>("i8 K=32, A Hankel no-swizzle", ctas);
    run<256, 0, 0>("i8 K=32, A Hankel no-swizzle", ctas);
    run<16, 0, 1>("i8 K=32, A/B swizzle-128B", ctas);
    run<48, 0, 1>("i8 K=32, A/B swizzle-128B", ctas);
    run<64, 0, 1>("i8 K=32, A/B swizzle-128B", ctas);
    run<128, 0, 1>("i8 K=32, A/B swizzle-128B", ctas);
    run<256, 0, 1>("i8 K=32, A/B swizzle-128B", ctas);
    run<64, 1, 1>("f16 K=16, A/B swizzle-128B", ctas);
    run<128, 1, 1>("f16 K=16, A/B swizzle-128B", ctas);
    run<256, 1, 1>("f16 K=16, A/B swizzle-128B", ctas);
    return 0;
}
