// mma_probe.cu -- how long does one tcgen05.mma kind::i8 (M = 128, K = 32) take as a function
// of N and of the shared-memory layout of its operands?  (development tool)
// One thread per SM issues `reps` MMAs back to back, commits and waits.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/mma_probe tools/labs/mma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../rtldavis_b200/csrc/davis_async.cuh"
using namespace davis;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mk_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
           (1ULL << 46) | ((uint64_t)layout << 61);
}
__global__ void k_probe(int n, int reps, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_layout, uint32_t a_step,
                        uint32_t b_lbo, uint32_t b_sbo, uint32_t b_layout, int d_stages, int d_stride,
                        int mode, unsigned long long *out) {
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
    if (threadIdx.x == 0) {
        const uint32_t a0 = (smem_addr(sm) + 1023u) & ~1023u, b0 = a0 + 64 * 1024;
        uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
        if (mode == 1) idesc = (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);          // f16 x f16 -> f32
        if (mode == 5) idesc = (2u << 4) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((64u >> 4) << 24);
        const uint32_t a_tm = tmem + 448;                       // 3 A tiles of 8 columns for the TS modes
        const uint64_t bd = mk_desc(b0, b_lbo, b_sbo, b_layout);
        unsigned long long best = ~0ull;
        for (int trial = 0; trial < 4; trial++) {
            const unsigned long long t0 = clock64();
            for (int r = 0; r < reps; r++) {
                const uint64_t ad = mk_desc(a0 + (r & 7) * a_step, a_lbo, a_sbo, a_layout);
                const uint32_t d = tmem + (uint32_t)((r % d_stages) * d_stride);
                if (mode == 0 || mode == 5)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                                 ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                else if (mode == 1)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                                 ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                else {
                    const uint32_t at = a_tm + 8u * (uint32_t)(r % 3);
                    if (mode == 3 || mode == 4)
                        asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(at), "l"(ad) : "memory");
                    if (mode == 2 || mode == 4)
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                     "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
                                     ::"r"(d), "r"(at), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                }
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(&bar)) : "memory");
            mbar_wait(smem_addr(&bar), trial & 1);
            const unsigned long long t1 = clock64();
            if (t1 - t0 < best) best = t1 - t0;
        }
        out[blockIdx.x] = best;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

static void run(const char *name, int n, int reps, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_layout, uint32_t a_step,
                uint32_t b_lbo, uint32_t b_sbo, uint32_t b_layout, int d_stages, int d_stride, int mode = 0) {
    unsigned long long *out; CK(cudaMalloc(&out, 148 * 8));
    const size_t smem = 97 * 1024 + 1024;
    CK(cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_probe<<<148, 32, smem>>>(n, reps, a_lbo, a_sbo, a_layout, a_step, b_lbo, b_sbo, b_layout, d_stages, d_stride, mode, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%-60s %s\n", name, cudaGetErrorString(e)); exit(2); }
    unsigned long long h[148]; CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
    double s = 0; for (int i = 0; i < 148; i++) s += (double)h[i];
    printf("%-64s N=%3d reps=%3d  %8.1f cycles per MMA (total %.0f)\n", name, n, reps, s / 148 / reps, s / 148);
    cudaFree(out);
}

int main() {
    const int R = 64;
    for (int n : {16, 32, 48, 64, 128, 256}) {
        const int st = n <= 64 ? 64 : n, ns = 512 / st;
        run("A no-swizzle Hankel (LBO 16, SBO 128), B no-swizzle", n, R, 16, 128, 0, 2048, n * 16, 128, 0, ns, st);
    }
    for (int n : {32, 48}) {
        run("A no-swizzle canonical [k][rows] (LBO 2048, SBO 128)", n, R, 2048, 128, 0, 4096, n * 16, 128, 0, 8, 64);
        run("A no-swizzle canonical [rows][k] (LBO 128, SBO 256)", n, R, 128, 256, 0, 4096, n * 16, 128, 0, 8, 64);
        run("A swizzle-32B (SBO 256)", n, R, 16, 256, 6, 4096, n * 16, 128, 0, 8, 64);
        run("A swizzle-64B (SBO 512)", n, R, 16, 512, 4, 8192, n * 16, 128, 0, 8, 64);
        run("A swizzle-128B (SBO 1024)", n, R, 16, 1024, 2, 8192, n * 16, 128, 0, 8, 64);
        run("A swizzle-32B, B swizzle-32B", n, R, 16, 256, 6, 4096, 16, 256, 6, 8, 64);
        run("A Hankel, B swizzle-32B", n, R, 16, 128, 0, 2048, 16, 256, 6, 8, 64);
        run("A Hankel, same D every time", n, R, 16, 128, 0, 2048, n * 16, 128, 0, 1, 64);
        run("A Hankel, 1 MMA (latency)", n, 1, 16, 128, 0, 2048, n * 16, 128, 0, 1, 64);
        run("A swizzle-128B, 1 MMA (latency)", n, 1, 16, 1024, 2, 8192, n * 16, 128, 0, 1, 64);
    }
    for (int n : {16, 48, 256}) {
        const int st = n <= 64 ? 64 : n, ns = 448 / st;
        run("kind::f16 SS, A Hankel", n, R, 16, 128, 0, 2048, n * 16, 128, 0, ns, st, 1);
        run("kind::i8 TS (A in tensor memory)", n, R, 16, 128, 0, 2048, n * 16, 128, 0, ns, st, 2);
        run("kind::i8 SS M=64", n, R, 16, 128, 0, 2048, n * 16, 128, 0, ns, st, 5);
    }
    for (int n : {192, 240, 256}) {
        run("kind::i8 SS, A canonical no-swizzle, 3 k-steps same D", n, 63, 2048, 128, 0, 4096, n * 16, 128, 0, 1, 256);
        run("kind::i8 SS, A canonical no-swizzle, 2 D stages", n, 64, 2048, 128, 0, 4096, n * 16, 128, 0, 2, 256);
    }
    run("tcgen05.cp 128x256b Hankel source only", 48, R, 16, 128, 0, 2048, 48 * 16, 128, 0, 7, 64, 3);
    run("tcgen05.cp 128x256b Hankel + kind::i8 TS", 48, R, 16, 128, 0, 2048, 48 * 16, 128, 0, 7, 64, 4);
    run("tcgen05.cp 128x256b Hankel + kind::i8 TS", 32, R, 16, 128, 0, 2048, 32 * 16, 128, 0, 7, 64, 4);
    run("tcgen05.cp + TS, 1 pair (latency)", 48, 1, 16, 128, 0, 2048, 48 * 16, 128, 0, 7, 64, 4);
    return 0;
}
