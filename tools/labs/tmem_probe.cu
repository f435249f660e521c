// tmem_probe.cu -- tensor-memory load/store throughput per SM (development tool).
// 16 warps (4 per lane quadrant) each issue `reps` tcgen05.ld (or .st) of 32 / 16 / 8 columns.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/labs/tmem_probe tools/labs/tmem_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

template <int COLS, bool STORE>
__global__ void k_tmem(int reps, int warps_used, unsigned long long *out, uint32_t *sink) {
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t taddr = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * 64);
    uint32_t acc = 0;
    uint32_t v[32];
#pragma unroll
    for (int i = 0; i < 32; i++) v[i] = threadIdx.x + i;
    __syncthreads();
    const unsigned long long t0 = clock64();
    if (warp < warps_used) {
        for (int r = 0; r < reps; r++) {
            if constexpr (STORE) {
                if constexpr (COLS == 32)
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
                else
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                                 ::"r"(taddr + (r & 7) * 8), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
                if ((r & 3) == 3) asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            } else {
                if constexpr (COLS == 32)
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                                 : "r"(taddr) : "memory");
                else
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                                 : "r"(taddr + (r & 7) * 8) : "memory");
                if ((r & 3) == 3) {
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    acc ^= v[0] ^ v[7];
                }
            }
        }
        if constexpr (STORE) asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        else asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    }
    __syncthreads();
    const unsigned long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    if (acc == 0x12345678u) sink[0] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

template <int COLS, bool STORE>
static void run(const char *name, int warps) {
    const int reps = 2048;
    unsigned long long *out; uint32_t *sink;
    CK(cudaMalloc(&out, 148 * 8)); CK(cudaMalloc(&sink, 4));
    k_tmem<COLS, STORE><<<148, 512>>>(reps, warps, out, sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); exit(2); }
    unsigned long long h[148]; CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
    double s = 0; for (int i = 0; i < 148; i++) s += (double)h[i];
    s /= 148;
    const double bytes = (double)warps * reps * 32 * COLS * 4;
    printf("%-28s %2d warps: %8.0f cycles, %7.1f B/clk/SM, %6.1f cycles per instruction and warp\n", name, warps, s, bytes / s, s / reps);
    cudaFree(out); cudaFree(sink);
}

int main() {
    for (int w : {1, 4, 8, 16}) {
        run<32, false>("tcgen05.ld 32x32b.x32", w);
        run<8, false>("tcgen05.ld 32x32b.x8", w);
        run<32, true>("tcgen05.st 32x32b.x32", w);
        run<8, true>("tcgen05.st 32x32b.x8", w);
    }
    return 0;
}
