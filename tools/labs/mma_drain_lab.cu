// mma_drain_lab.cu -- LAB, not product code.  VERDICT r02 item 1(b), first number: what do the
// copy + tcgen05.mma + tensor-memory drain of the Hankel-descriptor FIR (variant B1 of
// profiles/r02_mma_fir_lab.txt) cost for the bench workload when the issue path is cheap?
//
// Round 2's lab kernel waited on an mbarrier, rebuilt a descriptor and committed once per MMA and
// measured 1.83 ms "accumulators only drained" -- its control thread, not the tensor core
// (profiles/r03_mma_probe2.txt: an M128 x N48 x K32 kind::i8 instruction takes 44 cycles).  Here:
//   warp 0   producer: one elected lane streams 16 KiB pieces of the byte stream (+ 32 B of lead)
//            into a 4-slot shared-memory ring with cp.async.bulk (UBLKCP);
//   warp 1   MMA issuer: one elected lane; a slot = 8 MMAs = two accumulator stages of 4 MMAs
//            (4 x 48 = 192 tensor-memory columns each); per STAGE one tm_empty wait, four
//            back-to-back UTCIMMAs whose A descriptors are slot base + compile-time constant
//            (raw bytes through the LBO = 16 B / SBO = 128 B Hankel view: row r = bytes
//            [16 r, 16 r + 32), 1 024 samples per instruction), one commit;
//   warps 2..17  16 drain warps: (quadrant q = warp % 4, MMA i of the stage); each reads its
//            32 lanes x 48 columns with three tcgen05.ld.32x32b.x16, optionally re-arms them with
//            the bias pattern (tcgen05.st, what the real epilogue needs so that tcgen05.ld returns
//            floats), and XORs the values (kept so that nothing is optimised away).
// No filter arithmetic, no bitmap: this is the floor under a tcgen05 k_demod, to be compared with
// 2.21 ms (SIMT k_demod) and 0.84 ms (5.5 GB at the measured HBM peak).
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -DDAVIS_MBAR_WATCHDOG \
//             -o tools/labs/mma_drain_lab tools/labs/mma_drain_lab.cu
// Run:   tools/labs/mma_drain_lab [streams=1024] [blocks=328] [reps=5]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../rtldavis_b200/csrc/davis_async.cuh"
using namespace davis;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

#ifndef LAB_L
#define LAB_L 3
#endif
constexpr int L = LAB_L;                          // tap limbs (3: 24-bit taps, 2: 16-bit taps)
constexpr int N = 16 * L;                          // 16 columns per limb
constexpr int MMA_BYTES = 2048;                    // 1 024 samples
#ifndef LAB_STAGE_MMAS
#define LAB_STAGE_MMAS 4
#endif
#ifndef LAB_NSTAGE
#define LAB_NSTAGE 2
#endif
#ifndef LAB_NSLOT
#define LAB_NSLOT 4
#endif
constexpr int STAGE_MMAS = LAB_STAGE_MMAS, STAGE_COLS = STAGE_MMAS * N;        // 192
constexpr int NSTAGE = LAB_NSTAGE;
static_assert(STAGE_MMAS * LAB_NSTAGE == 8 && STAGE_COLS * NSTAGE <= 512, "a 16 KiB slot is 8 MMAs");
constexpr int SLOT_MMAS = STAGE_MMAS * NSTAGE;     // 8
constexpr int PIECE = SLOT_MMAS * MMA_BYTES;       // 16 384 bytes of new samples per slot
constexpr int LEAD = 32;
constexpr int SLOT_BYTES = PIECE + LEAD + 96;      // multiple of 128
constexpr int NSLOT = LAB_NSLOT;
constexpr int DRAIN_WARPS = 4 * STAGE_MMAS;
constexpr int THREADS = (2 + DRAIN_WARPS) * 32;

__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ULL << 46);
}
__device__ __forceinline__ void tc_mma_i8(uint32_t d, uint64_t ad, uint64_t bd, uint32_t acc) {
    constexpr uint32_t idesc = (2u << 4) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}

struct Smem {
    alignas(128) uint8_t slot[NSLOT][SLOT_BYTES];
    alignas(128) uint8_t bmat[N * 32];
    unsigned long long full[NSLOT], slot_free[NSLOT], tm_full[NSTAGE], tm_empty[NSTAGE];
    uint32_t tmem_base;
};

// units = pieces of 16 KiB; unit u = (stream u / per_row, piece u % per_row); CTA b takes b, b + grid, ...
// SKIP 0: everything; 1: drain warps do not read tensor memory; 2: no MMAs either (copy + handshakes)
template <bool BIAS, int SKIP, bool EPI>
__global__ void __launch_bounds__(THREADS, 1)
k_drain(const uint8_t *__restrict__ iq, size_t row_bytes, int per_row, long long units, uint32_t *__restrict__ sink) {
    extern __shared__ __align__(128) uint8_t raw[];
    Smem &sm = *reinterpret_cast<Smem *>(raw + ((128u - (smem_addr(raw) & 127u)) & 127u));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int e = threadIdx.x; e < N * 32; e += THREADS) sm.bmat[e] = (uint8_t)(int8_t)((e * 37 + 11) % 251 - 125);
    for (int e = threadIdx.x; e < NSLOT * LEAD; e += THREADS) sm.slot[e / LEAD][e % LEAD] = 127;   // lead of a row's first piece
    if (threadIdx.x == 0) {
        for (int i = 0; i < NSLOT; i++) { mbar_init(smem_addr(&sm.full[i]), 1); mbar_init(smem_addr(&sm.slot_free[i]), 1); }
        for (int i = 0; i < NSTAGE; i++) { mbar_init(smem_addr(&sm.tm_full[i]), 1); mbar_init(smem_addr(&sm.tm_empty[i]), DRAIN_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&sm.tmem_base)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = sm.tmem_base;
    if (BIAS) {
        // the first MMAs accumulate onto the bias: store it before anybody issues one
        if (warp >= 2) {
            uint32_t b0[16];
#pragma unroll
            for (int e = 0; e < 16; e++) b0[e] = 0x4B400000u - 77u;
            const uint32_t t = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(((warp - 2) >> 2) * N);
#pragma unroll
            for (int h = 0; h < NSTAGE; h++)
#pragma unroll
                for (int c = 0; c < L; c++) tc_st16(t + h * STAGE_COLS + 16 * c, b0);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    const long long first = blockIdx.x, step = gridDim.x;
    const int count = units > first ? (int)((units - first + step - 1) / step) : 0;

    if (warp == 0) {
        // ===== producer =====
        if (elect_one()) {
            for (int k = 0; k < count; k++) {
                const int s = k % NSLOT;
                mbar_wait(smem_addr(&sm.slot_free[s]), ((k / NSLOT) & 1) ^ 1);     // first round: free
                const long long u = first + (long long)k * step;
                const long long r = u / per_row, j = u % per_row;
                const uint8_t *src = iq + (size_t)r * row_bytes + (size_t)j * PIECE;
                const uint32_t dst = smem_addr(sm.slot[s]), bar = smem_addr(&sm.full[s]);
                if (j == 0) {
                    mbar_expect_tx(bar, PIECE);
                    bulk_g2s(dst + LEAD, src, PIECE, bar);
                } else {
                    mbar_expect_tx(bar, PIECE + LEAD);
                    bulk_g2s(dst, src - LEAD, PIECE + LEAD, bar);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            const uint64_t bd = smem_desc(smem_addr(sm.bmat), N * 16, 128);
            uint64_t ad[NSLOT];
#pragma unroll
            for (int s = 0; s < NSLOT; s++) ad[s] = smem_desc(smem_addr(sm.slot[s]) + 16, 16, 128);   // sample -8 of the piece
            for (int k0 = 0; k0 < count; k0 += NSLOT) {
                const uint32_t ph = (uint32_t)(k0 / NSLOT) & 1u;
#pragma unroll
                for (int s = 0; s < NSLOT; s++) {
                    if (k0 + s < count) {
                        mbar_wait(smem_addr(&sm.full[s]), ph);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                        for (int h = 0; h < NSTAGE; h++) {
                            // stage h is used once per slot: use number k0 + s, parity of that count
                            mbar_wait(smem_addr(&sm.tm_empty[h]), ((uint32_t)(k0 + s) & 1u) ^ 1u);
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                            for (int i = 0; i < STAGE_MMAS; i++)
                                if (SKIP < 2) tc_mma_i8(tmem + (uint32_t)(h * STAGE_COLS + i * N),
                                          ad[s] + (uint64_t)(((h * STAGE_MMAS + i) * MMA_BYTES) >> 4), bd, BIAS ? 1u : 0u);
                            tc_commit(smem_addr(&sm.tm_full[h]));
                        }
                        tc_commit(smem_addr(&sm.slot_free[s]));
                    }
                }
            }
        }
    } else {
        // ===== drain warps =====
        const int q = warp & 3, i = (warp - 2) >> 2;
        const uint32_t t0 = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(i * N);
        uint32_t bias[16];
#pragma unroll
        for (int e = 0; e < 16; e++) asm volatile("mov.b32 %0, %1;" : "=r"(bias[e]) : "r"(0x4B400000u - 77u));
        uint32_t x = 0;
        for (int k = 0; k < count; k++) {
#pragma unroll
            for (int h = 0; h < NSTAGE; h++) {
                mbar_wait(smem_addr(&sm.tm_full[h]), (uint32_t)k & 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                uint32_t v[L][16];
#pragma unroll
                for (int c = 0; c < L; c++) {
                    if (SKIP == 0) tc_ld16(t0 + h * STAGE_COLS + 16 * c, v[c]);
                    else
#pragma unroll
                        for (int e = 0; e < 16; e++) v[c][e] = k + e;
                }
                if (SKIP == 0) asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (BIAS) {
#pragma unroll
                    for (int c = 0; c < L; c++) tc_st16(t0 + h * STAGE_COLS + 16 * c, bias);
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                }
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_addr(&sm.tm_empty[h]));
                if (EPI) {
                    // The arithmetic of the real epilogue on these 48 values (cost model, not checked
                    // here: the B matrix is not the tap matrix).  Column = 24 half + 8 limb + 2 s' + comp,
                    // lane = operand row = 8 consecutive samples.  int -> float: one integer add of the
                    // magic constant per value when the accumulators were not pre-biased.
                    float2 z[9];
#pragma unroll
                    for (int sidx = 0; sidx < 8; sidx++) {
                        const int hh = sidx >> 2, sp = sidx & 3;
                        float f[2][L];
#pragma unroll
                        for (int comp = 0; comp < 2; comp++)
#pragma unroll
                            for (int l = 0; l < L; l++) {
                                const int col = 8 * L * hh + 8 * l + 2 * sp + comp;
                                const uint32_t w = v[col >> 4][col & 15];
                                f[comp][l] = __uint_as_float(BIAS ? w : w + (0x4B400000u - 77u));
                            }
                        if (L == 3) {
                            z[sidx].x = f[0][0] + fmaf(f[0][1], 256.f, fmaf(f[0][L - 1], 65536.f, -65793.f * 12582912.f));
                            z[sidx].y = f[1][0] + fmaf(f[1][1], 256.f, fmaf(f[1][L - 1], 65536.f, -65793.f * 12582912.f));
                        } else {
                            z[sidx].x = f[0][0] + fmaf(f[0][1], 256.f, -257.f * 12582912.f);
                            z[sidx].y = f[1][0] + fmaf(f[1][1], 256.f, -257.f * 12582912.f);
                        }
                    }
                    z[8].x = __shfl_down_sync(0xFFFFFFFFu, z[0].x, 1);
                    z[8].y = __shfl_down_sync(0xFFFFFFFFu, z[0].y, 1);
                    uint32_t acc = 0;
#pragma unroll
                    for (int sidx = 7; sidx >= 0; sidx--)
                        acc = __funnelshift_l(__float_as_uint(fmaf(-z[sidx].y, z[sidx + 1].y, -z[sidx].x * z[sidx + 1].x)), acc, 1);
                    x = (x << 8) ^ acc;
                } else {
#pragma unroll
                    for (int c = 0; c < L; c++)
#pragma unroll
                        for (int e = 0; e < 16; e++) x ^= v[c][e];
                }
            }
        }
        sink[(size_t)blockIdx.x * THREADS + threadIdx.x] = x;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

__global__ void k_fill(uint8_t *p, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        p[i] = (uint8_t)(127 + (int)((i * 2654435761u >> 13) % 41) - 20);
}

template <bool BIAS, int SKIP, bool EPI = false>
static void measure(const char *name, const uint8_t *iq, size_t row, int per_row, long long units, uint32_t *sink, int reps, int sms) {
    const size_t smem = sizeof(Smem) + 256;
    CK(cudaFuncSetAttribute(k_drain<BIAS, SKIP, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f, sum = 0;
    for (int r = 0; r < reps + 1; r++) {
        CK(cudaEventRecord(e0));
        k_drain<BIAS, SKIP, EPI><<<sms, THREADS, smem>>>(iq, row, per_row, units, sink);
        CK(cudaEventRecord(e1));
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); exit(2); }
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (r) { best = ms < best ? ms : best; sum += ms; }
    }
    const double bytes = (double)units * PIECE;
    printf("%-46s best %7.3f ms  mean %7.3f ms  %7.1f GB/s  (%lld MMAs per SM, %.1f SM cycles per MMA at 1965 MHz)\n",
           name, best, sum / reps, bytes / best / 1e6, units * SLOT_MMAS / sms, best * 1e-3 * 1.965e9 / (units * SLOT_MMAS / sms));
}

int main(int argc, char **argv) {
    const int n_streams = argc > 1 ? atoi(argv[1]) : 1024, n_blocks = argc > 2 ? atoi(argv[2]) : 328, reps = argc > 3 ? atoi(argv[3]) : 5;
    const size_t row = (size_t)n_blocks * 16384;
    const int per_row = (int)(row / PIECE);
    const long long units = (long long)n_streams * per_row;
    int sms = 148; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    uint8_t *iq; CK(cudaMalloc(&iq, row * n_streams + 256));
    uint32_t *sink; CK(cudaMalloc(&sink, (size_t)sms * THREADS * 4));
    k_fill<<<sms * 8, 256>>>(iq, row * n_streams); CK(cudaDeviceSynchronize());
    printf("mma_drain_lab: %d streams x %d blocks = %.2f GB, %lld pieces of %d B, %d SMs, smem %zu B\n",
           n_streams, n_blocks, row * n_streams / 1e9, units, PIECE, sms, sizeof(Smem) + 256);
    printf("L = %d limbs (N = %d); stages: %d x %d MMAs (%d columns each), %d slots, %d drain warps\n", L, N, NSTAGE, STAGE_MMAS, STAGE_COLS, NSLOT, DRAIN_WARPS);
    measure<false, 2>("copy + handshakes only (no MMA, no ld)", iq, row, per_row, units, sink, reps, sms);
    measure<false, 1>("copy + MMA stages, accumulators not read", iq, row, per_row, units, sink, reps, sms);
    measure<false, 0>("copy + MMA stages + tcgen05.ld drain", iq, row, per_row, units, sink, reps, sms);
    measure<true, 0>("same + bias re-arm (tcgen05.st) per stage", iq, row, per_row, units, sink, reps, sms);
    measure<false, 0, true>("drain + epilogue arithmetic (IADD magic)", iq, row, per_row, units, sink, reps, sms);
    measure<true, 0, true>("bias re-arm + epilogue arithmetic", iq, row, per_row, units, sink, reps, sms);
    measure<false, 1, true>("epilogue arithmetic on registers, no tcgen05.ld", iq, row, per_row, units, sink, reps, sms);
    return 0;
}
