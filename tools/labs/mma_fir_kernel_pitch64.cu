// tools/labs/mma_fir_kernel_pitch64.cu -- LAB RECORD, not product code: variant B2 of profiles/r02_mma_fir_lab.txt.
// Correct (z and bitmap checked by tools/labs/mma_fir_lab.cu) and slower than the SIMT kernel; not built into the library.
// davis_demod_mma.cu -- the streaming demodulation kernel on the 5th-generation tensor
// cores (uint8 IQ -> decision bitmap), SURVEY section 7 "hard part 1" contingency.
//
// Replaces, like davis_demod.cu, ByteToCmplxLUT.execute -> rotate_fs4 -> fir9 ->
// discriminate -> quantize of Demodulator.demodulate (src/rtldavis/dsp.py:139-166; Go
// dsp/dsp.go:274-292) for a whole batch of streams and blocks.  The SIMT kernel spends
// 23 of its 30 issue cycles per sample on widening the bytes and on the 9-tap filter
// (profiles/r01_step_v12_all_kernels_ncu_full.txt: FMA pipe 80 %, DRAM 33 %).  Here the
// filter runs on the tensor cores and no arithmetic instruction touches an input byte:
//
//   * Folded filter.  With z[n] = sum_m c[m] j^m x[n+m]  (x = unrotated sample, the Fs/4
//     rotation of dsp.py:42-49 folded into the taps) the reference's F[g] = j^(g-9) z[g-9]
//     and its discriminator numerator (dsp.py:88) is  -Re(conj(z[g-10]) z[g-9]).  Re z[n]
//     and Im z[n] are two real 9-term dot products over the 18 interleaved bytes
//     u[2n .. 2n+17]: a banded Toeplitz product  Z = U T  of the byte stream with a tap matrix.
//   * Shape.  Measured on this part (profiles/r02_mma_probe_*.txt): one tcgen05.mma costs
//     ~100 ns (200 SM cycles) whatever N (16..256), M (64/128), operand layout or kind --
//     cuBLAS' own bf16 rate is one 128 x 256 x 16 instruction per 98 ns.  So the cost of the
//     filter is the NUMBER of MMA instructions, and an operand row must yield as many
//     samples as the 256 accumulator columns hold.  Row r of the A operand is the 96-byte
//     window that starts 64 r bytes (32 samples) into the stream: three K = 32 steps
//     (kind::i8, M = 128, N = 256) give z for the row's 32 samples plus one more (the
//     partner of the last decision), 3 limbs x 2 components each: 4096 samples per three
//     instructions.  (A 16-byte row pitch needs no re-layout -- the raw stream through a
//     no-swizzle descriptor with a 16-byte K-chunk stride IS that Hankel matrix, and the
//     hardware accepts it: profiles/r02_mma_fir_lab_v1_pitch16.txt -- but yields only 1024
//     samples per instruction: 1.84 ms of tensor time per 2.75 Gsample.)
//   * Re-layout.  The bulk-copy engine lands the raw bytes in shared memory; four
//     formatter warps copy 16-byte cells (LDS.128 / STS.128, no arithmetic) into the canonical
//     no-swizzle K-major operand: cell (row r, K chunk c) at c * A_LBO + 16 r.
//   * Limbs.  The taps are the reference's decimal constants scaled to 24-bit integers
//     (relative tap error 5e-8; the scale is chosen so that all five distinct taps land
//     within 0.04 of an integer) and split into three balanced signed 8-bit digits; the
//     u8 x s8 -> s32 products are exact.  The accumulator columns are pre-loaded
//     (tcgen05.st) with 0x4B400000 - kappa, so that what tcgen05.ld returns is already the
//     IEEE float 1.5 * 2^23 + (limb sum - kappa): no I2F and no integer add in the
//     epilogue.  kappa (1 + 256 + 65536) is 127.4 * sum(signed taps), the LUT offset of
//     dsp.py:26 passed through the filter (the LUT scale 1/127.6 cannot change a sign).
//   * Epilogue (SIMT, ~12 issue cycles per sample instead of 30): 3 fused multiply-adds
//     per component put the limbs together, one multiply + one FMA give the discriminator
//     numerator, one funnel shift per sample packs its sign.  An epilogue warp owns a lane
//     quadrant and one of four column parts: 8 consecutive samples (+ the 9th z) of 32 rows.
//   * The 8-bit pieces go to a shared-memory bit ring and are re-cut into the bitmap format
//     of the search stage (word w = samples [28 w, 28 w + 28)).
//
// Decomposition.  One persistent CTA per SM walks a contiguous range of tiles (28672 samples
// = 7 MMA sets = 1024 bitmap words) of the flattened (stream, tile) space.  Roles: one
// control thread (bulk copies, MMAs), 4 formatter warps, 16 epilogue warps; raw slots,
// operand buffers and the two accumulator stages are handed over through mbarriers.
//
// Every decision that reaches a caller is re-derived by k_exact (davis_search.cu), as for
// the SIMT kernel; tests/test_guard_band.py holds the numpy model of this arithmetic.
#include <algorithm>

#include "../../rtldavis_b200/csrc/davis_async.cuh"
#include "../../rtldavis_b200/csrc/davis_kernels.cuh"

namespace davis {
namespace mma {

constexpr int EPI_WARPS = 16, FMT_WARPS = 4;
constexpr int CTL_WARP = EPI_WARPS + FMT_WARPS;
constexpr int THREADS = (CTL_WARP + 1) * 32;
constexpr int ROW_SAMPLES = 32, ROW_BYTES = 2 * ROW_SAMPLES;     // operand row pitch
constexpr int KSTEPS = 3;                                        // 96-byte windows
constexpr int SET_SAMPLES = 128 * ROW_SAMPLES, SET_BYTES = 2 * SET_SAMPLES;   // 4096 / 8192
constexpr int SETS_PER_TILE = TILE_SAMPLES / SET_SAMPLES;
static_assert(SETS_PER_TILE * SET_SAMPLES == TILE_SAMPLES, "tiles are whole MMA sets");
// raw slot: V bytes [2 (4096 k - 16), 2 (4096 k - 16) + RAW_BYTES) of set k; row r's window
// starts at slot offset 16 + 64 r (z index 4096 k - 8 + 32 r)
constexpr int RAW_LEAD = 32;
constexpr int RAW_BYTES = 16 + ROW_BYTES * 127 + 32 * KSTEPS;    // 8240
constexpr int RAW_PITCH = ((RAW_BYTES + 127) / 128) * 128;
constexpr int NRAW = 6;                          // raw sets in flight (HBM latency)
constexpr int A_LBO = 2048 + 32;                 // K-chunk plane stride (+32: conflict-free cell stores)
constexpr int A_BYTES = 2 * KSTEPS * A_LBO;
constexpr int A_PITCH = ((A_BYTES + 127) / 128) * 128;
constexpr int NABUF = 3;
constexpr int L = 3;                             // limbs
constexpr int PART_SAMPLES = 8, PARTS = ROW_SAMPLES / PART_SAMPLES;
constexpr int PART_COLS = 64;                    // samples 0..4 in columns 0..29, 5..8 in 32..55
constexpr int N_COLS = PARTS * PART_COLS;        // 256
constexpr int B_LBO = (N_COLS / 8) * 128, B_BYTES = 2 * B_LBO;   // per K step
constexpr int NSTAGE = 2, STAGE_COLS = 256, TMEM_COLS = 512;
constexpr int RING_BYTES = 4 * SET_SAMPLES / 8;  // four sets of decision bits
constexpr uint32_t MAGIC = 0x4B400000u;          // 1.5 * 2^23: ulp 1, +-2^22 of integer room

// ---- taps as integers ------------------------------------------------------------------
// round(S c[m]) for the five distinct taps of dsp.py:56-69; S = 3.5289e7 is the scale inside
// the digit range at which all five products are nearest to integers.
__host__ __device__ constexpr int tap5(int i) {
    return i == 0 ? 623995 : i == 1 ? 1699934 : i == 2 ? 4320285 : i == 3 ? 6966413 : 8068069;
}
__host__ __device__ constexpr int balanced_digit(long long v, int l) {
    int d = 0;
    for (int i = 0; i <= l; i++) {
        d = (int)(((v % 256) + 256 + 128) % 256) - 128;
        v = (v - d) / 256;
    }
    return d;
}
__host__ __device__ constexpr int tap_digit(int m, int l) { return balanced_digit(tap5(m < 5 ? m : 8 - m), l); }
// 127.4 * (sum of the taps with the rotation's signs) = 127.4 * (2 t0 - 2 t2 + t4), the same
// for both components.  It is subtracted as the SAME integer kappa on every limb
// (kappa * 65793 ~ offset), so that one set of eight registers re-arms every accumulator
// column; what is left over (55 of 2^31: below the rounding of the fp32 combine) is dropped.
__host__ __device__ constexpr int kappa() {
    const long long s = 2LL * tap5(0) - 2LL * tap5(2) + tap5(4);
    return (int)(((s * 1274 + 5) / 10 + 65793 / 2) / 65793);
}
// tap of byte offset o (0..17) of a sample's window for component comp, with the Fs/4
// rotation's sign: z = sum_m c[m] j^m (I + jQ)
__host__ __device__ constexpr int byte_tap(int o, int comp, int l) {
    if (o < 0 || o >= 18) return 0;
    const int m = o >> 1, q = o & 1, ph = m & 3;
    // Re: +I, -Q, -I, +Q     Im: +Q, +I, -Q, -I      (phase 0, 1, 2, 3)
    const bool uses_q = comp == 0 ? (ph & 1) : !(ph & 1);
    if ((q == 1) != uses_q) return 0;
    const int sign = comp == 0 ? ((ph == 0 || ph == 3) ? 1 : -1) : ((ph == 0 || ph == 1) ? 1 : -1);
    return sign * tap_digit(m, l);
}

// ---- tcgen05 wrappers ---------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t *v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld8(uint32_t taddr, uint32_t *v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld4(uint32_t taddr, uint32_t *v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld2(uint32_t taddr, uint32_t *v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0,%1}, [%2];"
                 : "=r"(v[0]), "=r"(v[1]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// K-major, no swizzle: 8 x 16-byte core matrices; LBO = stride between the two K chunks of an
// MMA, SBO = stride between 8-row groups (cute/arch/mma_sm100_desc.hpp SmemDescriptor)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) |
           ((uint64_t)(sbo_bytes >> 4) << 32) | (1ULL << 46);
}
// kind::i8: D = s32, A = u8, B = s8, both K-major, M = 128
__host__ __device__ constexpr uint32_t instr_desc(int n) {
    return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
}

// ---- shared memory ----------------------------------------------------------------------
struct Smem {
    alignas(128) uint8_t raw[NRAW][RAW_PITCH];
    alignas(128) uint8_t abuf[NABUF][A_PITCH];
    alignas(128) uint8_t bmat[KSTEPS][B_BYTES];
    alignas(16) uint8_t ring[RING_BYTES];
    unsigned long long raw_full[NRAW], raw_free[NRAW], a_ready[NABUF], a_free[NABUF];
    unsigned long long tm_full[NSTAGE], tm_empty[NSTAGE];
    uint32_t tmem_base;
};

struct SubSpan { const uint8_t *p0, *p1; uint32_t n0, n1, skip; };

// bytes [b_first, b_first + RAW_BYTES) of stream s's sequence V = carry || input, clipped to
// what exists: at most one piece of the carry buffer and one of the input
__device__ __forceinline__ SubSpan raw_span(const View &vw, int s, long long b_first) {
    const long long carry_bytes = 2LL * vw.n_carry;
    const long long total = carry_bytes + 2LL * vw.n_new;
    long long b0 = b_first, b1 = b_first + RAW_BYTES;
    SubSpan sp{nullptr, nullptr, 0u, 0u, 0u};
    if (b0 < 0) { sp.skip = (uint32_t)(-b0); b0 = 0; }
    if (b1 > total) b1 = total;
    if (b1 <= b0) return sp;
    if (b0 < carry_bytes) {
        const long long e = b1 < carry_bytes ? b1 : carry_bytes;
        sp.p0 = vw.carry + (size_t)s * vw.carry_stride + b0;
        sp.n0 = (uint32_t)(e - b0);
        b0 = e;
    }
    if (b0 < b1) {
        sp.p1 = vw.input + (size_t)s * vw.input_stride + (b0 - carry_bytes);
        sp.n1 = (uint32_t)(b1 - b0);
    }
    return sp;
}

// limb floats (1.5 * 2^23 + limb sum - kappa) -> z component scaled by S:  e0 + 256 e1 + 65536 e2
__device__ __forceinline__ float combine(float f0, float f1, float f2) {
    const float t2 = fmaf(f2, 65536.f, -65793.f * 12582912.f);       // exact: 2^16 (e2 - 49344)
    return f0 + fmaf(f1, 256.f, t2);
}
// sign bit set <=> Re(conj(a) b) > 0 <=> the reference's numerator Im F[g-1] Re F[g] - Re F[g-1] Im F[g] < 0
__device__ __forceinline__ uint32_t neg_dot_bits(float2 a, float2 b) {
    return __float_as_uint(fmaf(-a.y, b.y, -a.x * b.x));
}
// z[n] from the 18 bytes at p (the sample's I byte), with the tensor core's integer arithmetic
__device__ float2 z_from_bytes(const uint8_t *p) {
    int acc[2][L] = {{0, 0, 0}, {0, 0, 0}};
    for (int o = 0; o < 18; o++) {
        const int u = p[o];
        for (int comp = 0; comp < 2; comp++)
            for (int l = 0; l < L; l++) acc[comp][l] += byte_tap(o, comp, l) * u;
    }
    float f[2][L];
    for (int comp = 0; comp < 2; comp++)
        for (int l = 0; l < L; l++) f[comp][l] = __uint_as_float(MAGIC + (uint32_t)(acc[comp][l] - kappa()));
    return make_float2(combine(f[0][0], f[0][1], f[0][2]), combine(f[1][0], f[1][1], f[1][2]));
}

// position of a role in the (stream, tile, set) walk
struct Walk {
    int s, j, i;
    __device__ __forceinline__ void next(int tiles) {
        if (++i == SETS_PER_TILE) { i = 0; if (++j == tiles) { j = 0; s++; } }
    }
    __device__ __forceinline__ int set_in_stream() const { return j * SETS_PER_TILE + i; }
};

// LAB = 0: the product kernel.  LAB = 1 (tools/labs/mma_fir_lab.cu): z can be dumped.  LAB = 2: the
// epilogue only drains the accumulators (what copy + re-layout + MMA + tensor memory cost alone).
template <int LAB>
__global__ void __launch_bounds__(THREADS, 1)
k_demod_mma(View vw, uint32_t *__restrict__ words, int words_alloc, int tiles, long long units,
            float2 *__restrict__ z_dump, unsigned long long *prof) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    Smem &sm = *reinterpret_cast<Smem *>(smem_raw + ((128u - (smem_addr(smem_raw) & 127u)) & 127u));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // lab builds: cycles per phase and warp (prof != nullptr); PH(i) closes phase i
    unsigned long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = 0;
    auto PH = [&](int i) {
        if (LAB && prof) { const unsigned long long t = clock64(); pc[i] += t - tprev; tprev = t; }
    };

    // B operand of K step t: [K chunk][column group][8 columns][16 bytes]; column n = 64 part +
    // col(s') + 2 limb + comp, col(s') = 6 s' (s' < 5) or 32 + 6 (s' - 5), is sample slot
    // s = 8 part + s' (s' = 0..8) of the row; K index k of step t is byte 32 t + k of the row's window
    for (int e = threadIdx.x; e < KSTEPS * N_COLS * 32; e += THREADS) {
        const int t = e / (N_COLS * 32), n = (e / 32) % N_COLS, k = e & 31;
        const int part = n / PART_COLS, c64 = n % PART_COLS;
        int v = 0;
        if (c64 < 30 || (c64 >= 32 && c64 < 56)) {
            const int sp = c64 < 30 ? c64 / 6 : 5 + (c64 - 32) / 6, rem = c64 < 30 ? c64 % 6 : (c64 - 32) % 6;
            v = byte_tap(32 * t + k - 2 * (PART_SAMPLES * part + sp), rem & 1, rem >> 1);
        }
        sm.bmat[t][(k >> 4) * B_LBO + (n >> 3) * 128 + (n & 7) * 16 + (k & 15)] = (uint8_t)(int8_t)v;
    }
    for (int e = threadIdx.x; e < RING_BYTES / 4; e += THREADS) reinterpret_cast<uint32_t *>(sm.ring)[e] = 0u;
    if (threadIdx.x == 0) {
        for (int i = 0; i < NRAW; i++) {
            mbar_init(smem_addr(&sm.raw_full[i]), 1);
            mbar_init(smem_addr(&sm.raw_free[i]), FMT_WARPS);
        }
        for (int i = 0; i < NABUF; i++) {
            mbar_init(smem_addr(&sm.a_ready[i]), FMT_WARPS);
            mbar_init(smem_addr(&sm.a_free[i]), 1);
        }
        for (int i = 0; i < NSTAGE; i++) {
            mbar_init(smem_addr(&sm.tm_full[i]), 1);
            mbar_init(smem_addr(&sm.tm_empty[i]), EPI_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == CTL_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_addr(&sm.tmem_base)), "r"((uint32_t)TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // bmat is read by the tensor core
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sm.tmem_base;

    // this CTA's contiguous range of tiles of the flattened (stream, tile) space
    const long long u0 = units * blockIdx.x / gridDim.x, u1 = units * (blockIdx.x + 1) / gridDim.x;
    const int total = (int)(u1 - u0) * SETS_PER_TILE;            // MMA sets to process
    const Walk w0{(int)(u0 / tiles), (int)(u0 % tiles), 0};

    if (warp == CTL_WARP) {
        // ===== control thread: bulk copies and MMAs =====
        if (lane == 0 && total > 0) {
            constexpr uint32_t idesc = instr_desc(N_COLS);
            uint64_t b_desc[KSTEPS];
#pragma unroll
            for (int t = 0; t < KSTEPS; t++) b_desc[t] = smem_desc(smem_addr(sm.bmat[t]), B_LBO, 128);
            Walk pf = w0;                                        // next set to fetch
            auto fetch = [&](int slot) {
                const SubSpan sp = raw_span(vw, pf.s, (long long)pf.set_in_stream() * SET_BYTES - RAW_LEAD);
                const uint32_t dst = smem_addr(sm.raw[slot]), bar = smem_addr(&sm.raw_full[slot]);
                mbar_expect_tx(bar, sp.n0 + sp.n1);
                if (sp.n0) bulk_g2s(dst + sp.skip, sp.p0, sp.n0, bar);
                if (sp.n1) bulk_g2s(dst + sp.skip + sp.n0, sp.p1, sp.n1, bar);
                pf.next(tiles);
            };
            for (int i = 0; i < NRAW && i < total; i++) fetch(i);
#pragma unroll 1
            for (int m = 0; m < total; m++) {
                const int ab = m % NABUF, st = m % NSTAGE;
                PH(7);
                mbar_wait(smem_addr(&sm.a_ready[ab]), (m / NABUF) & 1);
                PH(0);
                mbar_wait(smem_addr(&sm.tm_empty[st]), (m / NSTAGE) & 1);
                PH(1);
                tc_fence_after();
                const uint32_t a0 = smem_addr(sm.abuf[ab]);
#pragma unroll
                for (int t = 0; t < KSTEPS; t++)
                    tc_mma_i8(tmem + st * STAGE_COLS, smem_desc(a0 + 2 * t * A_LBO, A_LBO, 128), b_desc[t], idesc, 1u);
                tc_commit(smem_addr(&sm.tm_full[st]));
                tc_commit(smem_addr(&sm.a_free[ab]));
                PH(2);
                if (m + NRAW < total) {                         // the formatters released this raw slot
                    mbar_wait(smem_addr(&sm.raw_free[m % NRAW]), (m / NRAW) & 1);
                    PH(3);
                    fetch(m % NRAW);
                    PH(4);
                }
            }
        }
    } else if (warp >= EPI_WARPS) {
        // ===== formatter warps: raw bytes -> canonical K-major operand, 16-byte cells =====
        const int ft = threadIdx.x - EPI_WARPS * 32;             // 0..127
#pragma unroll 1
        for (int m = 0; m < total; m++) {
            const int rs = m % NRAW, ab = m % NABUF;
            PH(7);
            mbar_wait(smem_addr(&sm.raw_full[rs]), (m / NRAW) & 1);
            PH(0);
            if (m >= NABUF) mbar_wait(smem_addr(&sm.a_free[ab]), ((m / NABUF) - 1) & 1);
            PH(1);
            const uint8_t *src = sm.raw[rs] + 16;
            uint8_t *dst = sm.abuf[ab];
            // source cell q = 4 r + c holds chunk c of row r, and (c < 2) chunk c + 4 of row r - 1
#pragma unroll
            for (int q = ft; q < 4 * 128 + 2; q += 128) {
                const uint4 cell = *reinterpret_cast<const uint4 *>(src + 16 * q);
                const int r = q >> 2, c = q & 3;
                if (r < 128) *reinterpret_cast<uint4 *>(dst + c * A_LBO + 16 * r) = cell;
                if (c < 2 && r >= 1) *reinterpret_cast<uint4 *>(dst + (c + 4) * A_LBO + 16 * (r - 1)) = cell;
            }
            if (m == 0 && w0.j > 0 && ft == 0) {
                // a range that starts inside a stream: the two decisions in front of its first
                // tile (V samples 28672 j, 28672 j + 1) pair z[-10], z[-9], z[-8] relative to the
                // tile; they are the top bits of the ring byte in front of the first set
                const uint8_t *p = sm.raw[rs] + RAW_LEAD;        // I byte of the tile's first sample
                const float2 za = z_from_bytes(p - 20), zb = z_from_bytes(p - 18), zc = z_from_bytes(p - 16);
                const int k = w0.set_in_stream();
                sm.ring[(((k + 3) & 3) * SET_SAMPLES / 8) + SET_SAMPLES / 8 - 1] =
                    (uint8_t)(((neg_dot_bits(za, zb) >> 31) << 6) | ((neg_dot_bits(zb, zc) >> 31) << 7));
            }
            PH(2);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(smem_addr(&sm.a_ready[ab]));
                mbar_arrive(smem_addr(&sm.raw_free[rs]));
            }
            PH(3);
        }
    } else {
        // ===== epilogue warps: quadrant q = TMEM lanes 32 q .. 32 q + 31, column part p =====
        const int q = warp & 3, part = warp >> 2, row = 32 * q + lane;
        const uint32_t tm_part = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(PART_COLS * part);
        // eight registers that stay put (opaque to the compiler, or it rebuilds them in front
        // of every tcgen05.st)
        uint32_t bias[8];
#pragma unroll
        for (int e = 0; e < 8; e++) asm volatile("mov.b32 %0, %1;" : "=r"(bias[e]) : "r"(MAGIC - (uint32_t)kappa()));
        if (total > 0) {
            for (uint32_t st = 0; st < NSTAGE; st++)
#pragma unroll
                for (int c = 0; c < 7; c++) tc_st8(tm_part + st * STAGE_COLS + 8 * c, bias);
            tc_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0)
                for (uint32_t st = 0; st < NSTAGE; st++) mbar_arrive(smem_addr(&sm.tm_empty[st]));
        }
        Walk w = w0;
#pragma unroll 1
        for (int m = 0; m < total; m++) {
            const uint32_t st = m % NSTAGE;
            const int k = w.set_in_stream();
            PH(7);
            mbar_wait(smem_addr(&sm.tm_full[st]), (m / NSTAGE) & 1);
            PH(0);
            tc_fence_after();
            const uint32_t ta = tm_part + st * STAGE_COLS;
            // columns 0..29: samples 0..4;  32..55: samples 5..8 (8 = the partner of decision 7)
            uint32_t va[30], vb[24];
            tc_ld16(ta, va); tc_ld8(ta + 16, va + 16); tc_ld4(ta + 24, va + 24); tc_ld2(ta + 28, va + 28);
            tc_ld16(ta + 32, vb); tc_ld8(ta + 48, vb + 16);
            tc_wait_ld();
            PH(1);
#pragma unroll
            for (int c = 0; c < 7; c++) tc_st8(ta + 8 * c, bias);               // hand the stage back, re-armed
            tc_wait_st();
            PH(2);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_addr(&sm.tm_empty[st]));
            PH(3);

            uint8_t *ring = sm.ring + (k & 3) * (SET_SAMPLES / 8);
            if constexpr (LAB == 2) {
                uint32_t x = 0;
#pragma unroll
                for (int e = 0; e < 30; e++) x ^= va[e];
#pragma unroll
                for (int e = 0; e < 24; e++) x ^= vb[e];
                ring[4 * row + part] = (uint8_t)x;
            } else {
                float2 z[9];
#pragma unroll
                for (int sp = 0; sp < 9; sp++) {
                    const uint32_t *v = sp < 5 ? va + 6 * sp : vb + 6 * (sp - 5);
                    z[sp].x = combine(__uint_as_float(v[0]), __uint_as_float(v[2]), __uint_as_float(v[4]));
                    z[sp].y = combine(__uint_as_float(v[1]), __uint_as_float(v[3]), __uint_as_float(v[5]));
                }
                uint32_t acc = 0;
#pragma unroll
                for (int sp = 7; sp >= 0; sp--)                 // bit sp = decision of z[sp], z[sp + 1]
                    acc = __funnelshift_l(neg_dot_bits(z[sp], z[sp + 1]), acc, 1);
                // decisions of V samples 4096 k + 2 + 32 row + 8 part + (0..7)
                ring[4 * row + part] = (uint8_t)acc;
                if (LAB == 1 && z_dump) {
                    // debug tap: z[n], n = 4096 k - 8 + 32 row + 8 part + sp, row-major float2 per stream
                    float2 *zd = z_dump + ((size_t)w.s * tiles * SETS_PER_TILE + k) * SET_SAMPLES + 32 * row + 8 * part;
#pragma unroll
                    for (int sp = 0; sp < 8; sp++) zd[sp] = z[sp];
                }
            }
            // all 4096 decisions of the set are in the ring: cut the complete words out.
            // Ring bit (v - 2) mod 16384 is the decision of V sample v (set k fills quarter k & 3).
            PH(4);
            asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");
            PH(5);
            const int w_lo = (int)((4096LL * k + 2) / LANE_SAMPLES), w_hi = (int)((4096LL * (k + 1) + 2) / LANE_SAMPLES);
            const int wi = w_lo + (int)threadIdx.x;
            if (wi < w_hi) {
                const uint32_t *rw = reinterpret_cast<const uint32_t *>(sm.ring);
                const uint32_t p0 = (uint32_t)(LANE_SAMPLES * wi + 4 * SET_SAMPLES - 2) & (4 * SET_SAMPLES - 1);
                const uint32_t lo = rw[(p0 >> 5) & (RING_BYTES / 4 - 1)], hi = rw[((p0 >> 5) + 1) & (RING_BYTES / 4 - 1)];
                words[(size_t)w.s * words_alloc + wi] = __funnelshift_r(lo, hi, p0 & 31u) & 0x0FFFFFFFu;
            }
            w.next(tiles);
            PH(6);
        }
    }

    if (LAB && prof && lane == 0)
        for (int i = 0; i < 8; i++) prof[((size_t)blockIdx.x * (THREADS / 32) + warp) * 8 + i] = pc[i];
    tc_fence_before();
    __syncthreads();
    if (warp == CTL_WARP) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)TMEM_COLS) : "memory");
    }
}

template <int LAB>
static void launch(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                   float2 *z_dump, unsigned long long *prof, cudaStream_t st) {
    const long long units = (long long)n_streams * tiles;
    if (units == 0) return;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)std::min<long long>(units, (long long)sms);
    const size_t smem = sizeof(Smem) + 128;
    cudaFuncSetAttribute(k_demod_mma<LAB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_demod_mma<LAB><<<grid, THREADS, smem, st>>>(vw, words, words_alloc, tiles, units, z_dump, prof);
}

}  // namespace mma

void launch_demod_mma(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                      cudaStream_t st) {
    mma::launch<0>(vw, words, n_streams, tiles, words_alloc, nullptr, nullptr, st);
}

#ifdef DAVIS_MMA_LAB
void launch_demod_mma_lab(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                          int drain_only, float2 *z_dump, unsigned long long *prof, cudaStream_t st) {
    if (drain_only) mma::launch<2>(vw, words, n_streams, tiles, words_alloc, nullptr, prof, st);
    else mma::launch<1>(vw, words, n_streams, tiles, words_alloc, z_dump, prof, st);
}
#endif

}  // namespace davis
