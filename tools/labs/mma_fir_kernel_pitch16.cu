// tools/labs/mma_fir_kernel_pitch16.cu -- LAB RECORD, not product code: variant B1 of profiles/r02_mma_fir_lab.txt
// (the raw byte stream as the tcgen05 A operand through an overlapping descriptor), as measured in commit
// 199c4da with that commit's tools/labs/mma_fir_lab.cu.  The current lab drives tools/labs/mma_fir_kernel_pitch64.cu.
// davis_demod_mma.cu -- the streaming demodulation kernel on the 5th-generation tensor
// cores (uint8 IQ -> decision bitmap), SURVEY section 7 "hard part 1" contingency.
//
// Replaces, like davis_demod.cu, ByteToCmplxLUT.execute -> rotate_fs4 -> fir9 ->
// discriminate -> quantize of Demodulator.demodulate (src/rtldavis/dsp.py:139-166; Go
// dsp/dsp.go:274-292) for a whole batch of streams and blocks.  The SIMT kernel spends
// 23 of its 30 issue cycles per sample on widening the bytes and on the 9-tap filter
// (profiles/r01_step_v12_all_kernels_ncu_full.txt: FMA pipe 80 %, DRAM 33 %).  Here the
// bytes are never touched by a SIMT instruction before they have been filtered:
//
//   * Folded filter.  With z[n] = sum_m c[m] j^m x[n+m]  (x = unrotated sample, the Fs/4
//     rotation of dsp.py:42-49 folded into the taps) the reference's F[g] = j^(g-9) z[g-9]
//     and its discriminator numerator (dsp.py:88) is  -Re(conj(z[g-10]) z[g-9]).  Re z[n]
//     and Im z[n] are two real 9-term dot products over the 18 interleaved bytes
//     u[2n .. 2n+17].
//   * Hankel view.  The raw byte stream, exactly as the bulk-copy engine lands it in shared
//     memory, is the A operand of tcgen05.mma kind::i8 through a no-swizzle K-major
//     descriptor with a 16-byte K-chunk stride (LBO) and a 128-byte 8-row-group stride
//     (SBO): row r of the 128 x 32 operand is the 32-byte window that starts 16 r bytes
//     (8 samples) into the buffer, so consecutive rows overlap by half.  One MMA
//     (M = 128, K = 32, N = 16 L) filters 1024 samples: column (s, comp, limb) of row r is
//     the limb's share of Re/Im z[8r + s].  No copy, no conversion, no shared-memory store.
//   * Limbs.  The taps are the reference's decimal constants scaled to 8 L-bit integers
//     (L = 3: 24 bits, relative tap error 5e-8; the scale is chosen so that all five
//     distinct taps land within 0.04 of an integer) and split into L balanced signed
//     8-bit digits; the u8 x s8 -> s32 products are exact.  The accumulator columns are
//     pre-loaded (tcgen05.st) with 0x4B400000 - K_l, so that what tcgen05.ld returns is
//     already the IEEE float 1.5 * 2^23 + (limb sum - K_l): no I2F and no integer add in the
//     epilogue.  K_l are the digits of 127.4 * sum(signed taps), the LUT offset of
//     dsp.py:26 passed through the filter (the LUT scale 1/127.6 cannot change a sign).
//   * Epilogue (SIMT, ~10 issue cycles per sample instead of 30): L fused multiply-adds
//     per component put the limbs together, one multiply + one FMA give the discriminator
//     numerator, one funnel shift per sample packs its sign.  Lane = one operand row = 8
//     consecutive samples; the neighbouring row's first z comes by warp shuffle, and
//     across warps / MMAs through shared memory (patched after a group barrier).
//   * The 8-bit pieces are re-cut into the bitmap format of the search stage (word w =
//     samples [28 w, 28 w + 28)) through a shared-memory bit stream.
//
// Decomposition.  A CTA (one per SM, persistent) holds GROUPS independent warpgroups.  A
// group = 4 epilogue warps (one per TMEM lane quadrant) + one control thread that issues
// the group's bulk copies and MMAs.  A group walks a contiguous range of sub-tiles
// (7168 samples = 7 MMAs = 256 bitmap words) of the flattened (stream, sub-tile) space, so
// the two decisions that straddle sub-tiles are carried in shared memory; only the first
// sub-tile of a range computes them from the bytes (once per launch and group).
//
// Every decision that reaches a caller is re-derived by k_exact (davis_search.cu), as for
// the SIMT kernel; tests/test_guard_band.py holds the numpy model of this arithmetic.
#include <algorithm>

#include "../../rtldavis_b200/csrc/davis_async.cuh"
#include "../../rtldavis_b200/csrc/davis_kernels.cuh"

namespace davis {
namespace mma {

constexpr int GROUPS = 4;                        // warpgroups per CTA
constexpr int EPI_WARPS = 4 * GROUPS;
constexpr int THREADS = (EPI_WARPS + GROUPS) * 32;
constexpr int MMA_SAMPLES = 1024, MMA_BYTES = 2 * MMA_SAMPLES;   // 128 rows x 8 samples
constexpr int SUB_MMAS = 7;
constexpr int SUB_SAMPLES = SUB_MMAS * MMA_SAMPLES;              // 7168 = 256 words x 28
constexpr int SUB_BYTES = 2 * SUB_SAMPLES;
constexpr int SUB_WORDS = SUB_SAMPLES / LANE_SAMPLES;
constexpr int SUBS_PER_TILE = TILE_SAMPLES / SUB_SAMPLES;
static_assert(SUB_WORDS * LANE_SAMPLES == SUB_SAMPLES && SUBS_PER_TILE * SUB_SAMPLES == TILE_SAMPLES,
              "sub-tiles cut the bitmap at word boundaries");
constexpr int LEAD_BYTES = 32;                   // 16 samples of history in front of a sub-tile
constexpr int SLOT_BYTES = ((LEAD_BYTES + SUB_BYTES + 127) / 128) * 128;
constexpr int NSLOT = 3;                         // sub-tiles in flight per group
constexpr int NSTAGE = 2;                        // accumulator stages per group
constexpr int STAGE_COLS = 64;
constexpr int TMEM_COLS = GROUPS * NSTAGE * STAGE_COLS;
static_assert(TMEM_COLS == 512, "the CTA owns the whole tensor memory of its SM");
constexpr int CHUNKS = 4 * SUB_MMAS;             // (MMA, quadrant) pieces of 256 decisions
constexpr int BITS_BYTES = 4 + SUB_SAMPLES / 8 + 12;             // lead word + 896 + slack
constexpr uint32_t MAGIC = 0x4B400000u;          // 1.5 * 2^23: ulp 1, +-2^22 of integer room

// ---- taps as integers ------------------------------------------------------------------
// round(S c[m]) for the five distinct taps of dsp.py:56-69; S (3.5289e7 / 1.38444e5) is the
// scale inside the digit range at which all five products are nearest to integers.
template <int L> __host__ __device__ constexpr int tap5(int i) {
    return L == 3 ? (i == 0 ? 623995 : i == 1 ? 1699934 : i == 2 ? 4320285 : i == 3 ? 6966413 : 8068069)
                  : (i == 0 ? 2448 : i == 1 ? 6669 : i == 2 ? 16949 : i == 3 ? 27330 : 31652);
}

__host__ __device__ constexpr int balanced_digit(long long v, int l) {
    int d = 0;
    for (int i = 0; i <= l; i++) {
        d = (int)(((v % 256) + 256 + 128) % 256) - 128;
        v = (v - d) / 256;
    }
    return d;
}
template <int L> __host__ __device__ constexpr int tap_of(int m) { return tap5<L>(m < 5 ? m : 8 - m); }
template <int L> __host__ __device__ constexpr int tap_digit(int m, int l) { return balanced_digit(tap_of<L>(m), l); }
// 127.4 * (sum of the taps with the rotation's signs) = 127.4 * (2 t0 - 2 t2 + t4), the same
// for both components; its fraction (< 1 unit at a scale of 2^31) is dropped
template <int L> __host__ __device__ constexpr long long lut_offset() {
    const long long s = 2LL * tap5<L>(0) - 2LL * tap5<L>(2) + tap5<L>(4);
    return (s * 1274 + 5) / 10;
}
// The offset is subtracted as the SAME integer kappa on every limb (kappa (1 + 256 (+ 65536)) ~ offset),
// so that one set of eight registers re-arms every accumulator column; what is left over
// (55 of 2^31 for L = 3, 88 of 2^26 for L = 2: below the rounding of the fp32 combine) is dropped.
template <int L> __host__ __device__ constexpr int offset_digit(int) {
    const long long w = L == 3 ? 65793 : 257;
    return (int)((lut_offset<L>() + w / 2) / w);
}
// tap of byte offset o (0..17) of a sample's window for component comp, with the Fs/4
// rotation's sign: z = sum_m c[m] j^m (I + jQ)
template <int L> __host__ __device__ constexpr int byte_tap(int o, int comp, int l) {
    if (o < 0 || o >= 18) return 0;
    const int m = o >> 1, q = o & 1, ph = m & 3;
    // Re: +I, -Q, -I, +Q     Im: +Q, +I, -Q, -I      (phase 0, 1, 2, 3)
    const bool uses_q = comp == 0 ? (ph & 1) : !(ph & 1);
    if ((q == 1) != uses_q) return 0;
    const int sign = comp == 0 ? ((ph == 0 || ph == 3) ? 1 : -1) : ((ph == 0 || ph == 1) ? 1 : -1);
    return sign * tap_digit<L>(m, l);
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
__device__ __forceinline__ void tc_ld8(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr) : "memory");
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
template <int L>
struct Smem {
    static constexpr int N = 16 * L;
    alignas(128) uint8_t slot[GROUPS][NSLOT][SLOT_BYTES];
    alignas(128) uint8_t bmat[N * 32];
    alignas(16) uint8_t bits[GROUPS][2][BITS_BYTES];
    float2 zfirst[GROUPS][2][CHUNKS], zlast[GROUPS][2][CHUNKS];
    float2 zhead[GROUPS][2];
    unsigned long long full[GROUPS][NSLOT], slot_free[GROUPS][NSLOT];
    unsigned long long tm_full[GROUPS][NSTAGE], tm_empty[GROUPS][NSTAGE], head_ready[GROUPS];
    uint32_t tmem_base;
};

struct SubSpan { const uint8_t *p0, *p1; uint32_t n0, n1, skip; };

// bytes [j SUB_BYTES - LEAD_BYTES, (j + 1) SUB_BYTES) of stream s's sequence V = carry || input,
// clipped to what exists: at most one piece of the carry buffer and one of the input
__device__ __forceinline__ SubSpan sub_span(const View &vw, int s, int j) {
    const long long carry_bytes = 2LL * vw.n_carry;
    const long long total = carry_bytes + 2LL * vw.n_new;
    long long b0 = (long long)j * SUB_BYTES - LEAD_BYTES;
    long long b1 = b0 + SUB_BYTES + LEAD_BYTES;
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

// limb floats (1.5 * 2^23 + limb sum - K_l) -> z component, scaled by S:  e0 + 256 e1 (+ 65536 e2)
template <int L>
__device__ __forceinline__ float combine(float f0, float f1, float f2) {
    if constexpr (L == 3) {
        const float t2 = fmaf(f2, 65536.f, -65793.f * 12582912.f);   // exact: 2^16 (e2 - 49344)
        return f0 + fmaf(f1, 256.f, t2);
    } else {
        return f0 + fmaf(f1, 256.f, -257.f * 12582912.f);            // exact: 256 (e1 - 49152)
    }
}
// sign bit set <=> Re(conj(a) b) > 0 <=> the reference's numerator Im F[g-1] Re F[g] - Re F[g-1] Im F[g] < 0
__device__ __forceinline__ uint32_t neg_dot_bits(float2 a, float2 b) {
    return __float_as_uint(fmaf(-a.y, b.y, -a.x * b.x));
}

// z[n] from the 18 bytes at p (the sample's I byte), with the tensor core's integer arithmetic
template <int L>
__device__ float2 z_from_bytes(const uint8_t *p) {
    int acc[2][3] = {{0, 0, 0}, {0, 0, 0}};
    for (int o = 0; o < 18; o++) {
        const int u = p[o];
        for (int comp = 0; comp < 2; comp++)
            for (int l = 0; l < L; l++) acc[comp][l] += byte_tap<L>(o, comp, l) * u;
    }
    float f[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
    for (int comp = 0; comp < 2; comp++)
        for (int l = 0; l < L; l++)
            f[comp][l] = __uint_as_float(MAGIC + (uint32_t)(acc[comp][l] - offset_digit<L>(l)));
    return make_float2(combine<L>(f[0][0], f[0][1], f[0][2]), combine<L>(f[1][0], f[1][1], f[1][2]));
}

// LAB = 0: the product kernel.  LAB = 1 (tools/labs/mma_fir_lab.cu): the operand strides come from
// the launch (to probe the descriptor semantics) and z can be dumped.  LAB = 2: the epilogue
// only drains the accumulators (what the copy + MMA + tensor-memory path costs alone).
template <int L, int LAB>
__global__ void __launch_bounds__(THREADS, 1)
k_demod_mma(View vw, uint32_t *__restrict__ words, int words_alloc, int n_sub, long long units,
            float2 *__restrict__ z_dump, uint32_t a_lbo, uint32_t a_sbo, unsigned long long *prof) {
    constexpr int N = 16 * L;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    Smem<L> &sm = *reinterpret_cast<Smem<L> *>(
        smem_raw + ((128u - (smem_addr(smem_raw) & 127u)) & 127u));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // lab builds: cycles spent per phase, per warp (prof != nullptr)
    unsigned long long pc[6] = {0, 0, 0, 0, 0, 0};
    auto tick = [&]() -> unsigned long long { return (LAB && prof) ? (unsigned long long)clock64() : 0ull; };

    // B operand: [K chunk][column group][8 columns][16 bytes]; column n = (half h, limb l,
    // sample s', component c) -> 8 L h + 8 l + 2 s' + c, sample s = 4 h + s'
    for (int e = threadIdx.x; e < N * 32; e += THREADS) {
        const int n = e >> 5, k = e & 31;
        const int h = n / (8 * L), l = (n % (8 * L)) >> 3, sp = (n & 7) >> 1, c = n & 1;
        int v = 0;
#pragma unroll
        for (int ll = 0; ll < L; ll++) if (ll == l) v = byte_tap<L>(k - 2 * (4 * h + sp), c, ll);
        sm.bmat[(k >> 4) * (N * 16) + (n >> 3) * 128 + (n & 7) * 16 + (k & 15)] = (uint8_t)(int8_t)v;
    }
    if (threadIdx.x == 0) {
        for (int g = 0; g < GROUPS; g++) {
            for (int i = 0; i < NSLOT; i++) {
                mbar_init(smem_addr(&sm.full[g][i]), 1);
                mbar_init(smem_addr(&sm.slot_free[g][i]), 1);
            }
            for (int i = 0; i < NSTAGE; i++) {
                mbar_init(smem_addr(&sm.tm_full[g][i]), 1);
                mbar_init(smem_addr(&sm.tm_empty[g][i]), 4);
            }
            mbar_init(smem_addr(&sm.head_ready[g]), 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == EPI_WARPS) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_addr(&sm.tmem_base)), "r"((uint32_t)TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // bmat is read by the tensor core
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sm.tmem_base;

    // this group's contiguous range of sub-tiles of the flattened (stream, sub-tile) space
    const int g = warp < EPI_WARPS ? (warp >> 2) : (warp - EPI_WARPS);
    const long long n_groups = (long long)gridDim.x * GROUPS, gi = (long long)blockIdx.x * GROUPS + g;
    const long long u0 = units * gi / n_groups, u1 = units * (gi + 1) / n_groups;
    const int count = (int)(u1 - u0);
    const int s0 = (int)(u0 / n_sub), j0 = (int)(u0 % n_sub);
    const uint32_t tm_group = tmem + (uint32_t)(g * NSTAGE * STAGE_COLS);

    if (warp >= EPI_WARPS) {
        // ===== control thread: bulk copies and MMAs of group g =====
        if (lane == 0 && count > 0) {
            const uint64_t b_desc = smem_desc(smem_addr(sm.bmat), N * 16, 128);
            constexpr uint32_t idesc = instr_desc(N);
            int ps = s0, pj = j0;                               // next sub-tile to fetch
            auto fetch = [&](int slot) {
                const SubSpan sp = sub_span(vw, ps, pj);
                const uint32_t dst = smem_addr(sm.slot[g][slot]), bar = smem_addr(&sm.full[g][slot]);
                mbar_expect_tx(bar, sp.n0 + sp.n1);
                if (sp.n0) bulk_g2s(dst + sp.skip, sp.p0, sp.n0, bar);
                if (sp.n1) bulk_g2s(dst + sp.skip + sp.n0, sp.p1, sp.n1, bar);
                if (++pj == n_sub) { pj = 0; ps++; }
            };
            for (int i = 0; i < NSLOT && i < count; i++) fetch(i);
            uint32_t m = 0;                                     // MMAs issued by this group
            for (int k = 0; k < count; k++) {
                const int slot = k % NSLOT;
                if (k >= 1 && k + NSLOT - 1 < count) {          // refill the slot sub-tile k-1 used
                    const int fs = (k - 1) % NSLOT;
                    const unsigned long long t1 = tick();
                    mbar_wait(smem_addr(&sm.slot_free[g][fs]), ((k - 1) / NSLOT) & 1);
                    pc[2] += tick() - t1;
                    fetch(fs);
                }
                unsigned long long t0 = tick();
                mbar_wait(smem_addr(&sm.full[g][slot]), (k / NSLOT) & 1);
                pc[0] += tick() - t0;
                if (k == 0 && j0 > 0) {
                    // the two decisions in front of a range that starts inside a stream need
                    // z[-10], z[-9] relative to the sub-tile: samples -10 .. -1 of the lead
                    const uint8_t *lead = sm.slot[g][slot] + LEAD_BYTES;
                    sm.zhead[g][0] = z_from_bytes<L>(lead - 20);
                    sm.zhead[g][1] = z_from_bytes<L>(lead - 18);
                    mbar_arrive(smem_addr(&sm.head_ready[g]));
                }
                tc_fence_after();
                const uint32_t a0 = smem_addr(sm.slot[g][slot]) + 16;   // sample -8 of the sub-tile
#pragma unroll 1
                for (int i = 0; i < SUB_MMAS; i++, m++) {
                    const uint32_t st = m % NSTAGE;
                    t0 = tick();
                    mbar_wait(smem_addr(&sm.tm_empty[g][st]), (m / NSTAGE) & 1);
                    pc[1] += tick() - t0;
                    tc_fence_after();
                    tc_mma_i8(tm_group + st * STAGE_COLS, smem_desc(a0 + i * MMA_BYTES, LAB ? a_lbo : 16u, LAB ? a_sbo : 128u), b_desc, idesc, 1u);
                    tc_commit(smem_addr(&sm.tm_full[g][st]));
                }
                tc_commit(smem_addr(&sm.slot_free[g][slot]));
            }
        }
    } else {
        // ===== epilogue warps of group g: quadrant q = TMEM lanes 32 q .. 32 q + 31 =====
        const int q = warp & 3, tg = threadIdx.x & 127;
        const uint32_t tm_lane = tm_group + ((uint32_t)(32 * q) << 16);
        // eight registers that stay put (opaque to the compiler, or it rebuilds them in front
        // of every tcgen05.st: 48 moves per row)
        uint32_t bias[8];
#pragma unroll
        for (int e = 0; e < 8; e++)
            asm volatile("mov.b32 %0, %1;" : "=r"(bias[e]) : "r"(MAGIC - (uint32_t)offset_digit<L>(0)));
        auto put_bias = [&](uint32_t st) {
#pragma unroll
            for (int c = 0; c < 2 * L; c++) tc_st8(tm_lane + st * STAGE_COLS + 8 * c, bias);
        };
        if (count > 0) {
            for (uint32_t st = 0; st < NSTAGE; st++) put_bias(st);
            tc_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0)
                for (uint32_t st = 0; st < NSTAGE; st++) mbar_arrive(smem_addr(&sm.tm_empty[g][st]));
        }
        int s = s0, j = j0;
        uint32_t m = 0;
        for (int k = 0; k < count; k++) {
            const int par = k & 1;
            uint8_t *bits = sm.bits[g][par];
#pragma unroll 1
            for (int i = 0; i < SUB_MMAS; i++, m++) {
                const uint32_t st = m % NSTAGE;
                unsigned long long t0 = tick();
                mbar_wait(smem_addr(&sm.tm_full[g][st]), (m / NSTAGE) & 1);
                unsigned long long t1 = tick();
                pc[0] += t1 - t0;
                tc_fence_after();
                uint32_t v[2][L][8];
#pragma unroll
                for (int h = 0; h < 2; h++)
#pragma unroll
                    for (int l = 0; l < L; l++) tc_ld8(tm_lane + st * STAGE_COLS + h * 8 * L + 8 * l, v[h][l]);
                tc_wait_ld();
                t0 = tick();
                pc[1] += t0 - t1;
                put_bias(st);                                   // hand the stage back, biased
                tc_wait_st();
                t1 = tick();
                pc[2] += t1 - t0;
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_addr(&sm.tm_empty[g][st]));
                pc[3] += tick() - t1;

                if constexpr (LAB == 2) {
                    uint32_t x = 0;
#pragma unroll
                    for (int h = 0; h < 2; h++)
#pragma unroll
                        for (int l = 0; l < L; l++)
#pragma unroll
                            for (int e = 0; e < 8; e++) x ^= v[h][l][e];
                    bits[4 + 128 * i + 32 * q + lane] = (uint8_t)x;
                    continue;
                }
                float2 z[9];
#pragma unroll
                for (int sidx = 0; sidx < 8; sidx++) {
                    const int h = sidx >> 2, sp = sidx & 3;
                    z[sidx].x = combine<L>(__uint_as_float(v[h][0][2 * sp]), __uint_as_float(v[h][1][2 * sp]),
                                           __uint_as_float(v[h][L - 1][2 * sp]));
                    z[sidx].y = combine<L>(__uint_as_float(v[h][0][2 * sp + 1]), __uint_as_float(v[h][1][2 * sp + 1]),
                                           __uint_as_float(v[h][L - 1][2 * sp + 1]));
                }
                z[8].x = __shfl_down_sync(0xFFFFFFFFu, z[0].x, 1);
                z[8].y = __shfl_down_sync(0xFFFFFFFFu, z[0].y, 1);
                uint32_t acc = 0;
#pragma unroll
                for (int sidx = 7; sidx >= 0; sidx--)           // bit sidx = decision of z[sidx], z[sidx + 1]
                    acc = __funnelshift_l(neg_dot_bits(z[sidx], z[sidx + 1]), acc, 1);
                const int row = 32 * q + lane, chunk = 4 * i + q;
                bits[4 + 128 * i + row] = (uint8_t)acc;         // lane 31's bit 7 is patched below
                if (lane == 0) sm.zfirst[g][par][chunk] = z[0];
                if (lane == 31) sm.zlast[g][par][chunk] = z[7];
                if (LAB == 1 && z_dump) {
                    // debug tap: z[n], n = 7168 j - 8 + 1024 i + 8 row + sidx, as row-major float2
                    float2 *zd = z_dump + ((size_t)s * n_sub + j) * SUB_SAMPLES + 1024 * i + 8 * row;
#pragma unroll
                    for (int sidx = 0; sidx < 8; sidx++) zd[sidx] = z[sidx];
                }
            }
            asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
            if (q == 0) {
                if (lane < CHUNKS - 1) {
                    // decision between the last z of chunk `lane` and the first of the next
                    const uint32_t d = neg_dot_bits(sm.zlast[g][par][lane], sm.zfirst[g][par][lane + 1]);
                    uint8_t *b = &bits[4 + 32 * lane + 31];
                    *b = (uint8_t)((*b & 0x7Fu) | ((d >> 31) << 7));
                } else if (lane == CHUNKS - 1) {
                    // the two decisions in front of the sub-tile (V samples 7168 j, 7168 j + 1)
                    uint32_t b6 = 0, b7 = 0;
                    if (j > 0) {
                        if (k == 0) {
                            mbar_wait(smem_addr(&sm.head_ready[g]), 0);
                            b6 = neg_dot_bits(sm.zhead[g][0], sm.zhead[g][1]) >> 31;
                            b7 = neg_dot_bits(sm.zhead[g][1], sm.zfirst[g][par][0]) >> 31;
                        } else {
                            b6 = (sm.bits[g][par ^ 1][4 + SUB_SAMPLES / 8 - 1] >> 6) & 1u;
                            b7 = neg_dot_bits(sm.zlast[g][par ^ 1][CHUNKS - 1], sm.zfirst[g][par][0]) >> 31;
                        }
                    }
                    bits[3] = (uint8_t)((b6 << 6) | (b7 << 7));
                }
            }
            asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
            // bit 32 + p of the stream is the decision of V sample 7168 j + 2 + p
            const uint32_t *bw = reinterpret_cast<const uint32_t *>(bits);
            uint32_t *out = words + (size_t)s * words_alloc + (size_t)j * SUB_WORDS;
#pragma unroll
            for (int w = tg; w < SUB_WORDS; w += 128) {
                const uint32_t p0 = 28u * w + 30u;
                out[w] = __funnelshift_r(bw[p0 >> 5], bw[(p0 >> 5) + 1], p0 & 31u) & 0x0FFFFFFFu;
            }
            if (++j == n_sub) { j = 0; s++; }
        }
    }

    if (LAB && prof && lane == 0)
        for (int i = 0; i < 6; i++) prof[((size_t)blockIdx.x * (THREADS / 32) + warp) * 6 + i] = pc[i];
    tc_fence_before();
    __syncthreads();
    if (warp == EPI_WARPS) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)TMEM_COLS) : "memory");
    }
}

template <int L, int LAB>
static void launch(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                   float2 *z_dump, uint32_t a_lbo, uint32_t a_sbo, unsigned long long *prof, cudaStream_t st) {
    const int n_sub = tiles * SUBS_PER_TILE;
    const long long units = (long long)n_streams * n_sub;
    if (units == 0) return;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)std::min<long long>((units + GROUPS - 1) / GROUPS, (long long)sms);
    const size_t smem = sizeof(Smem<L>) + 128;
    cudaFuncSetAttribute(k_demod_mma<L, LAB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_demod_mma<L, LAB><<<grid, THREADS, smem, st>>>(vw, words, words_alloc, n_sub, units, z_dump,
                                                      a_lbo, a_sbo, prof);
}

}  // namespace mma

void launch_demod_mma(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                      int limbs, cudaStream_t st) {
    if (limbs == 2) mma::launch<2, 0>(vw, words, n_streams, tiles, words_alloc, nullptr, 16, 128, nullptr, st);
    else mma::launch<3, 0>(vw, words, n_streams, tiles, words_alloc, nullptr, 16, 128, nullptr, st);
}

#ifdef DAVIS_MMA_LAB
void launch_demod_mma_lab(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                          int limbs, int drain_only, float2 *z_dump, uint32_t a_lbo, uint32_t a_sbo,
                          unsigned long long *prof, cudaStream_t st) {
    if (drain_only) {
        if (limbs == 2) mma::launch<2, 2>(vw, words, n_streams, tiles, words_alloc, nullptr, a_lbo, a_sbo, prof, st);
        else mma::launch<3, 2>(vw, words, n_streams, tiles, words_alloc, nullptr, a_lbo, a_sbo, prof, st);
    } else {
        if (limbs == 2) mma::launch<2, 1>(vw, words, n_streams, tiles, words_alloc, z_dump, a_lbo, a_sbo, prof, st);
        else mma::launch<3, 1>(vw, words, n_streams, tiles, words_alloc, z_dump, a_lbo, a_sbo, prof, st);
    }
}
#endif

}  // namespace davis
