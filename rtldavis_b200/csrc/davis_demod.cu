// davis_demod.cu -- the streaming demodulation kernel (uint8 IQ -> decision bitmap).
//
// Replaces, for a whole batch of streams and blocks at once, the per-block chain
// ByteToCmplxLUT.execute -> rotate_fs4 -> fir9 -> discriminate -> quantize of
// Demodulator.demodulate (src/rtldavis/dsp.py:139-166; Go dsp/dsp.go:274-292).
//
// Work decomposition
//   * one lane decides TWO runs of LANE_SAMPLES = 28 consecutive samples (two symbol
//     periods each): run A = word t and run B = word TILE_WORDS/2 + t of the CTA's
//     tile.  The two runs are held as (A, B) register pairs, so every floating-point
//     instruction of the path -- conversion, the 9-tap filter, the cross product --
//     is a packed f32x2 operation (FADD2/FFMA2/FMUL2) on two independent samples.
//     (The previous version packed (re, im) of one sample; its conversions and cross
//     products were scalar, and on this part every switch between scalar and packed
//     FMA-pipe instructions costs issue bubbles: profiles/r01_demod_lab.txt.)
//   * each run keeps a register sliding window, so the 9-sample FIR history is
//     re-read for only 9/28 of the work;
//   * the 28 sign bits of a run are shifted into one 32-bit word with a funnel shift
//     per sample (SHF, 1 instruction) -- no compare, no warp vote; word w of a
//     stream holds samples [28w, 28w+28), bit i = sample 28w+i.  Stepping by one
//     symbol (14 samples) is a half-word step, which the search kernel exploits;
//   * a CTA of 16 warps owns a tile of 1024 words = 28672 samples = 56 KiB of input.
//     CTAs are persistent (1 per SM; with 8 warps x 2 CTAs the second CTA was not always
//     co-resident and the kernel was 4 % slower): each walks tiles unit = blockIdx.x, +gridDim.x,
//     ... and one elected thread streams the next tiles into shared memory with the
//     bulk-copy engine (cp.async.bulk + mbarrier, SASS UBLKCP), three tiles in
//     flight.  The copy's address arithmetic is done by all threads on uniform
//     values, so it compiles to uniform-datapath instructions and only the copy
//     instructions themselves are predicated;
//   * lanes read their two 80-byte windows with ten 8-byte shared loads each (lane
//     stride 56 B = 7 x 8 B, odd, so the 16 lanes of a half-warp phase hit 16 distinct bank
//     pairs).  ncu still counts 6.4 % of the shared-memory wavefronts as conflicting
//     (profiles/r02_k_demod_ncu_full.txt: 4.2 M of 65 M) -- loads meeting the copy engine's
//     writes of the next tiles; LSU is 5.6 % busy, so it costs nothing measurable.
//
// Arithmetic (fp32, signs only)
//   x = (u - 127.4)/127.6 and the division by |F|^2 + eps are positive scalings
//   and are dropped.  A byte becomes a float without I2F: one PRMT builds the bit
//   pattern of 2^23 + u, and one packed add with -+(2^23 + 127) yields +-(u - 127)
//   exactly for both runs; the sign is the Fs/4 rotation's (which byte lands in
//   which component is its swap).  The remaining -0.4 of the LUT offset survives
//   the filter as a period-4 constant folded into the first FMA.  The rotation phase
//   is taken relative to the lane's window because a common unit factor cancels in
//   Im(conj(F[n-1]) F[n]).
//   Every decision that reaches a caller is re-derived by k_exact (davis_search.cu): in
//   fp32 with an error bound beside it, in the reference's float64 form where the
//   numerator lies inside the bound.  So an fp32 sign here that disagrees with the
//   reference's (only possible when the cross product is ~1e-6 of |F[n-1]||F[n]|)
//   cannot change a packet.
//
// Cost model (measured, profiles/r01_demod_lab.txt): FFMA2/FADD2/FMUL2 occupy the FMA
// pipe 2 cycles per warp instruction and there is one issue slot per cycle.  Per
// sample the FMA pipe needs (18 x 29/28 filter + 2 x 37/28 convert + 2 cross) = 23.3
// cycles, and each ALU instruction (2.6 PRMT + 1 SHF per sample) costs ~1.5 more
// in practice, so the kernel is bound by the FP32 pipe at ~29 cycles per sample and
// lane = ~40 % of the HBM roofline, not by HBM.
#include <algorithm>

#include "davis_kernels.cuh"
#include "davis_async.cuh"

namespace davis {

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 neg2(float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 fma2(float2 c, float2 a, float2 acc) { return __ffma2_rn(c, a, acc); }

// Rotated integer-valued sample rho_i = j^(i mod 4) * ((I - 127) + j(Q - 127)) of window
// index i for both runs of a lane: re = (re_A, re_B), im = (im_A, im_B).  wa / wb are the
// 32-bit words holding samples (2k, 2k+1) of run A / run B.
//   phase 0: (+I, +Q)   1: (-Q, +I)   2: (-I, -Q)   3: (+Q, -I)
template <int I>
__device__ __forceinline__ void unpack_rotated(uint32_t wa, uint32_t wb, float2 &re, float2 &im) {
    // bytes (u, 00, 00, 4B) = the float 2^23 + u
    constexpr uint32_t sel_i = (I & 1) ? 0x7442u : 0x7440u, sel_q = (I & 1) ? 0x7443u : 0x7441u;
    const float2 mi = f2(__uint_as_float(__byte_perm(wa, 0x4B000000u, sel_i)),
                         __uint_as_float(__byte_perm(wb, 0x4B000000u, sel_i)));
    const float2 mq = f2(__uint_as_float(__byte_perm(wa, 0x4B000000u, sel_q)),
                         __uint_as_float(__byte_perm(wb, 0x4B000000u, sel_q)));
    const float2 c = f2(8388735.0f, 8388735.0f);                 // 2^23 + 127
    constexpr int ph = I & 3;
    if (ph == 0) { re = add2(mi, neg2(c)); im = add2(mq, neg2(c)); }
    else if (ph == 1) { re = add2(c, neg2(mq)); im = add2(mi, neg2(c)); }
    else if (ph == 2) { re = add2(c, neg2(mi)); im = add2(c, neg2(mq)); }
    else { re = add2(mq, neg2(c)); im = add2(c, neg2(mi)); }
}

template <int I>
struct Unpack {
    static __device__ __forceinline__ void run(const uint32_t (&wa)[20], const uint32_t (&wb)[20],
                                               float2 (&re)[39], float2 (&im)[39]) {
        unpack_rotated<I>(wa[I >> 1], wb[I >> 1], re[I], im[I]);
        Unpack<I + 1>::run(wa, wb, re, im);
    }
};
template <>
struct Unpack<39> {
    static __device__ __forceinline__ void run(const uint32_t (&)[20], const uint32_t (&)[20],
                                               float2 (&)[39], float2 (&)[39]) {}
};

// What the filter makes of the -0.4 the exact centring leaves out (samples are
// +-(u - 127), the LUT subtracts 127.4), for output q (window taps q+2 .. q+10):
// a period-4 constant per component.
__host__ __device__ constexpr float lut_residual(int q, int comp) {
    const double c[9] = {DAVIS_C0, DAVIS_C1, DAVIS_C2, DAVIS_C3, DAVIS_C4,
                         DAVIS_C3, DAVIS_C2, DAVIS_C1, DAVIS_C0};
    const double sre[4] = {-0.4, 0.4, 0.4, -0.4}, sim[4] = {-0.4, -0.4, 0.4, 0.4};
    double acc = 0.0;
    for (int m = 0; m < 9; m++) acc += c[m] * (comp ? sim[(q + 2 + m) & 3] : sre[(q + 2 + m) & 3]);
    return (float)acc;
}

// one component of F[q] for both runs: fir9 (dsp.py:52-73) with the symmetric taps folded
__device__ __forceinline__ float2 fir9_pair(const float2 (&r)[39], int q, float off) {
    const float c0 = (float)DAVIS_C0, c1 = (float)DAVIS_C1, c2 = (float)DAVIS_C2,
                c3 = (float)DAVIS_C3, c4 = (float)DAVIS_C4;
    float2 acc = fma2(f2(c4, c4), r[q + 6], f2(off, off));
    acc = fma2(f2(c0, c0), add2(r[q + 2], r[q + 10]), acc);
    acc = fma2(f2(c1, c1), add2(r[q + 3], r[q + 9]), acc);
    acc = fma2(f2(c2, c2), add2(r[q + 4], r[q + 8]), acc);
    acc = fma2(f2(c3, c3), add2(r[q + 5], r[q + 7]), acc);
    return acc;
}

constexpr int DEMOD_THREADS = DEMOD_WARPS * 32;
static_assert(2 * DEMOD_THREADS == TILE_WORDS, "two words per thread");
constexpr int TILE_BYTES = 2 * TILE_SAMPLES;                 // 57344
constexpr int HALF_BYTES = TILE_BYTES / 2;                   // run B starts here
constexpr int LEAD_BYTES = 32;                               // history in front of a tile
constexpr int BUF_BYTES = TILE_BYTES + LEAD_BYTES + 96;      // padded to a multiple of 128
constexpr int DEMOD_STAGES = 3;                              // tiles in flight per CTA

// The copy of tile j of stream s: bytes [j*TILE_BYTES - 32, (j+1)*TILE_BYTES) of the
// stream's V sequence, clipped to what exists, as at most one piece of the carry buffer
// and one piece of the input.  Computed by every thread (uniform values).
struct TileSpan { const uint8_t *p0, *p1; uint32_t n0, n1, skip; };

__device__ __forceinline__ TileSpan tile_span(const View &vw, int s, int j) {
    const long long carry_bytes = 2LL * vw.n_carry;
    const long long total = carry_bytes + 2LL * vw.n_new;
    long long b0 = (long long)j * TILE_BYTES - LEAD_BYTES;
    long long b1 = b0 + TILE_BYTES + LEAD_BYTES;
    TileSpan sp{nullptr, nullptr, 0u, 0u, 0u};
    if (b0 < 0) { sp.skip = (uint32_t)(-b0); b0 = 0; }
    if (b1 > total) b1 = total;
    if (b1 <= b0) return sp;                                   // tile past the end: bits unused
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
__device__ __forceinline__ void issue_tile(const TileSpan &sp, uint32_t dst, uint32_t bar) {
    mbar_expect_tx(bar, sp.n0 + sp.n1);
    if (sp.n0) bulk_g2s(dst + sp.skip, sp.p0, sp.n0, bar);
    if (sp.n1) bulk_g2s(dst + sp.skip + sp.n0, sp.p1, sp.n1, bar);
}

template <bool WRITE_DISC>
__global__ void __launch_bounds__(DEMOD_THREADS, DEMOD_CTAS_PER_SM)
k_demod(View vw, uint32_t *__restrict__ words, int words_alloc, int tiles, int n_streams,
        float *__restrict__ disc_out, size_t disc_stride) {
    extern __shared__ __align__(128) uint8_t buf[];              // DEMOD_STAGES x BUF_BYTES
    __shared__ __align__(8) unsigned long long full_bar[DEMOD_STAGES], empty_bar[DEMOD_STAGES];

    if (threadIdx.x == 0) {
        for (int i = 0; i < DEMOD_STAGES; i++) {
            mbar_init(smem_addr(&full_bar[i]), 1);               // the producer's expect_tx arrival
            mbar_init(smem_addr(&empty_bar[i]), DEMOD_WARPS);    // one arrival per consumer warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Bytes a clipped copy does not write (before the start / past the end of a
    // stream) keep whatever the buffer held: any byte is a finite sample, and the
    // decisions they influence lie outside the reportable range.
    __syncthreads();

    // work unit = (stream s, tile j); this CTA walks units blockIdx.x, +gridDim.x, ...
    // Thread 0 is the producer: it keeps DEMOD_STAGES tiles in flight.  Consumers
    // release a buffer as soon as their windows are in registers (one mbarrier
    // arrival per warp), so warps drift freely instead of meeting at a CTA barrier.
    const int step_s = (int)(gridDim.x / (unsigned)tiles), step_j = (int)(gridDim.x % (unsigned)tiles);
    int s = (int)(blockIdx.x / (unsigned)tiles), j = (int)(blockIdx.x % (unsigned)tiles);
    int ps = s, pj = j;                                          // next unit to prefetch
    const bool producer = threadIdx.x == 0;
    for (int i = 0; i < DEMOD_STAGES && ps < n_streams; i++) {
        const TileSpan sp = tile_span(vw, ps, pj);
        if (producer) issue_tile(sp, smem_addr(buf + i * BUF_BYTES), smem_addr(&full_bar[i]));
        ps += step_s; pj += step_j;
        if (pj >= tiles) { pj -= tiles; ps++; }
    }

    int slot = 0;
    uint32_t parity = 0;
    for (; s < n_streams;) {
        const uint32_t full = smem_addr(&full_bar[slot]), empty = smem_addr(&empty_bar[slot]);
        mbar_wait(full, parity);

        // window of run A: samples 28*t - 12 .. 28*t + 27 (t = thread), i.e. bytes
        // 56*t + 8 .. 56*t + 88 of the buffer (the buffer starts 16 samples early);
        // run B is the same window half a tile further on
        const uint8_t *base = buf + slot * BUF_BYTES + 56 * threadIdx.x + 8;
        uint32_t wa[20], wb[20];
#pragma unroll
        for (int k = 0; k < 10; k++) {
            const uint2 ta = reinterpret_cast<const uint2 *>(base)[k];
            const uint2 tb = reinterpret_cast<const uint2 *>(base + HALF_BYTES)[k];
            wa[2 * k] = ta.x; wa[2 * k + 1] = ta.y;
            wb[2 * k] = tb.x; wb[2 * k + 1] = tb.y;
        }
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(empty);          // this warp is done with the buffer
        {                                                        // refill it for DEMOD_STAGES tiles ahead
            // only the producer's warp works out the span (a warp-uniform branch: the other
            // 15 warps skip ~60 index instructions per tile)
            if (threadIdx.x < 32) {
                if (ps < n_streams) {
                    const TileSpan sp = tile_span(vw, ps, pj);
                    if (producer) {
                        mbar_wait(empty, parity);
                        issue_tile(sp, smem_addr(buf + slot * BUF_BYTES), full);
                    }
                }
                ps += step_s; pj += step_j;
                if (pj >= tiles) { pj -= tiles; ps++; }
            }
        }

        // re/im[i] = rotated sample of window index i (V sample 28t - 12 + i), i = 2..38
        // F[q] (q = 0..28) = filter output for V sample 28t - 1 + q, from window 2+q .. 10+q
        float2 re[39], im[39];
        Unpack<2>::run(wa, wb, re, im);

        uint32_t bits_a = 0, bits_b = 0;
        float2 pre = f2(0.f, 0.f), pim = f2(0.f, 0.f);
#pragma unroll
        for (int q = 0; q < 29; q++) {
            const float2 fre = fir9_pair(re, q, lut_residual(q & 3, 0));
            const float2 fim = fir9_pair(im, q, lut_residual(q & 3, 1));
            if (q > 0) {
                // dsp.py:88: im[n-1] * re[n] - re[n-1] * im[n]  (sign only)
                const float2 num = fma2(pim, fre, neg2(__fmul2_rn(pre, fim)));
                bits_a = __funnelshift_l(__float_as_uint(num.x), bits_a, 1);
                bits_b = __funnelshift_l(__float_as_uint(num.y), bits_b, 1);
                if constexpr (WRITE_DISC) {
                    const float eps = 1e-10f * 127.6f * 127.6f;
                    const float2 den = add2(fma2(pre, pre, __fmul2_rn(pim, pim)), f2(eps, eps));
                    const long long v = (long long)j * TILE_SAMPLES + LANE_SAMPLES * threadIdx.x + (q - 1);
                    const long long v_end = (long long)vw.n_carry + vw.n_new;
                    float *row = disc_out + (size_t)s * disc_stride;
                    if (v < v_end) row[v] = num.x / den.x;
                    if (v + TILE_SAMPLES / 2 < v_end) row[v + TILE_SAMPLES / 2] = num.y / den.y;
                }
            }
            pre = fre; pim = fim;
        }
        // first sample was shifted in first: bit 27 - p holds sample p; store bit p = sample p
        uint32_t *out = words + (size_t)s * words_alloc + (size_t)j * TILE_WORDS + threadIdx.x;
        out[0] = __brev(bits_a) >> 4;
        out[TILE_WORDS / 2] = __brev(bits_b) >> 4;
        s += step_s; j += step_j;
        if (j >= tiles) { j -= tiles; s++; }
        if (++slot == DEMOD_STAGES) { slot = 0; parity ^= 1; }
    }
}

void launch_demod(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                  float *disc_out, size_t disc_stride, cudaStream_t st) {
    const long long units = (long long)n_streams * tiles;
    if (units == 0) return;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)std::min<long long>(units, (long long)sms * DEMOD_CTAS_PER_SM);
    const size_t smem = (size_t)DEMOD_STAGES * BUF_BYTES;
    if (disc_out)
        cudaFuncSetAttribute(k_demod<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    else
        cudaFuncSetAttribute(k_demod<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (disc_out)
        k_demod<true><<<grid, DEMOD_THREADS, smem, st>>>(vw, words, words_alloc, tiles, n_streams,
                                                         disc_out, disc_stride);
    else
        k_demod<false><<<grid, DEMOD_THREADS, smem, st>>>(vw, words, words_alloc, tiles, n_streams,
                                                          nullptr, 0);
}

}  // namespace davis
