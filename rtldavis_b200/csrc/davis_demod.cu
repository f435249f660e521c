// davis_demod.cu -- the streaming demodulation kernel (uint8 IQ -> decision bitmap).
//
// Replaces, for a whole batch of streams and blocks at once, the per-block chain
// ByteToCmplxLUT.execute -> rotate_fs4 -> fir9 -> discriminate -> quantize of
// Demodulator.demodulate (src/rtldavis/dsp.py:139-166; Go dsp/dsp.go:274-292).
//
// Work decomposition
//   * one lane decides LANE_SAMPLES = 28 consecutive samples (two symbol periods)
//     with a register sliding window, so the 9-sample FIR history is re-read for
//     only 9/28 of the work (the previous 14-sample version paid 9/14);
//   * the 28 sign bits are shifted into one 32-bit word per lane with a funnel
//     shift per sample (SHF, 1 instruction) -- no compare, no warp vote; word w
//     of a stream holds samples [28w, 28w+28), bit i = sample 28w+i.  Stepping by
//     one symbol (14 samples) is a half-word step, which the search kernel
//     exploits (davis_search.cu);
//   * a CTA of 8 warps owns a tile of 256 words = 7168 samples = 14 KiB of input.
//     CTAs are persistent: each walks tiles unit = blockIdx.x, +gridDim.x, ...
//     and an elected thread streams the next tile into shared memory with the
//     bulk-copy engine (cp.async.bulk + mbarrier, SASS UBLKCP) while the current
//     one is being computed: no per-thread staging instructions, and global
//     latency is hidden behind the arithmetic;
//   * lanes read their 80-byte window with ten 8-byte shared loads (lane stride
//     56 B = 7 x 8 B, odd, so the loads are bank-conflict free).
//
// Arithmetic (fp32, signs only)
//   x = (u - 127.4)/127.6 and the division by |F|^2 + eps are positive scalings
//   and are dropped; bytes are centred exactly (u - 127) while unpacking and the
//   remaining -0.4 of the LUT offset survives the filter as a period-4 constant
//   (0.4 * (2c0 - 2c2 + c4) * (1+j) * j^q) folded into the first FMA;
//   the rotation j^i is applied while unpacking (which byte lands in which
//   component, and with which sign); the rotation phase is taken relative to the
//   lane's window because a common unit factor cancels in Im(conj(F[n-1]) F[n]).
//   Bytes become floats without I2F: one PRMT builds the fp16 pair
//   (1024 + I, 1024 + Q) and a mixed-precision add (add.f32.f16, SASS FHADD)
//   per component yields +-(u - 127) exactly.
//   Every decision that reaches a caller is re-derived in fp64 by the confirm
//   kernel, so an fp32 sign that disagrees with the reference's float64 (only
//   possible when the cross product is ~1e-6 of |F[n-1]||F[n]|) cannot change a
//   packet.
//
// Cost model (measured pipe rates, profiles/r01_microbench2_pipes.txt: FFMA/FADD/
// FHADD 1 cycle per warp instruction per SM sub-partition, FFMA2/FADD2/PRMT/SHF
// 2 cycles, FMA and ALU pipes run concurrently, 1 issue slot per cycle):
//   FMA pipe per sample: FIR 9 packed ops x 2 x 29/28 + 2 FHADD x 37/28 + 2 = 23.3
//   cycles => 148 SM x 4 x 1.965 GHz x 32 / 23.3 = 1.6 Tsample/s = 3.2 TB/s of
//   input, i.e. this kernel is bound by the FP32 pipe at ~49 % of the HBM
//   roofline, not by HBM.
#include <algorithm>

#include "davis_kernels.cuh"

namespace davis {

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 cfma(float2 c, float2 a, float2 acc) { return __ffma2_rn(c, a, acc); }

// (u - 127) or (127 - u) from the fp16 value 1024 + u held in half `H` (0 or 1) of h2
template <int H, bool NEG>
__device__ __forceinline__ float centred(uint32_t h2) {
    float r;
    const unsigned short h = (unsigned short)(H ? (h2 >> 16) : (h2 & 0xFFFFu));
    if (NEG)
        asm("{ .reg .b16 t; neg.f16 t, %1; add.rn.f32.f16 %0, t, %2; }" : "=f"(r) : "h"(h), "f"(1151.0f));
    else
        asm("add.rn.f32.f16 %0, %1, %2;" : "=f"(r) : "h"(h), "f"(-1151.0f));
    return r;
}

// Rotated integer-valued sample rho_i = j^(i mod 4) * ((I - 127) + j(Q - 127)) of
// window index i, taken from the 32-bit word holding samples (2k, 2k+1).
template <int I>
__device__ __forceinline__ float2 unpack_rotated(uint32_t w) {
    // fp16 pair (1024 + I, 1024 + Q): bytes (I, 0x64, Q, 0x64)
    const uint32_t h2 = __byte_perm(w, 0x64646464u, (I & 1) ? 0x4342 : 0x4140);
    constexpr int ph = I & 3;
    // phase 0: (+I, +Q)   1: (-Q, +I)   2: (-I, -Q)   3: (+Q, -I)
    if (ph == 0) return f2(centred<0, false>(h2), centred<1, false>(h2));
    if (ph == 1) return f2(centred<1, true>(h2), centred<0, false>(h2));
    if (ph == 2) return f2(centred<0, true>(h2), centred<1, true>(h2));
    return f2(centred<1, false>(h2), centred<0, true>(h2));
}

template <int I>
struct Unpack {
    static __device__ __forceinline__ void run(const uint32_t (&w)[20], float2 (&r)[39]) {
        r[I] = unpack_rotated<I>(w[I >> 1]);
        Unpack<I + 1>::run(w, r);
    }
};
template <>
struct Unpack<39> {
    static __device__ __forceinline__ void run(const uint32_t (&)[20], float2 (&)[39]) {}
};

// ---- bulk copy + mbarrier (sm_90+ PTX; the copy engine writes shared memory) ----
__device__ __forceinline__ uint32_t smem_addr(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; "
                     "selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}

constexpr int DEMOD_THREADS = DEMOD_WARPS * 32;
constexpr int TILE_BYTES = 2 * TILE_SAMPLES;                 // 14336
constexpr int LEAD_BYTES = 32;                               // history in front of a tile
constexpr int BUF_BYTES = TILE_BYTES + LEAD_BYTES;           // 14368
constexpr int DEMOD_STAGES = 3;                              // tiles in flight per CTA

// Enqueue the copy of tile j of stream s into `dst` (bytes [j*TILE_BYTES - 32,
// (j+1)*TILE_BYTES) of the stream's V sequence, clipped to what exists).
__device__ __forceinline__ void issue_tile(const View &vw, int s, int j, uint32_t dst, uint32_t bar) {
    const long long carry_bytes = 2LL * vw.n_carry;
    const long long total = carry_bytes + 2LL * vw.n_new;
    long long b0 = (long long)j * TILE_BYTES - LEAD_BYTES;
    long long b1 = b0 + BUF_BYTES;
    uint32_t skip = 0;
    if (b0 < 0) { skip = (uint32_t)(-b0); b0 = 0; }
    if (b1 > total) b1 = total;
    if (b1 <= b0) { mbar_expect_tx(bar, 0); return; }          // tile past the end: bits unused
    const uint8_t *carry = vw.carry + (size_t)s * vw.carry_stride;
    const uint8_t *input = vw.input + (size_t)s * vw.input_stride;
    mbar_expect_tx(bar, (uint32_t)(b1 - b0));
    if (b0 < carry_bytes) {
        const long long e = b1 < carry_bytes ? b1 : carry_bytes;
        bulk_g2s(dst + skip, carry + b0, (uint32_t)(e - b0), bar);
        skip += (uint32_t)(e - b0);
        b0 = e;
    }
    if (b0 < b1) bulk_g2s(dst + skip, input + (b0 - carry_bytes), (uint32_t)(b1 - b0), bar);
}

template <bool WRITE_DISC>
__global__ void __launch_bounds__(DEMOD_THREADS, DEMOD_CTAS_PER_SM)
k_demod(View vw, uint32_t *__restrict__ words, int words_alloc, int tiles, int n_streams,
        float *__restrict__ disc_out, size_t disc_stride) {
    __shared__ __align__(128) uint8_t buf[DEMOD_STAGES][BUF_BYTES];
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
    // release a buffer as soon as their window is in registers (one mbarrier
    // arrival per warp), so warps drift freely instead of meeting at a CTA barrier.
    const int step_s = (int)(gridDim.x / (unsigned)tiles), step_j = (int)(gridDim.x % (unsigned)tiles);
    int s = (int)(blockIdx.x / (unsigned)tiles), j = (int)(blockIdx.x % (unsigned)tiles);
    int ps = s, pj = j;                                          // next unit to prefetch (thread 0)
    if (threadIdx.x == 0) {
        for (int i = 0; i < DEMOD_STAGES && ps < n_streams; i++) {
            issue_tile(vw, ps, pj, smem_addr(&buf[i][0]), smem_addr(&full_bar[i]));
            ps += step_s; pj += step_j;
            if (pj >= tiles) { pj -= tiles; ps++; }
        }
    }

    const float c0 = (float)DAVIS_C0, c1 = (float)DAVIS_C1, c2 = (float)DAVIS_C2,
                c3 = (float)DAVIS_C3, c4 = (float)DAVIS_C4;
    const float K = (float)(0.4 * (2.0 * DAVIS_C0 - 2.0 * DAVIS_C2 + DAVIS_C4));

    int slot = 0;
    uint32_t parity = 0;
    for (; s < n_streams;) {
        const uint32_t full = smem_addr(&full_bar[slot]), empty = smem_addr(&empty_bar[slot]);
        mbar_wait(full, parity);

        // window of this lane: samples 28*t - 12 .. 28*t + 27 (t = thread), i.e. bytes
        // 56*t + 8 .. 56*t + 88 of the buffer (the buffer starts 16 samples early)
        const uint2 *lw = reinterpret_cast<const uint2 *>(&buf[slot][0] + 56 * threadIdx.x + 8);
        uint32_t w[20];
#pragma unroll
        for (int k = 0; k < 10; k++) {
            const uint2 t = lw[k];
            w[2 * k] = t.x;
            w[2 * k + 1] = t.y;
        }
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(empty);          // this warp is done with the buffer
        if (threadIdx.x == 0 && ps < n_streams) {                // refill it for DEMOD_STAGES tiles ahead
            mbar_wait(empty, parity);
            issue_tile(vw, ps, pj, smem_addr(&buf[slot][0]), full);
            ps += step_s; pj += step_j;
            if (pj >= tiles) { pj -= tiles; ps++; }
        }

        // r[i] = rotated sample of window index i (V sample 28t - 12 + i), i = 2..38
        // F[q] (q = 0..28) = filter output for V sample 28t - 1 + q, from r[q+2 .. q+10]
        float2 r[39];
        Unpack<2>::run(w, r);

        uint32_t bits = 0;
        float2 prev = f2(0.f, 0.f);
#pragma unroll
        for (int q = 0; q < 29; q++) {
            // -0.4 * S * (1+j) * j^(q+2)  (samples are u - 127, the LUT subtracts 127.4)
            const int ph = (q + 2) & 3;
            const float2 off = ph == 0 ? f2(-K, -K) : ph == 1 ? f2(K, -K) : ph == 2 ? f2(K, K) : f2(-K, K);
            float2 acc = cfma(f2(c4, c4), r[q + 6], off);
            acc = cfma(f2(c0, c0), cadd(r[q + 2], r[q + 10]), acc);
            acc = cfma(f2(c1, c1), cadd(r[q + 3], r[q + 9]), acc);
            acc = cfma(f2(c2, c2), cadd(r[q + 4], r[q + 8]), acc);
            acc = cfma(f2(c3, c3), cadd(r[q + 5], r[q + 7]), acc);
            if (q > 0) {
                // dsp.py:88: im[n-1] * re[n] - re[n-1] * im[n]  (sign only)
                const float num = fmaf(prev.y, acc.x, -(prev.x * acc.y));
                bits = __funnelshift_l(__float_as_uint(num), bits, 1);
                if constexpr (WRITE_DISC) {
                    const long long v = (long long)j * TILE_SAMPLES + LANE_SAMPLES * threadIdx.x + (q - 1);
                    if (v < vw.n_carry + vw.n_new) {
                        const float den = fmaf(prev.x, prev.x, prev.y * prev.y) + 1e-10f * 127.6f * 127.6f;
                        disc_out[(size_t)s * disc_stride + v] = num / den;
                    }
                }
            }
            prev = acc;
        }
        // first sample was shifted in first: bit 27 - p holds sample p; store bit p = sample p
        words[(size_t)s * words_alloc + (size_t)j * TILE_WORDS + threadIdx.x] = __brev(bits) >> 4;
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
    if (disc_out)
        k_demod<true><<<grid, DEMOD_THREADS, 0, st>>>(vw, words, words_alloc, tiles, n_streams,
                                                      disc_out, disc_stride);
    else
        k_demod<false><<<grid, DEMOD_THREADS, 0, st>>>(vw, words, words_alloc, tiles, n_streams,
                                                       nullptr, 0);
}

}  // namespace davis
