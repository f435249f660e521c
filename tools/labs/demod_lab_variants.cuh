// demod_lab_variants.cuh -- experimental variants of k_demod (development tool).
#pragma once
#include <algorithm>
#include <cstdio>

#include "../../rtldavis_b200/csrc/davis_kernels.cuh"

namespace lab {
using namespace davis;

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 cfma(float2 c, float2 a, float2 acc) { return __ffma2_rn(c, a, acc); }

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
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

constexpr int TILE_BYTES = 2 * TILE_SAMPLES;
constexpr int LEAD_BYTES = 32;
constexpr int BUF_BYTES = TILE_BYTES + LEAD_BYTES;

__device__ __forceinline__ void issue_tile(const View &vw, int s, int j, uint32_t dst, uint32_t bar) {
    const long long carry_bytes = 2LL * vw.n_carry;
    const long long total = carry_bytes + 2LL * vw.n_new;
    long long b0 = (long long)j * TILE_BYTES - LEAD_BYTES;
    long long b1 = b0 + BUF_BYTES;
    uint32_t skip = 0;
    if (b0 < 0) { skip = (uint32_t)(-b0); b0 = 0; }
    if (b1 > total) b1 = total;
    if (b1 <= b0) { mbar_expect_tx(bar, 0); return; }
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

// ---- conversions -----------------------------------------------------------
// CONV 0: PRMT -> fp16 pair (1024+I, 1024+Q), FHADD with +-1151 (the product kernel's way)
// CONV 1: word ^ mask, then I2F.S8 with byte select; residual constants 0.6 / 0.4
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

template <int CONV, int I>
__device__ __forceinline__ float2 unpack_rotated(uint32_t w, float k1151) {
    constexpr int ph = I & 3;
    if constexpr (CONV == 0) {
        const uint32_t h2 = __byte_perm(w, 0x64646464u, (I & 1) ? 0x4342 : 0x4140);
        if (ph == 0) return f2(centred<0, false>(h2, k1151), centred<1, false>(h2, k1151));
        if (ph == 1) return f2(centred<1, true>(h2, k1151), centred<0, false>(h2, k1151));
        if (ph == 2) return f2(centred<0, true>(h2, k1151), centred<1, true>(h2, k1151));
        return f2(centred<1, false>(h2, k1151), centred<0, true>(h2, k1151));
    } else if constexpr (CONV == 2) {
        return f2(__uint_as_float((w >> (I & 1)) | 0x3F800000u), __uint_as_float((w >> (3 + (I & 1))) | 0x3F800000u));
    } else {
        constexpr int sh = (I & 1) ? 16 : 0;
        const float a = (float)(signed char)(w >> sh);          // +-(I - 127/128)
        const float b = (float)(signed char)(w >> (sh + 8));    // +-(Q - 127/128)
        if (ph == 0 || ph == 2) return f2(a, b);
        return f2(b, a);
    }
}

template <int CONV, int I>
struct Unpack {
    static __device__ __forceinline__ void run(const uint32_t (&w)[20], float2 (&r)[39], float k) {
        r[I] = unpack_rotated<CONV, I>(w[I >> 1], k);
        Unpack<CONV, I + 1>::run(w, r, k);
    }
};
template <int CONV>
struct Unpack<CONV, 39> {
    static __device__ __forceinline__ void run(const uint32_t (&)[20], float2 (&)[39], float) {}
};

// residual of the LUT offset after the filter, for output q (window taps q+2 .. q+10)
template <int CONV>
__host__ __device__ constexpr float residual(int q, int comp) {
    const double c[9] = {DAVIS_C0, DAVIS_C1, DAVIS_C2, DAVIS_C3, DAVIS_C4, DAVIS_C3, DAVIS_C2, DAVIS_C1, DAVIS_C0};
    double acc = 0.0;
    for (int m = 0; m < 9; m++) {
        const int ph = (q + 2 + m) & 3;
        double e = 0.0;
        if (CONV == 0) {
            // samples are +-(u - 127): true value is that -+ 0.4
            const double sre[4] = {-0.4, 0.4, 0.4, -0.4}, sim[4] = {-0.4, -0.4, 0.4, 0.4};
            e = comp ? sim[ph] : sre[ph];
        } else {
            // '+' components are (u - 128): +0.6 ; '-' components are (127 - u): +0.4
            const double sre[4] = {0.6, 0.4, 0.4, 0.6}, sim[4] = {0.6, 0.6, 0.4, 0.4};
            e = comp ? sim[ph] : sre[ph];
        }
        acc += c[m] * e;
    }
    return (float)acc;
}

template <int CONV, bool NOMEM, int WARPS, int CTAS, int STAGES, int SKIPX = 0>
__global__ void __launch_bounds__(WARPS * 32, CTAS)
k_lab(View vw, uint32_t *__restrict__ words, int words_alloc, int tiles, int n_streams) {
    static_assert(WARPS * 32 == TILE_WORDS, "one word per thread");
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t (*buf)[BUF_BYTES + 96] = reinterpret_cast<uint8_t (*)[BUF_BYTES + 96]>(smem);
    __shared__ __align__(8) unsigned long long full_bar[STAGES], empty_bar[STAGES];

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; i++) {
            mbar_init(smem_addr(&full_bar[i]), 1);
            mbar_init(smem_addr(&empty_bar[i]), WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int step_s = (int)(gridDim.x / (unsigned)tiles), step_j = (int)(gridDim.x % (unsigned)tiles);
    int s = (int)(blockIdx.x / (unsigned)tiles), j = (int)(blockIdx.x % (unsigned)tiles);
    int ps = s, pj = j;
    if (!NOMEM && threadIdx.x == 0) {
        for (int i = 0; i < STAGES && ps < n_streams; i++) {
            issue_tile(vw, ps, pj, smem_addr(&buf[i][0]), smem_addr(&full_bar[i]));
            ps += step_s; pj += step_j;
            if (pj >= tiles) { pj -= tiles; ps++; }
        }
    }

    const float c0 = (float)DAVIS_C0, c1 = (float)DAVIS_C1, c2 = (float)DAVIS_C2,
                c3 = (float)DAVIS_C3, c4 = (float)DAVIS_C4;
    float k1151;
    if constexpr (SKIPX >= 16) {
        __shared__ float kconst[32];
        if (threadIdx.x < 32) kconst[threadIdx.x] = 1151.0f;
        __syncthreads();
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(k1151) : "r"(smem_addr(&kconst[threadIdx.x & 31])));
    } else {
        asm volatile("mov.f32 %0, 0f448FE000;" : "=f"(k1151));     // opaque: keeps the constant in a register
    }

    int slot = 0;
    uint32_t parity = 0;
    for (; s < n_streams;) {
        const uint32_t full = smem_addr(&full_bar[slot]), empty = smem_addr(&empty_bar[slot]);
        if (!NOMEM) mbar_wait(full, parity);

        const uint2 *lw = reinterpret_cast<const uint2 *>(&buf[slot][0] + 56 * threadIdx.x + 8);
        uint32_t w[20];
#pragma unroll
        for (int k = 0; k < 10; k++) {
            const uint2 t = lw[k];
            w[2 * k] = t.x;
            w[2 * k + 1] = t.y;
        }
        if (!NOMEM) {
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive(empty);
            if (threadIdx.x == 0 && ps < n_streams) {
                mbar_wait(empty, parity);
                issue_tile(vw, ps, pj, smem_addr(&buf[slot][0]), full);
                ps += step_s; pj += step_j;
                if (pj >= tiles) { pj -= tiles; ps++; }
            }
        }
        if constexpr (CONV == 1) {
#pragma unroll
            for (int k = 0; k < 20; k++) w[k] ^= (k & 1) ? 0x807F7F7Fu : 0x7F808080u;
        }

        float2 r[39];
        Unpack<CONV, 2>::run(w, r, k1151);

        uint32_t bits = 0;
        float2 prev = f2(0.f, 0.f);
#pragma unroll
        for (int q = 0; q < 29; q++) {
            const float2 off = f2(residual<CONV>(q & 3, 0), residual<CONV>(q & 3, 1));
            float2 acc = cfma(f2(c4, c4), r[q + 6], off);
            acc = cfma(f2(c0, c0), cadd(r[q + 2], r[q + 10]), acc);
            acc = cfma(f2(c1, c1), cadd(r[q + 3], r[q + 9]), acc);
            acc = cfma(f2(c2, c2), cadd(r[q + 4], r[q + 8]), acc);
            acc = cfma(f2(c3, c3), cadd(r[q + 5], r[q + 7]), acc);
            if (q > 0) {
                if constexpr ((SKIPX & 15) == 0) {
                    const float num = fmaf(prev.y, acc.x, -(prev.x * acc.y));
                    bits = __funnelshift_l(__float_as_uint(num), bits, 1);
                } else if constexpr ((SKIPX & 15) == 1) {
                    bits = __funnelshift_l(__float_as_uint(acc.x) ^ __float_as_uint(acc.y), bits, 1);
                } else {
                    bits ^= __float_as_uint(acc.x) ^ __float_as_uint(acc.y);
                }
            }
            prev = acc;
        }
        words[(size_t)s * words_alloc + (size_t)j * TILE_WORDS + threadIdx.x] = __brev(bits) >> 4;
        s += step_s; j += step_j;
        if (j >= tiles) { j -= tiles; s++; }
        if (++slot == STAGES) { slot = 0; parity ^= 1; }
    }
}

template <int CONV, bool NOMEM, int WARPS, int CTAS, int STAGES, int SKIPX = 0>
void launch_k(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc) {
    const long long units = (long long)n_streams * tiles;
    const unsigned grid = (unsigned)std::min<long long>(units, 148LL * CTAS);
    const size_t smem = (size_t)STAGES * (BUF_BYTES + 96);
    auto kern = k_lab<CONV, NOMEM, WARPS, CTAS, STAGES, SKIPX>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<grid, WARPS * 32, smem>>>(vw, words, words_alloc, tiles, n_streams);
}

// ---------------------------------------------------------------------------
// k_warp: every warp runs its own bulk-copy pipeline (no CTA-level coupling).
// BODY 0: one 28-sample run per lane (warp tile 32 words), FHADD conversion
// BODY 1: two runs per lane (words lane and 32+lane of a 64-word warp tile) held as
//         (A, B) register pairs so conversion, FIR and cross product are all packed
// ---------------------------------------------------------------------------
__device__ __forceinline__ void issue_span(const View &vw, int s, long long b0, int nbytes,
                                           uint32_t dst, uint32_t bar) {
    const long long carry_bytes = 2LL * vw.n_carry;
    const long long total = carry_bytes + 2LL * vw.n_new;
    long long b1 = b0 + nbytes;
    uint32_t skip = 0;
    if (b0 < 0) { skip = (uint32_t)(-b0); b0 = 0; }
    if (b1 > total) b1 = total;
    if (b1 <= b0) { mbar_expect_tx(bar, 0); return; }
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
    const float c0 = (float)DAVIS_C0, c1 = (float)DAVIS_C1, c2 = (float)DAVIS_C2,
                c3 = (float)DAVIS_C3, c4 = (float)DAVIS_C4;
    float2 acc = cfma(f2(c4, c4), r[q + 6], f2(off, off));
    acc = cfma(f2(c0, c0), cadd(r[q + 2], r[q + 10]), acc);
    acc = cfma(f2(c1, c1), cadd(r[q + 3], r[q + 9]), acc);
    acc = cfma(f2(c2, c2), cadd(r[q + 4], r[q + 8]), acc);
    acc = cfma(f2(c3, c3), cadd(r[q + 5], r[q + 7]), acc);
    return acc;
}

template <int BODY, bool NOMEM, int WARPS, int CTAS, int STAGES>
__global__ void __launch_bounds__(WARPS * 32, CTAS)
k_warp(View vw, uint32_t *__restrict__ words, int words_alloc, int tiles, int n_streams) {
    constexpr int WT_WORDS = BODY == 1 ? 64 : 32;
    constexpr int WT_BYTES = WT_WORDS * LANE_SAMPLES * 2;           // 1792 or 3584
    constexpr int WBUF = WT_BYTES + LEAD_BYTES + 96;                // padded: windows overrun the tile
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) unsigned long long full_bar[WARPS][STAGES];
    __shared__ float kconst[32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *wbuf = smem + (size_t)warp * STAGES * WBUF;

    if (threadIdx.x < 32) kconst[threadIdx.x] = 1151.0f;
    if (lane == 0) {
        for (int i = 0; i < STAGES; i++) mbar_init(smem_addr(&full_bar[warp][i]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    float k1151;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(k1151) : "r"(smem_addr(&kconst[lane])));

    const int n_wt = tiles * (TILE_WORDS / WT_WORDS);               // warp tiles per stream
    const unsigned total_warps = gridDim.x * WARPS;
    const int step_s = (int)(total_warps / (unsigned)n_wt), step_j = (int)(total_warps % (unsigned)n_wt);
    const unsigned gw = blockIdx.x * WARPS + warp;
    int s = (int)(gw / (unsigned)n_wt), j = (int)(gw % (unsigned)n_wt);
    int ps = s, pj = j;
    if (!NOMEM && lane == 0) {
        for (int i = 0; i < STAGES && ps < n_streams; i++) {
            issue_span(vw, ps, (long long)pj * WT_BYTES - LEAD_BYTES, WT_BYTES + LEAD_BYTES,
                       smem_addr(wbuf + i * WBUF), smem_addr(&full_bar[warp][i]));
            ps += step_s; pj += step_j;
            if (pj >= n_wt) { pj -= n_wt; ps++; }
        }
    }

    int slot = 0;
    uint32_t parity = 0;
    for (; s < n_streams;) {
        const uint32_t full = smem_addr(&full_bar[warp][slot]);
        if (!NOMEM) mbar_wait(full, parity);
        const uint8_t *base = wbuf + slot * WBUF + 56 * lane + 8;
        uint32_t w[20], wb[BODY == 1 ? 20 : 1];
#pragma unroll
        for (int k = 0; k < 10; k++) {
            const uint2 t = reinterpret_cast<const uint2 *>(base)[k];
            w[2 * k] = t.x; w[2 * k + 1] = t.y;
            if constexpr (BODY == 1) {
                const uint2 u = reinterpret_cast<const uint2 *>(base + 1792)[k];
                wb[2 * k] = u.x; wb[2 * k + 1] = u.y;
            }
        }
        if (!NOMEM) {
            __syncwarp();
            if (lane == 0 && ps < n_streams) {
                issue_span(vw, ps, (long long)pj * WT_BYTES - LEAD_BYTES, WT_BYTES + LEAD_BYTES,
                           smem_addr(wbuf + slot * WBUF), full);
                ps += step_s; pj += step_j;
                if (pj >= n_wt) { pj -= n_wt; ps++; }
            }
        }
        uint32_t *out = words + (size_t)s * words_alloc + (size_t)j * WT_WORDS + lane;
        if constexpr (BODY == 0) {
            float2 r[39];
            Unpack<0, 2>::run(w, r, k1151);
            uint32_t bits = 0;
            float2 prev = f2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < 29; q++) {
                const float2 off = f2(residual<0>(q & 3, 0), residual<0>(q & 3, 1));
                const float c0 = (float)DAVIS_C0, c1 = (float)DAVIS_C1, c2 = (float)DAVIS_C2,
                            c3 = (float)DAVIS_C3, c4 = (float)DAVIS_C4;
                float2 acc = cfma(f2(c4, c4), r[q + 6], off);
                acc = cfma(f2(c0, c0), cadd(r[q + 2], r[q + 10]), acc);
                acc = cfma(f2(c1, c1), cadd(r[q + 3], r[q + 9]), acc);
                acc = cfma(f2(c2, c2), cadd(r[q + 4], r[q + 8]), acc);
                acc = cfma(f2(c3, c3), cadd(r[q + 5], r[q + 7]), acc);
                if (q > 0) {
                    const float num = fmaf(prev.y, acc.x, -(prev.x * acc.y));
                    bits = __funnelshift_l(__float_as_uint(num), bits, 1);
                }
                prev = acc;
            }
            out[0] = __brev(bits) >> 4;
        } else {
            float2 re[39], im[39];
            ConvPair<2>::run(w, wb, re, im);
            uint32_t bits_a = 0, bits_b = 0;
            float2 pre = f2(0.f, 0.f), pim = f2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < 29; q++) {
                const float2 are = fir_pair(re, q, residual<0>(q & 3, 0));
                const float2 aim = fir_pair(im, q, residual<0>(q & 3, 1));
                if (q > 0) {
                    const float2 t = __fmul2_rn(pre, aim);
                    const float2 num = cfma(pim, are, f2(-t.x, -t.y));
                    bits_a = __funnelshift_l(__float_as_uint(num.x), bits_a, 1);
                    bits_b = __funnelshift_l(__float_as_uint(num.y), bits_b, 1);
                }
                pre = are; pim = aim;
            }
            out[0] = __brev(bits_a) >> 4;
            out[32] = __brev(bits_b) >> 4;
        }
        s += step_s; j += step_j;
        if (j >= n_wt) { j -= n_wt; s++; }
        if (++slot == STAGES) { slot = 0; parity ^= 1; }
    }
}

template <int BODY, bool NOMEM, int WARPS, int CTAS, int STAGES>
void launch_w(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc) {
    constexpr int WT_WORDS = BODY == 1 ? 64 : 32;
    constexpr int WBUF = WT_WORDS * LANE_SAMPLES * 2 + LEAD_BYTES + 96;
    const size_t smem = (size_t)WARPS * STAGES * WBUF;
    auto kern = k_warp<BODY, NOMEM, WARPS, CTAS, STAGES>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<148 * CTAS, WARPS * 32, smem>>>(vw, words, words_alloc, tiles, n_streams);
}

// ---------------------------------------------------------------------------
// k_cta2: CTA-shared tile pipeline again, but the producer's address arithmetic is
// done by every thread on uniform values (no divergent 64-bit math, no R2UR waterfall);
// only the copy instructions themselves are predicated on one elected thread.
// BODY as in k_warp; PROD 0: thread 0 of warp 0 issues, PROD 1: a dedicated extra warp.
// ---------------------------------------------------------------------------
struct Span { const uint8_t *p0, *p1; uint32_t n0, n1, skip; };

__device__ __forceinline__ Span span_of(const View &vw, int s, long long b0, int nbytes) {
    const long long carry_bytes = 2LL * vw.n_carry;
    const long long total = carry_bytes + 2LL * vw.n_new;
    long long b1 = b0 + nbytes;
    Span sp{nullptr, nullptr, 0u, 0u, 0u};
    if (b0 < 0) { sp.skip = (uint32_t)(-b0); b0 = 0; }
    if (b1 > total) b1 = total;
    if (b1 <= b0) return sp;
    const uint8_t *carry = vw.carry + (size_t)s * vw.carry_stride;
    const uint8_t *input = vw.input + (size_t)s * vw.input_stride;
    if (b0 < carry_bytes) {
        const long long e = b1 < carry_bytes ? b1 : carry_bytes;
        sp.p0 = carry + b0; sp.n0 = (uint32_t)(e - b0);
        b0 = e;
    }
    if (b0 < b1) { sp.p1 = input + (b0 - carry_bytes); sp.n1 = (uint32_t)(b1 - b0); }
    return sp;
}
__device__ __forceinline__ void issue(const Span &sp, uint32_t dst, uint32_t bar) {
    mbar_expect_tx(bar, sp.n0 + sp.n1);
    if (sp.n0) bulk_g2s(dst + sp.skip, sp.p0, sp.n0, bar);
    if (sp.n1) bulk_g2s(dst + sp.skip + sp.n0, sp.p1, sp.n1, bar);
}

template <int BODY, int WARPS, int CTAS, int STAGES, int PROD>
__global__ void __launch_bounds__((WARPS + PROD) * 32, CTAS)
k_cta2(View vw, uint32_t *__restrict__ words, int words_alloc, int tiles, int n_streams) {
    constexpr int T_WORDS = WARPS * 32 * (BODY == 1 ? 2 : 1);       // words per CTA tile
    constexpr int T_BYTES = T_WORDS * LANE_SAMPLES * 2;
    constexpr int TBUF = T_BYTES + LEAD_BYTES + 96;
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) unsigned long long full_bar[STAGES], empty_bar[STAGES];
    __shared__ float kconst[32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x < 32) kconst[threadIdx.x] = 1151.0f;
    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; i++) {
            mbar_init(smem_addr(&full_bar[i]), 1);
            mbar_init(smem_addr(&empty_bar[i]), WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    float k1151;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(k1151) : "r"(smem_addr(&kconst[lane])));

    const int n_t = (tiles * TILE_WORDS + T_WORDS - 1) / T_WORDS;    // CTA tiles per stream
    const int step_s = (int)(gridDim.x / (unsigned)n_t), step_j = (int)(gridDim.x % (unsigned)n_t);
    int s = (int)(blockIdx.x / (unsigned)n_t), j = (int)(blockIdx.x % (unsigned)n_t);
    const bool leader = PROD ? (threadIdx.x == WARPS * 32) : (threadIdx.x == 0);

    if (PROD && warp == WARPS) {                                    // dedicated producer warp
        int slot = 0; uint32_t parity = 0; int it = 0;
        for (; s < n_streams; it++) {
            const Span sp = span_of(vw, s, (long long)j * T_BYTES - LEAD_BYTES, T_BYTES + LEAD_BYTES);
            if (leader) {
                if (it >= STAGES) mbar_wait(smem_addr(&empty_bar[slot]), parity ^ 1);
                issue(sp, smem_addr(smem + slot * TBUF), smem_addr(&full_bar[slot]));
            }
            s += step_s; j += step_j;
            if (j >= n_t) { j -= n_t; s++; }
            if (++slot == STAGES) { slot = 0; parity ^= 1; }
        }
        return;
    }

    int ps = s, pj = j;
    if (!PROD) {
        for (int i = 0; i < STAGES && ps < n_streams; i++) {
            const Span sp = span_of(vw, ps, (long long)pj * T_BYTES - LEAD_BYTES, T_BYTES + LEAD_BYTES);
            if (leader) issue(sp, smem_addr(smem + i * TBUF), smem_addr(&full_bar[i]));
            ps += step_s; pj += step_j;
            if (pj >= n_t) { pj -= n_t; ps++; }
        }
    }

    int slot = 0;
    uint32_t parity = 0;
    for (; s < n_streams;) {
        const uint32_t full = smem_addr(&full_bar[slot]), empty = smem_addr(&empty_bar[slot]);
        mbar_wait(full, parity);
        const uint8_t *base = smem + slot * TBUF + 56 * threadIdx.x + 8;
        uint32_t w[20], wb[BODY == 1 ? 20 : 1];
#pragma unroll
        for (int k = 0; k < 10; k++) {
            const uint2 t = reinterpret_cast<const uint2 *>(base)[k];
            w[2 * k] = t.x; w[2 * k + 1] = t.y;
            if constexpr (BODY == 1) {
                const uint2 u = reinterpret_cast<const uint2 *>(base + T_BYTES / 2)[k];
                wb[2 * k] = u.x; wb[2 * k + 1] = u.y;
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty);
        if (!PROD) {
            const bool more = ps < n_streams;
            const Span sp = span_of(vw, more ? ps : 0, (long long)pj * T_BYTES - LEAD_BYTES, T_BYTES + LEAD_BYTES);
            if (leader && more) {
                mbar_wait(empty, parity);
                issue(sp, smem_addr(smem + slot * TBUF), full);
            }
            ps += step_s; pj += step_j;
            if (pj >= n_t) { pj -= n_t; ps++; }
        }
        uint32_t *out = words + (size_t)s * words_alloc + (size_t)j * T_WORDS + threadIdx.x;
        if constexpr (BODY == 0) {
            float2 r[39];
            Unpack<0, 2>::run(w, r, k1151);
            uint32_t bits = 0;
            float2 prev = f2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < 29; q++) {
                const float2 acc = f2(0.f, 0.f);
                const float2 a2 = cadd(fir_pair(r, q, 0.f), f2(residual<0>(q & 3, 0), residual<0>(q & 3, 1)));
                (void)acc;
                if (q > 0) {
                    const float num = fmaf(prev.y, a2.x, -(prev.x * a2.y));
                    bits = __funnelshift_l(__float_as_uint(num), bits, 1);
                }
                prev = a2;
            }
            out[0] = __brev(bits) >> 4;
        } else {
            float2 re[39], im[39];
            ConvPair<2>::run(w, wb, re, im);
            uint32_t bits_a = 0, bits_b = 0;
            float2 pre = f2(0.f, 0.f), pim = f2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < 29; q++) {
                const float2 are = fir_pair(re, q, residual<0>(q & 3, 0));
                const float2 aim = fir_pair(im, q, residual<0>(q & 3, 1));
                if (q > 0) {
                    const float2 t = __fmul2_rn(pre, aim);
                    const float2 num = cfma(pim, are, f2(-t.x, -t.y));
                    bits_a = __funnelshift_l(__float_as_uint(num.x), bits_a, 1);
                    bits_b = __funnelshift_l(__float_as_uint(num.y), bits_b, 1);
                }
                pre = are; pim = aim;
            }
            const int wi = j * T_WORDS + (int)threadIdx.x;
            if (wi < tiles * TILE_WORDS) out[0] = __brev(bits_a) >> 4;
            if (wi + T_WORDS / 2 < tiles * TILE_WORDS) out[T_WORDS / 2] = __brev(bits_b) >> 4;
        }
        s += step_s; j += step_j;
        if (j >= n_t) { j -= n_t; s++; }
        if (++slot == STAGES) { slot = 0; parity ^= 1; }
    }
}

template <int BODY, int WARPS, int CTAS, int STAGES, int PROD>
void launch_c2(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc) {
    constexpr int T_WORDS = WARPS * 32 * (BODY == 1 ? 2 : 1);
    constexpr int TBUF = T_WORDS * LANE_SAMPLES * 2 + LEAD_BYTES + 96;
    const size_t smem = (size_t)STAGES * TBUF;
    auto kern = k_cta2<BODY, WARPS, CTAS, STAGES, PROD>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<148 * CTAS, (WARPS + PROD) * 32, smem>>>(vw, words, words_alloc, tiles, n_streams);
}

struct Variant { const char *name; void (*fn)(const View &, uint32_t *, int, int, int); };
static const Variant VARIANTS[] = {
    {"cta2 body1 8w x2cta 3stg prod0", launch_c2<1, 8, 2, 3, 0>},
    {"cta2 body1 6w x3cta 3stg prod0", launch_c2<1, 6, 3, 3, 0>},
    {"cta2 body1 6w x2cta 4stg prod0", launch_c2<1, 6, 2, 4, 0>},
    {"cta2 body1 16w x1cta 3stg prod0", launch_c2<1, 16, 1, 3, 0>},
    {"cta2 body1 12w x1cta 4stg prod0", launch_c2<1, 12, 1, 4, 0>},
    {"cta2 body0 8w x3cta 3stg prod0", launch_c2<0, 8, 3, 3, 0>},
    {"cta2 body0 8w x4cta 3stg prod0", launch_c2<0, 8, 4, 3, 0>},
    {"cta2 body0 16w x2cta 3stg prod0", launch_c2<0, 16, 2, 3, 0>},
};
inline int n_variants() { return (int)(sizeof(VARIANTS) / sizeof(VARIANTS[0])); }
inline const char *name(int v) { return VARIANTS[v].name; }
inline void launch(int v, const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc) {
    VARIANTS[v].fn(vw, words, n_streams, tiles, words_alloc);
}

// ---- pipe probes -------------------------------------------------------------
// NA independent chains of A and NB of B per group; prints cycles per group per SMSP.
enum Op { NONE, FFMA, FFMA2, I2FS8, I2FS16, PRMT, IADD3, LOP3, SHF, FHADD, DFMA, FADD2 };

template <int OP>
__device__ __forceinline__ void step(float &f, float &g, uint32_t &u, double &d, float c, uint32_t k) {
    if (OP == FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(f) : "f"(c));
    if (OP == FFMA2) {
        unsigned long long a, cc;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(f), "f"(g));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(a) : "l"(cc));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(f), "=f"(g) : "l"(a));
    }
    if (OP == FADD2) {
        unsigned long long a, cc;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(f), "f"(g));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
        asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(a) : "l"(cc));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(f), "=f"(g) : "l"(a));
    }
    if (OP == I2FS8) {       // u -> f -> (bits feed u) keeps a dependent chain through the converter
        asm volatile("{ .reg .b16 lo, hi; .reg .s8 dummy; .reg .b32 t; bfe.s32 t, %1, 8, 8; cvt.rn.f32.s32 %0, t; }"
                     : "=f"(f) : "r"(u));
        u ^= __float_as_uint(f);
    }
    if (OP == I2FS16) {
        asm volatile("{ .reg .b32 t; bfe.s32 t, %1, 16, 16; cvt.rn.f32.s32 %0, t; }" : "=f"(f) : "r"(u));
        u ^= __float_as_uint(f);
    }
    if (OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x5140;" : "+r"(u) : "r"(k));
    if (OP == IADD3) asm volatile("add.u32 %0, %0, %1;" : "+r"(u) : "r"(k));
    if (OP == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(k), "r"(k + 1));
    if (OP == SHF) asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(u) : "r"(k));
    if (OP == FHADD) {
        unsigned short h = (unsigned short)u;
        asm volatile("add.rn.f32.f16 %0, %1, %0;" : "+f"(f) : "h"(h));
    }
    if (OP == DFMA) asm volatile("fma.rn.f64 %0, %0, %1, %1;" : "+d"(d) : "d"((double)c));
}

template <int A, int NA, int B, int NB>
__global__ void probe(float *out, long long *cycles, float seed, uint32_t kk) {
    float fa[8], ga[8], fb[8], gb[8];
    uint32_t ua[8], ub[8];
    double da[8], db[8];
    for (int i = 0; i < 8; i++) {
        fa[i] = seed + i + threadIdx.x; fb[i] = seed * 2 + i; ga[i] = fa[i] * 0.5f; gb[i] = fb[i] * 0.25f;
        ua[i] = threadIdx.x * 2654435761u + i; ub[i] = threadIdx.x * 40503u + i;
        da[i] = fa[i]; db[i] = fb[i];
    }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < 1024; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < (NA > NB ? NA : NB); i++) {
                if (i < NA) step<A>(fa[(r * NA + i) & 7], ga[(r * NA + i) & 7], ua[(r * NA + i) & 7], da[(r * NA + i) & 7], seed, kk);
                if (i < NB) step<B>(fb[(r * NB + i) & 7], gb[(r * NB + i) & 7], ub[(r * NB + i) & 7], db[(r * NB + i) & 7], seed, kk);
            }
        }
    }
    const long long t1 = clock64();
    float sacc = 0; uint32_t v = 0; double dd = 0;
    for (int i = 0; i < 8; i++) { sacc += fa[i] + fb[i] + ga[i] + gb[i]; v ^= ua[i] ^ ub[i]; dd += da[i] + db[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = sacc + (float)v + (float)dd;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int A, int NA, int B, int NB>
void run_probe(const char *nm) {
    const int warps = 16;
    float *out; long long *cyc;
    cudaMalloc(&out, sizeof(float) * 148 * warps * 32);
    cudaMalloc(&cyc, sizeof(long long) * 148);
    for (int rep = 0; rep < 2; rep++) probe<A, NA, B, NB><<<148, warps * 32>>>(out, cyc, 1.0001f, 0x01020304u);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += (double)h[i]; avg /= 148;
    const double groups = 1024.0 * 4 * (warps / 4);
    printf("  %-30s %6.2f cycles per group per SMSP (A x%d, B x%d)\n", nm, avg / groups, NA, NB);
    cudaFree(out); cudaFree(cyc);
}

inline void microbench() {
    printf("pipe probes:\n");
    run_probe<FFMA, 1, NONE, 0>("FFMA");
    run_probe<FFMA2, 1, NONE, 0>("FFMA2");
    run_probe<I2FS8, 1, NONE, 0>("I2F.S8(+LOP)");
    run_probe<I2FS16, 1, NONE, 0>("I2F.S16(+LOP)");
    run_probe<LOP3, 1, NONE, 0>("LOP3");
    run_probe<FFMA2, 1, I2FS8, 1>("FFMA2 + I2F.S8(+LOP)");
    run_probe<FFMA2, 2, I2FS8, 1>("2 FFMA2 + I2F.S8(+LOP)");
    run_probe<FFMA2, 4, I2FS8, 1>("4 FFMA2 + I2F.S8(+LOP)");
    run_probe<FFMA2, 2, LOP3, 1>("2 FFMA2 + LOP3");
    run_probe<FFMA2, 1, LOP3, 1>("FFMA2 + LOP3");
    run_probe<FFMA2, 1, PRMT, 1>("FFMA2 + PRMT");
    run_probe<FFMA2, 1, SHF, 1>("FFMA2 + SHF");
    run_probe<FFMA2, 1, FHADD, 1>("FFMA2 + FHADD");
    run_probe<FFMA2, 1, FADD2, 1>("FFMA2 + FADD2");
    run_probe<DFMA, 1, NONE, 0>("DFMA");
    run_probe<FFMA2, 1, DFMA, 1>("FFMA2 + DFMA");
    run_probe<FFMA2, 2, DFMA, 1>("2 FFMA2 + DFMA");
    run_probe<FFMA, 2, DFMA, 1>("2 FFMA + DFMA");
    run_probe<FFMA2, 4, FFMA, 1>("4 FFMA2 + 1 FFMA");
    run_probe<FFMA2, 2, FFMA, 1>("2 FFMA2 + 1 FFMA");
    run_probe<FFMA2, 1, FFMA, 1>("1 FFMA2 + 1 FFMA");
    run_probe<FFMA2, 1, FFMA, 2>("1 FFMA2 + 2 FFMA");
    run_probe<FFMA2, 4, FHADD, 1>("4 FFMA2 + 1 FHADD");
    run_probe<FFMA2, 4, SHF, 1>("4 FFMA2 + 1 SHF");
    run_probe<FFMA2, 4, PRMT, 2>("4 FFMA2 + 2 PRMT");
    run_probe<FFMA2, 4, PRMT, 4>("4 FFMA2 + 4 PRMT");
    run_probe<IADD3, 1, NONE, 0>("IADD");
    run_probe<FFMA2, 1, IADD3, 1>("FFMA2 + IADD");
    run_probe<FFMA2, 1, IADD3, 2>("FFMA2 + 2 IADD");
}

}  // namespace lab
