// davis_demod.cu -- the streaming demodulation kernel (uint8 IQ -> decision-bit planes).
//
// Replaces, for a whole batch of streams and blocks at once, the per-block chain
// ByteToCmplxLUT.execute -> rotate_fs4 -> fir9 -> discriminate -> quantize of
// Demodulator.demodulate (src/rtldavis/dsp.py:139-166; Go dsp/dsp.go:274-292).
//
// Work decomposition
//   * a warp pass decides one TILE = 32 symbols = 448 consecutive samples;
//     lane l owns symbol l of the tile, i.e. 14 consecutive samples, and keeps a
//     sliding window of rotated samples in registers;
//   * for sample phase p (0..13) one __ballot_sync over the lanes' sign tests
//     yields a 32-symbol word of "phase plane" p -- exactly the strided view
//     quantized[p::14] that Demodulator._search walks (dsp.py:175-176);
//   * a CTA of DEMOD_WARPS warps stages its contiguous 7 KiB of input (+32 B of
//     history) in shared memory with 16-byte loads; lane stride is 28 B = 7
//     words, so the per-lane 32-bit shared loads are bank-conflict free.
//
// Arithmetic (fp32, signs only)
//   x = (u - 127.4)/127.6 and the division by |F|^2 + eps are positive scalings
//   and are dropped; bytes are centred exactly (u - 127) while unpacking and the
//   remaining -0.4 of the LUT offset survives the filter as a period-4 constant
//   (0.4 * (2c0 - 2c2 + c4) * (1+j) * j^q) folded into the first FMA;
//   the rotation j^i is applied while unpacking bytes (which byte lands in which
//   component, and with which sign, is a PRMT selector); the rotation phase is
//   taken relative to the lane's window because a common unit factor cancels in
//   Im(conj(F[n-1]) F[n]).  Bytes become floats without I2F: PRMT builds the
//   bit pattern of +-(2^23 + u) and one packed add of -+(2^23 + 127) leaves
//   +-(u - 127) exactly.
//   Every decision that reaches a caller is re-derived in fp64 by the confirm
//   kernel, so an fp32 sign that disagrees with the reference's float64 (only
//   possible when the cross product is ~1e-7 of |F|^2) cannot change a packet.
#include "davis_kernels.cuh"

namespace davis {

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }

template <bool PACKED>
__device__ __forceinline__ float2 cadd(float2 a, float2 b) {
    if constexpr (PACKED) return __fadd2_rn(a, b);
    else return f2(a.x + b.x, a.y + b.y);
}
template <bool PACKED>
__device__ __forceinline__ float2 cfma(float2 c, float2 a, float2 acc) {
    if constexpr (PACKED) return __ffma2_rn(c, a, acc);
    else return f2(fmaf(c.x, a.x, acc.x), fmaf(c.y, a.y, acc.y));
}

// Rotated integer-valued sample rho_i = j^(i mod 4) * (I + jQ) of local index i,
// taken from the 32-bit word holding samples (2k, 2k+1).
template <bool PACKED, int I>
__device__ __forceinline__ float2 unpack_rotated(uint32_t w) {
    constexpr uint32_t POOL = 0x00CB4B00u;  // bytes 4..7 of the PRMT pool: 00 4B CB 00
    constexpr int bi = (I & 1) ? 2 : 0;     // byte of I
    constexpr int bq = bi + 1;              // byte of Q
    constexpr int ph = I & 3;
    // component sources and signs per phase: 0:(+I,+Q) 1:(-Q,+I) 2:(-I,-Q) 3:(+Q,-I)
    constexpr int src_re = (ph == 0 || ph == 2) ? bi : bq;
    constexpr int src_im = (ph == 0 || ph == 2) ? bq : bi;
    constexpr bool neg_re = (ph == 1 || ph == 2);
    constexpr bool neg_im = (ph == 2 || ph == 3);
    constexpr uint32_t sel_re = src_re | 0x440u | ((neg_re ? 6u : 5u) << 12);
    constexpr uint32_t sel_im = src_im | 0x440u | ((neg_im ? 6u : 5u) << 12);
    float2 raw = f2(__uint_as_float(__byte_perm(w, POOL, sel_re)),
                    __uint_as_float(__byte_perm(w, POOL, sel_im)));
    constexpr float M = 8388608.0f + 127.0f;   // leaves +-(u - 127): keeps the FIR free of the DC pedestal
    return cadd<PACKED>(raw, f2(neg_re ? M : -M, neg_im ? M : -M));
}

template <bool PACKED, int I>
struct Unpack {
    static __device__ __forceinline__ void run(const uint32_t (&w)[12], float2 (&r)[23]) {
        r[I] = unpack_rotated<PACKED, I>(w[I >> 1]);
        Unpack<PACKED, I + 1>::run(w, r);
    }
};
template <bool PACKED>
struct Unpack<PACKED, 23> {
    static __device__ __forceinline__ void run(const uint32_t (&)[12], float2 (&)[23]) {}
};

template <int WARPS, bool PACKED, bool WRITE_DISC>
__global__ void __launch_bounds__(WARPS * 32)
k_demod(View vw, uint32_t *__restrict__ planes, int tiles_alloc, int groups,
        float *__restrict__ disc_out, size_t disc_stride) {
    constexpr int SM_BYTES = WARPS * TILE * 2 + 2 * HALO;
    constexpr int CHUNKS = SM_BYTES / 16;
    __shared__ uint4 sm[CHUNKS];

    const int s = blockIdx.x / groups;
    const int j0 = (blockIdx.x - s * groups) * WARPS;
    const long long byte0 = 2LL * ((long long)TILE * j0 - HALO);       // V byte offset of sm[0]
    const long long carry_bytes = 2LL * vw.n_carry;
    const long long total_bytes = carry_bytes + 2LL * vw.n_new;
    const uint8_t *carry = vw.carry + (size_t)s * vw.carry_stride;
    const uint8_t *input = vw.input + (size_t)s * vw.input_stride;

    for (int c = threadIdx.x; c < CHUNKS; c += WARPS * 32) {
        long long off = byte0 + 16LL * c;
        uint4 val = make_uint4(0u, 0u, 0u, 0u);
        if (off >= 0 && off < total_bytes) {
            const uint8_t *src = off < carry_bytes ? carry + off : input + (off - carry_bytes);
            val = __ldg(reinterpret_cast<const uint4 *>(src));
        }
        sm[c] = val;
    }
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t *lw = reinterpret_cast<const uint32_t *>(sm) + warp * (TILE / 2) + lane * 7 + 3;
    uint32_t w[12];
#pragma unroll
    for (int k = 0; k < 12; k++) w[k] = lw[k];

    float2 r[23];
    Unpack<PACKED, 0>::run(w, r);

    const float c0 = (float)DAVIS_C0, c1 = (float)DAVIS_C1, c2 = (float)DAVIS_C2,
                c3 = (float)DAVIS_C3, c4 = (float)DAVIS_C4;
    const float K = (float)(0.4 * (2.0 * DAVIS_C0 - 2.0 * DAVIS_C2 + DAVIS_C4));

    float2 F[15];
#pragma unroll
    for (int q = 0; q < 15; q++) {
        // -0.4 * S * (1+j) * j^q  (samples are u - 127, the LUT subtracts 127.4)
        const float2 off = (q & 3) == 0 ? f2(-K, -K) : (q & 3) == 1 ? f2(K, -K)
                         : (q & 3) == 2 ? f2(K, K) : f2(-K, K);
        float2 acc = cfma<PACKED>(f2(c4, c4), r[q + 4], off);
        acc = cfma<PACKED>(f2(c0, c0), cadd<PACKED>(r[q], r[q + 8]), acc);
        acc = cfma<PACKED>(f2(c1, c1), cadd<PACKED>(r[q + 1], r[q + 7]), acc);
        acc = cfma<PACKED>(f2(c2, c2), cadd<PACKED>(r[q + 2], r[q + 6]), acc);
        acc = cfma<PACKED>(f2(c3, c3), cadd<PACKED>(r[q + 3], r[q + 5]), acc);
        F[q] = acc;
    }

    uint32_t mine = 0;
    const int tile = j0 + warp;
#pragma unroll
    for (int p = 0; p < SPS; p++) {
        const float2 a = F[p], b = F[p + 1];
        const float num = fmaf(a.y, b.x, -(a.x * b.y));
        const uint32_t word = __ballot_sync(0xFFFFFFFFu, num < 0.0f);
        if (lane == p) mine = word;
        if constexpr (WRITE_DISC) {
            const long long v = (long long)TILE * tile + SPS * lane + p;
            if (v < vw.n_carry + vw.n_new) {
                const float den = fmaf(a.x, a.x, a.y * a.y) + 1e-10f * 127.6f * 127.6f;
                disc_out[(size_t)s * disc_stride + v] = num / den;
            }
        }
    }
    if (lane < PLANE_WORDS)
        planes[((size_t)s * tiles_alloc + tile) * PLANE_WORDS + lane] = mine;
}

void launch_demod(const View &vw, uint32_t *planes, int n_streams, int tiles_alloc, int variant,
                  float *disc_out, size_t disc_stride, cudaStream_t st) {
    const int groups = tiles_alloc / DEMOD_WARPS;
    const unsigned grid = (unsigned)((size_t)n_streams * groups);
    if (grid == 0) return;
    if (disc_out)
        k_demod<DEMOD_WARPS, true, true><<<grid, DEMOD_WARPS * 32, 0, st>>>(
            vw, planes, tiles_alloc, groups, disc_out, disc_stride);
    else if (variant == 1)
        k_demod<DEMOD_WARPS, false, false><<<grid, DEMOD_WARPS * 32, 0, st>>>(
            vw, planes, tiles_alloc, groups, nullptr, 0);
    else
        k_demod<DEMOD_WARPS, true, false><<<grid, DEMOD_WARPS * 32, 0, st>>>(
            vw, planes, tiles_alloc, groups, nullptr, 0);
}

}  // namespace davis
