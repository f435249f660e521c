// davis_device.cuh -- shared device-side definitions of the demodulation path.
//
// The arithmetic follows rtldavis: LUT (src/rtldavis/dsp.py:26), rotate_fs4
// (dsp.py:42-49), fir9 (dsp.py:52-73), discriminate (dsp.py:76-90), quantize
// (dsp.py:93-98).  Two evaluations exist:
//   * the fp32 streaming form used by the demodulation kernel (sign decisions
//     only; scale factors that cannot change a sign are dropped), and
//   * the fp64 form below, written in the reference's own operation order, used
//     wherever a result leaves the library (confirmed preambles, packet bits,
//     RSSI/SNR, discriminator means, debug taps).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace davis {

constexpr int SPS = 14;          // samples per symbol
constexpr int PRE_SYMS = 16;     // sync symbols
constexpr int PKT_SYMS = 80;     // sliced symbols
constexpr int PKT_BYTES = 10;
constexpr int HALO = 16;         // extra history samples kept in front of the carried blocks
constexpr int TILE = 32 * SPS;   // samples decided by one warp pass: 32 symbols
constexpr int PLANE_WORDS = 16;  // words stored per tile (14 phase planes + 2 pad)

// fir9 taps, src/rtldavis/dsp.py:56-69
#define DAVIS_C0 0.017682261285
#define DAVIS_C1 0.048171339939
#define DAVIS_C2 0.122424706672
#define DAVIS_C3 0.197408519126
#define DAVIS_C4 0.228626345955

// Where the samples of one call live.  The call's virtual sample sequence V of a
// stream is  carry[0 .. n_carry)  followed by  input[0 .. n_new);  V index v maps
// to global sample g = v - n_carry + blocks_done * block.  Samples with g < 0 are
// the zero history a fresh Demodulator starts from (dsp.py:129-134).
struct View {
    const uint8_t *carry;     // processed part of the carry buffer, per stream
    const uint8_t *input;     // new blocks, per stream (device memory)
    size_t carry_stride;      // bytes between streams in carry
    size_t input_stride;      // bytes between streams in input
    const int64_t *blocks_done;  // per stream: blocks consumed before this call
    int n_carry;              // samples: (W-1)*block + HALO
    int n_new;                // samples: n_blocks * block
    int block;                // block_size
    int window_blocks;        // W = buffer_length / block_size
    int stream0;              // first stream of this pipeline stage (for records)
    const double *lut;        // device copy of the float64 byte->sample table (256 entries)
};

// copy the table into shared memory (call from every thread of the block)
__device__ __forceinline__ void stage_lut(double *sm_lut, const double *lut) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sm_lut[i] = lut[i];
    __syncthreads();
}

__device__ __forceinline__ const uint8_t *sample_ptr(const View &vw, int s, int v) {
    return v < vw.n_carry ? vw.carry + (size_t)s * vw.carry_stride + 2 * (size_t)v
                          : vw.input + (size_t)s * vw.input_stride + 2 * (size_t)(v - vw.n_carry);
}

__device__ __forceinline__ int64_t global_of(const View &vw, int s, int v) {
    return (int64_t)v - vw.n_carry + vw.blocks_done[s] * (int64_t)vw.block;
}

// ---------------------------------------------------------------------------
// fp64, reference order.  No FMA contraction anywhere (explicit _rn intrinsics).
// ---------------------------------------------------------------------------
struct Cplx { double re, im; };

// rotated sample r[g] = lut(u) * j^(g mod 4); zero before the stream began
// `lut` is the reference's 256-entry table (u - 127.4) / 127.6 (dsp.py:26), built
// once on the host in float64 and staged in shared memory by every kernel that
// evaluates the exact form (a divide per sample would dominate those kernels).
__device__ __forceinline__ Cplx exact_rotated(const double *lut, const View &vw, int s, int v,
                                              int64_t g) {
    Cplx r{0.0, 0.0};
    if (g < 0 || v < 0) return r;
    const uint8_t *p = sample_ptr(vw, s, v);
    const uchar2 u = *reinterpret_cast<const uchar2 *>(p);
    double xr = lut[u.x];
    double xi = lut[u.y];
    switch ((int)(g & 3)) {
    case 0: r.re = xr;  r.im = xi;  break;
    case 1: r.re = -xi; r.im = xr;  break;
    case 2: r.re = -xr; r.im = -xi; break;
    default: r.re = xi; r.im = -xr; break;
    }
    return r;
}

// F[g] = sum_m taps[m] * r[g-9+m], summed m = 0..8
__device__ __forceinline__ Cplx exact_filtered(const double *lut, const View &vw, int s, int v,
                                               int64_t g) {
    const double taps[9] = {DAVIS_C0, DAVIS_C1, DAVIS_C2, DAVIS_C3, DAVIS_C4,
                            DAVIS_C3, DAVIS_C2, DAVIS_C1, DAVIS_C0};
    Cplx f{0.0, 0.0};
#pragma unroll
    for (int m = 0; m < 9; m++) {
        Cplx r = exact_rotated(lut, vw, s, v - 9 + m, g - 9 + m);
        f.re = __dadd_rn(f.re, __dmul_rn(taps[m], r.re));
        f.im = __dadd_rn(f.im, __dmul_rn(taps[m], r.im));
    }
    return f;
}

// D[g] = (Im a * Re b - Re a * Im b) / (|a|^2 + 1e-10), a = F[g-1], b = F[g]
__device__ __forceinline__ double exact_disc_from(const Cplx &a, const Cplx &b) {
    double num = __dsub_rn(__dmul_rn(a.im, b.re), __dmul_rn(a.re, b.im));
    double den = __dadd_rn(__dadd_rn(__dmul_rn(a.re, a.re), __dmul_rn(a.im, a.im)), 1e-10);
    return __ddiv_rn(num, den);
}

__device__ __forceinline__ double exact_disc(const double *lut, const View &vw, int s, int v) {
    int64_t g = global_of(vw, s, v);
    if (g < 0) return 0.0;                       // untouched zero buffer
    Cplx a = exact_filtered(lut, vw, s, v - 1, g - 1);
    Cplx b = exact_filtered(lut, vw, s, v, g);
    return exact_disc_from(a, b);
}

// quantize: IEEE sign bit
__device__ __forceinline__ uint32_t exact_bit(const double *lut, const View &vw, int s, int v) {
    return (uint32_t)((unsigned long long)__double_as_longlong(exact_disc(lut, vw, s, v)) >> 63);
}

}  // namespace davis
