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
// decision bitmap: word w of a stream holds samples [28w, 28w+28), bit i = sample 28w+i
constexpr int LANE_SAMPLES = 2 * SPS;                    // 28 decisions per 32-bit word
constexpr int TILE_WORDS = 1024;                         // words per demod tile (two per thread)
constexpr int TILE_SAMPLES = TILE_WORDS * LANE_SAMPLES;  // 28672
constexpr int SEARCH_AHEAD = 12;                         // words the search may read past the last tile (keeps rows 16-byte aligned)

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
    int elem;                 // bytes per sample: 2 (uint8 I,Q), 8 (complex64) or 16 (complex128)
    // davis_cfg.flavor: the Python discriminator adds 1e-10 to |a|^2 (dsp.py:88), the Go one
    // adds nothing (dsp/dsp.go:93); the parser's mean starts at index (protocol.py:304-308)
    // or 8 symbols later (protocol/protocol.go:156-158)
    double disc_eps;
    int disc_off;
};

// copy the table into shared memory (call from every thread of the block)
__device__ __forceinline__ void stage_lut(double *sm_lut, const double *lut) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sm_lut[i] = lut[i];
    __syncthreads();
}

__device__ __forceinline__ const uint8_t *sample_ptr(const View &vw, int s, int v) {
    return v < vw.n_carry ? vw.carry + (size_t)s * vw.carry_stride + vw.elem * (size_t)v
                          : vw.input + (size_t)s * vw.input_stride + vw.elem * (size_t)(v - vw.n_carry);
}

__device__ __forceinline__ int64_t global_of(const View &vw, int s, int v) {
    return (int64_t)v - vw.n_carry + vw.blocks_done[s] * (int64_t)vw.block;
}

// ---------------------------------------------------------------------------
// fp64, reference order.  No FMA contraction anywhere (explicit _rn intrinsics).
//
// The evaluators are templated on a sample source `Src` with
//     Cplx sample(lut, int v)     the unrotated sample of V index v (v >= 0)
//     int64_t g0                  global sample index of v = 0
// GlobalSrc reads the carry/input buffers directly; SpanSrc reads a span that a
// warp staged in shared memory with coalesced 16-byte loads.
// `lut` is the reference's 256-entry table (u - 127.4) / 127.6 (dsp.py:26), built
// once on the host in float64 and staged in shared memory by every kernel that
// evaluates the exact form (a divide per sample would dominate those kernels).
// ---------------------------------------------------------------------------
struct Cplx { double re, im; };

// One sample as the reference holds it in raw_samples (complex128, dsp.py:144-151):
// the LUT value of the two bytes (ByteToCmplxLUT.execute, dsp.py:28-39), or the
// complex input itself (dest[:] = input_data).  ES = bytes per sample.
template <int ES>
__device__ __forceinline__ Cplx load_sample(const double *lut, const uint8_t *p) {
    if constexpr (ES == 2) {
        const uchar2 u = *reinterpret_cast<const uchar2 *>(p);
        return Cplx{lut[u.x], lut[u.y]};
    } else if constexpr (ES == 8) {
        const float2 f = *reinterpret_cast<const float2 *>(p);
        return Cplx{(double)f.x, (double)f.y};
    } else {
        const double2 d = *reinterpret_cast<const double2 *>(p);
        return Cplx{d.x, d.y};
    }
}

template <int ES>
struct GlobalSrcT {
    static constexpr int elem_bytes = ES;
    const uint8_t *carry, *input;
    int n_carry;
    int64_t g0;
    __device__ __forceinline__ GlobalSrcT(const View &vw, int s)
        : carry(vw.carry + (size_t)s * vw.carry_stride),
          input(vw.input + (size_t)s * vw.input_stride),
          n_carry(vw.n_carry),
          g0(vw.blocks_done[s] * (int64_t)vw.block - vw.n_carry) {}
    __device__ __forceinline__ Cplx sample(const double *lut, int v) const {
        return load_sample<ES>(lut, v < n_carry ? carry + ES * (size_t)v : input + ES * (size_t)(v - n_carry));
    }
};
using GlobalSrc = GlobalSrcT<2>;

template <int ES>
struct SpanSrcT {
    static constexpr int elem_bytes = ES;
    const uint8_t *base;      // shared memory; byte 0 holds V index v0
    int v0;
    int64_t g0;
    __device__ __forceinline__ Cplx sample(const double *lut, int v) const {
        return load_sample<ES>(lut, base + ES * (v - v0));
    }
};
using SpanSrc = SpanSrcT<2>;

// A warp copies V samples [v_begin, v_begin + n) of stream s into `dst` (shared,
// 16-byte aligned, room for ES*n + 32 bytes) and returns the source describing it.
// Chunks before the start of V are zero-filled (they are only ever addressed for
// g < 0, where the evaluators substitute the zero history anyway).
template <int ES>
__device__ __forceinline__ SpanSrcT<ES> stage_span(const View &vw, int s, int v_begin, int n,
                                                   uint8_t *dst, int lane) {
    const long long b0 = ((long long)ES * v_begin) & ~15LL;    // floor to a 16-byte chunk
    const long long b1 = (long long)ES * (v_begin + n);
    const long long carry_bytes = (long long)ES * vw.n_carry, total = carry_bytes + (long long)ES * vw.n_new;
    const uint8_t *carry = vw.carry + (size_t)s * vw.carry_stride;
    const uint8_t *input = vw.input + (size_t)s * vw.input_stride;
    const int chunks = (int)((b1 - b0 + 15) >> 4);
    for (int c = lane; c < chunks; c += 32) {
        const long long off = b0 + 16LL * c;
        uint4 val = make_uint4(0u, 0u, 0u, 0u);
        if (off >= 0 && off < total)
            val = __ldg(reinterpret_cast<const uint4 *>(off < carry_bytes ? carry + off
                                                                         : input + (off - carry_bytes)));
        reinterpret_cast<uint4 *>(dst)[c] = val;
    }
    __syncwarp();
    SpanSrcT<ES> src;
    src.base = dst;
    // v0 may be fractional for ES = 16 only if b0 is not a multiple of ES: it always is
    src.v0 = (int)(b0 / ES);
    src.g0 = vw.blocks_done[s] * (int64_t)vw.block - vw.n_carry;
    return src;
}

// The same in two halves for spans of at most 32 chunks, so that the load of the next
// span can be in flight while the current one is being used: span_chunk returns
// this lane's 16 bytes, span_store puts them in shared memory.
__device__ __forceinline__ uint4 span_chunk(const View &vw, int s, int v_begin, int n, int lane) {
    const long long b0 = (2LL * v_begin) & ~15LL;
    const long long b1 = 2LL * (v_begin + n);
    const long long carry_bytes = 2LL * vw.n_carry, total = carry_bytes + 2LL * vw.n_new;
    const long long off = b0 + 16LL * lane;
    uint4 val = make_uint4(0u, 0u, 0u, 0u);
    if (off < b1 && off >= 0 && off < total) {
        const uint8_t *p = off < carry_bytes ? vw.carry + (size_t)s * vw.carry_stride + off
                                             : vw.input + (size_t)s * vw.input_stride + (off - carry_bytes);
        val = __ldg(reinterpret_cast<const uint4 *>(p));
    }
    return val;
}
__device__ __forceinline__ SpanSrc span_store(const View &vw, int s, int v_begin, uint4 val,
                                              uint8_t *dst, int lane) {
    reinterpret_cast<uint4 *>(dst)[lane] = val;
    __syncwarp();
    SpanSrc src;
    src.base = dst;
    src.v0 = (int)(((2LL * v_begin) & ~15LL) >> 1);
    src.g0 = vw.blocks_done[s] * (int64_t)vw.block - vw.n_carry;
    return src;
}

// rotated sample r[g] = lut(u) * j^(g mod 4); zero before the stream began
template <class Src>
__device__ __forceinline__ Cplx exact_rotated(const double *lut, const Src &src, int v) {
    Cplx r{0.0, 0.0};
    const int64_t g = src.g0 + v;
    if (g < 0 || v < 0) return r;
    const Cplx x = src.sample(lut, v);
    const double xr = x.re;
    const double xi = x.im;
    switch ((int)(g & 3)) {
    case 0: r.re = xr;  r.im = xi;  break;
    case 1: r.re = -xi; r.im = xr;  break;
    case 2: r.re = -xr; r.im = -xi; break;
    default: r.re = xi; r.im = -xr; break;
    }
    return r;
}

// F[g] = sum_m taps[m] * r[g-9+m], summed m = 0..8
template <class Src>
__device__ __forceinline__ Cplx exact_filtered(const double *lut, const Src &src, int v) {
    const double taps[9] = {DAVIS_C0, DAVIS_C1, DAVIS_C2, DAVIS_C3, DAVIS_C4,
                            DAVIS_C3, DAVIS_C2, DAVIS_C1, DAVIS_C0};
    Cplx f{0.0, 0.0};
#pragma unroll
    for (int m = 0; m < 9; m++) {
        const Cplx r = exact_rotated(lut, src, v - 9 + m);
        f.re = __dadd_rn(f.re, __dmul_rn(taps[m], r.re));
        f.im = __dadd_rn(f.im, __dmul_rn(taps[m], r.im));
    }
    return f;
}

// F[v_first + k] for k = 0..NF-1 from the NF + 8 samples they share (register
// sliding window); every sum runs m = 0..8 exactly as exact_filtered does.
template <int NF, class Src>
__device__ __forceinline__ void exact_filtered_run(const double *lut, const Src &src, int v_first,
                                                   Cplx (&F)[NF]) {
    const double taps[9] = {DAVIS_C0, DAVIS_C1, DAVIS_C2, DAVIS_C3, DAVIS_C4,
                            DAVIS_C3, DAVIS_C2, DAVIS_C1, DAVIS_C0};
    Cplx r[NF + 8];
#pragma unroll
    for (int i = 0; i < NF + 8; i++) r[i] = exact_rotated(lut, src, v_first - 9 + i);
#pragma unroll
    for (int k = 0; k < NF; k++) {
        Cplx f{0.0, 0.0};
#pragma unroll
        for (int m = 0; m < 9; m++) {
            f.re = __dadd_rn(f.re, __dmul_rn(taps[m], r[k + m].re));
            f.im = __dadd_rn(f.im, __dmul_rn(taps[m], r[k + m].im));
        }
        F[k] = f;
    }
}

// The same run with the rotation phase taken relative to the run's first sample
// instead of the global sample index, and no test for the zero history: no
// switch, no 64-bit index arithmetic.  Multiplying every sample by one power of j
// maps each F[k] to j^p F[k], i.e. each component to +- a component (exactly:
// rounding is symmetric), which leaves |F|^2 = re^2 + im^2 (addition commutes) and
// the cross product im*re' - re*im' of two outputs of the run unchanged bit for
// bit.  Valid only for quantities with that invariance and only when every
// sample of the run exists (global index >= 0): see filtered_run below.
template <int K>
__device__ __forceinline__ Cplx fixed_rotated(double x, double y) {
    if ((K & 3) == 0) return Cplx{x, y};
    if ((K & 3) == 1) return Cplx{-y, x};
    if ((K & 3) == 2) return Cplx{-x, -y};
    return Cplx{y, -x};
}
template <int K, int N, class Src>
struct FixedWindow {
    static __device__ __forceinline__ void run(const double *lut, const Src &src, int v, Cplx (&r)[N]) {
        const Cplx x = src.sample(lut, v + K);
        r[K] = fixed_rotated<K>(x.re, x.im);
        FixedWindow<K + 1, N, Src>::run(lut, src, v, r);
    }
};
template <int N, class Src>
struct FixedWindow<N, N, Src> {
    static __device__ __forceinline__ void run(const double *, const Src &, int, Cplx (&)[N]) {}
};

template <int NF, class Src>
__device__ __forceinline__ void fixed_filtered_run(const double *lut, const Src &src, int v_first,
                                                   Cplx (&F)[NF]) {
    const double taps[9] = {DAVIS_C0, DAVIS_C1, DAVIS_C2, DAVIS_C3, DAVIS_C4,
                            DAVIS_C3, DAVIS_C2, DAVIS_C1, DAVIS_C0};
    Cplx r[NF + 8];
    FixedWindow<0, NF + 8, Src>::run(lut, src, v_first - 9, r);
#pragma unroll
    for (int k = 0; k < NF; k++) {
        // m = 0..8 in order, each step one fused multiply-add: np.convolve hands the sum to
        // a BLAS dot product whose rounding order is unspecified, so the reference defines
        // F only to a few ulp; the fused form is inside that and halves the fp64 work.
        Cplx f{__dmul_rn(taps[0], r[k].re), __dmul_rn(taps[0], r[k].im)};
#pragma unroll
        for (int m = 1; m < 9; m++) {
            f.re = __fma_rn(taps[m], r[k + m].re, f.re);
            f.im = __fma_rn(taps[m], r[k + m].im, f.im);
        }
        F[k] = f;
    }
}

// F[v_first + k], k < NF, for rotation-invariant uses: the fixed-phase run when all
// its samples exist, the reference-phase run (zero history) otherwise.
template <int NF, class Src>
__device__ __forceinline__ void filtered_run(const double *lut, const Src &src, int v_first,
                                             Cplx (&F)[NF]) {
    // byte input only: the LUT never yields a zero, so no signed zero can make the two
    // forms differ; complex input (which may hold exact zeros) keeps the reference phase
    if (Src::elem_bytes == 2 && v_first - 9 >= 0 && src.g0 + v_first - 9 >= 0)
        fixed_filtered_run<NF>(lut, src, v_first, F);
    else
        exact_filtered_run<NF>(lut, src, v_first, F);
}

// numerator of D[g]: Im a * Re b - Re a * Im b,  a = F[g-1], b = F[g]
__device__ __forceinline__ double exact_cross(const Cplx &a, const Cplx &b) {
    return __dsub_rn(__dmul_rn(a.im, b.re), __dmul_rn(a.re, b.im));
}

// D[g] = cross / (|a|^2 + eps), eps = 1e-10 (Python) or 0 (Go: 0/0 stays a NaN)
__device__ __forceinline__ double exact_disc_from(const Cplx &a, const Cplx &b, double eps) {
    const double den = __dadd_rn(__dadd_rn(__dmul_rn(a.re, a.re), __dmul_rn(a.im, a.im)), eps);
    return __ddiv_rn(exact_cross(a, b), den);
}

// a = F[g-1], b = F[g] from one pass over the ten samples they share; each sum
// runs m = 0..8 exactly as exact_filtered does.
template <class Src>
__device__ __forceinline__ void exact_pair(const double *lut, const Src &src, int v, Cplx &a,
                                           Cplx &b) {
    const double taps[9] = {DAVIS_C0, DAVIS_C1, DAVIS_C2, DAVIS_C3, DAVIS_C4,
                            DAVIS_C3, DAVIS_C2, DAVIS_C1, DAVIS_C0};
    a = Cplx{0.0, 0.0};
    b = Cplx{0.0, 0.0};
    Cplx prev = exact_rotated(lut, src, v - 10);
#pragma unroll
    for (int m = 0; m < 9; m++) {
        const Cplx cur = exact_rotated(lut, src, v - 9 + m);
        a.re = __dadd_rn(a.re, __dmul_rn(taps[m], prev.re));
        a.im = __dadd_rn(a.im, __dmul_rn(taps[m], prev.im));
        b.re = __dadd_rn(b.re, __dmul_rn(taps[m], cur.re));
        b.im = __dadd_rn(b.im, __dmul_rn(taps[m], cur.im));
        prev = cur;
    }
}

template <class Src>
__device__ __forceinline__ double exact_disc(const double *lut, const Src &src, int v, double eps) {
    if (src.g0 + v < 0) return 0.0;              // untouched zero buffer
    Cplx a, b;
    exact_pair(lut, src, v, a, b);
    return exact_disc_from(a, b, eps);
}

// quantize: IEEE sign bit of D[g].  The denominator is positive and finite, so
// the quotient (even when it underflows to a signed zero) carries the sign bit
// of the numerator: the division is skipped.
template <class Src>
__device__ __forceinline__ uint32_t exact_bit(const double *lut, const Src &src, int v) {
    if (src.g0 + v < 0) return 0u;               // +0.0 in the zero-initialised buffer
    Cplx a, b;
    exact_pair(lut, src, v, a, b);
    return (uint32_t)((unsigned long long)__double_as_longlong(exact_cross(a, b)) >> 63);
}

}  // namespace davis
