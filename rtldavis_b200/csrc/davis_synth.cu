// davis_synth.cu -- device-side generator of synthetic Davis ISS captures.
//
// Not part of the demodulation path: it only produces the benchmark/test INPUT
// (uint8 interleaved I/Q, RTL-SDR format) at sizes numpy cannot generate in
// reasonable time (1,024+ streams x 10 s).  Signal model = SURVEY.md 8(d) /
// rtldavis_b200/synth.py: 2-FSK, Gaussian BT 0.5, 14 samples/symbol, +-4.8 kHz,
// carrier at -Fs/4 (+ per-stream offset), sync AA AA AA AA CB 89, message bytes
// LSB-first with CRC-16/0x1021/init 0, AWGN at a stated SNR, quantised with the
// inverse of the reference LUT (src/rtldavis/dsp.py:26).
// Everything is a pure function of (seed, stream, packet, sample), so the host
// can recompute the embedded messages (davis_synth_truth) without a copy.
#include <math.h>
#include <stdint.h>

#include <cuda_runtime.h>

#include "../../include/davis_b200.h"

namespace {

constexpr int SPS = 14;
constexpr int BURST_BITS = 112;
constexpr double FS = 268800.0;
constexpr double DEV_HZ = 4800.0;
constexpr float AMP = 0.6f;
constexpr int GSPAN = 4 * SPS;            // Gaussian taps - 1

__host__ __device__ inline uint64_t mix(uint64_t x) {           // splitmix64 finaliser
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__host__ __device__ inline uint64_t key(uint64_t seed, uint64_t stream, uint64_t a, uint64_t b) {
    return mix(mix(mix(seed ^ 0xD1B54A32D192ED03ull) + stream) + (a << 40) + b);
}
__host__ __device__ inline double unit(uint64_t h) { return (double)(h >> 11) * (1.0 / 9007199254740992.0); }

struct Burst {
    long long start;
    uint8_t msg[8];
    double phase0;
};

__host__ __device__ inline uint16_t crc16_bits(const uint8_t *p, int n) {
    uint32_t reg = 0;
    for (int i = 0; i < n; i++) {
        reg ^= (uint32_t)p[i] << 8;
        for (int b = 0; b < 8; b++) reg = (reg & 0x8000u) ? ((reg << 1) ^ 0x1021u) : (reg << 1);
        reg &= 0xFFFFu;
    }
    return (uint16_t)reg;
}

__host__ __device__ inline Burst burst_of(uint64_t seed, uint64_t stream, int k, size_t n_samples,
                                          int n_packets) {
    Burst b;
    const long long gap = (long long)(n_samples / (size_t)n_packets);
    const uint64_t h = key(seed, stream, 1, (uint64_t)k);
    b.start = (long long)k * gap + gap / 4 + (long long)(h % SPS);
    const uint64_t h2 = key(seed, stream, 2, (uint64_t)k);
    const int sensor = 2 + (int)((h >> 8) % 9);
    b.msg[0] = (uint8_t)((sensor << 4) | (int)(stream & 7));
    for (int i = 0; i < 5; i++) b.msg[1 + i] = (uint8_t)(h2 >> (8 * i));
    const uint16_t c = crc16_bits(b.msg, 6);
    b.msg[6] = (uint8_t)(c >> 8);
    b.msg[7] = (uint8_t)(c & 0xFF);
    b.phase0 = 6.283185307179586 * unit(key(seed, stream, 3, (uint64_t)k));
    return b;
}

__host__ __device__ inline int air_bit(const Burst &b, int i) {      // i in [0, 112)
    const uint8_t sync[6] = {0xAA, 0xAA, 0xAA, 0xAA, 0xCB, 0x89};
    if (i < 48) return (sync[i >> 3] >> (7 - (i & 7))) & 1;
    const int j = i - 48;
    return (b.msg[j >> 3] >> (j & 7)) & 1;                          // LSB first on air
}

__device__ inline float2 noise_at(uint64_t seed, uint64_t stream, uint64_t n, float sigma) {
    const uint64_t h = key(seed, stream, 4, n);
    const float u1 = ((float)(uint32_t)(h >> 32) + 1.0f) * (1.0f / 4294967296.0f);
    const float u2 = (float)(uint32_t)h * (1.0f / 4294967296.0f);
    const float r = sigma * sqrtf(-2.0f * __logf(u1));
    float s, c;
    __sincosf(6.2831853f * u2, &s, &c);
    return make_float2(r * c, r * s);
}

__device__ inline uint8_t quant(float x) {
    return (uint8_t)fminf(fmaxf(rintf(x * 127.6f + 127.4f), 0.0f), 255.0f);
}

__global__ void k_noise(uint8_t *out, size_t stride, size_t n_samples, uint64_t seed, float sigma,
                        size_t n_streams) {
    const size_t per = (n_samples + 7) / 8;
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= per * n_streams) return;
    const size_t s = t / per, c = t - s * per;
    uint8_t bytes[16];
    for (int i = 0; i < 8; i++) {
        const float2 z = noise_at(seed, s, c * 8 + i, sigma);
        bytes[2 * i] = quant(z.x);
        bytes[2 * i + 1] = quant(z.y);
    }
    uint8_t *dst = out + s * stride + c * 16;
    if (c * 8 + 8 <= n_samples && ((uintptr_t)dst & 15) == 0)
        *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<uint4 *>(bytes);
    else
        for (size_t i = 0; i < 16 && c * 16 + i < 2 * n_samples; i++) dst[i] = bytes[i];
}

__global__ void k_bursts(uint8_t *out, size_t stride, size_t n_samples, uint64_t seed, float sigma,
                         size_t n_streams, int n_packets, float df_range, float ppm_range) {
    __shared__ float gauss[GSPAN + 1];
    if (threadIdx.x <= GSPAN) {
        const float sg = 0.8325546f / (6.2831853f * 0.5f);           // sqrt(ln 2) / (2 pi BT)
        const float t = ((float)threadIdx.x - GSPAN / 2) / SPS;
        gauss[threadIdx.x] = __expf(-0.5f * (t / sg) * (t / sg));
    }
    __syncthreads();
    float norm = 0.f;
    for (int i = 0; i <= GSPAN; i++) norm += gauss[i];

    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_streams * (size_t)n_packets) return;
    const size_t s = t / n_packets;
    const int k = (int)(t - s * n_packets);
    const Burst b = burst_of(seed, s, k, n_samples, n_packets);
    const double df = df_range * (2.0 * unit(key(seed, s, 5, 0)) - 1.0);
    const double ppm = ppm_range * (2.0 * unit(key(seed, s, 6, 0)) - 1.0);
    const double sps = SPS * (1.0 + ppm * 1e-6);
    const int len = (int)ceil(BURST_BITS * sps);
    double phase = b.phase0;
    for (int n = 0; n < len; n++) {
        float shaped = 0.f;
        for (int j = 0; j <= GSPAN; j++) {
            int m = n + j - GSPAN / 2;
            m = m < 0 ? 0 : (m >= len ? len - 1 : m);
            int bit = (int)((double)m / sps);
            bit = bit >= BURST_BITS ? BURST_BITS - 1 : bit;
            shaped += gauss[j] * (air_bit(b, bit) ? 1.f : -1.f);
        }
        shaped /= norm;
        phase += 6.283185307179586 * (DEV_HZ * shaped + df - FS / 4.0) / FS;
        if (phase > 3.141592653589793) phase -= 6.283185307179586 * floor((phase + 3.141592653589793) / 6.283185307179586);
        if (phase < -3.141592653589793) phase += 6.283185307179586 * floor((3.141592653589793 - phase) / 6.283185307179586);
        const long long g = b.start + n;
        if (g < 0 || g >= (long long)n_samples) continue;
        float sn, cs;
        sincosf((float)phase, &sn, &cs);
        const float2 z = noise_at(seed, s, (uint64_t)g, sigma);
        out[s * stride + 2 * g] = quant(AMP * cs + z.x);
        out[s * stride + 2 * g + 1] = quant(AMP * sn + z.y);
    }
}

}  // namespace

extern "C" {

// Fill d_out (device memory, n_streams rows of 2*n_samples bytes at `stride`) with
// seeded captures holding n_packets bursts each.  df_range_hz / ppm_range: per-
// stream carrier and sample-clock offsets drawn uniformly in +-range.
DAVIS_API int davis_synth_fill(uint8_t *d_out, size_t n_streams, size_t stride, size_t n_samples,
                               uint64_t seed, int n_packets, float snr_db, float df_range_hz,
                               float ppm_range, int device, void *cuda_stream) {
    if (!d_out || !n_streams || !n_samples || n_packets < 0) return DAVIS_E_INVAL;
    if (cudaSetDevice(device) != cudaSuccess) return DAVIS_E_CUDA;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const float sigma = AMP / sqrtf(2.0f) * powf(10.0f, -snr_db / 20.0f);
    const size_t per = (n_samples + 7) / 8;
    const size_t threads = per * n_streams;
    k_noise<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(d_out, stride, n_samples, seed, sigma,
                                                              n_streams);
    if (n_packets > 0) {
        const size_t nb = n_streams * (size_t)n_packets;
        k_bursts<<<(unsigned)((nb + 63) / 64), 64, 0, st>>>(d_out, stride, n_samples, seed, sigma,
                                                            n_streams, n_packets, df_range_hz,
                                                            ppm_range);
    }
    return cudaGetLastError() == cudaSuccess ? DAVIS_OK : DAVIS_E_CUDA;
}

// Host-side recomputation of what davis_synth_fill embedded: msgs[n_streams][n_packets][8]
// (host bit order) and starts[n_streams][n_packets] (first sample of the burst).
DAVIS_API int davis_synth_truth(size_t n_streams, size_t n_samples, uint64_t seed, int n_packets,
                                float /*snr_db*/, uint8_t *msgs, int64_t *starts) {
    if (!msgs || !starts || n_packets < 1) return DAVIS_E_INVAL;
    for (size_t s = 0; s < n_streams; s++)
        for (int k = 0; k < n_packets; k++) {
            const Burst b = burst_of(seed, s, k, n_samples, n_packets);
            for (int i = 0; i < 8; i++) msgs[(s * n_packets + k) * 8 + i] = b.msg[i];
            starts[s * n_packets + k] = b.start;
        }
    return DAVIS_OK;
}

}  // extern "C"
