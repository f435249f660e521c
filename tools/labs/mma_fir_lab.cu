// mma_fir_lab.cu -- lab for the tensor-core demodulation kernel (development tool).
//   1. small cases: the filter outputs z[n] of mma_fir_kernel_pitch64.cu against a float64 evaluation
//      on the host, and its decision bitmap against the SIMT kernel's (davis_demod.cu);
//   2. full size (1024 streams x 328 blocks): time and bit difference against the SIMT kernel, the
//      drain-only form, and the cycles every warp role spends per phase.
//   (Commit 199c4da holds the version of this file that drove mma_fir_kernel_pitch16.cu, with the
//   2-limb variant and the check of which descriptor stride is which.)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -DDAVIS_MMA_LAB \
//        -DDAVIS_MBAR_WATCHDOG -o tools/labs/mma_fir_lab tools/labs/mma_fir_lab.cu \
//        rtldavis_b200/csrc/davis_demod.cu tools/labs/mma_fir_kernel_pitch64.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../rtldavis_b200/csrc/davis_kernels.cuh"

using namespace davis;

namespace davis {
void launch_demod_mma_lab(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                          int drain_only, float2 *z_dump, unsigned long long *prof, cudaStream_t st);
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void fill_random(uint8_t *p, size_t n, uint32_t seed) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n / 4; i += stride) {
        uint32_t x = (uint32_t)i * 2654435761u + seed;
        x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
        uint32_t out = 0;
        for (int b = 0; b < 4; b++) {
            int v = 127 + (int)((x >> (8 * b)) & 0x1F) - 16 + (int)(40.f * __sinf(0.05f * (float)(4 * i + b)));
            out |= (uint32_t)(v & 0xFF) << (8 * b);
        }
        reinterpret_cast<uint32_t *>(p)[i] = out;
    }
}

struct Case {
    int n_streams, n_blocks, B = 8192, W = 2;
    int n_carry, n_new, v_len, tiles, words_alloc;
    size_t carry_stride, input_stride;
    uint8_t *carry = nullptr, *input = nullptr;
    int64_t *done = nullptr;
    View vw{};
    Case(int ns, int nb) : n_streams(ns), n_blocks(nb) {
        n_carry = (W - 1) * B + HALO; n_new = n_blocks * B; v_len = n_carry + n_new;
        tiles = (v_len + TILE_SAMPLES - 1) / TILE_SAMPLES;
        words_alloc = tiles * TILE_WORDS + SEARCH_AHEAD;
        carry_stride = 2 * (size_t)n_carry; input_stride = 2 * (size_t)n_new;
        CK(cudaMalloc(&carry, carry_stride * n_streams));
        CK(cudaMalloc(&input, input_stride * n_streams));
        CK(cudaMalloc(&done, sizeof(int64_t) * n_streams));
        CK(cudaMemset(done, 0, sizeof(int64_t) * n_streams));
        fill_random<<<148 * 8, 256>>>(carry, carry_stride * n_streams, 1u);
        fill_random<<<148 * 8, 256>>>(input, input_stride * n_streams, 77u);
        CK(cudaDeviceSynchronize());
        vw.carry = carry; vw.input = input; vw.carry_stride = carry_stride; vw.input_stride = input_stride;
        vw.blocks_done = done; vw.n_carry = n_carry; vw.n_new = n_new; vw.block = B; vw.window_blocks = W;
        vw.stream0 = 0; vw.lut = nullptr; vw.elem = 2;
    }
    ~Case() { cudaFree(carry); cudaFree(input); cudaFree(done); }
};

static const int T3[5] = {623995, 1699934, 4320285, 6966413, 8068069};
static const int T2[5] = {2448, 6669, 16949, 27330, 31652};

// z[n] = sum_m T[m] j^m ((I - 127.4) + j (Q - 127.4)) in float64
static void host_z(const uint8_t *v, int n, const int *T5, double &zr, double &zi) {
    zr = zi = 0.0;
    for (int m = 0; m < 9; m++) {
        const double t = T5[m < 5 ? m : 8 - m], xr = v[2 * (n + m)] - 127.4, xi = v[2 * (n + m) + 1] - 127.4;
        switch (m & 3) {
        case 0: zr += t * xr; zi += t * xi; break;
        case 1: zr -= t * xi; zi += t * xr; break;
        case 2: zr -= t * xr; zi -= t * xi; break;
        default: zr += t * xi; zi -= t * xr; break;
        }
    }
}

static size_t diff_words(const std::vector<uint32_t> &a, const std::vector<uint32_t> &b, const Case &c,
                         size_t *bits, int first_word) {
    size_t dw = 0; *bits = 0;
    const int full_words = c.v_len / LANE_SAMPLES;
    for (int s = 0; s < c.n_streams; s++)
        for (int w = first_word; w < full_words; w++) {
            const uint32_t x = (a[(size_t)s * c.words_alloc + w] ^ b[(size_t)s * c.words_alloc + w]) & 0x0FFFFFFFu;
            if (x) { dw++; *bits += __builtin_popcount(x); }
        }
    return dw;
}

static bool small_case(int n_streams, int n_blocks) {
    Case c(n_streams, n_blocks);
    const int limbs = 3;
    const size_t per_stream = (size_t)c.tiles * TILE_SAMPLES;
    const size_t zn = (size_t)c.n_streams * per_stream;
    float2 *zd; uint32_t *w_ref, *w_mma;
    const size_t nw = (size_t)c.words_alloc * c.n_streams;
    CK(cudaMalloc(&zd, zn * sizeof(float2))); CK(cudaMemset(zd, 0, zn * sizeof(float2)));
    CK(cudaMalloc(&w_ref, nw * 4)); CK(cudaMalloc(&w_mma, nw * 4));
    CK(cudaMemset(w_ref, 0, nw * 4)); CK(cudaMemset(w_mma, 0, nw * 4));
    launch_demod(c.vw, w_ref, c.n_streams, c.tiles, c.words_alloc, nullptr, 0, 0);
    launch_demod_mma_lab(c.vw, w_mma, c.n_streams, c.tiles, c.words_alloc, 0, zd, nullptr, 0);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("small case %d x %d: %s\n", n_streams, n_blocks, cudaGetErrorString(e)); exit(2); }
    std::vector<float2> z(zn);
    std::vector<uint8_t> V((size_t)2 * c.v_len + 64);
    std::vector<uint32_t> hr(nw), hm(nw);
    CK(cudaMemcpy(z.data(), zd, zn * sizeof(float2), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hr.data(), w_ref, nw * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hm.data(), w_mma, nw * 4, cudaMemcpyDeviceToHost));
    double worst = 0.0; size_t bad = 0, checked = 0;
    for (int s = 0; s < c.n_streams; s++) {
        CK(cudaMemcpy(V.data(), c.carry + s * c.carry_stride, c.carry_stride, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(V.data() + c.carry_stride, c.input + s * c.input_stride, c.input_stride, cudaMemcpyDeviceToHost));
        for (int n = 0; n + 9 <= c.v_len; n++) {
            double zr, zi; host_z(V.data(), n, limbs == 3 ? T3 : T2, zr, zi);
            const float2 got = z[(size_t)s * per_stream + n + 8];
            const double scale = (limbs == 3 ? 8068069.0 : 31652.0) * 128.0;
            const double err = fmax(fabs(got.x - zr), fabs(got.y - zi)) / scale;
            if (err > worst) worst = err;
            if (err > 1e-5) { if (bad < 5) printf("  z mismatch s=%d n=%d got (%g, %g) want (%g, %g)\n", s, n, got.x, got.y, zr, zi); bad++; }
            checked++;
        }
    }
    size_t bits; const size_t dw = diff_words(hr, hm, c, &bits, 1);
    printf("small case %d streams x %d blocks: z checked %zu, bad %zu, worst err %.3g of full scale; bitmap vs SIMT: %zu bits in %zu words differ\n",
           n_streams, n_blocks, checked, bad, worst, bits, dw);
    cudaFree(zd); cudaFree(w_ref); cudaFree(w_mma);
    return bad == 0;
}

int main(int argc, char **argv) {
    const int n_streams = argc > 1 ? atoi(argv[1]) : 1024;
    const int n_blocks = argc > 2 ? atoi(argv[2]) : 328;
    const int reps = argc > 3 ? atoi(argv[3]) : 5;
    // 3 x 4: one CTA per tile; 200 x 9: ranges that start inside streams (head decisions)
    if (!small_case(3, 4) || !small_case(200, 9)) return 3;

    Case c(n_streams, n_blocks);
    const size_t nw = (size_t)c.words_alloc * c.n_streams;
    uint32_t *words[2];
    for (int i = 0; i < 2; i++) CK(cudaMalloc(&words[i], nw * 4));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const double samples = (double)n_streams * c.n_new;
    int clock_khz = 0;
    CK(cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0));
    std::vector<uint32_t> ref(nw), got(nw);
    const char *names[3] = {"SIMT k_demod (davis_demod.cu)", "mma, 64-byte row pitch, 3 limbs", "mma, drain only"};
    for (int v = 0; v < 3; v++) {
        uint32_t *out = words[v == 0 ? 0 : 1];
        CK(cudaMemset(out, 0, nw * 4));
        float best = 1e30f, sum = 0.f;
        for (int r = 0; r < reps + 1; r++) {
            CK(cudaEventRecord(e0));
            if (v == 0) launch_demod(c.vw, out, n_streams, c.tiles, c.words_alloc, nullptr, 0, 0);
            else launch_demod_mma_lab(c.vw, out, n_streams, c.tiles, c.words_alloc, v == 2, nullptr, nullptr, 0);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            if (r > 0) { sum += ms; if (ms < best) best = ms; }
        }
        CK(cudaGetLastError());
        size_t bits = 0, dw = 0;
        if (v == 0) CK(cudaMemcpy(ref.data(), out, nw * 4, cudaMemcpyDeviceToHost));
        else if (v < 2) { CK(cudaMemcpy(got.data(), out, nw * 4, cudaMemcpyDeviceToHost)); dw = diff_words(ref, got, c, &bits, 1); }
        printf("%-32s best %7.3f ms  mean %7.3f ms  %7.1f GB/s  %6.2f cyc/sample/lane@%d MHz  diff_bits=%zu (words %zu) of %.0f\n",
               names[v], best, sum / reps, 2.0 * samples / best * 1e-6,
               best * 1e-3 * clock_khz * 1e3 * 148 * 128 / samples, clock_khz / 1000, bits, dw, samples);
        fflush(stdout);
        if (v > 0) {
            // where the warps of the kernel spend their cycles (one extra, instrumented launch)
            unsigned long long *prof; const size_t np = (size_t)148 * 21 * 8;
            CK(cudaMalloc(&prof, np * 8)); CK(cudaMemset(prof, 0, np * 8));
            launch_demod_mma_lab(c.vw, out, n_streams, c.tiles, c.words_alloc, v == 2, nullptr, prof, 0);
            CK(cudaDeviceSynchronize());
            std::vector<unsigned long long> hp(np);
            CK(cudaMemcpy(hp.data(), prof, np * 8, cudaMemcpyDeviceToHost));
            double epi[8] = {0}, fmt[8] = {0}, ctl[8] = {0};
            for (int b = 0; b < 148; b++)
                for (int w = 0; w < 21; w++)
                    for (int i = 0; i < 8; i++) (w < 16 ? epi : w < 20 ? fmt : ctl)[i] += (double)hp[((size_t)b * 21 + w) * 8 + i];
            const double sets = samples / 4096.0;
            printf("    cycles per set: epilogue warp: wait_full %.0f ld %.0f st %.0f arrive %.0f math %.0f barrier %.0f repack %.0f loop %.0f\n",
                   epi[0] / (16 * sets), epi[1] / (16 * sets), epi[2] / (16 * sets), epi[3] / (16 * sets), epi[4] / (16 * sets), epi[5] / (16 * sets), epi[6] / (16 * sets), epi[7] / (16 * sets));
            printf("                    formatter warp: wait_raw %.0f wait_a_free %.0f copy %.0f arrive %.0f loop %.0f | control: wait_a_ready %.0f wait_tm_empty %.0f issue %.0f wait_raw_free %.0f fetch %.0f loop %.0f\n",
                   fmt[0] / (4 * sets), fmt[1] / (4 * sets), fmt[2] / (4 * sets), fmt[3] / (4 * sets), fmt[7] / (4 * sets),
                   ctl[0] / sets, ctl[1] / sets, ctl[2] / sets, ctl[3] / sets, ctl[4] / sets, ctl[7] / sets);
            cudaFree(prof);
        }
    }
    return 0;
}
