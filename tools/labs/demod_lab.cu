// demod_lab.cu -- bench for variants of the demodulation kernel (development tool).
// Links the product kernel (davis_demod.cu) as the baseline and times experimental
// variants against it on the same random input, comparing the decision bitmaps.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo \
//        -o tools/labs/demod_lab tools/labs/demod_lab.cu rtldavis_b200/csrc/davis_demod.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../rtldavis_b200/csrc/davis_kernels.cuh"
#include "demod_lab_variants.cuh"

using namespace davis;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void fill_random(uint8_t *p, size_t n, uint32_t seed) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n / 4; i += stride) {
        uint32_t x = (uint32_t)i * 2654435761u + seed;
        x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
        // noise around mid-scale with a slow tone so that signs are not pure coin flips
        uint32_t out = 0;
        for (int b = 0; b < 4; b++) {
            int v = 127 + (int)((x >> (8 * b)) & 0x1F) - 16 + (int)(40.f * __sinf(0.05f * (float)(4 * i + b)));
            out |= (uint32_t)(v & 0xFF) << (8 * b);
        }
        reinterpret_cast<uint32_t *>(p)[i] = out;
    }
}

int main(int argc, char **argv) {
    const int n_streams = argc > 1 ? atoi(argv[1]) : 1024;
    const int n_blocks = argc > 2 ? atoi(argv[2]) : 328;
    const int reps = argc > 3 ? atoi(argv[3]) : 5;
    const int B = 8192, W = 2;
    const int n_carry = (W - 1) * B + HALO, n_new = n_blocks * B;
    const int v_len = n_carry + n_new;
    const int tiles = (v_len + TILE_SAMPLES - 1) / TILE_SAMPLES;
    const int words_alloc = tiles * TILE_WORDS + SEARCH_AHEAD;
    const size_t carry_stride = 2 * (size_t)n_carry, input_stride = 2 * (size_t)n_new;

    uint8_t *carry, *input;
    int64_t *done;
    uint32_t *words[2];
    CK(cudaMalloc(&carry, carry_stride * n_streams));
    CK(cudaMalloc(&input, input_stride * n_streams));
    CK(cudaMalloc(&done, sizeof(int64_t) * n_streams));
    CK(cudaMemset(done, 0, sizeof(int64_t) * n_streams));
    for (int i = 0; i < 2; i++) CK(cudaMalloc(&words[i], sizeof(uint32_t) * (size_t)words_alloc * n_streams));
    fill_random<<<148 * 8, 256>>>(carry, carry_stride * n_streams, 1u);
    fill_random<<<148 * 8, 256>>>(input, input_stride * n_streams, 77u);
    CK(cudaDeviceSynchronize());

    View vw{};
    vw.carry = carry; vw.input = input; vw.carry_stride = carry_stride; vw.input_stride = input_stride;
    vw.blocks_done = done; vw.n_carry = n_carry; vw.n_new = n_new; vw.block = B; vw.window_blocks = W;
    vw.stream0 = 0; vw.lut = nullptr;

    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const double samples = (double)n_streams * n_new;
    int clock_khz = 0;
    CK(cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0));

    std::vector<uint32_t> ref((size_t)words_alloc * n_streams), got((size_t)words_alloc * n_streams);
    const int n_var = lab::n_variants();
    for (int v = -1; v < n_var; v++) {
        uint32_t *out = words[v < 0 ? 0 : 1];
        CK(cudaMemset(out, 0, sizeof(uint32_t) * (size_t)words_alloc * n_streams));
        float best = 1e30f;
        for (int r = 0; r < reps + 1; r++) {
            CK(cudaEventRecord(e0));
            if (v < 0) launch_demod(vw, out, n_streams, tiles, words_alloc, nullptr, 0, 0);
            else lab::launch(v, vw, out, n_streams, tiles, words_alloc);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            if (r > 0 && ms < best) best = ms;
        }
        CK(cudaGetLastError());
        size_t diff_bits = 0, diff_words = 0;
        if (v < 0) CK(cudaMemcpy(ref.data(), out, ref.size() * 4, cudaMemcpyDeviceToHost));
        else {
            CK(cudaMemcpy(got.data(), out, got.size() * 4, cudaMemcpyDeviceToHost));
            // compare words whose 28 samples all exist (skip the first word of a stream: zero history)
            const int full_words = v_len / LANE_SAMPLES;
            for (int s = 0; s < n_streams; s++)
                for (int w = 1; w < full_words; w++) {
                    const uint32_t x = (ref[(size_t)s * words_alloc + w] ^ got[(size_t)s * words_alloc + w]) & 0x0FFFFFFFu;
                    if (x) { diff_words++; diff_bits += __builtin_popcount(x); }
                }
        }
        printf("%-34s %8.3f ms  %7.1f GB/s  %6.2f cyc/sample/lane@%d MHz  diff_bits=%zu (words %zu)\n",
               v < 0 ? "baseline (davis_demod.cu)" : lab::name(v), best, 2.0 * samples / best * 1e-6,
               best * 1e-3 * clock_khz * 1e3 * 148 * 128 / samples, clock_khz / 1000, diff_bits, diff_words);
        fflush(stdout);
    }
    lab::microbench();
    return 0;
}
