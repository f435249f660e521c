// microbench.cu -- issue-rate probes for the instruction mix of the demod kernel
// on sm_100a (B200).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
// Prints warp-instructions per clock per SM for each probe (clock64-based).
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 4096
#define UNROLL 8

template <int KIND>
__global__ void probe(float *out, long long *cycles, float seed) {
    float2 a[UNROLL], b = make_float2(seed, seed * 0.5f), c = make_float2(0.999f, 1.001f);
    unsigned u[UNROLL];
    for (int i = 0; i < UNROLL; i++) { a[i] = make_float2(threadIdx.x + i, i * seed); u[i] = threadIdx.x * 2654435761u + i; }
    unsigned acc = 0;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < UNROLL; i++) {
            if (KIND == 0) { a[i].x = fmaf(a[i].x, c.x, b.x); }                       // FFMA
            if (KIND == 1) { a[i] = __ffma2_rn(a[i], c, b); }                         // FFMA2
            if (KIND == 2) { a[i] = __fadd2_rn(a[i], b); }                            // FADD2
            if (KIND == 3) { a[i].x = a[i].x + b.x; }                                 // FADD
            if (KIND == 4) { u[i] = __byte_perm(u[i], 0x00CB4B00u, 0x5140 + (it & 3)); }     // PRMT
            if (KIND == 5) { acc += __ballot_sync(0xFFFFFFFFu, a[i].x < b.x); a[i].x += 1.0f; } // FSETP+VOTE+FADD
            if (KIND == 6) { a[i].x = fmaf(a[i].x, c.x, b.x); a[i].y = a[i].y + b.y; } // FFMA + FADD
            if (KIND == 7) { a[i] = __ffma2_rn(a[i], c, b); u[i] = __byte_perm(u[i], 0x00CB4B00u, 0x5140 + (it & 3)); } // FFMA2 + PRMT
            if (KIND == 8) { a[i].x = (float)(u[i] & 0xFF); u[i] += 3; }               // I2F
            if (KIND == 9) { a[i] = __ffma2_rn(a[i], c, b); a[i].x = fmaf(a[i].x, c.x, b.x); } // FFMA2+FFMA
            if (KIND == 10) { u[i] = __funnelshift_r(u[i], acc, it & 31) & (u[i] ^ acc); } // SHF + LOP3
        }
    }
    long long t1 = clock64();
    float s = 0; unsigned v = acc;
    for (int i = 0; i < UNROLL; i++) { s += a[i].x + a[i].y; v ^= u[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)v;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int KIND>
void run(const char *name, int instr_per_iter, int warps) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float *out; long long *cyc;
    cudaMalloc(&out, sizeof(float) * sms * warps * 32);
    cudaMalloc(&cyc, sizeof(long long) * sms);
    probe<KIND><<<sms, warps * 32>>>(out, cyc, 1.0001f);
    probe<KIND><<<sms, warps * 32>>>(out, cyc, 1.0001f);
    cudaDeviceSynchronize();
    long long h[1024]; cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < sms; i++) avg += h[i]; avg /= sms;
    double winst = (double)ITER * UNROLL * instr_per_iter * warps;
    printf("%-22s warps/SM=%2d  warp-instr/clk/SM = %.3f  (cycles %.0f)\n", name, warps, winst / avg, avg);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int warps : {8, 16, 32}) {
        run<0>("FFMA", 1, warps);
        run<1>("FFMA2", 1, warps);
        run<2>("FADD2", 1, warps);
        run<3>("FADD", 1, warps);
        run<4>("PRMT", 1, warps);
        run<5>("FSETP+VOTE+FADD", 3, warps);
        run<6>("FFMA+FADD", 2, warps);
        run<7>("FFMA2+PRMT", 2, warps);
        run<8>("I2F(+LOP+IADD)", 3, warps);
        run<9>("FFMA2+FFMA", 2, warps);
        run<10>("SHF+LOP3(+LOP3)", 3, warps);
    }
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
