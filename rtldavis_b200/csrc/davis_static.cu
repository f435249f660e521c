// davis_static.cu -- davis_disc32_static(): the fp32 discriminator tap for callers that
// must not allocate device memory.
//
// Who that is.  The rest of the library sizes its buffers with cudaMalloc as it goes
// (davis_create, the first call of a new shape, davis_debug_disc32 on every call).  Some
// hosts of this code cannot live with that:
//   * a receive thread that is fed by the dongle's USB callback has a deadline per block;
//     cudaMalloc and cudaFree take a process-wide driver lock and may wait for the whole
//     device to drain, so one call can cost more than the block it was meant to serve;
//   * a plug-in inside a process whose device memory already belongs to somebody else's
//     pool (torch's caching allocator, an SDR framework's arena) -- a cudaMalloc there
//     either fails or takes memory the owner believes it has;
//   * a process started with a fixed device-memory budget, where everything a library
//     will ever hold has to be visible the moment the library is loaded.
// For them this one entry point keeps its working memory in a file-scope __device__
// array.  The price: the array exists on every GPU that loads the library, for the life
// of the process, whether the entry point is ever called or not, and it caps the size of
// a call (DAVIS_STATIC_MAX_SAMPLES).
//
// What it computes is what davis_debug_disc32 computes -- the streaming kernel's own
// fp32 value of Demodulator.discriminated (src/rtldavis/dsp.py:76-90,162) -- for the first
// n_samples samples of a fresh stream, with the same k_demod<true> launch.  That kernel
// needs three pieces of memory the caller never sees:
//   history   the 16 samples in front of the input (a fresh stream has none: byte 127, the
//             nearest a uint8 sample gets to zero, exactly as davis_create fills them);
//   bitmap    the decision words k_demod always stores (nobody reads them here);
//   staging   the discriminator of history + input; the caller gets the input part.
// All three live in g_scratch below.
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "davis_kernels.cuh"

using namespace davis;

namespace {

constexpr int HIST = HALO;                                         // samples of history
constexpr size_t MAX_SAMPLES = DAVIS_STATIC_MAX_SAMPLES;
constexpr size_t MAX_TILES = (HIST + MAX_SAMPLES + TILE_SAMPLES - 1) / TILE_SAMPLES;
constexpr size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t OFF_HIST = 0;
constexpr size_t OFF_BITMAP = align256(2 * HIST);
constexpr size_t OFF_DISC = OFF_BITMAP + align256(MAX_TILES * TILE_WORDS * sizeof(uint32_t));
constexpr size_t SCRATCH_BYTES = OFF_DISC + align256((HIST + MAX_SAMPLES) * sizeof(float));
static_assert(MAX_SAMPLES % 8 == 0, "the bulk copies move 16-byte pieces");
static_assert(SCRATCH_BYTES < (2u << 20), "the header promises less than 2 MiB per GPU");

// The working memory.  In host code this name is only a handle for cudaGetSymbolAddress:
// it is NOT the device address (a kernel handed `g_scratch` from the host would fault, or
// on a machine where the GPU can reach host memory quietly use the wrong bytes).
static __device__ __align__(256) unsigned char g_scratch[SCRATCH_BYTES];

// Device address of g_scratch on each GPU, looked up the first time the GPU is used.
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned char *g_base[MAX_DEVICES];

struct DeviceGuard {                     // leave the caller's current device as it was
    int prev = -1;
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int say(char *err, size_t cap, int code, const char *fmt, ...) __attribute__((format(printf, 4, 5)));
int say(char *err, size_t cap, int code, const char *fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

// true if `p` is memory the GPU `device` can use as a kernel argument
bool on_device(const void *p, int device) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    if (a.type == cudaMemoryTypeManaged) return true;
    return a.type == cudaMemoryTypeDevice && a.device == device;
}

}  // namespace

extern "C" int davis_disc32_static(const uint8_t *iq_dev, size_t n_samples, float *out_dev,
                                   int device, void *cuda_stream, char *err, size_t err_cap) {
    if (err && err_cap) err[0] = 0;
    if (n_samples > MAX_SAMPLES)
        return say(err, err_cap, DAVIS_E_SIZE,
                   "davis_disc32_static: %zu samples exceed DAVIS_STATIC_MAX_SAMPLES = %zu, the size its "
                   "fixed scratch array was built for; split the input or use davis_debug_disc32",
                   n_samples, MAX_SAMPLES);
    if (n_samples % 8)
        return say(err, err_cap, DAVIS_E_SIZE, "davis_disc32_static: n_samples = %zu is not a multiple of 8",
                   n_samples);
    if (n_samples == 0) return DAVIS_OK;
    if (!iq_dev || !out_dev) return say(err, err_cap, DAVIS_E_INVAL, "davis_disc32_static: null buffer");
    if ((uintptr_t)iq_dev & 15)
        return say(err, err_cap, DAVIS_E_INVAL, "davis_disc32_static: iq_dev must be 16-byte aligned");
    if ((uintptr_t)out_dev & 3)
        return say(err, err_cap, DAVIS_E_INVAL, "davis_disc32_static: out_dev must be 4-byte aligned");
    if (device < 0 || device >= MAX_DEVICES)
        return say(err, err_cap, DAVIS_E_INVAL, "davis_disc32_static: device %d outside 0..%d", device,
                   MAX_DEVICES - 1);

    DeviceGuard guard;
    cudaError_t e = cudaGetDevice(&guard.prev);
    if (e == cudaSuccess) e = cudaSetDevice(device);
    if (e != cudaSuccess)
        return say(err, err_cap, DAVIS_E_CUDA, "davis_disc32_static: cudaSetDevice(%d): %s", device,
                   cudaGetErrorString(e));
    if (!on_device(iq_dev, device) || !on_device(out_dev, device))
        return say(err, err_cap, DAVIS_E_INVAL,
                   "davis_disc32_static: iq_dev and out_dev must be device memory of GPU %d", device);

    // Set-up, once per GPU, with that GPU current: ask the runtime where g_scratch lives there.
    unsigned char *base = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_mu);
        if (!g_base[device]) {
            unsigned char *ptr = nullptr;
            e = cudaGetSymbolAddress((void **)&ptr, g_scratch);
            if (e != cudaSuccess || !ptr)
                return say(err, err_cap, DAVIS_E_CUDA, "davis_disc32_static: cudaGetSymbolAddress on GPU %d: %s",
                           device, cudaGetErrorString(e));
            g_base[device] = ptr;
        }
        base = g_base[device];
    }
    uint8_t *hist = base + OFF_HIST;
    uint32_t *bitmap = reinterpret_cast<uint32_t *>(base + OFF_BITMAP);
    float *disc = reinterpret_cast<float *>(base + OFF_DISC);

    cudaStream_t st = (cudaStream_t)cuda_stream;
    // The array still holds the previous call's bytes.  The bitmap and the staging area are
    // overwritten for every sample this call covers before anything reads them; the history
    // is the one part that must hold a known value, so it is written again every time (in
    // stream order, whichever stream the previous call used).
    e = cudaMemsetAsync(hist, 127, 2 * HIST, st);
    if (e != cudaSuccess)
        return say(err, err_cap, DAVIS_E_CUDA, "davis_disc32_static: cudaMemsetAsync: %s", cudaGetErrorString(e));

    View vw{};
    vw.carry = hist;
    vw.input = iq_dev;
    vw.n_carry = HIST;
    vw.n_new = (int)n_samples;
    vw.block = (int)n_samples;
    vw.window_blocks = 1;
    vw.elem = 2;
    const int tiles = (int)((HIST + n_samples + TILE_SAMPLES - 1) / TILE_SAMPLES);
    launch_demod(vw, bitmap, 1, tiles, tiles * TILE_WORDS, disc, 0, st);
    e = cudaGetLastError();
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(out_dev, disc + HIST, n_samples * sizeof(float), cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess)
        return say(err, err_cap, DAVIS_E_CUDA, "davis_disc32_static: %s", cudaGetErrorString(e));
    return DAVIS_OK;
}
