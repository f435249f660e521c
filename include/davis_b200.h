/*
 * davis_b200.h -- C ABI of the B200 demodulation library (libdavis_b200.so).
 *
 * This is the drop-in boundary for rtldavis's demodulation hot path.  Each entry
 * point names the reference interface it stands in for (paths relative to the
 * rtldavis source tree).  Plain pointers and sizes only; no C++/torch types; no
 * exception crosses the boundary.  Every function returns 0 (DAVIS_OK) or a
 * negative DAVIS_E_* code; davis_last_error() gives the text for the last
 * failure on a handle.  The caller owns every buffer it passes; the library
 * owns the opaque handle and all device memory behind it.
 *
 * Batched model: one handle demodulates `n_streams` independent IQ streams.
 * One davis_demod_u8() call feeds `n_blocks` consecutive blocks of every stream
 * and yields exactly the packets that `n_blocks` successive calls of the
 * reference's Demodulator.demodulate() would have returned for each stream,
 * tagged with the call number.  A batch of one stream and one block is the
 * reference's own call shape.
 */
#ifndef DAVIS_B200_H
#define DAVIS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define DAVIS_API __attribute__((visibility("default")))
#else
#define DAVIS_API
#endif

#define DAVIS_ABI_VERSION 3

enum {
    DAVIS_OK = 0,
    DAVIS_E_INVAL = -1,     /* bad argument / unsupported configuration            */
    DAVIS_E_SIZE = -2,      /* "Incompatible array sizes" (dsp.py:32-36,145-149)   */
    DAVIS_E_CUDA = -3,      /* a CUDA runtime call failed                          */
    DAVIS_E_NOMEM = -4,     /* host or device allocation failed                    */
    DAVIS_E_OVERFLOW = -5,  /* hit or candidate buffer too small: nothing dropped
                               silently, the call's results are withheld           */
    DAVIS_E_STATE = -6      /* call sequence error (e.g. fetch with nothing queued) */
};

/* PacketConfig, src/rtldavis/dsp.py:101-125 (Go: dsp/dsp.go:172-218), as built by
 * protocol.new_packet_config, src/rtldavis/protocol.py:68-76.  Derived sizes
 * (sample_rate, preamble_length, packet_length, buffer_length) are recomputed by
 * the library exactly as dsp.py:118-125 does.  Supported: symbol_length 14,
 * preamble_symbols 16, packet_symbols 80, block_size a multiple of 8 and >= 256. */
typedef struct davis_cfg {
    int32_t bit_rate;          /* 19200                                             */
    int32_t symbol_length;     /* samples per symbol, 14                            */
    int32_t preamble_symbols;  /* 16                                                */
    int32_t packet_symbols;    /* 80                                                */
    uint32_t preamble;         /* sync pattern, first symbol in the MSB: 0xCB89     */
    int32_t block_size;        /* samples per block: 8192 (Python) or 512 (Go)      */
    int32_t max_blocks;        /* largest n_blocks a call will pass (>= 1)          */
    int32_t max_hits;          /* most packets one call may return (first fetch: DAVIS_E_OVERFLOW beyond); 0 = no limit */
    int32_t host_slots;        /* pipeline depth for host-memory input; 0 = auto    */
    int32_t host_group;        /* streams per pipeline stage for host input; 0=auto */
    int32_t flavor;            /* DAVIS_FLAVOR_PYTHON (0) or DAVIS_FLAVOR_GO (1), ABI >= 2 */
} davis_cfg;

/* Which of the reference's two implementations the numbers follow where they differ.
 * Decision bits, packets, the CRC verdict, rssi and snr are the same in both.
 *   PYTHON: discriminator denominator |a|^2 + 1e-10 (src/rtldavis/dsp.py:88);
 *           davis_hit.disc_mean = mean(discriminated[index : index+224])
 *           (src/rtldavis/protocol.py:304-310).
 *   GO:     denominator |a|^2, no epsilon (dsp/dsp.go:93; 0/0 gives NaN);
 *           davis_hit.disc_mean = mean(Discriminated[Idx+8*14 : Idx+24*14])
 *           (protocol/protocol.go:156-164; the caller adds the 9600 Hz bias of
 *           protocol.go:168).  Use block_size 512 (dsp/dsp.go:212).  Two bits are not
 *           Go's on amd64: the decisions of global samples 0 and 1 of a fresh stream are
 *           0/0 in Go, a NaN whose sign bit is platform-defined (set on amd64, clear on
 *           arm64); this library keeps 0 there, as Python and arm64 do. */
enum { DAVIS_FLAVOR_PYTHON = 0, DAVIS_FLAVOR_GO = 1 };

/* dsp.Packet, src/rtldavis/dsp.py:12-17 (Go: dsp/dsp.go:133-136), plus the
 * verdict of the CRC gate in Parser.parse, src/rtldavis/protocol.py:290-300, and
 * the mean the parser takes over `discriminated` (protocol.py:304-311). */
typedef struct davis_hit {
    int64_t position;     /* global sample index of the preamble in its stream      */
    uint32_t stream;      /* stream the packet belongs to                           */
    uint32_t call;        /* which demodulate() call of that stream returns it      */
    int32_t index;        /* Packet.index (0..block_size inclusive)                 */
    uint8_t data[10];     /* Packet.data, on-air bit order, MSB-first               */
    uint8_t crc_ok;       /* 1 iff CRC-16(0x1021, init 0) of swapped bytes 2..9 = 0 */
    uint8_t flags;        /* bit0: a decision bit was corrected by the fp64 re-check */
    double rssi;          /* Packet.rssi, dsp.py:207-229                            */
    double snr;           /* Packet.snr,  dsp.py:230-236                            */
    double disc_mean;     /* mean(discriminated[index:index+224]), protocol.py:304-310
                             (flavor GO: [index+112:index+336], protocol.go:156-164);
                             NaN when crc_ok == 0 (the reference takes it after the CRC gate) */
} davis_hit;

typedef struct davis_handle davis_handle;

/* Counters of the most recent davis_demod_u8() call (all pipeline stages summed). */
typedef struct davis_stats {
    uint64_t samples;         /* complex samples demodulated (new blocks only)       */
    uint64_t launches;        /* kernels launched                                    */
    uint64_t candidates;      /* near-matches of the sync word in the fp32 bits      */
    uint64_t confirmed;       /* preamble positions confirmed in fp64                */
    uint64_t records;         /* (call, index) records produced, before dedupe       */
    float demod_ms;           /* device time in the demodulation kernel              */
    float search_ms;          /* device time in the sync search kernel               */
    float finish_ms;          /* device time in confirm + slice/CRC + carry kernels  */
    float total_ms;           /* first kernel start to last kernel end               */
    uint64_t h2d_bytes;       /* bytes copied host->device by the call               */
    uint64_t d2h_bytes;       /* bytes copied device->host by the fetch              */
    float host_submit_ms;     /* host time inside davis_demod_u8 (enqueue only)      */
    float host_wait_ms;       /* host time the fetch spent waiting for the device    */
    float host_copy_ms;       /* host time copying packet records back               */
    float host_sort_ms;       /* always 0: ordering (dsp.py:175-186) is done on the device */
    uint64_t reruns;          /* chains run again with larger lists (sync-word-dense input) */
    uint64_t graph_replays;   /* small calls whose chain went out as one cached CUDA graph   */
} davis_stats;

DAVIS_API int davis_abi_version(void);

/* Demodulator.__init__, src/rtldavis/dsp.py:127-137 (Go NewDemodulator,
 * dsp/dsp.go:220-241), for n_streams independent streams on CUDA device `device`. */
DAVIS_API int davis_create(const davis_cfg *cfg, int n_streams, int device, davis_handle **out);
DAVIS_API int davis_destroy(davis_handle *h);

/* Demodulator.reset, src/rtldavis/dsp.py:248-253 (Go: dsp/dsp.go:294-310).
 * stream = -1 resets every stream (history, call counter). */
DAVIS_API int davis_reset(davis_handle *h, int stream);

/* Demodulator.demodulate, src/rtldavis/dsp.py:139-169 (Go: dsp/dsp.go:274-292),
 * uint8 entry (ByteToCmplxLUT.execute, dsp.py:28-39), for n_blocks blocks of every
 * stream.  `iq` is interleaved I,Q bytes; stream s starts at iq + s*stream_stride
 * and supplies bytes_per_stream = n_blocks * 2 * block_size bytes (anything else
 * is DAVIS_E_SIZE).  `iq` may be device memory (16-byte aligned, stride a multiple
 * of 16; used in place) or host memory (pinned for overlap; staged through the handle's
 * pipeline).  Either kind must stay valid AND unmodified until the fetch: device input is
 * read in place by the whole chain, and host input may be copied a second time at the
 * fetch when a pipeline stage's lists overflowed after a later group of streams had
 * taken over its staging buffer (davis_stats.reruns).  Work is enqueued
 * asynchronously; `cuda_stream` (a cudaStream_t, may be NULL) is the stream the
 * caller's device input is ordered on. */
DAVIS_API int davis_demod_u8(davis_handle *h, const uint8_t *iq, size_t bytes_per_stream,
                   size_t stream_stride, int n_blocks, void *cuda_stream);

/* Same for complex input, the live pyrtlsdr entry, src/rtldavis/dsp.py:144-151
 * (`dest[:] = input_data`: the samples are used as they are, no LUT).  complex64 is
 * interleaved float re, im; complex128 interleaved double re, im (what pyrtlsdr's
 * read_samples returns and what worker.py:49 passes on).  samples_per_stream =
 * n_blocks * block_size.  Every decision of this entry is made in float64 in the
 * reference's operation order (there is no fp32 streaming form for arbitrary complex
 * samples).  The carried tail is kept in the input's element type, so the input type
 * may change only on a fresh or reset handle (DAVIS_E_STATE otherwise). */
DAVIS_API int davis_demod_c64(davis_handle *h, const float *iq, size_t samples_per_stream,
                    size_t stream_stride_bytes, int n_blocks, void *cuda_stream);
DAVIS_API int davis_demod_c128(davis_handle *h, const double *iq, size_t samples_per_stream,
                     size_t stream_stride_bytes, int n_blocks, void *cuda_stream);

/* Waits for the queued call and returns its packets: the concatenation, stream by
 * stream and call by call, of what the reference returns (search order of
 * dsp.py:171-188, the index <= block_size filter and per-call dedupe of
 * dsp.py:190-205).  *n receives the number of packets; if it exceeds `cap` the
 * function returns DAVIS_E_OVERFLOW with *n set and nothing copied, and may be
 * called again with a larger buffer.
 * The device-side lists are sized for the sync-word matches of noise; an input that
 * matches far more often (a transmitter repeating the sync word) makes the library run
 * the affected chain again with larger lists (davis_stats.reruns) -- the reference
 * answers such input (dsp.py:171-205), so does this.  The one remaining overflow error
 * is a limit the caller set with davis_cfg.max_hits > 0: the fetch that finds more
 * packets than that returns DAVIS_E_OVERFLOW once, with *n set and nothing copied or
 * truncated.  The packets are kept (the streams have advanced past the call, so they
 * could not be produced again): the next fetch hands them out. */
DAVIS_API int davis_fetch_hits(davis_handle *h, davis_hit *out, size_t cap, size_t *n);

/* Same, without the copy: *out points at the handle's own (pinned) result buffer,
 * valid until the next davis_demod_*, davis_reset or davis_destroy on the handle. */
DAVIS_API int davis_fetch_hits_view(davis_handle *h, const davis_hit **out, size_t *n);

DAVIS_API int davis_get_stats(davis_handle *h, davis_stats *out);

/* Verification taps (not on the product path).
 * davis_debug_bits: the fp32 kernel's decision bits (Demodulator.quantized,
 *   dsp.py:163-166) of the last call for one stream, one 0/1 byte per sample,
 *   starting at global sample *first.
 * davis_debug_disc: Demodulator.discriminated (dsp.py:133,162), the 2*block_size
 *   float64 values the reference holds after the last call, recomputed in fp64.
 * davis_debug_disc32: the fp32 kernel arithmetic's discriminator for the same
 *   2*block_size samples (for the 1e-5 tolerance count). */
DAVIS_API int davis_debug_bits(davis_handle *h, int stream, uint8_t *out, size_t cap,
                     int64_t *first, size_t *n);
DAVIS_API int davis_debug_disc(davis_handle *h, int stream, double *out, size_t cap);
DAVIS_API int davis_debug_disc32(davis_handle *h, int stream, float *out, size_t cap);

/* No counterpart in the reference: rtldavis never has to think about device memory.
 * davis_disc32_static is davis_debug_disc32 for callers that must not allocate on the
 * GPU -- a receive thread on the dongle's callback deadline (cudaMalloc/cudaFree take a
 * process-wide lock and can wait for the device to drain), a plug-in in a process whose
 * device memory belongs to another allocator's pool, a process held to a fixed memory
 * budget.  It needs no handle and calls neither cudaMalloc nor cudaFree nor anything that
 * does: its working memory (16 samples of history, the decision bitmap the kernel always
 * stores, a staging row) is a fixed array of just over 1 MiB inside the library image,
 * present on every GPU that loads the library until the process exits.
 *   in   iq_dev    device memory, 16-byte aligned: interleaved I,Q bytes, the FIRST n_samples
 *                  samples of a stream (zero history in front, like a fresh Demodulator)
 *   out  out_dev   device memory: n_samples floats, the fp32 streaming kernel's
 *                  discriminator of those samples (the values davis_debug_disc32 reports)
 *   n_samples      a multiple of 8, at most DAVIS_STATIC_MAX_SAMPLES (about one second of
 *                  one channel); more is DAVIS_E_SIZE, nothing is run
 *   device, cuda_stream   the GPU both buffers live on; the stream the work is put on
 *                  (NULL = the default stream).  The call returns once the work is
 *                  enqueued: out_dev is complete when that stream has caught up.
 *   err, err_cap   optional buffer for the reason of a failure (there is no handle to ask)
 * The first call on each GPU looks up where the array lives there (cudaGetSymbolAddress);
 * make that call before any stream capture is begun.  Every call on a GPU uses the same
 * array and finds the previous call's bytes in it; the entry point rewrites what it relies
 * on, so calls that follow one another on a stream are independent, but two calls that can
 * run on the GPU at the same time (different streams, different threads) overwrite each
 * other's working memory and both results are garbage: serialise them. */
#define DAVIS_STATIC_MAX_SAMPLES 262144
DAVIS_API int davis_disc32_static(const uint8_t *iq_dev, size_t n_samples, float *out_dev,
                        int device, void *cuda_stream, char *err, size_t err_cap);

/* Pinned host memory for pipelined host input (cudaHostAlloc / cudaFreeHost). */
DAVIS_API int davis_host_alloc(void **p, size_t bytes);
DAVIS_API int davis_host_free(void *p);

/* CRC.checksum, src/rtldavis/crc.py:19-26 (Go: crc/crc.go:37-44), and
 * protocol.swap_bit_order, src/rtldavis/protocol.py:79-83, run on the device for
 * `n` independent byte strings of `len` bytes each (parity taps for the CRC
 * kernel; host buffers in and out). */
DAVIS_API int davis_crc16_batch(const uint8_t *data, size_t n, size_t len, uint16_t init,
                      uint16_t poly, int swap_bits, int device, uint16_t *out);

DAVIS_API const char *davis_last_error(const davis_handle *h);
DAVIS_API const char *davis_strerror(int code);

#ifdef __cplusplus
}
#endif
#endif /* DAVIS_B200_H */
