/*
 * davis_oracle.c -- plain C restatement of the reference demodulation path.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Parity status: PINNED --
 * tests/test_oracle_golden.py checks this file against the tests/golden fixtures,
 * which were produced by the unmodified Python reference
 * (oracle/gen_golden.py).  It exists because the numpy restatement is too slow
 * for the multi-thousand-stream parity cases and for a meaningful multi-core
 * CPU baseline.
 *
 * All arithmetic is IEEE double, evaluated in the order the reference's
 * formulas are written:
 *   LUT            src/rtldavis/dsp.py:26        (u - 127.4) / 127.6
 *   rotate_fs4     src/rtldavis/dsp.py:42-49     x * j^(i mod 4)
 *   fir9           src/rtldavis/dsp.py:52-73     9 symmetric taps, "valid"
 *   discriminate   src/rtldavis/dsp.py:76-90     cross product / (|a|^2 + 1e-10)
 *   quantize       src/rtldavis/dsp.py:93-98     IEEE sign bit (so -0.0 -> 1)
 *   _search        src/rtldavis/dsp.py:171-188   all 14 phases, every occurrence
 *   _slice         src/rtldavis/dsp.py:190-205   index <= B, 80 bits MSB-first, dedupe
 *   swap + CRC     src/rtldavis/protocol.py:79-83,290-297 ; src/rtldavis/crc.py:19-42
 *
 * Build: make -C oracle   (gcc -O2 -pthread -shared)
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SPS 14
#define PRE_SYMS 16
#define PKT_SYMS 80
#define PKT_BYTES 10

static const double TAPS[9] = {
    0.017682261285, 0.048171339939, 0.122424706672, 0.197408519126,
    0.228626345955,
    0.197408519126, 0.122424706672, 0.048171339939, 0.017682261285};

typedef struct {
    uint32_t stream;
    uint32_t call;
    int32_t index;
    uint8_t data[PKT_BYTES];
    uint8_t crc_ok;
    uint8_t pad;
} dor_packet;

/* ---- CRC-16, table built exactly like crc.py:28-42 ---- */
static uint16_t g_tbl[256];
static int g_tbl_ready = 0;

static void build_table(void) {
    for (int i = 0; i < 256; i++) {
        uint16_t reg = (uint16_t)(i << 8);
        for (int b = 0; b < 8; b++)
            reg = (reg & 0x8000) ? (uint16_t)((reg << 1) ^ 0x1021) : (uint16_t)(reg << 1);
        g_tbl[i] = reg;
    }
    g_tbl_ready = 1;
}

uint16_t dor_crc16(const uint8_t *p, size_t n, uint16_t init) {
    if (!g_tbl_ready) build_table();
    uint16_t reg = init;
    for (size_t i = 0; i < n; i++)
        reg = (uint16_t)((reg << 8) ^ g_tbl[(reg >> 8) ^ p[i]]);
    return reg;
}

uint8_t dor_swap_bits(uint8_t b) {
    b = (uint8_t)(((b & 0xF0) >> 4) | ((b & 0x0F) << 4));
    b = (uint8_t)(((b & 0xCC) >> 2) | ((b & 0x33) << 2));
    b = (uint8_t)(((b & 0xAA) >> 1) | ((b & 0x55) << 1));
    return b;
}

int dor_crc_gate(const uint8_t air[PKT_BYTES]) {
    uint8_t host[PKT_BYTES];
    for (int i = 0; i < PKT_BYTES; i++) host[i] = dor_swap_bits(air[i]);
    return dor_crc16(host + 2, 8, 0) == 0;
}

/* ---- demodulation of one whole stream into decision bits ---- */
static inline void rotated(const uint8_t *iq, long g, double *re, double *im) {
    if (g < 0) { *re = 0.0; *im = 0.0; return; }     /* zero history before the first block */
    double xr = ((double)iq[2 * g] - 127.4) / 127.6;
    double xi = ((double)iq[2 * g + 1] - 127.4) / 127.6;
    switch (g & 3) {                                   /* (xr + j xi) * j^(g mod 4) */
    case 0: *re = xr;  *im = xi;  break;
    case 1: *re = -xi; *im = xr;  break;
    case 2: *re = -xr; *im = -xi; break;
    default: *re = xi; *im = -xr; break;
    }
}

/* bits[g] (0/1 bytes) and optionally disc[g], filt[2g..2g+1] for g < n */
void dor_stream_bits(const uint8_t *iq, size_t n, uint8_t *bits, double *disc, double *filt) {
    double pr = 0.0, pi_ = 0.0;                        /* F[g-1]; F[-1] = 0 */
    double wr[9], wi[9];
    for (int m = 0; m < 9; m++) rotated(iq, (long)m - 9, &wr[m], &wi[m]);
    for (size_t g = 0; g < n; g++) {
        double fr = 0.0, fi = 0.0;
        for (int m = 0; m < 9; m++) { fr += TAPS[m] * wr[m]; fi += TAPS[m] * wi[m]; }
        double d = (pi_ * fr - pr * fi) / (pr * pr + pi_ * pi_ + 1e-10);
        bits[g] = (uint8_t)(signbit(d) ? 1 : 0);
        if (disc) disc[g] = d;
        if (filt) { filt[2 * g] = fr; filt[2 * g + 1] = fi; }
        pr = fr; pi_ = fi;
        memmove(wr, wr + 1, 8 * sizeof(double));
        memmove(wi, wi + 1, 8 * sizeof(double));
        rotated(iq, (long)g, &wr[8], &wi[8]);          /* F[g+1] uses samples g-8..g */
    }
}

/* ---- search + call mapping + slice + dedupe ---- */
typedef struct { int phase; int index; long g; } cand;

static int cand_cmp(const void *a, const void *b) {
    const cand *x = (const cand *)a, *y = (const cand *)b;
    if (x->phase != y->phase) return x->phase - y->phase;
    return x->index - y->index;
}

/* Packets of calls [0, n_calls) of one stream whose bits are given.  Returns the
 * number written, or -1 if `cap` is too small. */
long dor_packets_from_bits(const uint8_t *bits, size_t n, int block, uint32_t preamble,
                           uint32_t stream, dor_packet *out, size_t cap) {
    const long B = block;
    const long W = (PKT_SYMS * SPS) / B + 2;
    const long n_calls = (long)(n / (size_t)B);
    long written = 0;
    uint8_t pat[PRE_SYMS];
    for (int i = 0; i < PRE_SYMS; i++) pat[i] = (uint8_t)((preamble >> (PRE_SYMS - 1 - i)) & 1);

    for (long k = 0; k < n_calls; k++) {
        long g0 = (k - W + 1) * B;                     /* global index of window position 0 */
        long lo = g0 < 0 ? 0 : g0, hi = g0 + B;        /* positions kept: index in [0, B] */
        cand list[512];
        int nc = 0;
        for (long g = lo; g <= hi; g++) {
            if (g + (PKT_SYMS - 1) * SPS >= (long)n) break;
            int ok = 1;
            for (int i = 0; i < PRE_SYMS && ok; i++) ok = bits[g + i * SPS] == pat[i];
            if (!ok) continue;
            if (nc == 512) return -1;
            list[nc].index = (int)(g - g0);
            list[nc].phase = list[nc].index % SPS;
            list[nc].g = g;
            nc++;
        }
        qsort(list, (size_t)nc, sizeof(cand), cand_cmp);
        long first = written;
        for (int c = 0; c < nc; c++) {
            uint8_t data[PKT_BYTES] = {0};
            for (int i = 0; i < PKT_SYMS; i++)
                data[i >> 3] = (uint8_t)((data[i >> 3] << 1) | bits[list[c].g + i * SPS]);
            int dup = 0;
            for (long j = first; j < written && !dup; j++) dup = !memcmp(out[j].data, data, PKT_BYTES);
            if (dup) continue;
            if ((size_t)written >= cap) return -1;
            out[written].stream = stream;
            out[written].call = (uint32_t)k;
            out[written].index = list[c].index;
            memcpy(out[written].data, data, PKT_BYTES);
            out[written].crc_ok = (uint8_t)dor_crc_gate(data);
            out[written].pad = 0;
            written++;
        }
    }
    return written;
}

/* Whole path for a batch of independent streams laid out at `stride` bytes.
 * per_stream_cap slots are reserved per stream in `out`; counts[s] receives the
 * number used (or -1 on overflow).  Streams are handed out to n_threads POSIX
 * threads from a shared counter (the image's gcc has no OpenMP runtime). */
typedef struct {
    const uint8_t *iq;
    size_t n_streams, stride, usable, per_stream_cap;
    int block;
    uint32_t preamble;
    dor_packet *out;
    long *counts;
    long next;
    pthread_mutex_t lock;
} batch_job;

static void *batch_worker(void *arg) {
    batch_job *j = (batch_job *)arg;
    uint8_t *bits = (uint8_t *)malloc(j->usable ? j->usable : 1);
    for (;;) {
        pthread_mutex_lock(&j->lock);
        long s = j->next++;
        pthread_mutex_unlock(&j->lock);
        if (s >= (long)j->n_streams) break;
        dor_stream_bits(j->iq + (size_t)s * j->stride, j->usable, bits, NULL, NULL);
        j->counts[s] = dor_packets_from_bits(bits, j->usable, j->block, j->preamble, (uint32_t)s,
                                             j->out + (size_t)s * j->per_stream_cap,
                                             j->per_stream_cap);
    }
    free(bits);
    return NULL;
}

void dor_batch(const uint8_t *iq, size_t n_streams, size_t stride, size_t n_samples, int block,
               uint32_t preamble, dor_packet *out, size_t per_stream_cap, long *counts,
               int n_threads) {
    if (!g_tbl_ready) build_table();
    batch_job j;
    j.iq = iq; j.n_streams = n_streams; j.stride = stride;
    j.usable = (n_samples / (size_t)block) * (size_t)block;
    j.per_stream_cap = per_stream_cap; j.block = block; j.preamble = preamble;
    j.out = out; j.counts = counts; j.next = 0;
    pthread_mutex_init(&j.lock, NULL);
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t tid[256];
    int started = 0;
    for (int t = 1; t < n_threads; t++)
        if (pthread_create(&tid[started], NULL, batch_worker, &j) == 0) started++;
    batch_worker(&j);
    for (int t = 0; t < started; t++) pthread_join(tid[t], NULL);
    pthread_mutex_destroy(&j.lock);
}
