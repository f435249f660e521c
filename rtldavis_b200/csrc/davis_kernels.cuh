// davis_kernels.cuh -- kernel declarations and the records they exchange.
#pragma once
#include "davis_device.cuh"
#include "../../include/davis_b200.h"

namespace davis {

// A position where the fp32 decision bits match the sync word with at most one
// disagreeing symbol.  `miss` is the disagreeing symbol (0..15) or -1.
struct Cand {
    uint32_t stream;   // local to the pipeline stage
    int32_t v;         // index into the call's virtual sample sequence
    int32_t miss;
    uint32_t pad;
};

// One (call, index) report of a confirmed preamble before per-call dedupe
// (Demodulator._slice, dsp.py:190-205, up to the `seen` check).
struct alignas(16) Rec {
    int64_t g;                 // global sample index of the preamble
    int32_t v;                 // index into the call's virtual sample sequence
    uint32_t stream;           // local to the pipeline stage
    uint32_t call;
    int32_t index;             // Packet.index
    uint8_t data[10];          // Packet.data (on-air order)
    uint8_t crc_ok;
    uint8_t flags;
    uint32_t pad[3];
};
static_assert(sizeof(Rec) == 48, "Rec layout");

// Device-side counters of one pipeline stage (zeroed before every call).
struct Counters {
    unsigned int n_cand;       // entries appended to the candidate list
    unsigned int n_conf;       // entries appended to the confirmed list
    unsigned int n_rec;        // records produced (before dedupe)
    unsigned int n_hit;        // packets written to the hit buffer (after dedupe)
    unsigned int overflow;     // bit0 cand, bit1 conf, bit2 rec/hit, bit3 group list
    unsigned int n_grp;        // bitmap word groups queued for the fp64 pass
    unsigned int n_fix;        // bitmap words the fp64 pass changed
    unsigned int pad;
};

struct SearchParams {
    uint32_t pat_rev;          // sync word with symbol 0 in bit 0
    int v_lo, v_hi;            // inclusive range of reportable positions
    int words;                 // words per stream that hold reportable positions
    int words_alloc;           // words per stream in the bitmap
    int n_streams;
    unsigned int cand_cap, conf_cap, hit_cap;
    unsigned int tbl_mask;     // dedupe hash table size - 1 (power of two)
    int mark_words;            // 32-bit words of group marks per stream
    unsigned int grp_cap;      // capacity of the group list
};

// The fp64 pass works on groups of GROUP_WORDS bitmap words (224 samples = 7 per lane).
constexpr int GROUP_WORDS = 8;
constexpr int GROUP_SAMPLES = GROUP_WORDS * LANE_SAMPLES;
constexpr unsigned int FIX_CAP = 1024;

// A bitmap word whose fp32 decisions differed from the reference's float64 ones.
struct Fix {
    uint32_t stream, word, diff, pad;
};

constexpr int DEMOD_WARPS = 16;
constexpr int DEMOD_CTAS_PER_SM = 1;

void launch_demod(const View &vw, uint32_t *words, int n_streams, int tiles, int words_alloc,
                  float *disc_out, size_t disc_stride, cudaStream_t st);
void launch_search(const uint32_t *planes, const SearchParams &sp, Cand *cand, Counters *ctr,
                   cudaStream_t st);
void launch_precheck(const View &vw, const SearchParams &sp, const Cand *cand, Cand *conf,
                     uint32_t *marks, uint2 *glist, Counters *ctr, bool bitmap_is_exact,
                     cudaStream_t st);
void launch_exact_all(const View &vw, const SearchParams &sp, uint32_t *planes, cudaStream_t st);
void launch_exact(const View &vw, const SearchParams &sp, uint32_t *planes, const uint2 *glist,
                  Fix *fixes, Counters *ctr, cudaStream_t st);
void launch_slice(const View &vw, const SearchParams &sp, const uint32_t *planes,
                  const Cand *conf, const Fix *fixes, Rec *recs, uint32_t *tbl, uint8_t *lost,
                  Counters *ctr, int n_blocks, cudaStream_t st);
void launch_finalize(const View &vw, const SearchParams &sp, const Rec *recs, const uint8_t *lost,
                     davis_hit *hits, unsigned int *per_stream, Counters *ctr, cudaStream_t st);
void launch_order(const davis_hit *tmp, const Counters *ctr, int stream0, int n_streams,
                  unsigned int *per_stream, unsigned int *start, unsigned int *fill,
                  unsigned int *bucket, davis_hit *out, cudaStream_t st);
void launch_carry(const uint8_t *carry_old, uint8_t *carry_new, size_t carry_stride, int keep,
                  const uint8_t *input, size_t input_stride, int n_new, int n_streams,
                  int64_t *blocks_done, int n_blocks, int elem, cudaStream_t st);
void launch_rewind(int64_t *blocks_done, int n_streams, int n_blocks, cudaStream_t st);
void launch_exact_disc(const View &vw, int stream, int v0, int n, double *out, cudaStream_t st);
void launch_crc16(const uint8_t *data, size_t n, size_t len, uint16_t init, uint16_t poly,
                  int swap_bits, uint16_t *out, cudaStream_t st);

}  // namespace davis
