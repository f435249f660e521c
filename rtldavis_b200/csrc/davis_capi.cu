// davis_capi.cu -- host side of libdavis_b200.so: the C ABI of include/davis_b200.h.
//
// Owns per-stream state (the carried tail of every stream and its call counter) and the
// pipeline stages used for host-memory input.  Ordering (Demodulator._search,
// dsp.py:171-188) and per-call dedupe (_slice, dsp.py:203-205) happen on the device; the host
// only sizes the lists, and runs a group's chain again with larger ones when an input matches
// the sync word more often than noise does (the reference never fails on such input).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "davis_kernels.cuh"

using namespace davis;

namespace {

struct Stage {
    cudaStream_t st = nullptr;
    uint8_t *staging = nullptr;  size_t staging_bytes = 0;
    uint32_t *planes = nullptr;  size_t planes_words = 0;
    Cand *cand = nullptr;        unsigned int cand_cap = 0;
    Cand *conf = nullptr;        unsigned int conf_cap = 0;
    uint32_t *marks = nullptr;   size_t marks_cap = 0;       // one bit per bitmap word group
    uint2 *glist = nullptr;      size_t glist_cap = 0;       // groups queued for the fp64 pass
    Fix *fixes = nullptr;
    int mark_words = 0;
    Rec *recs = nullptr;         size_t recs_cap = 0;
    uint32_t *tbl = nullptr;     size_t tbl_size = 0;        // dedupe table, power of two
    uint8_t *lost = nullptr;     size_t lost_cap = 0;
    davis_hit *tmp = nullptr;    size_t tmp_cap = 0;         // packets before ordering
    unsigned int *bucket = nullptr;  size_t bucket_cap = 0;
    unsigned int *per_stream = nullptr;  size_t per_stream_cap = 0;   // count | start | fill
    Counters *ctr = nullptr;
    int last_group = -1;
    int tiles = 0, words_alloc = 0;
};

struct Group {
    int stage = 0, s0 = 0, ns = 0;
    size_t hit_off = 0;  unsigned int hit_cap = 0, n_hit = 0;
    uint32_t *planes = nullptr;          // bitmap rows of this group's streams
    davis_hit *hits = nullptr;           // where the chain writes this group's packets
    const uint8_t *dev_input = nullptr;  // first stream of the group on the device
    size_t dev_stride = 0;
    bool shared_demod = false;           // the bitmap came from a launch that covers several groups
    bool graph_events = false;           // ev[] belong to a cached CUDA graph, not to the event pool
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
};

}  // namespace

struct davis_handle {
    davis_cfg cfg{};
    int n_streams = 0, device = 0;
    int B = 0, W = 0, n_carry = 0, n_keep = 0;       // samples
    size_t carry_stride = 0;
    int elem = 2;                                     // bytes per sample of the current input type
    uint8_t *carry[2] = {nullptr, nullptr};
    int cur = 0;
    int64_t *d_blocks_done = nullptr;
    double *d_lut = nullptr;                          // dsp.py:26 table, float64
    std::vector<int64_t> h_blocks_done;
    uint32_t pat_rev = 0;

    std::vector<Stage> stages;
    std::vector<Group> groups;            // groups of the queued call
    std::vector<cudaEvent_t> event_pool;
    davis_hit *d_hits = nullptr;  size_t d_hits_cap = 0;
    Counters *h_ctr = nullptr;    size_t h_ctr_cap = 0;     // pinned, one per group
    cudaEvent_t demod_ev[2] = {nullptr, nullptr};           // around a demodulation launch shared by groups

    bool queued = false;
    int queued_blocks = 0;
    // the queued call, kept for a second run of a group's chain (list overflow)
    const uint8_t *q_iq = nullptr;  size_t q_bytes = 0, q_stride = 0;  bool q_on_device = false;
    std::vector<davis_hit *> retry_hits;                    // larger packet buffers of re-run groups
    // Small calls (the single-stream, one-block drop-in above all) are launch-bound: five
    // memsets + nine kernels + the counter copy.  Their chain is captured once per
    // (blocks, carry buffer, input pointer) into a CUDA graph and replayed.
    struct ChainGraph {
        int n_blocks = 0, cur = 0;
        const uint8_t *dev_input = nullptr;  size_t dev_stride = 0;
        uint64_t epoch = 0, launches = 0;
        cudaGraphExec_t exec = nullptr;
        cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    };
    std::vector<ChainGraph> graphs;
    int graph_misses = 0;                                   // cache misses in a row with the cache full
    uint64_t alloc_epoch = 0;                               // bumped whenever a device buffer moves
    bool have_results = false;
    size_t n_results = 0;
    davis_hit *h_hits = nullptr;  size_t h_hits_cap = 0;       // pinned; the fetched packets
    davis_stats stats{};
    std::string err;
};

namespace {

// davis_cfg.flavor -> the two places where the Go mirror's numbers differ (include/davis_b200.h)
void set_flavor(const davis_handle *h, View *vw) {
    const bool go = h->cfg.flavor == DAVIS_FLAVOR_GO;
    vw->disc_eps = go ? 0.0 : 1e-10;                 // dsp/dsp.go:93 vs dsp.py:88
    vw->disc_off = go ? 8 * SPS : 0;                 // protocol/protocol.go:156 vs protocol.py:305
}

double now_ms() {
    return std::chrono::duration<double, std::milli>(
               std::chrono::steady_clock::now().time_since_epoch()).count();
}

int fail(davis_handle *h, int code, const std::string &msg) {
    if (h) h->err = msg;
    return code;
}

#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail(h, e_ == cudaErrorMemoryAllocation ? DAVIS_E_NOMEM : DAVIS_E_CUDA, \
                        std::string(#call) + ": " + cudaGetErrorString(e_));              \
    } while (0)

template <typename T>
int grow(davis_handle *h, T **p, size_t *have, size_t want) {
    if (*have >= want) return DAVIS_OK;
    h->alloc_epoch++;
    if (*p) CU(cudaFree(*p));
    *p = nullptr; *have = 0;
    CU(cudaMalloc((void **)p, want * sizeof(T)));
    *have = want;
    return DAVIS_OK;
}

int tiles_needed(int v_len) { return (v_len + TILE_SAMPLES - 1) / TILE_SAMPLES; }

int prepare_stage(davis_handle *h, Stage &st, int ns, int n_blocks, size_t staging_bytes,
                  size_t rec_cap, size_t min_cand = 0, size_t min_conf = 0, size_t min_grp = 0) {
    const int v_len = h->n_carry + n_blocks * h->B;
    st.tiles = tiles_needed(v_len);
    st.words_alloc = st.tiles * TILE_WORDS + SEARCH_AHEAD;
    int rc;
    if ((rc = grow(h, &st.planes, &st.planes_words, (size_t)ns * st.words_alloc))) return rc;
    if (staging_bytes && (rc = grow(h, &st.staging, &st.staging_bytes, staging_bytes))) return rc;
    // expected near-matches: 17/65536 per sample position; allow 4x plus slack
    const double positions = (double)ns * ((double)n_blocks + h->W) * h->B;
    size_t want = (size_t)(positions * 17.0 / 65536.0 * 4.0) + 4096;
    size_t have = st.cand_cap;
    if ((rc = grow(h, &st.cand, &have, std::max(want, min_cand)))) return rc;
    st.cand_cap = (unsigned int)std::min<size_t>(have, 0xFFFFFFFFu);
    have = st.conf_cap;
    if ((rc = grow(h, &st.conf, &have, std::max(want / 2, min_conf)))) return rc;
    st.conf_cap = (unsigned int)std::min<size_t>(have, 0xFFFFFFFFu);
    const int groups = (st.words_alloc + GROUP_WORDS - 1) / GROUP_WORDS;
    st.mark_words = (groups + 31) / 32;
    if ((rc = grow(h, &st.marks, &st.marks_cap, (size_t)ns * st.mark_words))) return rc;
    const int per_pos = (SPS * (PKT_SYMS - 1)) / GROUP_SAMPLES + 2;        // groups one position touches
    if ((rc = grow(h, &st.glist, &st.glist_cap,
                   std::max(std::min((size_t)ns * groups, (size_t)st.conf_cap * per_pos),
                            std::min((size_t)ns * groups, min_grp))))) return rc;
    if (!st.fixes) CU(cudaMalloc((void **)&st.fixes, sizeof(Fix) * FIX_CAP));
    if ((rc = grow(h, &st.recs, &st.recs_cap, rec_cap))) return rc;
    if ((rc = grow(h, &st.lost, &st.lost_cap, rec_cap))) return rc;
    if ((rc = grow(h, &st.tmp, &st.tmp_cap, rec_cap))) return rc;
    if ((rc = grow(h, &st.bucket, &st.bucket_cap, rec_cap))) return rc;
    if ((rc = grow(h, &st.per_stream, &st.per_stream_cap, 3 * (size_t)ns + 4))) return rc;
    size_t tbl = 1024;
    while (tbl < 2 * rec_cap) tbl <<= 1;                       // load factor <= 0.5
    if (st.tbl_size < tbl) {
        size_t have_t = st.tbl_size;
        if ((rc = grow(h, &st.tbl, &have_t, tbl))) return rc;
        st.tbl_size = tbl;
    }
    if (!st.ctr) CU(cudaMalloc((void **)&st.ctr, sizeof(Counters)));
    if (!st.st) CU(cudaStreamCreateWithFlags(&st.st, cudaStreamNonBlocking));
    return DAVIS_OK;
}

int get_event(davis_handle *h, cudaEvent_t *e) {
    if (!h->event_pool.empty()) {
        *e = h->event_pool.back();
        h->event_pool.pop_back();
        return DAVIS_OK;
    }
    CU(cudaEventCreate(e));
    return DAVIS_OK;
}

void release_groups(davis_handle *h) {
    for (auto &e : h->demod_ev)
        if (e) { h->event_pool.push_back(e); e = nullptr; }
    for (auto &g : h->groups)
        for (auto &e : g.ev) {
            if (e && !g.graph_events) h->event_pool.push_back(e);
            e = nullptr;
        }
    h->groups.clear();
    for (auto p : h->retry_hits) cudaFree(p);
    h->retry_hits.clear();
}

// Enqueue every kernel of one group of streams on `st`.  `dev_input` points at the group's
// first stream.  With `demod_done` set, the bitmap rows of the group (g.planes) are being
// written by a demodulation launch on another stream that covers several groups: this
// chain then starts at the search and waits for that event first (its memsets do not).
int run_group(davis_handle *h, Stage &stage, Group &g, int gi, const uint8_t *dev_input,
              size_t input_stride, int n_blocks, cudaStream_t st, cudaEvent_t demod_done = nullptr,
              int cur = -1, bool timed = true) {
    // timed = false while the chain is being captured into a graph: an event whose last record was
    // captured cannot be waited on or read; run_group_cached records around the graph launch
    auto mark = [&](int i) -> cudaError_t { return timed ? cudaEventRecord(g.ev[i], st) : cudaSuccess; };
    if (cur < 0) cur = h->cur;           // which carry buffer holds the state BEFORE this call
    g.dev_input = dev_input;
    g.dev_stride = input_stride;
    g.shared_demod = demod_done != nullptr;
    if (!g.hits) g.hits = h->d_hits + g.hit_off;
    View vw;
    vw.carry = h->carry[cur] + (size_t)g.s0 * h->carry_stride + h->elem * (size_t)(h->n_keep - h->n_carry);
    vw.input = dev_input;
    vw.carry_stride = h->carry_stride;
    vw.input_stride = input_stride;
    vw.blocks_done = h->d_blocks_done + g.s0;
    vw.n_carry = h->n_carry;
    vw.n_new = n_blocks * h->B;
    vw.block = h->B;
    vw.window_blocks = h->W;
    vw.stream0 = g.s0;
    vw.lut = h->d_lut;
    vw.elem = h->elem;
    set_flavor(h, &vw);
    const bool bytes_in = h->elem == 2;

    SearchParams sp;
    sp.pat_rev = h->pat_rev;
    sp.v_lo = HALO;
    sp.v_hi = HALO + n_blocks * h->B;
    sp.words = sp.v_hi / LANE_SAMPLES + 1;
    sp.words_alloc = stage.words_alloc;
    sp.n_streams = g.ns;
    sp.cand_cap = stage.cand_cap;
    sp.conf_cap = stage.conf_cap;
    sp.hit_cap = g.hit_cap;
    sp.tbl_mask = (unsigned int)(stage.tbl_size - 1);
    sp.mark_words = stage.mark_words;
    sp.grp_cap = (unsigned int)std::min<size_t>(stage.glist_cap, 0xFFFFFFFFu);

    CU(cudaMemsetAsync(stage.ctr, 0, sizeof(Counters), st));
    CU(cudaMemsetAsync(stage.tbl, 0xFF, stage.tbl_size * sizeof(uint32_t), st));
    CU(cudaMemsetAsync(stage.lost, 0, g.hit_cap, st));
    CU(cudaMemsetAsync(stage.marks, 0, sizeof(uint32_t) * (size_t)g.ns * stage.mark_words, st));
    if (!g.planes) g.planes = stage.planes;
    if (demod_done) {
        CU(cudaStreamWaitEvent(st, demod_done, 0));
        CU(mark(0));
    } else {
        CU(mark(0));
        if (bytes_in) launch_demod(vw, g.planes, g.ns, stage.tiles, stage.words_alloc, nullptr, 0, st);
        else launch_exact_all(vw, sp, g.planes, st);               // complex input: fp64 decisions throughout
    }
    CU(mark(1));
    launch_search(g.planes, sp, stage.cand, stage.ctr, st);
    CU(mark(2));
    launch_precheck(vw, sp, stage.cand, stage.conf, stage.marks, stage.glist, stage.ctr, !bytes_in, st);
    if (bytes_in) launch_exact(vw, sp, g.planes, stage.glist, stage.fixes, stage.ctr, st);
    launch_slice(vw, sp, g.planes, stage.conf, stage.fixes, stage.recs, stage.tbl, stage.lost,
                 stage.ctr, n_blocks, st);
    unsigned int *count = stage.per_stream, *start = count + g.ns, *fill = start + g.ns + 1;
    CU(cudaMemsetAsync(count, 0, sizeof(unsigned int) * g.ns, st));
    launch_finalize(vw, sp, stage.recs, stage.lost, stage.tmp, count, stage.ctr, st);
    launch_order(stage.tmp, stage.ctr, g.s0, g.ns, count, start, fill, stage.bucket, g.hits, st);
    launch_carry(h->carry[cur] + (size_t)g.s0 * h->carry_stride,
                 h->carry[cur ^ 1] + (size_t)g.s0 * h->carry_stride, h->carry_stride, h->n_keep,
                 dev_input, input_stride, n_blocks * h->B, g.ns, h->d_blocks_done + g.s0, n_blocks,
                 h->elem, st);
    CU(mark(3));
    CU(cudaMemcpyAsync(&h->h_ctr[gi], stage.ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
    CU(mark(4));
    CU(cudaGetLastError());
    stage.last_group = gi;
    h->stats.launches += (demod_done ? 8 : 9) + (uint64_t)((g.ns + 65534) / 65535);
    return DAVIS_OK;
}

// A small call's chain through a cached CUDA graph: captured from run_group the first time a
// (blocks, carry buffer, input pointer) combination is seen, replayed afterwards.
int run_group_cached(davis_handle *h, Stage &stage, Group &g, int gi, const uint8_t *dev_input,
                     size_t input_stride, int n_blocks, cudaStream_t st) {
    davis_handle::ChainGraph *cg = nullptr;
    for (auto &c : h->graphs)
        if (c.n_blocks == n_blocks && c.cur == h->cur && c.dev_input == dev_input && c.dev_stride == input_stride) {
            cg = &c;
            break;
        }
    if (cg) h->graph_misses = 0;
    if (!cg) {
        // A caller that hands in a fresh device buffer every time (buffers straight from an
        // allocator) never hits the cache: capture + instantiate + destroy per call costs more
        // than the chain launched directly.  Once the cache is full and nothing but misses has
        // been seen for a while, small calls go out as plain launches until a known buffer returns.
        // (every 64th miss is still captured, so a new working set of buffers finds its way in)
        if (h->graphs.size() >= 16 && ++h->graph_misses > 16 && h->graph_misses % 64)
            return run_group(h, stage, g, gi, dev_input, input_stride, n_blocks, st);
        if (h->graphs.size() >= 16) {                      // callers cycling through many input buffers
            auto &old = h->graphs.front();
            if (old.exec) cudaGraphExecDestroy(old.exec);
            for (auto e : old.ev) if (e) h->event_pool.push_back(e);
            h->graphs.erase(h->graphs.begin());
        }
        h->graphs.emplace_back();
        cg = &h->graphs.back();
        cg->n_blocks = n_blocks; cg->cur = h->cur; cg->dev_input = dev_input; cg->dev_stride = input_stride;
        for (auto &e : cg->ev) { const int rc = get_event(h, &e); if (rc) return rc; }
    }
    if (cg->exec && cg->epoch != h->alloc_epoch) {          // a buffer the graph points at has moved
        cudaGraphExecDestroy(cg->exec);
        cg->exec = nullptr;
    }
    for (auto &e : g.ev) { if (e && !g.graph_events) h->event_pool.push_back(e); e = nullptr; }
    for (int i = 0; i < 5; i++) g.ev[i] = cg->ev[i];
    g.graph_events = true;
    if (!cg->exec) {
        const uint64_t before = h->stats.launches;
        // captured on the stage's own stream, never the caller's (which may itself be in a
        // capture, or carry work the graph must not swallow); the graph is launched on `st`
        cudaStream_t cap = stage.st;
        CU(cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal));
        const int rc = run_group(h, stage, g, gi, dev_input, input_stride, n_blocks, cap, nullptr, -1, false);
        cudaGraph_t graph = nullptr;
        const cudaError_t e = cudaStreamEndCapture(cap, &graph);
        if (rc || e != cudaSuccess) {
            if (graph) cudaGraphDestroy(graph);
            cudaGetLastError();
            return rc ? rc : fail(h, DAVIS_E_CUDA, std::string("stream capture: ") + cudaGetErrorString(e));
        }
        const cudaError_t ei = cudaGraphInstantiate(&cg->exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ei != cudaSuccess) { cg->exec = nullptr; return fail(h, DAVIS_E_CUDA, cudaGetErrorString(ei)); }
        cg->launches = h->stats.launches - before;
        cg->epoch = h->alloc_epoch;
    } else {                                                // what run_group notes on the host
        g.dev_input = dev_input;
        g.dev_stride = input_stride;
        g.shared_demod = false;
        g.hits = h->d_hits + g.hit_off;
        g.planes = stage.planes;
        stage.last_group = gi;
        h->stats.launches += cg->launches;
    }
    // per-kernel times are not taken in this mode: the whole chain counts as "finish"
    for (int i = 0; i < 3; i++) CU(cudaEventRecord(g.ev[i], st));
    CU(cudaGraphLaunch(cg->exec, st));
    for (int i = 3; i < 5; i++) CU(cudaEventRecord(g.ev[i], st));
    h->stats.graph_replays += 1;
    return DAVIS_OK;
}

}  // namespace

extern "C" {

int davis_abi_version(void) { return DAVIS_ABI_VERSION; }

const char *davis_strerror(int code) {
    switch (code) {
    case DAVIS_OK: return "ok";
    case DAVIS_E_INVAL: return "invalid argument or unsupported configuration";
    case DAVIS_E_SIZE: return "Incompatible array sizes";
    case DAVIS_E_CUDA: return "CUDA runtime error";
    case DAVIS_E_NOMEM: return "out of memory";
    case DAVIS_E_OVERFLOW: return "result buffer too small";
    case DAVIS_E_STATE: return "call sequence error";
    default: return "unknown error";
    }
}

const char *davis_last_error(const davis_handle *h) { return h ? h->err.c_str() : ""; }

int davis_create(const davis_cfg *cfg, int n_streams, int device, davis_handle **out) {
    if (!cfg || !out || n_streams < 1) return DAVIS_E_INVAL;
    *out = nullptr;
    if (cfg->symbol_length != SPS || cfg->preamble_symbols != PRE_SYMS ||
        cfg->packet_symbols != PKT_SYMS || cfg->block_size < 256 || cfg->block_size % 8 ||
        cfg->max_blocks < 1 || cfg->preamble > 0xFFFFu ||
        (cfg->flavor != DAVIS_FLAVOR_PYTHON && cfg->flavor != DAVIS_FLAVOR_GO))
        return DAVIS_E_INVAL;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count)
        return DAVIS_E_CUDA;
    davis_handle *h = new (std::nothrow) davis_handle;
    if (!h) return DAVIS_E_NOMEM;
    h->cfg = *cfg;
    h->n_streams = n_streams;
    h->device = device;
    h->B = cfg->block_size;
    h->W = (PKT_SYMS * SPS) / h->B + 2;                          // dsp.py:123-125
    h->n_carry = (h->W - 1) * h->B + HALO;
    h->n_keep = std::max(h->W - 1, 2) * h->B + HALO;             // also covers `discriminated`
    h->carry_stride = 2 * (size_t)h->n_keep;
    for (int i = 0; i < PRE_SYMS; i++)                           // symbol 0 is the MSB
        h->pat_rev |= ((cfg->preamble >> (PRE_SYMS - 1 - i)) & 1u) << i;
    h->h_blocks_done.assign(n_streams, 0);

    auto bail = [&](int rc) { davis_destroy(h); return rc; };
    if (cudaSetDevice(device) != cudaSuccess) return bail(DAVIS_E_CUDA);
    for (int i = 0; i < 2; i++) {
        if (cudaMalloc((void **)&h->carry[i], h->carry_stride * n_streams) != cudaSuccess)
            return bail(DAVIS_E_NOMEM);
        // history of a fresh stream: the reference starts from zero samples (dsp.py:129-134); byte
        // 127 is the nearest a uint8 sample gets to zero (-0.003).  Only the streaming kernel's
        // candidate bits of global samples < 10 see it; every exact path substitutes true zeros.
        cudaMemset(h->carry[i], 127, h->carry_stride * n_streams);
    }
    if (cudaMalloc((void **)&h->d_blocks_done, sizeof(int64_t) * n_streams) != cudaSuccess)
        return bail(DAVIS_E_NOMEM);
    cudaMemset(h->d_blocks_done, 0, sizeof(int64_t) * n_streams);
    {
        double lut[256];                                         // dsp.py:26, same float64 ops
        for (int i = 0; i < 256; i++) lut[i] = ((double)i - 127.4) / 127.6;
        if (cudaMalloc((void **)&h->d_lut, sizeof lut) != cudaSuccess) return bail(DAVIS_E_NOMEM);
        if (cudaMemcpy(h->d_lut, lut, sizeof lut, cudaMemcpyHostToDevice) != cudaSuccess)
            return bail(DAVIS_E_CUDA);
    }
    int slots = cfg->host_slots > 0 ? cfg->host_slots : 3;
    h->stages.resize(slots);
    if (cudaDeviceSynchronize() != cudaSuccess) return bail(DAVIS_E_CUDA);
    *out = h;
    return DAVIS_OK;
}

int davis_destroy(davis_handle *h) {
    if (!h) return DAVIS_OK;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    release_groups(h);
    for (auto &c : h->graphs) {
        if (c.exec) cudaGraphExecDestroy(c.exec);
        for (auto e : c.ev) if (e) cudaEventDestroy(e);
    }
    h->graphs.clear();
    for (auto e : h->event_pool) cudaEventDestroy(e);
    for (auto &s : h->stages) {
        if (s.staging) cudaFree(s.staging);
        if (s.planes) cudaFree(s.planes);
        if (s.cand) cudaFree(s.cand);
        if (s.conf) cudaFree(s.conf);
        if (s.marks) cudaFree(s.marks);
        if (s.glist) cudaFree(s.glist);
        if (s.fixes) cudaFree(s.fixes);
        if (s.recs) cudaFree(s.recs);
        if (s.tbl) cudaFree(s.tbl);
        if (s.lost) cudaFree(s.lost);
        if (s.tmp) cudaFree(s.tmp);
        if (s.bucket) cudaFree(s.bucket);
        if (s.per_stream) cudaFree(s.per_stream);
        if (s.ctr) cudaFree(s.ctr);
        if (s.st) cudaStreamDestroy(s.st);
    }
    for (int i = 0; i < 2; i++) if (h->carry[i]) cudaFree(h->carry[i]);
    if (h->d_blocks_done) cudaFree(h->d_blocks_done);
    if (h->d_lut) cudaFree(h->d_lut);
    if (h->d_hits) cudaFree(h->d_hits);
    if (h->h_ctr) cudaFreeHost(h->h_ctr);
    if (h->h_hits) cudaFreeHost(h->h_hits);
    delete h;
    return DAVIS_OK;
}

int davis_reset(davis_handle *h, int stream) {
    if (!h || stream < -1 || stream >= h->n_streams) return fail(h, DAVIS_E_INVAL, "bad stream");
    CU(cudaSetDevice(h->device));
    CU(cudaDeviceSynchronize());
    // the carried tail goes back to "no history" too (what is left of the previous stream would
    // only reach the streaming kernel's candidate bits of the first ten samples, but it costs
    // nothing to keep a reset stream identical to a fresh one)
    const int fill = h->elem == 2 ? 127 : 0;
    if (stream < 0) {
        std::fill(h->h_blocks_done.begin(), h->h_blocks_done.end(), 0);
        CU(cudaMemset(h->d_blocks_done, 0, sizeof(int64_t) * h->n_streams));
        for (int i = 0; i < 2; i++) CU(cudaMemset(h->carry[i], fill, h->carry_stride * h->n_streams));
    } else {
        h->h_blocks_done[stream] = 0;
        CU(cudaMemset(h->d_blocks_done + stream, 0, sizeof(int64_t)));
        for (int i = 0; i < 2; i++)
            CU(cudaMemset(h->carry[i] + (size_t)stream * h->carry_stride, fill, h->carry_stride));
    }
    h->queued = false;
    h->have_results = false;
    release_groups(h);
    return DAVIS_OK;
}

}  // extern "C"

// The carried tail is kept in the input's own element type.  Changing the type is
// allowed only on a fresh (or reset) handle: the reference keeps one complex128
// buffer for both entries (dsp.py:129,144-151), this library keeps raw input.
static int set_elem(davis_handle *h, int elem) {
    if (h->elem == elem) return DAVIS_OK;
    for (int64_t b : h->h_blocks_done)
        if (b != 0)
            return fail(h, DAVIS_E_STATE, "input type changed in mid-stream: davis_reset(h, -1) first");
    CU(cudaSetDevice(h->device));
    CU(cudaDeviceSynchronize());
    const size_t stride = (size_t)elem * h->n_keep;
    uint8_t *fresh[2] = {nullptr, nullptr};
    for (int i = 0; i < 2; i++) {
        if (cudaMalloc((void **)&fresh[i], stride * h->n_streams) != cudaSuccess) {
            if (fresh[0]) cudaFree(fresh[0]);
            cudaGetLastError();
            return fail(h, DAVIS_E_NOMEM, "carry buffers for the new input type");
        }
        cudaMemset(fresh[i], elem == 2 ? 127 : 0, stride * h->n_streams);
    }
    for (int i = 0; i < 2; i++) { cudaFree(h->carry[i]); h->carry[i] = fresh[i]; }
    h->alloc_epoch++;
    h->carry_stride = stride;
    h->elem = elem;
    h->cur = 0;
    return DAVIS_OK;
}

// Demodulator.demodulate for n_blocks blocks of every stream; `elem` = bytes per sample.
static int demod_common(davis_handle *h, const uint8_t *iq, size_t bytes_per_stream,
                        size_t stream_stride, int n_blocks, void *cuda_stream, int elem) {
    if (!h) return DAVIS_E_INVAL;
    if (!iq || n_blocks < 1) return fail(h, DAVIS_E_INVAL, "null input or n_blocks < 1");
    if (n_blocks > h->cfg.max_blocks)
        return fail(h, DAVIS_E_INVAL, "n_blocks exceeds davis_cfg.max_blocks");
    if (bytes_per_stream != (size_t)n_blocks * elem * h->B)
        return fail(h, DAVIS_E_SIZE, "Incompatible array sizes");
    if (h->queued) return fail(h, DAVIS_E_STATE, "previous call not fetched");
    {
        const int rc_e = set_elem(h, elem);
        if (rc_e) return rc_e;
    }
    if (h->n_streams > 1 && stream_stride < bytes_per_stream)
        return fail(h, DAVIS_E_SIZE, "stream_stride smaller than bytes_per_stream");
    if (h->queued) return fail(h, DAVIS_E_STATE, "previous call not fetched");
    const double t_submit = now_ms();
    CU(cudaSetDevice(h->device));

    cudaPointerAttributes attr;
    bool on_device = false;
    if (cudaPointerGetAttributes(&attr, iq) == cudaSuccess)
        on_device = attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
    else
        cudaGetLastError();
    if (on_device && (((uintptr_t)iq & 15) || (h->n_streams > 1 && (stream_stride & 15))))
        return fail(h, DAVIS_E_INVAL, "device input must be 16-byte aligned with a 16-byte stride");

    release_groups(h);
    h->have_results = false;
    h->stats = davis_stats{};
    h->stats.samples = (uint64_t)h->n_streams * n_blocks * h->B;
    h->q_iq = iq; h->q_bytes = bytes_per_stream; h->q_stride = stream_stride; h->q_on_device = on_device;

    // split into groups
    int per_group = h->n_streams;
    // Device-resident byte input: ONE demodulation launch over all streams, then the search
    // and finish kernels of two halves of the streams on two streams.  Those kernels are
    // latency- or issue-bound and leave the GPU half idle when they run one after another;
    // two independent chains fill each other's gaps (3.34 -> ~3.1 ms per 1 024 x 328 blocks).
    const bool split_finish = on_device && elem == 2 && h->n_streams >= 64 && h->stages.size() >= 2;
    if (split_finish) {
        const int chains = 2;        // measured: 3 chains are slower again (more small launches)
        per_group = (h->n_streams + chains - 1) / chains;
    }
    if (!on_device) {
        per_group = h->cfg.host_group > 0 ? h->cfg.host_group : 0;
        if (per_group <= 0) {                       // ~64 MiB of input per pipeline stage
            per_group = (int)std::max<size_t>(1, ((size_t)64 << 20) / bytes_per_stream);
        }
        per_group = std::min(per_group, h->n_streams);
    }
    const int n_groups = (h->n_streams + per_group - 1) / per_group;
    h->groups.resize(n_groups);
    size_t hit_total = 0;
    for (int gi = 0; gi < n_groups; gi++) {
        Group &g = h->groups[gi];
        g.stage = on_device ? (split_finish ? gi : 0) : gi % (int)h->stages.size();
        g.planes = nullptr;
        g.s0 = gi * per_group;
        g.ns = std::min(per_group, h->n_streams - g.s0);
        // records BEFORE the per-call dedupe (a burst matches at up to ~14 adjacent sample phases);
        // davis_cfg.max_hits bounds the packets handed to the caller, not this list
        size_t cap = 4096 + (size_t)g.ns * ((size_t)n_blocks + h->W) * 6;
        g.hit_cap = (unsigned int)std::min<size_t>(cap, 0x7FFFFFFFu);
        g.hit_off = hit_total;
        hit_total += g.hit_cap;
        for (auto &e : g.ev) { int rc = get_event(h, &e); if (rc) return rc; }
    }
    int rc;
    if ((rc = grow(h, &h->d_hits, &h->d_hits_cap, hit_total))) return rc;
    if (h->h_ctr_cap < (size_t)n_groups) {
        if (h->h_ctr) CU(cudaFreeHost(h->h_ctr));
        h->h_ctr = nullptr; h->h_ctr_cap = 0;
        h->alloc_epoch++;
        CU(cudaHostAlloc((void **)&h->h_ctr, sizeof(Counters) * n_groups, cudaHostAllocDefault));
        h->h_ctr_cap = n_groups;
    }

    if (split_finish) {
        cudaStream_t main_st = cuda_stream ? (cudaStream_t)cuda_stream : h->stages[0].st;
        // stage 0 holds the bitmap of every stream (and group 0's lists), stage g group g's lists
        if ((rc = prepare_stage(h, h->stages[0], h->n_streams, n_blocks, 0, h->groups[0].hit_cap))) return rc;
        for (int gi = 1; gi < n_groups; gi++)
            if ((rc = prepare_stage(h, h->stages[gi], h->groups[gi].ns, n_blocks, 0, h->groups[gi].hit_cap)))
                return rc;
        if (!cuda_stream) main_st = h->stages[0].st;
        Stage &s0 = h->stages[0];
        View vw{};
        vw.carry = h->carry[h->cur] + h->elem * (size_t)(h->n_keep - h->n_carry);
        vw.input = iq;
        vw.carry_stride = h->carry_stride;
        vw.input_stride = stream_stride;
        vw.blocks_done = h->d_blocks_done;
        vw.n_carry = h->n_carry;
        vw.n_new = n_blocks * h->B;
        vw.block = h->B;
        vw.window_blocks = h->W;
        vw.stream0 = 0;
        vw.lut = h->d_lut;
        vw.elem = h->elem;
        set_flavor(h, &vw);
        cudaEvent_t t0, t1;
        if ((rc = get_event(h, &t0)) || (rc = get_event(h, &t1))) return rc;
        CU(cudaEventRecord(t0, main_st));
        launch_demod(vw, s0.planes, h->n_streams, s0.tiles, s0.words_alloc, nullptr, 0, main_st);
        CU(cudaEventRecord(t1, main_st));
        h->stats.launches += 1;
        h->demod_ev[0] = t0;
        h->demod_ev[1] = t1;
        for (int gi = 0; gi < n_groups; gi++) {
            Group &g = h->groups[gi];
            g.planes = s0.planes + (size_t)g.s0 * s0.words_alloc;
            cudaStream_t st = gi == 0 ? main_st : h->stages[g.stage].st;
            if ((rc = run_group(h, h->stages[g.stage], g, gi, iq + (size_t)g.s0 * stream_stride,
                                stream_stride, n_blocks, st, t1)))
                return rc;
            // whatever the caller enqueues next on its stream comes after every chain
            if (st != main_st) CU(cudaStreamWaitEvent(main_st, g.ev[4], 0));
        }
    } else {
    // every stage is sized before the first kernel is enqueued: a failure here leaves the
    // stream state untouched
    const size_t stage_bytes = on_device ? 0 : (size_t)per_group * bytes_per_stream;
    for (int si = 0; si < std::min(n_groups, (int)h->stages.size()); si++)
        if ((rc = prepare_stage(h, h->stages[h->groups[si].stage], on_device ? h->n_streams : per_group,
                                n_blocks, stage_bytes, h->groups[0].hit_cap)))
            return rc;
    for (int gi = 0; gi < n_groups; gi++) {
        Group &g = h->groups[gi];
        Stage &stage = h->stages[g.stage];
        cudaStream_t st = stage.st;
        const uint8_t *dev_in;
        size_t dev_stride;
        if (on_device) {
            if (cuda_stream) st = (cudaStream_t)cuda_stream;
            dev_in = iq;
            dev_stride = stream_stride;
        } else {
            CU(cudaMemcpy2DAsync(stage.staging, bytes_per_stream, iq + (size_t)g.s0 * stream_stride,
                                 stream_stride, bytes_per_stream, g.ns, cudaMemcpyHostToDevice, st));
            h->stats.h2d_bytes += (uint64_t)bytes_per_stream * g.ns;
            dev_in = stage.staging;
            dev_stride = bytes_per_stream;
        }
        const bool small = n_groups == 1 && (size_t)h->n_streams * (size_t)n_blocks <= 512;
        if ((rc = small ? run_group_cached(h, stage, g, gi, dev_in, dev_stride, n_blocks, st)
                        : run_group(h, stage, g, gi, dev_in, dev_stride, n_blocks, st)))
            return rc;
    }
    }
    h->cur ^= 1;
    for (auto &b : h->h_blocks_done) b += n_blocks;
    h->queued = true;
    h->queued_blocks = n_blocks;
    h->stats.host_submit_ms = (float)(now_ms() - t_submit);
    return DAVIS_OK;
}

extern "C" {

int davis_demod_u8(davis_handle *h, const uint8_t *iq, size_t bytes_per_stream,
                   size_t stream_stride, int n_blocks, void *cuda_stream) {
    return demod_common(h, iq, bytes_per_stream, stream_stride, n_blocks, cuda_stream, 2);
}

int davis_demod_c64(davis_handle *h, const float *iq, size_t samples_per_stream,
                    size_t stream_stride_bytes, int n_blocks, void *cuda_stream) {
    return demod_common(h, reinterpret_cast<const uint8_t *>(iq), samples_per_stream * 8,
                        stream_stride_bytes, n_blocks, cuda_stream, 8);
}

int davis_demod_c128(davis_handle *h, const double *iq, size_t samples_per_stream,
                     size_t stream_stride_bytes, int n_blocks, void *cuda_stream) {
    return demod_common(h, reinterpret_cast<const uint8_t *>(iq), samples_per_stream * 16,
                        stream_stride_bytes, n_blocks, cuda_stream, 16);
}

// Second (third, ...) run of one group's chain after a list overflow, with every list that
// overflowed resized from the device counters.  The stream state before the call is still in
// the other carry buffer; the call counters are rewound for the group's streams first.
static int rerun_group(davis_handle *h, int gi) {
    Group &g = h->groups[gi];
    Stage &stage = h->stages[g.stage];
    const Counters c = h->h_ctr[gi];
    const int n_blocks = h->queued_blocks;
    h->stats.reruns += 1;
    if (h->stats.reruns > 16 * (uint64_t)h->groups.size())
        return fail(h, DAVIS_E_OVERFLOW, "device lists keep overflowing");
    CU(cudaStreamSynchronize(stage.st));                    // later groups of this stage are done with its lists
    const size_t min_cand = (c.overflow & 1u) ? (size_t)c.n_cand + c.n_cand / 4 + 1024 : 0;
    const size_t min_conf = (c.overflow & 2u) ? (size_t)c.n_conf + c.n_conf / 4 + 1024 : 0;
    const size_t min_grp = (c.overflow & 8u) ? (size_t)c.n_grp + c.n_grp / 4 + 1024 : 0;
    size_t rec_cap = g.hit_cap;
    if (c.overflow & 4u) {
        rec_cap = std::min<size_t>((size_t)c.n_rec + c.n_rec / 4 + 1024, 0x7FFFFFFFu);
        davis_hit *bigger = nullptr;
        CU(cudaMalloc((void **)&bigger, rec_cap * sizeof(davis_hit)));
        h->retry_hits.push_back(bigger);
        g.hits = bigger;
        g.hit_cap = (unsigned int)rec_cap;
    }
    // the lists of the stage may be shared with the bitmap of every stream (stage 0 of a
    // shared demodulation launch): keep its row count
    const int ns_alloc = (int)std::max<size_t>((size_t)g.ns, stage.planes_words / (size_t)std::max(stage.words_alloc, 1));
    uint32_t *planes_before = stage.planes;
    int rc = prepare_stage(h, stage, ns_alloc, n_blocks, stage.staging_bytes, rec_cap, min_cand, min_conf, min_grp);
    if (rc) return rc;
    if (stage.planes != planes_before) return fail(h, DAVIS_E_STATE, "bitmap moved during a re-run");
    cudaStream_t st = stage.st;
    launch_rewind(h->d_blocks_done + g.s0, g.ns, n_blocks, st);
    const int cur_before = h->cur ^ 1;                      // demod_common flipped it after enqueueing
    if (!h->q_on_device) {
        // The staging buffer still holds this group's bytes unless a later group of the same
        // pipeline stage has been copied over them; only then is the caller's host buffer read a
        // second time (which is why davis_b200.h asks for it to stay untouched until the fetch).
        if (stage.last_group != gi) {
            CU(cudaMemcpy2DAsync(stage.staging, h->q_bytes, h->q_iq + (size_t)g.s0 * h->q_stride, h->q_stride,
                                 h->q_bytes, g.ns, cudaMemcpyHostToDevice, st));
            h->stats.h2d_bytes += (uint64_t)h->q_bytes * g.ns;
        }
        g.planes = nullptr;
        return run_group(h, stage, g, gi, stage.staging, h->q_bytes, n_blocks, st, nullptr, cur_before);
    }
    return run_group(h, stage, g, gi, g.dev_input, g.dev_stride, n_blocks, st,
                     g.shared_demod ? h->demod_ev[1] : nullptr, cur_before);
}

// Wait for the queued call and bring its packets into the pinned result buffer.
static int complete_call(davis_handle *h) {
    CU(cudaSetDevice(h->device));
    if (h->queued) {
        h->n_results = 0;
        float demod = 0, search = 0, finish = 0;
        double t_wait = 0;
        for (size_t gi = 0; gi < h->groups.size(); gi++) {
            Group &g = h->groups[gi];
            const double t0 = now_ms();
            CU(cudaEventSynchronize(g.ev[4]));
            t_wait += now_ms() - t0;
            const Counters c = h->h_ctr[gi];
            if (c.overflow) {
                // An input that matches the sync word far more often than noise does (a
                // transmitter repeating it, a test pattern) overflows lists sized for noise.  The
                // reference answers such input (dsp.py:171-205), so the group's chain runs again
                // with lists sized from the counters, which hold the true counts.
                const int rc_r = rerun_group(h, (int)gi);
                if (rc_r) { h->queued = false; return rc_r; }
                gi--;                               // look at the same group again
                continue;
            }
            // counters and times of the pass that is kept (a pass discarded for an overflow shows
            // up in davis_stats.reruns only)
            float ms;
            CU(cudaEventElapsedTime(&ms, g.ev[0], g.ev[1])); demod += ms;
            CU(cudaEventElapsedTime(&ms, g.ev[1], g.ev[2])); search += ms;
            CU(cudaEventElapsedTime(&ms, g.ev[2], g.ev[3])); finish += ms;
            h->stats.candidates += c.n_cand;
            h->stats.confirmed += c.n_conf;
            h->stats.records += c.n_rec;
            g.n_hit = c.n_hit;
            h->n_results += c.n_hit;
            h->stats.d2h_bytes += sizeof(Counters);
        }
        // the one overflow that stays an error: a limit the CALLER set.  It is reported by this
        // fetch, after the packets are safe in the result buffer: the next fetch hands them out.
        const bool over_max = h->cfg.max_hits > 0 && h->n_results > (size_t)h->cfg.max_hits;
        if (h->demod_ev[0]) {
            float ms;
            CU(cudaEventElapsedTime(&ms, h->demod_ev[0], h->demod_ev[1]));
            demod += ms;
        }
        h->stats.demod_ms = demod;
        h->stats.search_ms = search;
        h->stats.finish_ms = finish;
        if (!h->groups.empty()) {
            float first_to_last = 0;
            // stages overlap; report the span of the slowest single group chain instead
            for (auto &g : h->groups) {
                float ms;
                CU(cudaEventElapsedTime(&ms, g.ev[0], g.ev[3]));
                first_to_last += ms;
            }
            h->stats.total_ms = first_to_last;
        }
        // Nothing to do on the host: the device applied the per-call dedupe
        // (dsp.py:203-205) and delivered every group in (stream, call, index mod 14,
        // index) order (dsp.py:175-186); groups hold ascending stream ranges.
        h->stats.host_wait_ms = (float)t_wait;
        h->stats.host_sort_ms = 0.0f;
        // one DMA per group into the handle's pinned result buffer
        const double t1 = now_ms();
        if (h->h_hits_cap < h->n_results) {
            if (h->h_hits) CU(cudaFreeHost(h->h_hits));
            h->h_hits = nullptr; h->h_hits_cap = 0;
            const size_t want = h->n_results + h->n_results / 4 + 1024;
            CU(cudaHostAlloc((void **)&h->h_hits, want * sizeof(davis_hit), cudaHostAllocDefault));
            h->h_hits_cap = want;
        }
        size_t at = 0;
        for (const Group &g : h->groups) {
            if (!g.n_hit) continue;
            CU(cudaMemcpyAsync(h->h_hits + at, g.hits, sizeof(davis_hit) * g.n_hit,
                               cudaMemcpyDeviceToHost, h->stages[g.stage].st));
            at += g.n_hit;
        }
        for (auto &st : h->stages) if (st.st) CU(cudaStreamSynchronize(st.st));
        h->stats.host_copy_ms = (float)(now_ms() - t1);
        h->stats.d2h_bytes += sizeof(davis_hit) * at;
        h->queued = false;
        h->have_results = true;
        if (over_max) {
            char msg[200];
            snprintf(msg, sizeof msg,
                     "%zu packets exceed max_hits = %d; they are kept: fetch again (with room for "
                     "them) to take them, or go on without",
                     h->n_results, h->cfg.max_hits);
            return fail(h, DAVIS_E_OVERFLOW, msg);
        }
    } else if (!h->have_results) {
        return fail(h, DAVIS_E_STATE, "nothing queued");
    }
    return DAVIS_OK;
}

int davis_fetch_hits(davis_handle *h, davis_hit *out, size_t cap, size_t *n) {
    if (!h || !n) return DAVIS_E_INVAL;
    const int rc = complete_call(h);
    if (rc == DAVIS_E_OVERFLOW && h->have_results) *n = h->n_results;      // max_hits: see complete_call
    if (rc) return rc;
    *n = h->n_results;
    if (*n > cap) return fail(h, DAVIS_E_OVERFLOW, "caller buffer smaller than the packet count");
    if (*n && !out) return fail(h, DAVIS_E_INVAL, "null output buffer");
    if (*n) memcpy(out, h->h_hits, sizeof(davis_hit) * *n);
    return DAVIS_OK;
}

int davis_fetch_hits_view(davis_handle *h, const davis_hit **out, size_t *n) {
    if (!h || !out || !n) return DAVIS_E_INVAL;
    *out = nullptr;
    *n = 0;
    const int rc = complete_call(h);
    if (rc == DAVIS_E_OVERFLOW && h->have_results) *n = h->n_results;
    if (rc) return rc;
    *out = h->n_results ? h->h_hits : nullptr;
    *n = h->n_results;
    return DAVIS_OK;
}

int davis_get_stats(davis_handle *h, davis_stats *out) {
    if (!h || !out) return DAVIS_E_INVAL;
    *out = h->stats;
    return DAVIS_OK;
}

int davis_debug_bits(davis_handle *h, int stream, uint8_t *out, size_t cap, int64_t *first,
                     size_t *n) {
    if (!h || !out || !first || !n || stream < 0 || stream >= h->n_streams)
        return fail(h, DAVIS_E_INVAL, "bad argument");
    if (h->queued || h->groups.empty()) return fail(h, DAVIS_E_STATE, "fetch the call first");
    CU(cudaSetDevice(h->device));
    for (size_t gi = 0; gi < h->groups.size(); gi++) {
        const Group &g = h->groups[gi];
        if (stream < g.s0 || stream >= g.s0 + g.ns) continue;
        const Stage &st = h->stages[g.stage];
        if (st.last_group != (int)gi)
            return fail(h, DAVIS_E_STATE, "planes of that stream were recycled by a later group");
        const int v_len = h->n_carry + h->queued_blocks * h->B;
        const size_t count = (size_t)v_len - HALO;
        *n = count;
        *first = (h->h_blocks_done[stream] - h->queued_blocks) * (int64_t)h->B - h->n_carry + HALO;
        if (count > cap) return fail(h, DAVIS_E_OVERFLOW, "bit buffer too small");
        std::vector<uint32_t> words((size_t)st.words_alloc);
        CU(cudaMemcpy(words.data(),
                      (g.planes ? g.planes : st.planes) + (size_t)(stream - g.s0) * st.words_alloc,
                      words.size() * 4, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < count; i++) {
            const size_t v = i + HALO;
            out[i] = (uint8_t)((words[v / LANE_SAMPLES] >> (v % LANE_SAMPLES)) & 1u);
        }
        return DAVIS_OK;
    }
    return fail(h, DAVIS_E_STATE, "stream not in the last call");
}

static int debug_view(davis_handle *h, View *vw) {
    vw->carry = h->carry[h->cur];
    vw->input = nullptr;
    vw->carry_stride = h->carry_stride;
    vw->input_stride = 0;
    vw->blocks_done = h->d_blocks_done;
    vw->n_carry = h->n_keep;
    vw->n_new = 0;
    vw->block = h->B;
    vw->window_blocks = h->W;
    vw->stream0 = 0;
    vw->lut = h->d_lut;
    vw->elem = h->elem;
    set_flavor(h, vw);
    return DAVIS_OK;
}

int davis_debug_disc(davis_handle *h, int stream, double *out, size_t cap) {
    if (!h || !out || stream < 0 || stream >= h->n_streams) return fail(h, DAVIS_E_INVAL, "bad argument");
    if (cap < (size_t)2 * h->B) return fail(h, DAVIS_E_OVERFLOW, "need 2*block_size doubles");
    if (h->queued) return fail(h, DAVIS_E_STATE, "fetch the call first");
    CU(cudaSetDevice(h->device));
    CU(cudaDeviceSynchronize());
    View vw;
    debug_view(h, &vw);
    double *d = nullptr;
    CU(cudaMalloc((void **)&d, sizeof(double) * 2 * h->B));
    launch_exact_disc(vw, stream, h->n_keep - 2 * h->B, 2 * h->B, d, 0);
    cudaError_t e = cudaMemcpy(out, d, sizeof(double) * 2 * h->B, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) return fail(h, DAVIS_E_CUDA, cudaGetErrorString(e));
    return DAVIS_OK;
}

int davis_debug_disc32(davis_handle *h, int stream, float *out, size_t cap) {
    if (!h || !out || stream < 0 || stream >= h->n_streams) return fail(h, DAVIS_E_INVAL, "bad argument");
    if (cap < (size_t)2 * h->B) return fail(h, DAVIS_E_OVERFLOW, "need 2*block_size floats");
    if (h->elem != 2) return fail(h, DAVIS_E_INVAL, "the fp32 kernel exists for uint8 input only");
    if (h->queued) return fail(h, DAVIS_E_STATE, "fetch the call first");
    CU(cudaSetDevice(h->device));
    CU(cudaDeviceSynchronize());
    View vw;
    debug_view(h, &vw);
    vw.carry += (size_t)stream * h->carry_stride;
    vw.blocks_done += stream;
    const int tiles = tiles_needed(h->n_keep);
    const int words_alloc = tiles * TILE_WORDS;
    uint32_t *planes = nullptr;
    float *d = nullptr;
    const size_t n_disc = (size_t)tiles * TILE_SAMPLES;
    CU(cudaMalloc((void **)&planes, sizeof(uint32_t) * words_alloc));
    if (cudaMalloc((void **)&d, sizeof(float) * n_disc) != cudaSuccess) {
        cudaFree(planes);
        return fail(h, DAVIS_E_NOMEM, "cudaMalloc");
    }
    launch_demod(vw, planes, 1, tiles, words_alloc, d, n_disc, 0);
    cudaError_t e = cudaMemcpy(out, d + (h->n_keep - 2 * h->B), sizeof(float) * 2 * h->B,
                               cudaMemcpyDeviceToHost);
    cudaFree(planes);
    cudaFree(d);
    if (e != cudaSuccess) return fail(h, DAVIS_E_CUDA, cudaGetErrorString(e));
    return DAVIS_OK;
}

int davis_host_alloc(void **p, size_t bytes) {
    if (!p) return DAVIS_E_INVAL;
    return cudaHostAlloc(p, bytes, cudaHostAllocDefault) == cudaSuccess ? DAVIS_OK : DAVIS_E_NOMEM;
}

int davis_host_free(void *p) {
    return cudaFreeHost(p) == cudaSuccess ? DAVIS_OK : DAVIS_E_CUDA;
}

int davis_crc16_batch(const uint8_t *data, size_t n, size_t len, uint16_t init, uint16_t poly,
                      int swap_bits, int device, uint16_t *out) {
    if ((!data && n * len) || !out) return DAVIS_E_INVAL;
    if (n == 0) return DAVIS_OK;
    if (cudaSetDevice(device) != cudaSuccess) return DAVIS_E_CUDA;
    uint8_t *d_in = nullptr;
    uint16_t *d_out = nullptr;
    int rc = DAVIS_OK;
    if (cudaMalloc((void **)&d_in, n * len + 1) != cudaSuccess ||
        cudaMalloc((void **)&d_out, n * sizeof(uint16_t)) != cudaSuccess)
        rc = DAVIS_E_NOMEM;
    if (!rc && n * len && cudaMemcpy(d_in, data, n * len, cudaMemcpyHostToDevice) != cudaSuccess)
        rc = DAVIS_E_CUDA;
    if (!rc) {
        launch_crc16(d_in, n, len, init, poly, swap_bits, d_out, 0);
        if (cudaMemcpy(out, d_out, n * sizeof(uint16_t), cudaMemcpyDeviceToHost) != cudaSuccess)
            rc = DAVIS_E_CUDA;
    }
    if (d_in) cudaFree(d_in);
    if (d_out) cudaFree(d_out);
    return rc;
}

}  // extern "C"
