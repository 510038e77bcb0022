// cudasw.cu - host side of libcudasw.so: engine handle, batch pipeline, C ABI (include/cudasw.h).
//
// Pipeline of one batch (all on one stream, no host round trip per pair):
//   H2D blobs+offsets -> encode (clean(): SswJniAligner.java:71-79) -> bin pairs by query rows
//   (device radix sort) -> forward score pass per variant -> bin by read_end -> reverse score pass
//   -> banded traceback + CIGAR emission -> D2H results.
// There is no CPU alignment path in this file or anywhere in the library.
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <condition_variable>
#include <atomic>
#include <mutex>
#include <shared_mutex>
#include <thread>
#include <unordered_set>
#include <string>
#include <unordered_map>
#include <vector>

#define CUSW_MAIN_TU 1  // non-template kernels are emitted in this translation unit only
#include "sw_kernels.cuh"
#include "rev_band.cuh"
#include "band_filter.cuh"
#include "ja_mode.cuh"
#include "variants.h"

using namespace cusw;

namespace {

thread_local std::string g_create_error;
std::vector<cusw::Variant> g_var;
int g_nvar = 0;
std::once_flag g_var_once;

#define CK(expr)                                                                              \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                   \
            return CUDASW_ECUDA;                                                              \
        }                                                                                     \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

}  // namespace

// Queries named as segments instead of shipped as bytes: pair i = segs[seg_off[i] .. seg_off[i+1]) concatenated (or
// slots [i*stride, (i+1)*stride) when stride > 0; empty slots have start == end); a segment with contig < 0 is
// bytes [start, end) of `lit`. w_stride > 0: the WINDOW descriptors handed to run() are fixed-stride too.
struct QuerySegs {
    const WindowSeg* segs;
    const int64_t* seg_off;
    int stride;
    const uint8_t* lit;
    int64_t lit_bytes;
    int w_stride;
};

// Host inputs packed as 2 bits per base + an optional N mask (cudasw_align_batch_packed); NULL masks = no N anywhere.
struct PackedIn {
    const uint32_t *seq2, *seqN, *ref2, *refN;
};

// pinned host staging that grows on demand (cudasw_breakpoint_homology)
template <typename T>
struct PinBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMallocHost(reinterpret_cast<void**>(&p), n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct cudasw {
    int device = 0;
    cudasw_scoring scoring{};
    ScoreConst sc{};
    TbConst tc{};
    VariantTable vt{};
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;  // H2D of the next chunk while the current one computes
    std::vector<cudaEvent_t> h2d_done;
    static constexpr int kAux = 4;  // variant kernels of one pass run side by side (mixed-length batches)
    // A phase's launches: its main stream plus auxiliary streams forked/joined by events. Two sets, because the
    // forward pass of chunk i (compute) runs concurrently with the reverse pass and traceback of chunk i-1 (post).
    struct StreamSet {
        cudaStream_t main = nullptr;
        cudaStream_t aux[kAux]{};
        cudaEvent_t fork_ev = nullptr, join_ev[kAux]{};
    };
    StreamSet fwd_set, post_set;  // fwd_set.main == stream
    cudaStream_t tb_streams[5]{};  // the one-thread traceback classes run side by side (post_chunk)
    cudaEvent_t tb_join[5]{};
    cudaStream_t pack_stream = nullptr;  // encode / bin of chunk i+1 while chunk i is in its forward pass
    struct ChunkEvents { cudaEvent_t pack0, enc1, pack1, fwd0, fwd1; };
    std::vector<ChunkEvents> chunk_ev;
    unsigned long long* d_strip = nullptr;  // [0], [1]: bump counters of the strip-mining boundary buffers (forward, reverse); [2], [3]: cells, padded cells
    unsigned long long* h_cells = nullptr;  // pinned copy of d_strip[2..3]
    DevBuf<uint8_t> strip_fwd, strip_rev;
    DevBuf<uint32_t> rev_ring[2 * kRevVariants];  // banded reverse pass: per-group boundary rings of every (variant, flavour) launch
    int* d_rev_slots = nullptr;                    // (start, count) per slot of that pass, filled on the device
    int64_t batch_max_q = 0, batch_max_w = 0;      // longest query / window of the batch being run (host offsets)
    int launch_variants(StreamSet& ss, bool reverse, const int* order, PairIO io, const int4* fwd_c, int4* out_c, StripPool spool, uint64_t* launches,
                        const int* counts, const int* dev_slots, bool maybe_x);
    std::vector<std::vector<int>> fwd_counts;  // per chunk: pairs per query-length class (variant), counted on the host
    DevBuf<int> fwd_slots;                     // per chunk: (start, count) of every forward slot, filled on the device
    DevBuf<uint8_t> cub_tmp_post;              // sort workspace of the post stream (it may run beside a later chunk's pack)
    bool batch_maybe_n = true;                 // false when the caller's packed inputs carry no N mask for the windows
    std::vector<int16_t> var_by_q;  // host copy of variant_of() for the stats
    cudaEvent_t ev[8]{};
    std::mutex mu;
    std::string error;
    cudasw_stats stats{};

    // staged raw inputs (host API) and encoded copies
    DevBuf<uint8_t> raw_seq, raw_ref, enc_seq, enc_ref;
    DevBuf<uint32_t> pk_seq2, pk_seqN, pk_ref2, pk_refN;  // packed host inputs as copied (cudasw_align_batch_packed)
    int64_t pk_word0_q = 0, pk_word0_w = 0;               // first 16-base word of the batch held in pk_* (even)
    bool pk_has_qn = false, pk_has_wn = false;
    DevBuf<int64_t> seq_off, ref_off;
    DevBuf<uint32_t> keys, keys_alt;
    DevBuf<int> vals, vals_alt, retry_a, retry_b;
    DevBuf<int4> fwd, rev;
    DevBuf<cudasw_result> res;
    DevBuf<uint32_t> cig;
    DevBuf<uint8_t> cub_tmp;
    // k-mer filter in front of the forward pass (band_filter.cuh): per-pair column range + certificate data, the pairs whose
    // sub-window result is proven, and the second (full-window) pass's own bins
    FilterConst fc{};
    bool use_filter = false;
    bool filter_always = false;         // CUDASW_FILTER=2: also for chunks of less than two waves of forward CTAs (tests)
    std::vector<char> chunk_filter;     // per chunk: the filter runs (scan_batch)
    DevBuf<FilterRec> frec;
    DevBuf<uint32_t> fb_ring[kBandVariants];  // banded forward pass: per-group boundary rings of every strip-height launch
    int* d_band_slots = nullptr;              // (start, count) per slot of that pass
    DevBuf<uint32_t> bkeys, bkeys_alt;        // its sort keys / job lists (per chunk range, like keys / vals)
    DevBuf<int> bvals, bvals_alt;
    DevBuf<uint32_t> fhc;  // per pair: hits of the core bins extended by j bins either side, j < kFltHc
    DevBuf<uint8_t> fdone;
    DevBuf<int> fwd_slots2;
    DevBuf<uint8_t> cub_tmp_fwd2;
    int* d_hist2 = nullptr;
    unsigned long long* d_fcount = nullptr;  // [0] pairs given a band, [1] pairs the filter passed on, [2] certificates that failed
    DevBuf<uint16_t> anchor;  // per pair: seed estimate of the alignment's first window column / 4 (scheduling hint)
    DevBuf<uint8_t> wflags;  // per pair: window contains N (xor-selector flavour)
    DevBuf<uint8_t> tb_pool;
    DevBuf<int> tb_esc;   // escalation lists between the traceback frame classes
    DevBuf<int> tb_band;  // per pair: band the traceback reached so far (0 = not started)
    // device-resident reference genome (cudasw_reference_load) and the window descriptors of the current batch
    DevBuf<uint32_t> ref2bit, refN;
    DevBuf<int64_t> d_contig_off, d_seg_off;
    DevBuf<WindowSeg> d_segs;
    // queries assembled on the device (cudasw_breakpoint_homology): their segments and the literal bytes between them
    DevBuf<WindowSeg> d_qsegs;
    DevBuf<int64_t> d_qseg_off;
    DevBuf<uint8_t> d_lit;
    PinBuf<WindowSeg> hom_qs, hom_ws;
    PinBuf<int64_t> hom_seq_off, hom_ref_off;
    PinBuf<uint8_t> hom_lit;
    PinBuf<cudasw_result> hom_res;
    std::mutex hom_mu;  // one cudasw_breakpoint_homology call at a time owns the staging above
    std::vector<int64_t> contig_off;  // n_contigs + 1, linear base coordinates
    int* d_hist = nullptr;  // 256 ints: pairs per (flavour, variant) slot
    uint32_t* d_pool = nullptr;  // {poolA, poolB}
    unsigned long long* d_counters = nullptr;  // [0] pool used, [1] cigar used, [2] retry count(int)
    int* h_hist = nullptr;                     // pinned
    unsigned long long* h_counters = nullptr;  // pinned
    size_t tb_pool_bytes = 0;

    // async trio state
    std::vector<uint8_t> pend_seq, pend_ref;
    std::vector<int64_t> pend_seq_off{0}, pend_ref_off{0};
    std::vector<int64_t> pend_tickets;
    int64_t next_ticket = 1;
    struct Done {
        cudasw_result r;
        std::vector<uint32_t> cigar;
    };
    std::unordered_map<int64_t, Done> done;
    std::unordered_set<int64_t> live;  // submitted and not yet delivered (pending, on the GPU, or done)
    std::mutex qmu;                    // guards the queue state above; `mu` guards the engine itself
    std::condition_variable cv_work, cv_done;
    std::thread disp;
    bool disp_run = false;
    int64_t disp_max_pairs = 0, disp_delay_us = 0;
    std::chrono::steady_clock::time_point first_pending;

    std::mutex err_mu;           // guards `error` (written by whichever thread fails, read by cudasw_last_error)
    std::shared_mutex ref_mu;    // the loaded reference (contig_off, ref2bit, refN): shared by batches, exclusive for a (re)load
    void set_error(const std::string& s) { std::lock_guard<std::mutex> l(err_mu); error = s; }
    void set_error_locked(const std::string& s) { set_error(s); }
    int pack_chunk(size_t ci, int64_t first, int64_t cnt, const uint8_t* d_seq, const int64_t* d_so, const uint8_t* d_ref,
                   const int64_t* d_ro, int64_t q_lo, int64_t q_hi, int64_t w_lo, int64_t w_hi, bool gather,
                   uint64_t* launches_out, bool qgather, int q_stride, int w_stride, bool packed, const int64_t* h_seq_off,
                   const int64_t* h_ref_off);
    int forward_chunk(size_t ci, int64_t first, int64_t cnt, const int64_t* d_so, const int64_t* d_ro, uint64_t* launches_out);
    int post_chunk(size_t ci, int64_t first, int64_t n, const int64_t* d_so, const int64_t* d_ro, int64_t cigar_cap, uint64_t* launches_out,
                   bool ends_only = false);
    int post_chunk_ja(size_t ci, int64_t first, int64_t n, const int64_t* d_so, const int64_t* d_ro, int64_t cigar_cap, uint64_t* launches_out);
    int tie_break = CUDASW_TIEBREAK_SSW;
    int size_strip_buffers(unsigned long long need_f, unsigned long long need_r);
    void scan_batch(const std::vector<std::pair<int64_t, int64_t>>& ranges, const int64_t* h_seq_off, const int64_t* h_ref_off);
    std::vector<unsigned long long> strip_need;  // per chunk: boundary-row bytes of its strip-mined pairs
    float last_post_ms = 0.f;
    int run(int64_t n, const uint8_t* d_seq, const int64_t* d_seq_off, const uint8_t* d_ref,
            const int64_t* d_ref_off, int64_t total_q, int64_t total_w, cudasw_result* out,
            uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used, bool inputs_on_host,
            const uint8_t* h_seq, const int64_t* h_seq_off, const uint8_t* h_ref,
            const int64_t* h_ref_off, const WindowSeg* h_segs = nullptr, const int64_t* h_seg_off = nullptr,
            const QuerySegs* qsegs = nullptr, bool ends_only = false, const PackedIn* packed = nullptr);
};

static uint32_t pack2(int v) { return ((uint32_t)v & 0xFFFFu) | (((uint32_t)v & 0xFFFFu) << 16); }

extern "C" int cudasw_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" const char* cudasw_version(void) { return "cudasw 0.1 (sm_100a)"; }

extern "C" const char* cudasw_last_error(cudasw_t* h) {
    if (!h) return g_create_error.c_str();
    // a copy per calling thread: the handle's string may be rewritten by another thread's batch at any time
    thread_local std::string copy;
    { std::lock_guard<std::mutex> l(h->err_mu); copy = h->error; }
    return copy.c_str();
}

extern "C" int cudasw_create(const cudasw_scoring* s, int device, cudasw_t** out) {
    return cudasw_create_ex(s, device, CUDASW_TIEBREAK_SSW, out);
}

extern "C" int cudasw_create_ex(const cudasw_scoring* s, int device, int tie_break, cudasw_t** out) {
    if (!out) return CUDASW_EINVAL;
    *out = nullptr;
    if (tie_break != CUDASW_TIEBREAK_SSW && tie_break != CUDASW_TIEBREAK_JALIGNER) {
        g_create_error = "unknown tie-break mode";
        return CUDASW_EINVAL;
    }
    cudasw_scoring sc = {1, -4, -4, 6, 1};  // AlignerFactory.java:90
    if (s) sc = *s;
    // ssw.Aligner.align argument checks (jar bytecode)
    if (sc.gap_open < 0 || sc.gap_open > 255 || sc.gap_extend < 0 || sc.gap_extend > 255) {
        g_create_error = "Gap open and extension penalties must be in [0, 255]";
        return CUDASW_EINVAL;
    }
    if (std::abs(sc.match) > 127 || std::abs(sc.mismatch) > 127 || std::abs(sc.ambiguous) > 127) {
        g_create_error = "Scoring matrix values must fit into signed 8-bit integer";
        return CUDASW_EINVAL;
    }
    if (sc.mismatch > 0 || sc.ambiguous > 0 || sc.match < 0) {
        g_create_error = "cudasw requires match >= 0 and mismatch, ambiguous <= 0";
        return CUDASW_EINVAL;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) +
                         " (libcudasw has no CPU fallback)";
        return CUDASW_ECUDA;
    }
    if (device < 0 || device >= ndev) {
        g_create_error = "device index out of range";
        return CUDASW_EINVAL;
    }
    std::call_once(g_var_once, [] { g_var = cusw::collect_variants(); g_nvar = (int)g_var.size(); });

    cudasw_t* h = new cudasw();
    h->device = device;
    h->scoring = sc;
    h->tie_break = tie_break;
    auto fail = [&](const std::string& m) {
        g_create_error = m;
        delete h;
        return CUDASW_ECUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(cudaGetErrorString(e));
    if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess)
        return fail(cudaGetErrorString(e));
    if ((e = cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking)) != cudaSuccess)
        return fail(cudaGetErrorString(e));
    for (auto& ev : h->ev)
        if ((e = cudaEventCreate(&ev)) != cudaSuccess) return fail(cudaGetErrorString(e));
    {
        // pack and post work is latency-bound and small next to the forward pass it overlaps: give it priority so its
        // CTAs take the SM slots that the forward kernel frees
        int lo_pri = 0, hi_pri = 0;
        cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri);
        h->fwd_set.main = h->stream;
        if ((e = cudaStreamCreateWithPriority(&h->post_set.main, cudaStreamNonBlocking, hi_pri)) != cudaSuccess) return fail(cudaGetErrorString(e));
        if ((e = cudaStreamCreateWithPriority(&h->pack_stream, cudaStreamNonBlocking, hi_pri)) != cudaSuccess) return fail(cudaGetErrorString(e));
        for (auto* ss : {&h->fwd_set, &h->post_set}) {
            for (auto& a : ss->aux) {
                e = ss == &h->post_set ? cudaStreamCreateWithPriority(&a, cudaStreamNonBlocking, hi_pri) : cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking);
                if (e != cudaSuccess) return fail(cudaGetErrorString(e));
            }
            if ((e = cudaEventCreateWithFlags(&ss->fork_ev, cudaEventDisableTiming)) != cudaSuccess) return fail(cudaGetErrorString(e));
            for (auto& j : ss->join_ev)
                if ((e = cudaEventCreateWithFlags(&j, cudaEventDisableTiming)) != cudaSuccess) return fail(cudaGetErrorString(e));
        }
    }
    {
        int lo_pri = 0, hi_pri = 0;
        cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri);
        for (int k = 0; k < 5; k++) {
            if ((e = cudaStreamCreateWithPriority(&h->tb_streams[k], cudaStreamNonBlocking, hi_pri)) != cudaSuccess) return fail(cudaGetErrorString(e));
            if ((e = cudaEventCreateWithFlags(&h->tb_join[k], cudaEventDisableTiming)) != cudaSuccess) return fail(cudaGetErrorString(e));
        }
    }
    if ((e = cudaMalloc(&h->d_strip, 4 * sizeof(unsigned long long))) != cudaSuccess) return fail(cudaGetErrorString(e));
    if ((e = cudaMalloc(&h->d_hist, 256 * sizeof(int))) != cudaSuccess) return fail(cudaGetErrorString(e));
    if ((e = cudaMalloc(&h->d_rev_slots, 64 * sizeof(int))) != cudaSuccess) return fail(cudaGetErrorString(e));
    if ((e = cudaMalloc(&h->d_hist2, 256 * sizeof(int))) != cudaSuccess) return fail(cudaGetErrorString(e));
    if ((e = cudaMalloc(&h->d_band_slots, 32 * sizeof(int))) != cudaSuccess) return fail(cudaGetErrorString(e));
    if ((e = cudaMalloc(&h->d_fcount, 4 * sizeof(unsigned long long))) != cudaSuccess) return fail(cudaGetErrorString(e));
    cudaMemset(h->d_fcount, 0, 4 * sizeof(unsigned long long));
    h->fc = make_filter_const(sc.match, sc.mismatch, sc.ambiguous, sc.gap_open, sc.gap_extend);
    if (const char* fe = getenv("CUDASW_FILTER_EST")) h->fc.est_pct = std::min(std::max(atoi(fe), 10), 200);
    {
        const char* fe = getenv("CUDASW_FILTER");  // 0: every pair sweeps its whole window (the pre-filter pipeline)
        h->use_filter = h->fc.ok && !(fe && atoi(fe) == 0);
        h->filter_always = fe && atoi(fe) == 2;
        // CUDASW_BAND=1 (experiment, measured slower): certifiable pairs sweep their diagonal band with the strip
        // kernels of the reverse pass run forwards, instead of a column range with the wavefront kernels
        const char* be = getenv("CUDASW_BAND");
        h->fc.band = be && atoi(be) == 1;
        if ((e = cudaFuncSetAttribute(filter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kFltWarps * kFltSmemPerWarp))) != cudaSuccess)
            return fail(cudaGetErrorString(e));
    }
    if ((e = cudaMalloc(&h->d_counters, 8 * sizeof(unsigned long long))) != cudaSuccess)
        return fail(cudaGetErrorString(e));
    if ((e = cudaMallocHost(&h->h_hist, 256 * sizeof(int))) != cudaSuccess) return fail(cudaGetErrorString(e));
    if ((e = cudaMallocHost(&h->h_cells, 8 * sizeof(unsigned long long))) != cudaSuccess) return fail(cudaGetErrorString(e));
    if ((e = cudaMallocHost(&h->h_counters, 8 * sizeof(unsigned long long))) != cudaSuccess)
        return fail(cudaGetErrorString(e));

    // PRMT pool: x = qcode ^ wcode; 0 -> match, 1..3 -> mismatch, 4..7 -> ambiguous
    auto b = [](int v) { return (uint32_t)(uint8_t)(int8_t)v; };
    h->sc.poolA = b(sc.match) | (b(sc.mismatch) << 8) | (b(sc.mismatch) << 16) | (b(sc.mismatch) << 24);
    h->sc.poolB = b(sc.ambiguous) * 0x01010101u;
    h->sc.B = sc.gap_open;
    h->sc.BB = pack2(sc.gap_open);
    h->sc.GO2 = pack2(sc.gap_open);
    h->sc.GOBB = pack2(2 * sc.gap_open);
    h->sc.NGE2 = pack2(-sc.gap_extend);
    h->sc.ge = sc.gap_extend;
    h->sc.match = sc.match;
    {
        // per lane: the packed constants, then the query profile words: byte x of profile[c] = score of query code c
        // (A0 C1 G2 T3, N4, pads 5..7) against window base x (SswJniAligner.createMatrix, SswJniAligner.java:19-33)
        uint32_t pool[32 * 16];
        uint32_t one[16] = {h->sc.poolA, h->sc.poolB, h->sc.BB, h->sc.GO2, h->sc.NGE2, h->sc.GOBB, 0u, 0u};
        for (int c = 0; c < 8; c++) {
            uint32_t wd = 0;
            for (int x = 0; x < 4; x++) wd |= b(c < 4 ? (c == x ? sc.match : sc.mismatch) : sc.ambiguous) << (8 * x);
            one[8 + c] = wd;
        }
        for (int l = 0; l < 32; l++) memcpy(pool + 16 * l, one, sizeof one);
        if ((e = cudaMalloc(&h->d_pool, sizeof pool + 64)) != cudaSuccess) return fail(cudaGetErrorString(e));  // + 8 step-statistics words (CUSW_STEP_STATS builds)
        cudaMemset(h->d_pool, 0, sizeof pool + 64);
        if ((e = cudaMemcpy(h->d_pool, pool, sizeof pool, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(cudaGetErrorString(e));
        h->sc.pool_dev = h->d_pool;
    }
    h->tc = TbConst{sc.match, sc.mismatch, sc.ambiguous, sc.gap_open, sc.gap_extend, h->sc.poolA, h->sc.poolB};
    if (2 * g_nvar + 1 > 127) { cudasw_destroy(h); g_create_error = "too many variants"; return CUDASW_EINVAL; }
    h->vt.n = g_nvar;
    h->vt.n_single = 0;
    for (int i = 0; i < g_nvar; i++) {
        h->vt.rows[i] = g_var[i].rows;
        h->vt.lanes[i] = g_var[i].G;
        if (!g_var[i].multi) h->vt.n_single = i + 1;
    }

    h->var_by_q.assign(32768, 0);
    for (int q = 1; q < 32768; q++) h->var_by_q[(size_t)q] = (int16_t)variant_of(h->vt, q);

    // self-test, like AlignerFactory's "A" vs "A" probe (AlignerFactory.java:74-84)
#ifndef CUSW_EXPERIMENT_NO_SELFTEST
    {
        const uint8_t a[1] = {'A'};
        const int64_t off[2] = {0, 1};
        cudasw_result r;
        uint32_t cg[4];
        int64_t used = 0;
        int rc = cudasw_align_batch(h, 1, a, off, a, off, &r, cg, 4, &used);
        if (rc != CUDASW_OK || r.status != 0 || r.score != sc.match || r.cigar_len != 1 || cg[0] != (1u << 4)) {
            std::string m = std::string("self-test failed: ") + cudasw_last_error(h);
            cudasw_destroy(h);
            g_create_error = m;
            return CUDASW_ECUDA;
        }
    }
#endif
    *out = h;
    return CUDASW_OK;
}

extern "C" void cudasw_destroy(cudasw_t* h) {
    if (!h) return;
    cudasw_dispatcher_stop(h);
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    h->raw_seq.release(); h->raw_ref.release(); h->enc_seq.release(); h->enc_ref.release();
    h->pk_seq2.release(); h->pk_seqN.release(); h->pk_ref2.release(); h->pk_refN.release();
    h->seq_off.release(); h->ref_off.release(); h->keys.release(); h->keys_alt.release();
    h->vals.release(); h->vals_alt.release(); h->retry_a.release(); h->retry_b.release();
    h->fwd.release(); h->rev.release(); h->res.release(); h->cig.release();
    h->cub_tmp.release(); h->cub_tmp_post.release(); h->fwd_slots.release(); h->tb_pool.release(); h->tb_band.release(); h->tb_esc.release(); h->wflags.release(); h->anchor.release();
    h->ref2bit.release(); h->refN.release(); h->d_contig_off.release(); h->d_seg_off.release(); h->d_segs.release();
    h->d_qsegs.release(); h->d_qseg_off.release(); h->d_lit.release();
    h->hom_qs.release(); h->hom_ws.release(); h->hom_seq_off.release(); h->hom_ref_off.release(); h->hom_lit.release(); h->hom_res.release();
    if (h->d_hist) cudaFree(h->d_hist);
    if (h->d_hist2) cudaFree(h->d_hist2);
    if (h->d_band_slots) cudaFree(h->d_band_slots);
    for (auto& r : h->fb_ring) r.release();
    if (h->d_fcount) cudaFree(h->d_fcount);
    h->frec.release(); h->fhc.release(); h->fdone.release(); h->bkeys.release(); h->bkeys_alt.release(); h->bvals.release(); h->bvals_alt.release(); h->fwd_slots2.release(); h->cub_tmp_fwd2.release();
    if (h->d_rev_slots) cudaFree(h->d_rev_slots);
    for (auto& r : h->rev_ring) r.release();
    if (h->d_pool) cudaFree(h->d_pool);
    if (h->d_counters) cudaFree(h->d_counters);
    if (h->h_hist) cudaFreeHost(h->h_hist);
    if (h->h_counters) cudaFreeHost(h->h_counters);
    for (auto& ev : h->ev)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : h->h2d_done) cudaEventDestroy(ev);
    for (auto* ss : {&h->fwd_set, &h->post_set}) {
        for (auto& a : ss->aux) if (a) { cudaStreamSynchronize(a); cudaStreamDestroy(a); }
        if (ss->fork_ev) cudaEventDestroy(ss->fork_ev);
        for (auto& j : ss->join_ev) if (j) cudaEventDestroy(j);
    }
    for (int k = 0; k < 5; k++) {
        if (h->tb_streams[k]) { cudaStreamSynchronize(h->tb_streams[k]); cudaStreamDestroy(h->tb_streams[k]); }
        if (h->tb_join[k]) cudaEventDestroy(h->tb_join[k]);
    }
    if (h->post_set.main) { cudaStreamSynchronize(h->post_set.main); cudaStreamDestroy(h->post_set.main); }
    if (h->pack_stream) { cudaStreamSynchronize(h->pack_stream); cudaStreamDestroy(h->pack_stream); }
    for (auto& c : h->chunk_ev) { cudaEventDestroy(c.pack0); cudaEventDestroy(c.enc1); cudaEventDestroy(c.pack1); cudaEventDestroy(c.fwd0); cudaEventDestroy(c.fwd1); }
    if (h->d_strip) cudaFree(h->d_strip);
    if (h->h_cells) cudaFreeHost(h->h_cells);
    h->strip_fwd.release(); h->strip_rev.release();
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

// Launches the score kernel of every variant that has pairs (sorted order in `order`). With several variants the
// launches are spread over auxiliary streams so that small launches overlap (fork/join by events).
// dev_slots != nullptr (forward pass): `counts[v]` = pairs whose query length falls into variant v, counted by the host
// from the offsets it already holds; it sizes the grids of both flavours of the variant, while the kernels read the true
// (start, count) of their list from dev_slots (filled by slot_scan_kernel), so the host never waits for the device bins.
// dev_slots == nullptr (CUDASW_REV_FULL reverse pass): counts[slot] are the exact per-slot counts read back from the device.
int cudasw::launch_variants(StreamSet& ss, bool reverse, const int* order, PairIO io, const int4* fwd_c, int4* out_c,
                            StripPool spool, uint64_t* launches, const int* counts, const int* dev_slots, bool maybe_x) {
    cudaStream_t st = ss.main;
    int nonempty = 0;
    for (int v = 0; v < (dev_slots ? g_nvar : 2 * g_nvar); v++) nonempty += counts[v] > 0;
    static const bool no_spread = getenv("CUDASW_NO_SPREAD") != nullptr;  // experiment hook: variant launches one after the other
    const bool spread = nonempty > 1 && !no_spread;
    // Quiet-chain flavour of the forward kernels: faster launch by launch (C2 -11 %, C3's variants -2.5 % serialised),
    // slower when many different variants share the SMs (C3 with 52 variants side by side: +7.5 %; larger code and
    // register footprints next to each other, profiles/c3_probe_r01j.txt). CUDASW_CHAIN=0/1 forces either flavour.
    static const char* chain_env = getenv("CUDASW_CHAIN");
    const bool chained = chain_env ? atoi(chain_env) != 0 : nonempty <= kAux;
    if (spread) CK(cudaEventRecord(ss.fork_ev, st));
    bool used[kAux] = {};
    int64_t start = 0;
    int k = 0;
    for (int slot = 0; slot < 2 * g_nvar; slot++) {  // [0, nvar): query-profile flavour, [nvar, 2 nvar): xor-selector flavour
        const bool xr = slot >= g_nvar;
        const int c = dev_slots ? ((xr && !maybe_x) ? 0 : counts[slot % g_nvar]) : counts[slot];
        if (c > 0) {
            const cusw::Variant& V = g_var[slot % g_nvar];
            cudaStream_t s = st;
            if (spread) {
                const int a = k++ % kAux;
                s = ss.aux[a];
                if (!used[a]) { CK(cudaStreamWaitEvent(s, ss.fork_ev, 0)); used[a] = true; }
            }
            (reverse ? (xr ? V.revx : V.rev) : (xr ? V.fwdx : chained ? V.fwdc : V.fwd))(dev_slots ? order : order + start, c, io, fwd_c, out_c, sc, spool, s,
                                                                                      dev_slots ? dev_slots + 2 * slot : nullptr);
            (*launches)++;
        }
        if (!dev_slots) start += c;
    }
    for (int a = 0; a < kAux; a++)
        if (used[a]) {
            CK(cudaEventRecord(ss.join_ev[a], ss.aux[a]));
            CK(cudaStreamWaitEvent(st, ss.join_ev[a], 0));
        }
    return CUDASW_OK;
}

// Chunk pipeline. Blob offsets are absolute, so a chunk is a shifted view of the per-pair arrays; every per-pair
// scratch array (keys, order, results) is used at offset `first`, which keeps concurrently running chunks apart.
//   pack_chunk    (pack stream)    encode -> window flags / gather -> seed anchors -> bin; ends with a host sync
//   forward_chunk (compute stream) forward score pass of every variant that has pairs
//   post_chunk    (post stream)    bin -> reverse pass -> traceback + CIGARs; host-synchronised
// The host issues pack(i), forward(i), post(i-1): the ALU-bound forward kernels of chunk i then share the SMs with
// the memory-bound pack of chunk i+1 and the latency-bound reverse/traceback kernels of chunk i-1. pack and post
// never overlap each other (both are host-synchronised), so they share d_hist, h_hist and the sort workspace.
int cudasw::pack_chunk(size_t ci, int64_t first, int64_t cnt, const uint8_t* d_seq, const int64_t* d_so,
                       const uint8_t* d_ref, const int64_t* d_ro, int64_t q_lo, int64_t q_hi,
                       int64_t w_lo, int64_t w_hi, bool gather, uint64_t* launches_out, bool qgather, int q_stride, int w_stride,
                       bool packed, const int64_t* h_seq_off, const int64_t* h_ref_off) {
    cudaStream_t st = pack_stream;
    uint64_t launches = 0;
    const int64_t n = cnt;
    int* any_n = reinterpret_cast<int*>(d_counters + 5);  // set by encode_kernel when a window holds an N
    CK(cudaEventRecord(chunk_ev[ci].pack0, st));
    {
        const int64_t tq = q_hi - q_lo, tw = w_hi - w_lo;
        if (packed) {
            // 2 bit + N mask words -> the encoded byte layout, 16 bases per thread
            int64_t thr = (tq + 15) / 16 + 1;
            int blocks = (int)std::min<int64_t>((thr + 255) / 256, 148 * 16);
            if (tq > 0) { unpack_kernel<<<std::max(blocks, 1), 256, 0, st>>>(pk_seq2.p, pk_has_qn ? pk_seqN.p : nullptr, pk_word0_q, q_lo, q_hi, enc_seq.p, 0u, nullptr); launches++; }
            thr = (tw + 15) / 16 + 1;
            blocks = (int)std::min<int64_t>((thr + 255) / 256, 148 * 16);
            if (tw > 0) { unpack_kernel<<<std::max(blocks, 1), 256, 0, st>>>(pk_ref2.p, pk_has_wn ? pk_refN.p : nullptr, pk_word0_w, w_lo, w_hi, enc_ref.p, 0x80u, any_n); launches++; }
        }
        int64_t thr = (tq + 31) / 32;
        int blocks = (int)std::min<int64_t>((thr + 255) / 256, 148 * 16);
        if (packed) {}
        else if (tq > 0 && !qgather) { encode_kernel<<<std::max(blocks, 1), 256, 0, st>>>(d_seq + q_lo, enc_seq.p + q_lo, tq, 0u, nullptr); launches++; }
        thr = (tw + 31) / 32;
        blocks = (int)std::min<int64_t>((thr + 255) / 256, 148 * 16);
        if (tw > 0 && !gather && !packed) { encode_kernel<<<std::max(blocks, 1), 256, 0, st>>>(d_ref + w_lo, enc_ref.p + w_lo, tw, 0x80u, any_n); launches++; }
    }
    PairIO io{enc_seq.p, enc_ref.p, d_so + first, d_ro + first};
    if (qgather) {
        // queries assembled from reference segments and literal bytes, in the query encoding
        window_gather_kernel<<<(int)((n * 32 + 255) / 256), 256, 0, st>>>((int)n, d_qsegs.p + (q_stride ? first * q_stride : 0),
                                                                         q_stride ? nullptr : d_qseg_off.p + first, io.seq_off,
                                                                         d_contig_off.p, ref2bit.p, refN.p, enc_seq.p,
                                                                         nullptr, nullptr, 0, d_lit.p, 0u, q_stride);
        launches++;
    }
    if (gather) {
        // windows come out of the device-resident 2-bit reference already encoded, with their N flags
        window_gather_kernel<<<(int)((n * 32 + 255) / 256), 256, 0, st>>>((int)n, d_segs.p + (w_stride ? first * w_stride : 0),
                                                                         w_stride ? nullptr : d_seg_off.p + first, io.ref_off,
                                                                         d_contig_off.p, ref2bit.p, refN.p, enc_ref.p,
                                                                         wflags.p + first, any_n, 0, nullptr, 0x80u, w_stride);
    } else {
        CK(cudaEventRecord(chunk_ev[ci].enc1, st));
        window_flags_kernel<<<(int)((n * 32 + 255) / 256), 256, 0, st>>>((int)n, enc_ref.p, io.ref_off, any_n, wflags.p + first);
    }
    if (gather) CK(cudaEventRecord(chunk_ev[ci].enc1, st));
    // 1 byte read + 1 written per base; a gathered window reads 3 bits and writes 1 byte per base
    if (packed) stats.encode_bytes += ((uint64_t)(q_hi - q_lo) + (uint64_t)(w_hi - w_lo)) * 5ull / 4ull + (pk_has_qn ? (uint64_t)(q_hi - q_lo) / 8 : 0) + (pk_has_wn ? (uint64_t)(w_hi - w_lo) / 8 : 0);
    else stats.encode_bytes += 2ull * (uint64_t)(q_hi - q_lo) + (gather ? (uint64_t)(w_hi - w_lo) * 11ull / 8ull : 2ull * (uint64_t)(w_hi - w_lo));
    launches++;
    const bool have_multi = g_var[g_nvar - 1].multi;
    const int max_rows = std::min(have_multi ? 32767 : g_var[g_nvar - 1].rows,
                                  scoring.match > 0 ? (32767 - scoring.gap_open) / scoring.match : 32767);
    const int max_cols = (1 << 24) - 2;
    size_t cub_bytes = cub_tmp.cap;
    CK(cudaMemsetAsync(d_hist, 0, 256 * sizeof(int), st));
    const bool flt = use_filter && chunk_filter[ci];
    if (use_filter && !flt)  // every record names the whole window (the reverse pass reads the records of the whole batch)
        CK(cudaMemsetAsync(frec.p + first, 0xFF, (size_t)n * sizeof(FilterRec), st));
    if (flt) {
        // per pair: hits per diagonal bin -> the column range the forward pass sweeps first (band_filter.cuh)
        const FilterSmem fsm = make_filter_smem(std::min<int64_t>(batch_max_q, kFltMaxQ), std::min<int64_t>(batch_max_w, (int64_t)kFltBinW * kFltMaxBins));
        filter_kernel<<<(int)std::min<int64_t>((n + kFltWarps - 1) / kFltWarps, 148 * 2 * 8), 32 * kFltWarps, kFltWarps * flt_smem_per_warp(fsm), st>>>(
            (int)n, io, any_n, wflags.p + first, fc, frec.p + first, fhc.p + (size_t)first * kFltHc, d_fcount, fsm);
        launches++;
    }
    seed_anchor_kernel<<<(int)std::min<int64_t>((n + kSeedWarps - 1) / kSeedWarps, 148 * 5 * 4), 32 * kSeedWarps, 0, st>>>(
        (int)n, enc_seq.p, enc_ref.p, io.seq_off, io.ref_off, anchor.p + first, flt ? frec.p + first : nullptr);
    launches++;
    fwd_keys_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>((int)n, io.seq_off, io.ref_off, vt, max_rows, max_cols,
                                                           keys.p + first, vals.p + first, d_hist, res.p + first, any_n,
                                                           wflags.p + first, anchor.p + first, d_strip + 2,
                                                           flt ? frec.p + first : nullptr, nullptr, 1, flt && fc.band ? 1 : 0);
    launches++;
    CK(cub::DeviceRadixSort::SortPairs(cub_tmp.p, cub_bytes, keys.p + first, keys_alt.p + first, vals.p + first,
                                       vals_alt.p + first, (int)n, 0, 32, st));
    launches += 3;
    // where every (flavour, variant) slot starts in the sorted order: stays on the device (the forward kernels read it)
    slot_scan_kernel<<<1, 32, 0, st>>>(d_hist, 2 * g_nvar + 1, fwd_slots.p + ci * (size_t)(2 * (2 * g_nvar + 1)));
    launches++;
    CK(cudaEventRecord(chunk_ev[ci].pack1, st));
    CK(cudaGetLastError());
    (void)h_seq_off; (void)h_ref_off;  // (the per-class pair counts come from scan_batch)
    *launches_out += launches;
    return CUDASW_OK;
}

// Forward score pass of one packed chunk (fwd_counts / fwd_slots hold its bins). Waits for the pack by event.
int cudasw::forward_chunk(size_t ci, int64_t first, int64_t cnt, const int64_t* d_so, const int64_t* d_ro, uint64_t* launches_out) {
    (void)cnt;
    cudaStream_t st = stream;
    uint64_t launches = 0;
    PairIO io{enc_seq.p, enc_ref.p, d_so + first, d_ro + first};
    CK(cudaStreamWaitEvent(st, chunk_ev[ci].pack1, 0));
    CK(cudaEventRecord(chunk_ev[ci].fwd0, st));
    CK(cudaMemsetAsync(d_strip, 0, sizeof(unsigned long long), st));
    StripPool spool{strip_fwd.p, (unsigned long long)strip_fwd.cap, d_strip};
    const bool flt = use_filter && chunk_filter[ci];
    if (flt) io.sub = frec.p + first;
    if (flt && fc.band) {
        // Pairs with a diagonal band: binned by strip height on the device, swept by the persistent banded kernels (the
        // forward flavour of the reverse pass's strips) side by side on the auxiliary streams. Uses the post stage's key
        // arrays of this chunk's range (free until the reverse pass) and the second pass's sort workspace.
        const int64_t n = cnt;
        CK(cudaMemsetAsync(d_hist2, 0, 256 * sizeof(int), st));
        band_keys_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>((int)n, io.seq_off, io.ref_off, frec.p + first, n < 4096 ? 1 : 0,
                                                                bkeys.p + first, bvals.p + first, d_hist2, d_strip + 2);
        size_t cub_bytes = cub_tmp_fwd2.cap;
        CK(cub::DeviceRadixSort::SortPairs(cub_tmp_fwd2.p, cub_bytes, bkeys.p + first, bkeys_alt.p + first, bvals.p + first,
                                           bvals_alt.p + first, (int)n, 0, 24, st));
        slot_scan_kernel<<<1, 32, 0, st>>>(d_hist2, kBandVariants + 1, d_band_slots);
        launches += 5;
        CK(cudaEventRecord(fwd_set.fork_ev, st));
        const long long kBandMaxWd = kFltBandMaxWd;
        for (int v = 0; v < kBandVariants; v++) {
            const int gpb = 128 / rev_variant_lanes(v);
            long long half = std::min<long long>(band_variant_rows(v) + kBandMaxWd + 8, batch_max_w + 8);
            long long blocks = std::min<long long>((n + gpb - 1) / gpb, 148 * 4);
            const long long per_block = (long long)gpb * 2 * half * 4;
            blocks = std::max<long long>(1, std::min<long long>(blocks, ((long long)1 << 30) / std::max<long long>(per_block, 1)));
            CK(fb_ring[v].ensure((size_t)(blocks * gpb * 2 * half)));
            cudaStream_t sx = fwd_set.aux[v % kAux];
            if (v < kAux) CK(cudaStreamWaitEvent(sx, fwd_set.fork_ev, 0));
            cusw::launch_fwd_band(v, (int)blocks, bvals_alt.p + first, d_band_slots + 2 * v, io, fwd.p + first, sc, fb_ring[v].p, (int)half, sx);
            launches++;
        }
        for (int a = 0; a < kAux; a++) {
            CK(cudaEventRecord(fwd_set.join_ev[a], fwd_set.aux[a]));
            CK(cudaStreamWaitEvent(st, fwd_set.join_ev[a], 0));
        }
    }
    { int rc = launch_variants(fwd_set, false, vals_alt.p + first, io, nullptr, fwd.p + first, spool, &launches, fwd_counts[ci].data(),
                               fwd_slots.p + ci * (size_t)(2 * (2 * g_nvar + 1)), batch_maybe_n); if (rc != CUDASW_OK) return rc; }
    if (flt) {
        // Which sub-window results are proven (band_filter.cuh)? The others sweep their whole window now: binned again
        // (everything proven lands in the "nothing to do" slot) and launched over the same host-side upper bounds.
        const int64_t n = cnt;
        const bool have_multi = g_var[g_nvar - 1].multi;
        const int max_rows = std::min(have_multi ? 32767 : g_var[g_nvar - 1].rows,
                                      scoring.match > 0 ? (32767 - scoring.gap_open) / scoring.match : 32767);
        const int max_cols = (1 << 24) - 2;
        int* any_n = reinterpret_cast<int*>(d_counters + 5);
        int* slots2 = fwd_slots2.p + ci * (size_t)(2 * (2 * g_nvar + 1));
        filter_certify_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>((int)n, io.seq_off, io.ref_off, fwd.p + first, fc, frec.p + first,
                                                                     fhc.p + (size_t)first * kFltHc, fdone.p + first, d_fcount);
        CK(cudaMemsetAsync(d_hist2, 0, 256 * sizeof(int), st));
        fwd_keys_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>((int)n, io.seq_off, io.ref_off, vt, max_rows, max_cols,
                                                               keys.p + first, vals.p + first, d_hist2, res.p + first, any_n,
                                                               wflags.p + first, anchor.p + first, d_strip + 2, frec.p + first, fdone.p + first, 0, 0);
        size_t cub_bytes = cub_tmp_fwd2.cap;
        CK(cub::DeviceRadixSort::SortPairs(cub_tmp_fwd2.p, cub_bytes, keys.p + first, keys_alt.p + first, vals.p + first,
                                           vals_alt.p + first, (int)n, 0, 32, st));
        slot_scan_kernel<<<1, 32, 0, st>>>(d_hist2, 2 * g_nvar + 1, slots2);
        launches += 6;
        { int rc = launch_variants(fwd_set, false, vals_alt.p + first, io, nullptr, fwd.p + first, spool, &launches, fwd_counts[ci].data(),
                                   slots2, batch_maybe_n); if (rc != CUDASW_OK) return rc; }
    }
    CK(cudaEventRecord(chunk_ev[ci].fwd1, st));
    CK(cudaGetLastError());
    *launches_out += launches;
    return CUDASW_OK;
}

// Reverse pass + traceback of one chunk whose forward pass has been queued (waits for it by event).
int cudasw::post_chunk(size_t ci, int64_t first, int64_t n, const int64_t* d_so, const int64_t* d_ro, int64_t cigar_cap, uint64_t* launches_out,
                       bool ends_only) {
    cudaStream_t st = post_set.main;
    CK(cudaStreamWaitEvent(st, chunk_ev[ci].fwd1, 0));
    cudaStream_t* aux = post_set.aux;
    cudaEvent_t fork_ev = post_set.fork_ev;
    cudaEvent_t* join_ev = post_set.join_ev;
    uint64_t launches = 0;
    PairIO io{enc_seq.p, enc_ref.p, d_so + first, d_ro + first};
    if (use_filter) io.sub = frec.p + first;  // the banded reverse pass reads the certified gap bounds
    int4* fwd_c = fwd.p + first;
    int4* rev_c = rev.p + first;
    cudasw_result* res_c = res.p + first;
    DevBuf<uint32_t> keys = this->keys, keys_alt = this->keys_alt;  // shifted views (never released here)
    DevBuf<int> vals = this->vals, vals_alt = this->vals_alt, retry_a = this->retry_a, retry_b = this->retry_b, tb_band = this->tb_band;
    keys.p += first; keys_alt.p += first; vals.p += first; vals_alt.p += first; retry_a.p += first; retry_b.p += first; tb_band.p += first;
    const uint8_t* wflags_c = wflags.p + first;
    size_t cub_bytes = cub_tmp_post.cap;
    StripPool spool{strip_rev.p, (unsigned long long)strip_rev.cap, d_strip + 1};
    CK(cudaEventRecord(ev[3], st));

    // ---- reverse score pass
    const bool rev_full = getenv("CUDASW_REV_FULL") != nullptr;  // experiment hook: round 1's full-rectangle kernels
    if (!rev_full) {
        // Banded, strip-mined, persistent kernels (rev_band.cuh): binned on the device, job list bounds read from device
        // memory, so nothing here waits for the host.
        int* dhr = d_hist + 160;  // this pass's own histogram (the pack of a later chunk may be using d_hist[0 ..))
        CK(cudaMemsetAsync(dhr, 0, 32 * sizeof(int), st));
        const int few_pairs = n < 4096 ? 1 : 0;
        rev_band_keys_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>((int)n, fwd_c, res_c, sc.match, sc.B, sc.ge, few_pairs, keys.p, vals.p, dhr,
                                                                    reinterpret_cast<int*>(d_counters + 5), wflags_c, io.sub);
        launches++;
        CK(cub::DeviceRadixSort::SortPairs(cub_tmp_post.p, cub_bytes, keys.p, keys_alt.p, vals.p, vals_alt.p, (int)n, 0, 24, st));
        launches += 3;
        slot_scan_kernel<<<1, 32, 0, st>>>(dhr, 2 * kRevVariants + 1, d_rev_slots);
        launches++;
        CK(cudaEventRecord(fork_ev, st));
        for (int s = 0; s < 2 * kRevVariants; s++) {
            const int v = s % kRevVariants;
            const bool prof = s < kRevVariants;
            const int gpb = 128 / rev_variant_lanes(v);
            // persistent grid: what fits the machine (4 CTAs per SM), fewer for the rare N flavour, never more than the
            // batch could fill; and no more ring memory than 1 GB per launch
            const long long half = rev_ring_half(v, batch_max_q, batch_max_w, sc.match, sc.B, sc.ge);
            long long blocks = std::min<long long>((n + gpb - 1) / gpb, prof ? 148 * 4 : 148);
            const long long per_block = (long long)gpb * 2 * half * 4;
            blocks = std::max<long long>(1, std::min<long long>(blocks, ((long long)1 << 30) / std::max<long long>(per_block, 1)));
            CK(rev_ring[s].ensure((size_t)(blocks * gpb * 2 * half)));
            cudaStream_t sx = aux[s % kAux];
            if (s < kAux) CK(cudaStreamWaitEvent(sx, fork_ev, 0));
            cusw::launch_rev_band(v, prof, (int)blocks, vals_alt.p, d_rev_slots + 2 * s, io, fwd_c, rev_c, sc, rev_ring[s].p, (int)half, sx);
            launches++;
        }
        for (int a = 0; a < kAux; a++) {
            CK(cudaEventRecord(join_ev[a], aux[a]));
            CK(cudaStreamWaitEvent(st, join_ev[a], 0));
        }
        if (getenv("CUDASW_REV_STATS")) {  // experiment hook: pairs per strip-height variant (synchronises)
            int hh[2 * kRevVariants + 1];
            CK(cudaStreamSynchronize(st));
            CK(cudaMemcpy(hh, dhr, sizeof hh, cudaMemcpyDeviceToHost));
            fprintf(stderr, "[cudasw reverse] pairs %lld: S = 32/64/128/256/1024: %d %d %d %d %d; with N: %d %d %d %d %d; none %d\n", (long long)n,
                    hh[0], hh[1], hh[2], hh[3], hh[4], hh[5], hh[6], hh[7], hh[8], hh[9], hh[10]);
        }
    } else {
    CK(cudaMemsetAsync(d_strip + 1, 0, sizeof(unsigned long long), st));
    CK(cudaMemsetAsync(d_hist, 0, 256 * sizeof(int), st));
    rev_keys_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>((int)n, fwd_c, res_c, vt, keys.p, vals.p, d_hist,
                                                           reinterpret_cast<int*>(d_counters + 5), wflags_c);
    launches++;
    CK(cub::DeviceRadixSort::SortPairs(cub_tmp_post.p, cub_bytes, keys.p, keys_alt.p, vals.p, vals_alt.p, (int)n, 0, 27, st));
    launches += 3;
    CK(cudaMemcpyAsync(h_hist, d_hist, 256 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    { int rc = launch_variants(post_set, true, vals_alt.p, io, fwd_c, rev_c, spool, &launches, h_hist, nullptr, true); if (rc != CUDASW_OK) return rc; }
    }
    CK(cudaGetLastError());
    CK(cudaEventRecord(ev[4], st));

    if (ends_only) {  // the caller reads clip lengths only: no traceback, no CIGAR
        ends_only_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>((int)n, fwd_c, rev_c, res_c);
        launches++;
        CK(cudaEventRecord(ev[5], st));
        CK(cudaMemcpyAsync(h_counters, d_counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    // ---- traceback (+ retry rounds when the scratch pool runs out); the CIGAR counter [1] runs on across rounds.
    // Pairs are sorted by estimated banded cells: the widest go to the cooperative kernels (a CTA or a warp each, on
    // auxiliary streams), the rest to the one-thread-per-pair kernels in descending cost order. A round is queued
    // without the host knowing how the pairs fall into these lists: every kernel is launched over an upper-bound grid
    // and reads its list bounds from device memory (tb_slots_kernel); the one synchronisation is at the end of the round.
    if (!ends_only) {
        // default: the one-thread kernel keeps bands up to kTbThreadBand in registers, wider pairs take a warp each;
        // CUDASW_TB_WIDE_CELLS additionally sends every pair above that many banded cells to the warp kernel (0 = all)
        long long wide_cells = -1;
        if (const char* e = getenv("CUDASW_TB_WIDE_CELLS")) wide_cells = atoll(e);
        // widest starting band the one-thread kernels take (experiment hook; they hold up to kTbThreadBand in registers)
        int thread_band = kTbThreadBand;
        if (const char* e = getenv("CUDASW_TB_THREAD_BAND")) thread_band = std::min(std::max(atoi(e), 0), kTbThreadBand);
        const int cta_wave = getenv("CUDASW_TB_NO_BLOCK") ? -1 : 148 * 8;  // big pairs take a CTA each while they fit one wave
        CK(cudaMemsetAsync(tb_band.p, 0, (size_t)n * sizeof(int), st));
        TbPool pool{tb_pool.p, (unsigned long long)tb_pool_bytes, d_counters};
        int* retry_count = reinterpret_cast<int*>(d_counters + 2);
        int* wide_count = reinterpret_cast<int*>(d_counters + 3);
        int* dh = d_hist + 128;            // [0, 8): pairs per frame class, [8, 16): escalation list counters
        int* tb_slots = d_rev_slots + 32;  // (start, count) of the kTbSlots lists of this round
        const int* list = nullptr;  // first round: every pair
        int c = (int)n;
        int* retry_out = retry_a.p;
        bool stuck = false;
        for (int round = 0; round < 64 && c > 0; round++) {
            CK(cudaMemsetAsync(d_counters, 0, sizeof(unsigned long long), st));      // pool
            CK(cudaMemsetAsync(d_counters + 2, 0, 2 * sizeof(unsigned long long), st));  // retry + wide counters
            CK(cudaMemsetAsync(dh, 0, 16 * sizeof(int), st));                            // class and escalation counters
            CK(cudaMemsetAsync(d_counters + 7, 0, sizeof(unsigned long long), st));      // scratch estimate
            // few pairs cannot fill the machine with one thread each: long ones then take a warp / CTA each
            // (and run all their band passes in one kernel: a kernel boundary between passes costs the sum of the slowest
            // pairs of each stage instead of the slowest pair overall)
            // (measured on C1: 10 k pairs 1.31 -> 0.76 ms with the 1024-cell rule, neutral at 50 k, slower from 400 k up)
            const long long wc = wide_cells >= 0 ? wide_cells : (c < 8192 ? 2048 : c < 32768 ? 1024 : -1);
            tb_keys_kernel<<<(c + 255) / 256, 256, 0, st>>>(list, c, fwd_c, rev_c, res_c, wc, c < 8192 ? 1 : 0, thread_band, tb_band.p, keys.p, vals.p, wide_count,
                                                            dh, d_counters + 7);
            launches++;
            CK(cub::DeviceRadixSort::SortPairs(cub_tmp_post.p, cub_bytes, keys.p, keys_alt.p, vals.p, vals_alt.p, c, 0, 23, st));
            launches += 3;
            tb_slots_kernel<<<1, 32, 0, st>>>(wide_count, dh, cta_wave, tb_slots);
            launches++;
            // escalation lists: kTbClasses + 2 arrays of c ints (list k = input of thread class k, list kTbClasses = input
            // of the small warp kernel, list kTbClasses + 1 = input of the big one); their counters live in dh[8 ..]
            CK(tb_esc.ensure((size_t)(kTbClasses + 2) * (size_t)c));
            int* esc_small = tb_esc.p + (size_t)kTbClasses * c;
            int* esc_big = tb_esc.p + (size_t)(kTbClasses + 1) * c;
            int* esc_small_count = dh + 8 + kTbClasses;
            int* esc_big_count = dh + 8 + kTbClasses + 1;
            const unsigned long long ccap = (unsigned long long)cigar_cap;
            {
                CK(cudaEventRecord(fork_ev, st));
                // big pairs: a CTA each while they fit one wave (a quarter of the per-pair latency of a lone warp; anything
                // wider than its register frames is handed to the one-warp kernel's list), else a warp each
                CK(cudaStreamWaitEvent(aux[0], fork_ev, 0));
                sw_traceback_block_kernel<<<std::min(c, 148 * 8), 128, 0, aux[0]>>>(vals_alt.p, 0, esc_big, esc_big_count, io, fwd_c, rev_c, tc, pool, res_c,
                                                                                   cig.p, ccap, d_counters + 1, retry_out, retry_count, tb_band.p, tb_slots + 0);
                sw_traceback_warp_kernel<32><<<std::min((c + 3) / 4, 148 * 8), 128, 0, aux[0]>>>(vals_alt.p, 0, nullptr, nullptr, nullptr, io, fwd_c, rev_c, tc,
                                                                                                pool, res_c, cig.p, ccap, d_counters + 1, retry_out,
                                                                                                retry_count, tb_band.p, tb_slots + 2);
                launches += 2;
                CK(cudaEventRecord(join_ev[0], aux[0]));
                CK(cudaStreamWaitEvent(aux[1], fork_ev, 0));
                sw_traceback_warp_kernel<8><<<std::min((c + 3) / 4, 148 * 16), 128, 0, aux[1]>>>(vals_alt.p, 0, nullptr, esc_big, esc_big_count, io, fwd_c, rev_c,
                                                                                                tc, pool, res_c, cig.p, ccap, d_counters + 1, retry_out,
                                                                                                retry_count, tb_band.p, tb_slots + 4);
                launches++;
                CK(cudaEventRecord(join_ev[1], aux[1]));
            }
            {
                // One launch per frame class over its own (sorted) list, all classes side by side on their own streams: each is
                // latency-bound (one thread per pair, a few hundred registers' worth of frame) and leaves most of the machine
                // idle on its own. A pair that outgrows its frame is appended to the escalation list of the next wider class;
                // those lists are drained afterwards, narrowest first (each drain may append to the next), and the widest
                // class hands over to the warp kernels. Class slot 5 = pairs with nothing to trace.
                const int grid = std::min((c + 127) / 128, 148 * 8);
#define CUSW_TB_LAUNCH(B, STREAM, SLOT, ESC_IN, ESC_IN_COUNT)                                                                                    \
    sw_traceback_kernel<B><<<grid, 128, 0, STREAM>>>(vals_alt.p, 0, ESC_IN, ESC_IN_COUNT, esc_out, esc_out_count, io, fwd_c, rev_c, tc, pool, res_c, \
                                                     cig.p, ccap, d_counters + 1, retry_out, retry_count, tb_band.p, SLOT)
                // own lists: classes 1.. on their streams first, class 0 on the main stream
                for (int kk = 1; kk <= kTbClasses + 1; kk++) {
                    const int k = kk <= kTbClasses ? kk : 0;
                    const bool none = k == kTbClasses;  // score 0 / failed pairs: "1M(q-1)S" or nothing; any class will do
                    int* esc_out = none ? tb_esc.p : tb_esc.p + (size_t)(k + 1) * c;
                    int* esc_out_count = none ? dh + 8 : dh + 8 + k + 1;
                    cudaStream_t sk = k == 0 ? st : tb_streams[k - 1];
                    if (k > 0) CK(cudaStreamWaitEvent(sk, fork_ev, 0));
                    switch (k) {
                        case 1: CUSW_TB_LAUNCH(2, sk, tb_slots + 6 + 2 * k, nullptr, nullptr); break;
                        case 2: CUSW_TB_LAUNCH(4, sk, tb_slots + 6 + 2 * k, nullptr, nullptr); break;
                        case 3: CUSW_TB_LAUNCH(8, sk, tb_slots + 6 + 2 * k, nullptr, nullptr); break;
                        case 4: CUSW_TB_LAUNCH(16, sk, tb_slots + 6 + 2 * k, nullptr, nullptr); break;
                        default: CUSW_TB_LAUNCH(1, sk, tb_slots + 6 + 2 * k, nullptr, nullptr); break;
                    }
                    launches++;
                    if (k > 0) CK(cudaEventRecord(tb_join[k - 1], sk));
                }
                // drain the escalation lists on the main stream, narrowest first: class k takes what class k-1 (own list and
                // drain) handed over as soon as its own list is through, while the wider classes' own lists are still running
                const int egrid = std::min((c + 127) / 128, 148 * 2);
                for (int k = 1; k < kTbClasses; k++) {
                    const int* esc_in = tb_esc.p + (size_t)k * c;
                    const int* esc_in_count = dh + 8 + k;
                    int* esc_out = tb_esc.p + (size_t)(k + 1) * c;
                    int* esc_out_count = dh + 8 + k + 1;
                    const int grid = egrid;
                    CK(cudaStreamWaitEvent(st, tb_join[k - 1], 0));
                    switch (k) {
                        case 1: CUSW_TB_LAUNCH(2, st, tb_slots + 2 * kTbSlots, esc_in, esc_in_count); break;
                        case 2: CUSW_TB_LAUNCH(4, st, tb_slots + 2 * kTbSlots, esc_in, esc_in_count); break;
                        case 3: CUSW_TB_LAUNCH(8, st, tb_slots + 2 * kTbSlots, esc_in, esc_in_count); break;
                        default: CUSW_TB_LAUNCH(16, st, tb_slots + 2 * kTbSlots, esc_in, esc_in_count); break;
                    }
                    launches++;
                }
                CK(cudaStreamWaitEvent(st, tb_join[kTbClasses - 1], 0));  // the "nothing to trace" class
#undef CUSW_TB_LAUNCH
                // pairs that outgrew the widest frame, then those that outgrew the small warp kernel (from either launch)
                sw_traceback_warp_kernel<8><<<148 * 4, 128, 0, st>>>(esc_small, 0, esc_small_count, esc_big, esc_big_count, io, fwd_c,
                                                                      rev_c, tc, pool, res_c, cig.p, ccap, d_counters + 1, retry_out, retry_count, tb_band.p);
                launches++;
                CK(cudaStreamWaitEvent(st, join_ev[1], 0));
                CK(cudaStreamWaitEvent(st, join_ev[0], 0));  // the CTA-per-pair kernel may append to esc_big too
                sw_traceback_warp_kernel<32><<<148 * 3, 128, 0, st>>>(esc_big, 0, esc_big_count, nullptr, nullptr, io, fwd_c,
                                                                      rev_c, tc, pool, res_c, cig.p, ccap, d_counters + 1, retry_out, retry_count, tb_band.p);
                launches++;
            }
            CK(cudaEventRecord(ev[5], st));
            CK(cudaMemcpyAsync(h_counters, d_counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));  // the one host round trip of the round
            if (getenv("CUDASW_TB_STATS")) {  // experiment hook: how the pairs fell into the traceback lists
                int hh[16];
                CK(cudaMemcpy(hh, dh, sizeof hh, cudaMemcpyDeviceToHost));
                const int* wc2 = reinterpret_cast<const int*>(h_counters + 3);
                fprintf(stderr, "[cudasw tb round %d] pairs %d: wide %d (big %d), classes 1/2/4/8/16/none = %d %d %d %d %d %d; escalated into 2/4/8/16/warp8/warp32 = %d %d %d %d %d %d; retry %d\n",
                        round, c, wc2[0], wc2[1], hh[0], hh[1], hh[2], hh[3], hh[4], hh[5], hh[9], hh[10], hh[11], hh[12], hh[13], hh[14],
                        *reinterpret_cast<int*>(h_counters + 2));
            }
            int nretry = *reinterpret_cast<int*>(h_counters + 2);
            if (nretry == 0) { c = 0; break; }
            // a whole round may legitimately come back once (every pair handed from the thread kernel to the warp
            // kernel); twice in a row means no pair fits the scratch pool
            if (nretry == c && stuck) { set_error("traceback scratch pool too small"); return CUDASW_ENOMEM; }
            stuck = nretry == c;
            {
                // the scratch pool ran out: let it follow the demand (first band passes of the round + a margin for doublings)
                const unsigned long long est = 3ull * h_counters[7] + ((unsigned long long)256 << 20);
                if (est > tb_pool_bytes && !getenv("CUDASW_TB_POOL_MB")) {
                    size_t free_b = 0, total_b = 0;
                    CK(cudaMemGetInfo(&free_b, &total_b));
                    const size_t want = (size_t)std::min<unsigned long long>(est, (unsigned long long)((free_b + tb_pool_bytes) / 2));
                    if (want > tb_pool_bytes) {
                        tb_pool.release();
                        tb_pool_bytes = 0;
                        CK(tb_pool.ensure(want));
                        tb_pool_bytes = want;
                        pool = TbPool{tb_pool.p, (unsigned long long)tb_pool_bytes, d_counters};
                        stuck = false;
                    }
                }
            }
            list = retry_out;
            retry_out = (retry_out == retry_a.p) ? retry_b.p : retry_a.p;
            c = nretry;
        }
        if (c > 0) { set_error("traceback did not converge"); return CUDASW_ENOMEM; }
    }
    CK(cudaGetLastError());
    // the stream is idle here (synchronised above): fold this chunk's phase times into the batch stats
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev[3], ev[4]); stats.ms_reverse += ms; last_post_ms = ms;
    cudaEventElapsedTime(&ms, ev[4], ev[5]); stats.ms_traceback += ms; last_post_ms += ms;
    *launches_out += launches;
    return CUDASW_OK;
}

// JAligner-compatible mode (ja_mode.cuh): no reverse pass; a banded pointer DP from row 0 to the end cell and jaligner's
// walk, one CTA per pair, with the same scratch pool / retry rounds as the SSW traceback.
int cudasw::post_chunk_ja(size_t ci, int64_t first, int64_t n, const int64_t* d_so, const int64_t* d_ro, int64_t cigar_cap,
                          uint64_t* launches_out) {
    cudaStream_t st = post_set.main;
    CK(cudaStreamWaitEvent(st, chunk_ev[ci].fwd1, 0));
    uint64_t launches = 0;
    PairIO io{enc_seq.p, enc_ref.p, d_so + first, d_ro + first};
    int4* fwd_c = fwd.p + first;
    cudasw_result* res_c = res.p + first;
    uint32_t* keys_c = keys.p + first; uint32_t* keys_alt_c = keys_alt.p + first;
    int* vals_c = vals.p + first; int* vals_alt_c = vals_alt.p + first;
    int* retry_a_c = retry_a.p + first; int* retry_b_c = retry_b.p + first;
    size_t cub_bytes = cub_tmp_post.cap;
    CK(cudaEventRecord(ev[3], st));
    CK(cudaEventRecord(ev[4], st));  // (no reverse pass in this mode)
    TbPool pool{tb_pool.p, (unsigned long long)tb_pool_bytes, d_counters};
    int* retry_count = reinterpret_cast<int*>(d_counters + 2);
    int* dh = d_hist + 128;
    int* slots = d_rev_slots + 32;
    const int* list = nullptr;
    int c = (int)n;
    int* retry_out = retry_a_c;
    bool stuck = false;
    for (int round = 0; round < 256 && c > 0; round++) {
        CK(cudaMemsetAsync(d_counters, 0, sizeof(unsigned long long), st));          // pool
        CK(cudaMemsetAsync(d_counters + 2, 0, sizeof(unsigned long long), st));      // retry counter
        CK(cudaMemsetAsync(d_counters + 7, 0, sizeof(unsigned long long), st));      // scratch estimate
        CK(cudaMemsetAsync(dh, 0, 8 * sizeof(int), st));
        ja_keys_kernel<<<(c + 255) / 256, 256, 0, st>>>(list, c, io, fwd_c, res_c, tc, keys_c, vals_c, dh, d_counters + 7);
        CK(cub::DeviceRadixSort::SortPairs(cub_tmp_post.p, cub_bytes, keys_c, keys_alt_c, vals_c, vals_alt_c, c, 0, 22, st));
        slot_scan_kernel<<<1, 32, 0, st>>>(dh, 3, slots);
        launches += 5;
        sw_ja_traceback_kernel<kJaSmallD><<<std::min(c, 148 * 8), 128, 0, st>>>(vals_alt_c, slots + 0, io, fwd_c, tc, pool, res_c, cig.p,
                                                                               (unsigned long long)cigar_cap, d_counters + 1, retry_out, retry_count);
        sw_ja_traceback_kernel<kJaBigD><<<std::min(c, 148 * 2), 128, 0, st>>>(vals_alt_c, slots + 2, io, fwd_c, tc, pool, res_c, cig.p,
                                                                             (unsigned long long)cigar_cap, d_counters + 1, retry_out, retry_count);
        launches += 2;
        CK(cudaEventRecord(ev[5], st));
        CK(cudaMemcpyAsync(h_counters, d_counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        const int nretry = *reinterpret_cast<int*>(h_counters + 2);
        if (nretry == 0) { c = 0; break; }
        if (nretry == c && stuck) { set_error("traceback scratch pool too small"); return CUDASW_ENOMEM; }
        stuck = nretry == c;
        {
            // every pair keeps a pointer byte per band cell: let the pool grow towards what a round asks for
            const unsigned long long est = h_counters[7] + ((unsigned long long)64 << 20);
            if (est > tb_pool_bytes && !getenv("CUDASW_TB_POOL_MB")) {
                size_t free_b = 0, total_b = 0;
                CK(cudaMemGetInfo(&free_b, &total_b));
                const size_t want = (size_t)std::min<unsigned long long>(est, (unsigned long long)((free_b + tb_pool_bytes) / 2));
                if (want > tb_pool_bytes) {
                    tb_pool.release();
                    tb_pool_bytes = 0;
                    CK(tb_pool.ensure(want));
                    tb_pool_bytes = want;
                    pool = TbPool{tb_pool.p, (unsigned long long)tb_pool_bytes, d_counters};
                    stuck = false;
                }
            }
        }
        list = retry_out;
        retry_out = (retry_out == retry_a_c) ? retry_b_c : retry_a_c;
        c = nretry;
    }
    if (c > 0) { set_error("traceback did not converge"); return CUDASW_ENOMEM; }
    CK(cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev[4], ev[5]); stats.ms_traceback += ms; last_post_ms = ms;
    *launches_out += launches;
    return CUDASW_OK;
}

// Boundary rows of the strip-mined score variants (queries above the tallest single-strip variant): two rows of one
// 32-bit word per window column and pair. The forward buffer is sized for the chunk [first, first + cnt) whose forward
// pass is about to be queued; the reverse buffer for `rev_first`/`rev_cnt`, the range ONE post_chunk call walks (the whole
// batch unless CUDASW_POST_PER_CHUNK is set), so the kernels' bump allocation cannot run out.
int cudasw::size_strip_buffers(unsigned long long need_f, unsigned long long need_r) {
    if (need_f > strip_fwd.cap || need_r > strip_rev.cap) {
        // the buffers may still be in use by the chunk before: let it finish before they are replaced
        CK(cudaStreamSynchronize(stream));
        CK(cudaStreamSynchronize(post_set.main));
        if (need_f > strip_fwd.cap) CK(strip_fwd.ensure((size_t)need_f + 256));
        if (need_r > strip_rev.cap) CK(strip_rev.ensure((size_t)need_r + 256));
    }
    return CUDASW_OK;
}

// One pass over the host's copy of the offsets (several threads for large batches), while the first H2D copies are in
// flight: per chunk the pairs per query-length class (the grids of the forward launches) and the boundary-row memory its
// strip-mined pairs need; for the batch the longest query / window (the reverse pass's rings).
void cudasw::scan_batch(const std::vector<std::pair<int64_t, int64_t>>& ranges, const int64_t* h_seq_off, const int64_t* h_ref_off) {
    const bool have_multi = g_var[g_nvar - 1].multi;
    const int max_rows = std::min(have_multi ? 32767 : g_var[g_nvar - 1].rows,
                                  scoring.match > 0 ? (32767 - scoring.gap_open) / scoring.match : 32767);
    const int max_cols = (1 << 24) - 2;
    const int64_t single_rows = vt.n_single > 0 ? vt.rows[vt.n_single - 1] : 0;
    fwd_counts.assign(ranges.size(), std::vector<int>((size_t)g_nvar, 0));
    strip_need.assign(ranges.size(), 0ull);
    struct Part { std::vector<int> counts; unsigned long long need = 0; int64_t mq = 0, mw = 0; double cells = 0.0, sq = 0.0, sw = 0.0; };
    std::vector<double> chunk_cells(ranges.size(), 0.0), chunk_q(ranges.size(), 0.0), chunk_w(ranges.size(), 0.0);
    int64_t mq = 0, mw = 0;
    for (size_t ci = 0; ci < ranges.size(); ci++) {
        const int64_t first = ranges[ci].first, cnt = ranges[ci].second;
        const int nt = (int)std::min<int64_t>(8, std::max<int64_t>(1, cnt / 65536));
        std::vector<Part> parts((size_t)nt);
        auto work = [&](int t) {
            Part& p = parts[(size_t)t];
            p.counts.assign((size_t)g_nvar, 0);
            const int64_t a = first + cnt * t / nt, b = first + cnt * (t + 1) / nt;
            for (int64_t i = a; i < b; i++) {
                const int64_t q = h_seq_off[i + 1] - h_seq_off[i], w = h_ref_off[i + 1] - h_ref_off[i];
                p.mq = std::max(p.mq, q);
                p.mw = std::max(p.mw, w);
                if (q > 0 && w > 0 && q <= max_rows && w <= max_cols) {
                    p.counts[(size_t)var_by_q[(size_t)q]]++;
                    p.cells += (double)q * (double)w;
                    p.sq += (double)q;
                    p.sw += (double)w;
                    if (q > single_rows) p.need += ((((unsigned long long)w * 4ull + 15ull) & ~15ull) * 2ull);
                }
            }
        };
        if (nt == 1) work(0);
        else {
            std::vector<std::thread> th;
            for (int t = 1; t < nt; t++) th.emplace_back(work, t);
            work(0);
            for (auto& x : th) x.join();
        }
        for (auto& p : parts) {
            for (int v = 0; v < g_nvar; v++) fwd_counts[ci][(size_t)v] += p.counts[(size_t)v];
            strip_need[ci] += p.need;
            chunk_cells[ci] += p.cells;
            chunk_q[ci] += p.sq;
            chunk_w[ci] += p.sw;
            mq = std::max(mq, p.mq);
            mw = std::max(mw, p.mw);
        }
    }
    // The filter pays once the forward pass is throughput-bound: a small chunk runs as long as its longest CTA, and a
    // sub-window pass followed by a full-window pass for the few unproven pairs is then longer than the full-window pass
    // alone (C1 10 k pairs = 3 G cells: 1.26 against 0.76 ms; C2 5 k pairs = 6 G cells: 2.13 against 1.37 ms; C3 10 k
    // pairs = 21 G cells: 15.4 against 33.4 ms).
    chunk_filter.assign(ranges.size(), 0);
    // Nor does the column-range flavour (CUDASW_BAND=0) pay when the windows are hardly longer than the queries
    // (BreakpointHomology's 600 x 620: the columns a band can reach are the whole window); the diagonal-band kernel does.
    for (size_t ci = 0; ci < ranges.size(); ci++)
        chunk_filter[ci] = use_filter && (filter_always || (chunk_cells[ci] >= 12e9 && (fc.band || chunk_w[ci] >= 1.5 * chunk_q[ci])));
    batch_max_q = std::min<int64_t>(std::max<int64_t>(mq, 1), 32767);
    batch_max_w = std::min<int64_t>(std::max<int64_t>(mw, 1), (1 << 24));
}

int cudasw::run(int64_t n, const uint8_t* d_seq_in, const int64_t* d_seq_off_in,
                const uint8_t* d_ref_in, const int64_t* d_ref_off_in, int64_t total_q,
                int64_t total_w, cudasw_result* out, uint32_t* cigar_blob, int64_t cigar_cap,
                int64_t* cigar_used, bool inputs_on_host, const uint8_t* h_seq,
                const int64_t* h_seq_off, const uint8_t* h_ref, const int64_t* h_ref_off,
                const WindowSeg* h_segs, const int64_t* h_seg_off, const QuerySegs* qsegs, bool ends_only,
                const PackedIn* packed) {
    std::lock_guard<std::mutex> lock(mu);
    const bool gather = h_segs != nullptr;
    const bool qgather = qsegs != nullptr;
    const int q_stride = qsegs ? qsegs->stride : 0, w_stride = qsegs ? qsegs->w_stride : 0;
    set_error("");
    stats = cudasw_stats{};
    if (cigar_used) *cigar_used = 0;
    if (n < 0 || n > 0x7fffff00ll) { set_error("n_pairs out of range"); return CUDASW_EINVAL; }
    if (n == 0) return CUDASW_OK;
    const int64_t n_wsegs = !gather ? 0 : (w_stride ? n * w_stride : h_seg_off[n]);
    const int64_t n_qsegs = !qgather ? 0 : (q_stride ? n * q_stride : qsegs->seg_off[n]);
    if (!out || (!cigar_blob && cigar_cap > 0) || cigar_cap < 0) { set_error("null output"); return CUDASW_EINVAL; }
    if (total_q < 0 || total_w < 0) { set_error("offsets must be non-decreasing from 0"); return CUDASW_EINVAL; }
    CK(cudaSetDevice(device));
    cudaStream_t st = stream;
    uint64_t launches = 0;

    // ---- chunk plan: pair ranges of about chunk_bytes of sequence each. With host inputs the H2D copy of
    // chunk i+1 (copy stream) overlaps the kernels of chunk i (compute stream).
    struct Chunk { int64_t first, cnt, q_lo, q_hi, w_lo, w_hi; };
    std::vector<Chunk> chunks;
    {
        // Host inputs: chunk sizes grow geometrically (64 MB, 160 MB, 400 MB, ... capped at 4 GB; a pinned copy runs about
        // 2.5x faster than the kernels consume the bytes): the first copy is the only one
        // the kernels have to wait for, and each later copy hides behind the (slower) kernels of the chunk before it.
        // Resident inputs: one chunk. (Splitting them so that the pack of chunk i+1 runs beside the forward pass of chunk i
        // was measured neutral on B200: the forward kernel fills the register file, so whatever runs beside it takes SM
        // slots from it. The split is kept only for the CUDASW_POST_PER_CHUNK experiment.)
        const int64_t total_bytes = total_q + total_w;
        const bool ppc = getenv("CUDASW_POST_PER_CHUNK") != nullptr;
        const bool pipelined = ppc && !getenv("CUDASW_NO_OVERLAP") && total_bytes >= ((int64_t)256 << 20);
        int64_t target = inputs_on_host ? (int64_t)64 << 20 : (pipelined ? std::max<int64_t>(total_bytes / 16, 1) : (int64_t)1 << 60);
        int64_t target_max = inputs_on_host ? (int64_t)4 << 30 : (pipelined && ppc ? std::max<int64_t>(5 * total_bytes / 16, 1) : (int64_t)1 << 60);
        const int64_t tail_bytes = pipelined && ppc ? total_bytes / 16 : 0;
        // windows gathered on the device: only queries and descriptors cross PCIe (~1/12 of the bytes), so the copy of
        // everything behind the small leading chunk hides behind that chunk
        if (gather && inputs_on_host && !pipelined) { target = (int64_t)128 << 20; target_max = (int64_t)1 << 60; }
        if (gather && pipelined) { target = std::max<int64_t>(total_bytes / 16, 1); target_max = ppc ? std::max<int64_t>(5 * total_bytes / 16, 1) : (int64_t)1 << 60; }
        bool fixed_target = false;
        if (const char* e = getenv("CUDASW_CHUNK_MB")) { target = target_max = std::max<int64_t>(1, atoll(e)) << 20; fixed_target = true; }
        const int64_t max_pairs = 8 << 20;  // keeps the per-chunk sort/bin arrays small
        int64_t first = 0;
        while (first < n) {
            int64_t last = first;
            if (h_seq_off && h_ref_off) {
                // largest `last` with bytes(first..last) <= target (offsets are monotone: binary search)
                int64_t lo = first + 1, hi = std::min(n, first + max_pairs);
                const int64_t base = h_seq_off[first] + h_ref_off[first];
                int64_t want_bytes = target;
                if (!fixed_target && tail_bytes > 0) {
                    // keep a small trailing chunk: whatever is left after this one is either >= 1.5 tails or exactly one
                    const int64_t left = total_bytes - base;
                    if (left <= tail_bytes + tail_bytes / 2) want_bytes = left;                 // this is the trailing chunk
                    else if (left - want_bytes < tail_bytes + tail_bytes / 2) want_bytes = left - tail_bytes;
                }
                const int64_t target = want_bytes;
                while (lo < hi) {
                    int64_t mid = lo + (hi - lo + 1) / 2;
                    if (h_seq_off[mid] + h_ref_off[mid] - base <= target) lo = mid; else hi = mid - 1;
                }
                last = lo;
            } else {
                last = n;
            }
            // an offset array that is not monotone is caught per pair by fwd_keys_kernel (EINVAL); clamp here
            Chunk c{first, last - first, 0, 0, 0, 0};
            if (h_seq_off && h_ref_off) {
                c.q_lo = h_seq_off[first]; c.q_hi = h_seq_off[last];
                c.w_lo = h_ref_off[first]; c.w_hi = h_ref_off[last];
                if (c.q_lo < 0 || c.q_hi < c.q_lo || c.q_hi > total_q || c.w_lo < 0 || c.w_hi < c.w_lo || c.w_hi > total_w) {
                    set_error("offsets must be non-decreasing from 0");
                    return CUDASW_EINVAL;
                }
            } else {
                c.q_hi = total_q; c.w_hi = total_w;
            }
            chunks.push_back(c);
            first = last;
            target = fixed_target ? target : (inputs_on_host && !gather ? std::min(target * 5 / 2, std::max(target, target_max)) : target_max);
        }
    }
    int64_t max_cnt = 0;
    for (auto& c : chunks) max_cnt = std::max(max_cnt, c.cnt);

    // ---- capacity
    CK(enc_seq.ensure((size_t)total_q + 16));
    CK(enc_ref.ensure((size_t)total_w + 16));
    CK(keys.ensure(n)); CK(keys_alt.ensure(n)); CK(vals.ensure(n)); CK(vals_alt.ensure(n));
    CK(retry_a.ensure(n)); CK(retry_b.ensure(n));
    CK(fwd.ensure(n)); CK(rev.ensure(n)); CK(res.ensure(n)); CK(wflags.ensure(n)); CK(anchor.ensure(n)); CK(tb_band.ensure(n));
    CK(cig.ensure((size_t)std::max<int64_t>(cigar_cap, 1)));
    {
        size_t cub_bytes = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, keys.p, keys_alt.p, vals.p, vals_alt.p, (int)n, 0, 27, st));
        CK(cub_tmp.ensure(cub_bytes + 16));
        CK(cub_tmp_post.ensure(cub_bytes + 16));
        if (use_filter) CK(cub_tmp_fwd2.ensure(cub_bytes + 16));
    }
    CK(fwd_slots.ensure(chunks.size() * (size_t)(2 * (2 * g_nvar + 1))));
    if (use_filter) { CK(frec.ensure(n)); CK(fhc.ensure((size_t)n * kFltHc)); CK(fdone.ensure(n)); CK(bkeys.ensure(n)); CK(bkeys_alt.ensure(n)); CK(bvals.ensure(n)); CK(bvals_alt.ensure(n)); CK(fwd_slots2.ensure(chunks.size() * (size_t)(2 * (2 * g_nvar + 1)))); }
    batch_maybe_n = !(packed && packed->refN == nullptr);
    if (!tb_pool.p) {
        size_t free_b = 0, total_b = 0;
        CK(cudaMemGetInfo(&free_b, &total_b));
        size_t want = (size_t)8 << 30;
        if (const char* e = getenv("CUDASW_TB_POOL_MB")) want = (size_t)atoll(e) << 20;
        want = std::min(want, free_b / 4);
        want = std::max(want, (size_t)64 << 20);
        CK(tb_pool.ensure(want));
        tb_pool_bytes = want;
    }
    while (h2d_done.size() < chunks.size()) {
        cudaEvent_t e;
        CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h2d_done.push_back(e);
    }
    while (chunk_ev.size() < chunks.size()) {
        ChunkEvents ce{};
        CK(cudaEventCreate(&ce.pack0)); CK(cudaEventCreate(&ce.enc1)); CK(cudaEventCreate(&ce.pack1)); CK(cudaEventCreate(&ce.fwd0)); CK(cudaEventCreate(&ce.fwd1));
        chunk_ev.push_back(ce);
    }

    CK(cudaEventRecord(ev[0], st));
    CK(cudaMemsetAsync(d_counters, 0, 8 * sizeof(unsigned long long), st));
    CK(cudaMemsetAsync(d_strip + 2, 0, 2 * sizeof(unsigned long long), st));
    CK(cudaMemsetAsync(d_fcount, 0, 4 * sizeof(unsigned long long), st));
    CK(cudaEventRecord(ev[1], st));
    CK(cudaStreamWaitEvent(pack_stream, ev[1], 0));  // the counters (any-N flag) are reset before the first encode
    // ---- H2D: everything is queued on the copy stream now, one event per chunk
    const uint8_t* d_seq = d_seq_in;
    const uint8_t* d_ref = d_ref_in;
    const int64_t* d_so = d_seq_off_in;
    const int64_t* d_ro = d_ref_off_in;
    if (packed) {
        // packed host inputs: words [word0, ...) of each blob; word0 = the batch's first base rounded down to 32 bases
        pk_word0_q = (h_seq_off[0] >> 5) << 1;
        pk_word0_w = (h_ref_off[0] >> 5) << 1;
        pk_has_qn = packed->seqN != nullptr;
        pk_has_wn = packed->refN != nullptr;
        CK(pk_seq2.ensure((size_t)(((total_q + 31) >> 5) * 2 - pk_word0_q) + 4));
        CK(pk_ref2.ensure((size_t)(((total_w + 31) >> 5) * 2 - pk_word0_w) + 4));
        if (pk_has_qn) CK(pk_seqN.ensure((size_t)(((total_q + 31) >> 5) - (pk_word0_q >> 1)) + 4));
        if (pk_has_wn) CK(pk_refN.ensure((size_t)(((total_w + 31) >> 5) - (pk_word0_w >> 1)) + 4));
    }
    if (inputs_on_host) {
        if (!qgather && !packed) CK(raw_seq.ensure((size_t)total_q + 16));
        if (!gather && !packed) CK(raw_ref.ensure((size_t)total_w + 16));
        CK(seq_off.ensure(n + 1));
        CK(ref_off.ensure(n + 1));
        CK(cudaEventRecord(ev[7], st));
        CK(cudaStreamWaitEvent(copy_stream, ev[7], 0));  // the previous batch's kernels no longer read raw_*
        CK(cudaMemcpyAsync(seq_off.p, h_seq_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, copy_stream));
        CK(cudaMemcpyAsync(ref_off.p, h_ref_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, copy_stream));
        if (gather) {
            const int64_t nsegs = w_stride ? n * w_stride : h_seg_off[n];
            CK(d_segs.ensure((size_t)std::max<int64_t>(nsegs, 1)));
            if (!w_stride) {
                CK(d_seg_off.ensure(n + 1));
                CK(cudaMemcpyAsync(d_seg_off.p, h_seg_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, copy_stream));
            }
            if (nsegs > 0) CK(cudaMemcpyAsync(d_segs.p, h_segs, (size_t)nsegs * sizeof(WindowSeg), cudaMemcpyHostToDevice, copy_stream));
        }
        if (qgather) {
            const int64_t nsegs = q_stride ? n * q_stride : qsegs->seg_off[n];
            CK(d_qsegs.ensure((size_t)std::max<int64_t>(nsegs, 1)));
            CK(d_lit.ensure((size_t)qsegs->lit_bytes + 16));
            if (!q_stride) {
                CK(d_qseg_off.ensure(n + 1));
                CK(cudaMemcpyAsync(d_qseg_off.p, qsegs->seg_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, copy_stream));
            }
            if (nsegs > 0) CK(cudaMemcpyAsync(d_qsegs.p, qsegs->segs, (size_t)nsegs * sizeof(WindowSeg), cudaMemcpyHostToDevice, copy_stream));
            if (qsegs->lit_bytes > 0) CK(cudaMemcpyAsync(d_lit.p, qsegs->lit, (size_t)qsegs->lit_bytes, cudaMemcpyHostToDevice, copy_stream));
        }
        uint64_t packed_bytes = 0;
        for (size_t i = 0; packed && i < chunks.size(); i++) {
            const Chunk& c = chunks[i];
            // whole 32-base groups (two 2-bit words + one mask word) covering the chunk's bases; neighbouring chunks may
            // copy the same boundary group twice
            auto up = [&](DevBuf<uint32_t>& d2, DevBuf<uint32_t>& dN, const uint32_t* h2, const uint32_t* hN, int64_t word0, int64_t lo, int64_t hi) -> cudaError_t {
                if (hi <= lo) return cudaSuccess;
                const int64_t g0 = lo >> 5, g1 = (hi + 31) >> 5;  // groups of 32 bases
                cudaError_t e = cudaMemcpyAsync(d2.p + (2 * g0 - word0), h2 + 2 * g0, (size_t)(g1 - g0) * 8, cudaMemcpyHostToDevice, copy_stream);
                packed_bytes += (uint64_t)(g1 - g0) * 8;
                if (e == cudaSuccess && hN) {
                    e = cudaMemcpyAsync(dN.p + (g0 - (word0 >> 1)), hN + g0, (size_t)(g1 - g0) * 4, cudaMemcpyHostToDevice, copy_stream);
                    packed_bytes += (uint64_t)(g1 - g0) * 4;
                }
                return e;
            };
            CK(up(pk_seq2, pk_seqN, packed->seq2, packed->seqN, pk_word0_q, c.q_lo, c.q_hi));
            CK(up(pk_ref2, pk_refN, packed->ref2, packed->refN, pk_word0_w, c.w_lo, c.w_hi));
            CK(cudaEventRecord(h2d_done[i], copy_stream));
        }
        for (size_t i = 0; !packed && i < chunks.size(); i++) {
            const Chunk& c = chunks[i];
            if (c.q_hi > c.q_lo && !qgather) CK(cudaMemcpyAsync(raw_seq.p + c.q_lo, h_seq + c.q_lo, (size_t)(c.q_hi - c.q_lo), cudaMemcpyHostToDevice, copy_stream));
            if (c.w_hi > c.w_lo && !gather) CK(cudaMemcpyAsync(raw_ref.p + c.w_lo, h_ref + c.w_lo, (size_t)(c.w_hi - c.w_lo), cudaMemcpyHostToDevice, copy_stream));
            CK(cudaEventRecord(h2d_done[i], copy_stream));
        }
        d_seq = raw_seq.p; d_ref = raw_ref.p; d_so = seq_off.p; d_ro = ref_off.p;
        stats.h2d_bytes = (qgather ? (q_stride ? 0 : (uint64_t)(n + 1) * 8) + (uint64_t)n_qsegs * sizeof(WindowSeg) + (uint64_t)qsegs->lit_bytes
                                   : (uint64_t)total_q) + (uint64_t)(n + 1) * 16 +
                          (gather ? (w_stride ? 0 : (uint64_t)(n + 1) * 8) + (uint64_t)n_wsegs * sizeof(WindowSeg) : (uint64_t)total_w);
        if (packed) stats.h2d_bytes = packed_bytes + (uint64_t)(n + 1) * 16;
    }

    {
        // (runs on the host while the first copies are on their way)
        std::vector<std::pair<int64_t, int64_t>> ranges;
        for (auto& c : chunks) ranges.emplace_back(c.first, c.cnt);
        scan_batch(ranges, h_seq_off, h_ref_off);
    }
    // ---- the chunk pipeline: pack(i) | forward(i) | post(i-1), see pack_chunk
    auto drain = [&]() {  // nothing may stay in flight over the caller's buffers or this handle's scratch
        cudaStreamSynchronize(copy_stream); cudaStreamSynchronize(pack_stream); cudaStreamSynchronize(st);
        cudaStreamSynchronize(post_set.main);
    };
    const bool serial = getenv("CUDASW_NO_OVERLAP") != nullptr;
    const bool post_per_chunk = getenv("CUDASW_POST_PER_CHUNK") != nullptr;
    for (size_t i = 0; i < chunks.size(); i++) {
        const Chunk& c = chunks[i];
        if (inputs_on_host) CK(cudaStreamWaitEvent(pack_stream, h2d_done[i], 0));
        int rc = pack_chunk(i, c.first, c.cnt, d_seq, d_so, d_ref, d_ro, c.q_lo, c.q_hi, c.w_lo, c.w_hi, gather, &launches, qgather, q_stride, w_stride, packed != nullptr, h_seq_off, h_ref_off);
        // reverse-pass strips: per chunk when post runs per chunk, else once for the whole batch (sized with the first chunk)
        if (rc == CUDASW_OK && g_var[g_nvar - 1].multi) {
            unsigned long long need_all = 0;
            for (auto x : strip_need) need_all += x;
            rc = size_strip_buffers(strip_need[i], post_per_chunk ? strip_need[i] : need_all);
        }
        if (rc == CUDASW_OK) rc = forward_chunk(i, c.first, c.cnt, d_so, d_ro, &launches);
        if (rc == CUDASW_OK && serial) rc = cudaStreamSynchronize(st) == cudaSuccess ? CUDASW_OK : CUDASW_ECUDA;
        if (rc == CUDASW_OK && post_per_chunk && tie_break == CUDASW_TIEBREAK_SSW && i > 0) rc = post_chunk(i - 1, chunks[i - 1].first, chunks[i - 1].cnt, d_so, d_ro, cigar_cap, &launches, ends_only);
        if (rc != CUDASW_OK) { drain(); return rc; }
    }
    {
        // Reverse pass + traceback: once over the whole batch. Running them per chunk beside the next chunk's forward
        // pass was measured slower on B200 (C2: 253 vs 237 ms per 1 M pairs): their latency-bound CTAs displace
        // forward CTAs that fill the register file. CUDASW_POST_PER_CHUNK=1 keeps that experiment available.
        const Chunk& c = chunks.back();
        int rc = tie_break == CUDASW_TIEBREAK_JALIGNER ? post_chunk_ja(chunks.size() - 1, 0, n, d_so, d_ro, cigar_cap, &launches)
                 : post_per_chunk ? post_chunk(chunks.size() - 1, c.first, c.cnt, d_so, d_ro, cigar_cap, &launches, ends_only)
                                  : post_chunk(chunks.size() - 1, 0, n, d_so, d_ro, cigar_cap, &launches, ends_only);
        if (rc != CUDASW_OK) { drain(); return rc; }
    }
    CK(cudaEventRecord(ev[5], st));
#ifdef CUSW_STEP_STATS
    {
        unsigned long long ss[8];
        CK(cudaStreamSynchronize(st));
        CK(cudaMemcpy(ss, d_pool + 512, sizeof ss, cudaMemcpyDeviceToHost));
        CK(cudaMemset(d_pool + 512, 0, sizeof ss));
        fprintf(stderr, "[cudasw step stats, forward+reverse warp-steps] quiet %llu, quiet attempt failed %llu, full %llu, chained quiet %llu, chain step failed %llu, chains %llu\n",
                ss[0], ss[1], ss[2], ss[3], ss[4], ss[5]);
    }
#endif

    // ---- D2H
    unsigned long long used = h_counters[1];
    if (cigar_used) *cigar_used = (int64_t)used;
    CK(cudaMemcpyAsync(out, res.p, (size_t)n * sizeof(cudasw_result), cudaMemcpyDeviceToHost, st));
    unsigned long long ncopy = std::min<unsigned long long>(used, (unsigned long long)cigar_cap);
    if (ncopy) CK(cudaMemcpyAsync(cigar_blob, cig.p, (size_t)ncopy * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(h_cells, d_strip + 2, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(h_cells + 4, d_fcount, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(ev[6], st));
    CK(cudaStreamSynchronize(st));
    stats.d2h_bytes = (uint64_t)n * sizeof(cudasw_result) + ncopy * 4;
    stats.filter_banded = use_filter ? h_cells[4] : 0;
    stats.filter_failed = use_filter ? h_cells[6] : 0;
    if (use_filter && getenv("CUDASW_FILTER_STATS"))  // experiment hook
        fprintf(stderr, "[cudasw filter] pairs %lld: banded %llu, passed on %llu (ineligible pairs not counted), certificates failed %llu\n",
                (long long)n, h_cells[4], h_cells[5], h_cells[6]);

    stats.pairs = (uint64_t)n;
    stats.kernel_launches = launches;
    cudaEventElapsedTime(&stats.ms_d2h, ev[5], ev[6]);
    cudaEventElapsedTime(&stats.ms_total, ev[0], ev[6]);
    // per-phase device time summed over the chunks. The phases of neighbouring chunks overlap (pack and post run beside a
    // forward pass), so they add up to more than ms_total; ms_h2d = the part of ms_total that is none of: first pack,
    // forward passes, last post, D2H (i.e. copy waits and pipeline bubbles)
    float pack_first = 0.f;
    for (size_t i = 0; i < chunks.size(); i++) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, chunk_ev[i].pack0, chunk_ev[i].enc1); stats.ms_encode += ms;
        cudaEventElapsedTime(&ms, chunk_ev[i].pack0, chunk_ev[i].pack1); stats.ms_pack += ms;
        if (i == 0) pack_first = ms;
        cudaEventElapsedTime(&ms, chunk_ev[i].fwd0, chunk_ev[i].fwd1); stats.ms_forward += ms;
    }
    stats.ms_h2d = std::max(0.f, stats.ms_total - pack_first - stats.ms_forward - last_post_ms - stats.ms_d2h);
    stats.cells = h_cells[0];                 // counted by fwd_keys_kernel while binning (no per-pair host loop)
    stats.forward_cells_padded = h_cells[1];
    if (used > (unsigned long long)cigar_cap) {
        set_error("cigar buffer too small");
        return CUDASW_ENOMEM;
    }
    return CUDASW_OK;
}

extern "C" int cudasw_align_batch(cudasw_t* h, int64_t n, const uint8_t* seq_blob,
                                  const int64_t* seq_off, const uint8_t* ref_blob,
                                  const int64_t* ref_off, cudasw_result* out, uint32_t* cigar_blob,
                                  int64_t cigar_cap, int64_t* cigar_used) {
    if (!h) return CUDASW_EINVAL;
    if (n > 0 && (!seq_blob || !seq_off || !ref_blob || !ref_off)) { h->set_error("null input"); return CUDASW_EINVAL; }
    if (n > 0 && (seq_off[0] != 0 || ref_off[0] != 0)) { h->set_error("offsets must start at 0"); return CUDASW_EINVAL; }
    int64_t tq = n > 0 ? seq_off[n] : 0, tw = n > 0 ? ref_off[n] : 0;
    return h->run(n, nullptr, nullptr, nullptr, nullptr, tq, tw, out, cigar_blob, cigar_cap, cigar_used, true,
                  seq_blob, seq_off, ref_blob, ref_off);
}

extern "C" int cudasw_align_batch_device(cudasw_t* h, int64_t n, const uint8_t* d_seq_blob,
                                         const int64_t* d_seq_off, const uint8_t* d_ref_blob,
                                         const int64_t* d_ref_off, const int64_t* h_seq_off,
                                         const int64_t* h_ref_off, cudasw_result* out,
                                         uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used) {
    if (!h) return CUDASW_EINVAL;
    if (n > 0 && (!d_seq_blob || !d_seq_off || !d_ref_blob || !d_ref_off || !h_seq_off || !h_ref_off)) {
        h->set_error("null input");
        return CUDASW_EINVAL;
    }
    int64_t tq = n > 0 ? h_seq_off[n] : 0, tw = n > 0 ? h_ref_off[n] : 0;
    return h->run(n, d_seq_blob, d_seq_off, d_ref_blob, d_ref_off, tq, tw, out, cigar_blob, cigar_cap, cigar_used,
                  false, nullptr, h_seq_off, nullptr, h_ref_off);
}

// ---- packed host inputs (north_star (1): "packs 2-bit or byte sequences ... into pinned host buffers")
// Host-side packer for the shim: bases [base0, base0 + n) of a blob from ASCII. A/C/G/T in either case -> 0..3; every
// other byte is what clean() (SswJniAligner.java:71-79) turns into N: stored as 0 with its mask bit set. The words that
// hold the range keep their bits below base0 and lose everything from base0 + n up, so a blob can be appended to pair
// by pair without clearing it first. Returns the number of ambiguous bases, or CUDASW_EINVAL.
extern "C" int64_t cudasw_pack_bases(const uint8_t* ascii, int64_t n, int64_t base0, uint32_t* two_bit, uint32_t* nmask) {
    if (n < 0 || base0 < 0 || (n > 0 && (!ascii || !two_bit))) return CUDASW_EINVAL;
    static const struct Lut {
        uint8_t c[256];
        Lut() {
            for (int i = 0; i < 256; i++) c[i] = 4;
            c['A'] = c['a'] = 0; c['C'] = c['c'] = 1; c['G'] = c['g'] = 2; c['T'] = c['t'] = 3;
        }
    } lut;
    int64_t amb = 0;
    int64_t b = base0;
    const int64_t end = base0 + n;
    const uint8_t* src = ascii;
    // head: up to the next 32-base boundary, bit by bit into the partly filled words
    if (b < end && (b & 31)) {
        const int64_t stop = std::min(end, (b | 31) + 1);
        uint32_t w0 = two_bit[(b >> 5) * 2], w1 = two_bit[(b >> 5) * 2 + 1], nm = nmask ? nmask[b >> 5] : 0u;
        const int k0 = (int)(b & 31);
        // keep the bits below base0, clear the rest
        if (k0 < 16) { w0 &= (1u << (2 * k0)) - 1u; w1 = 0u; } else w1 &= k0 == 16 ? 0u : (1u << (2 * (k0 - 16))) - 1u;
        nm &= (1u << k0) - 1u;
        for (; b < stop; b++) {
            const uint32_t c = lut.c[*src++];
            const int k = (int)(b & 31);
            if (k < 16) w0 |= (c & 3u) << (2 * k); else w1 |= (c & 3u) << (2 * (k - 16));
            nm |= (c >> 2) << k;
            amb += c >> 2;
        }
        two_bit[((b - 1) >> 5) * 2] = w0; two_bit[((b - 1) >> 5) * 2 + 1] = w1;
        if (nmask) nmask[(b - 1) >> 5] = nm;
    }
    // body + tail: whole groups of 32 bases (the last one zero-padded)
    while (b < end) {
        const int cnt = (int)std::min<int64_t>(32, end - b);
        uint32_t w0 = 0, w1 = 0, nm = 0;
        for (int k = 0; k < cnt; k++) {
            const uint32_t c = lut.c[src[k]];
            if (k < 16) w0 |= (c & 3u) << (2 * k); else w1 |= (c & 3u) << (2 * (k - 16));
            nm |= (c >> 2) << k;
        }
        amb += __builtin_popcount(nm);
        two_bit[(b >> 5) * 2] = w0; two_bit[(b >> 5) * 2 + 1] = w1;
        if (nmask) nmask[b >> 5] = nm;
        src += cnt; b += cnt;
    }
    return amb;
}

extern "C" int cudasw_align_batch_packed(cudasw_t* h, int64_t n, const uint32_t* seq_two_bit, const uint32_t* seq_nmask,
                                         const int64_t* seq_off, const uint32_t* ref_two_bit, const uint32_t* ref_nmask,
                                         const int64_t* ref_off, cudasw_result* out, uint32_t* cigar_blob,
                                         int64_t cigar_cap, int64_t* cigar_used) {
    if (!h) return CUDASW_EINVAL;
    if (n > 0 && (!seq_two_bit || !seq_off || !ref_two_bit || !ref_off)) { h->set_error_locked("null input"); return CUDASW_EINVAL; }
    if (n > 0 && (seq_off[0] < 0 || ref_off[0] < 0)) { h->set_error_locked("offsets must be non-negative"); return CUDASW_EINVAL; }
    const int64_t tq = n > 0 ? seq_off[n] : 0, tw = n > 0 ? ref_off[n] : 0;
    const PackedIn pk{seq_two_bit, seq_nmask, ref_two_bit, ref_nmask};
    return h->run(n, nullptr, nullptr, nullptr, nullptr, tq, tw, out, cigar_blob, cigar_cap, cigar_used, true,
                  nullptr, seq_off, nullptr, ref_off, nullptr, nullptr, nullptr, false, &pk);
}

// ---- device-resident reference + window descriptors
static_assert(sizeof(cudasw_window) == sizeof(WindowSeg), "cudasw_window and the device descriptor share one layout");

extern "C" int cudasw_reference_load(cudasw_t* h, int32_t n_contigs, const uint8_t* bases, const int64_t* contig_off) {
    if (!h) return CUDASW_EINVAL;
    std::unique_lock<std::shared_mutex> ref_lock(h->ref_mu);  // no batch may be reading the reference being replaced
    std::lock_guard<std::mutex> lock(h->mu);
    h->set_error("");
    auto set_error = [&](const std::string& m) { h->set_error(m); };
    if (n_contigs < 0 || (n_contigs > 0 && (!bases || !contig_off))) { h->set_error("null reference"); return CUDASW_EINVAL; }
    if (n_contigs > 0 && contig_off[0] != 0) { h->set_error("contig offsets must start at 0"); return CUDASW_EINVAL; }
    for (int32_t c = 0; c < n_contigs; c++)
        if (contig_off[c + 1] < contig_off[c]) { h->set_error("contig offsets must be non-decreasing"); return CUDASW_EINVAL; }
    CK(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    CK(cudaStreamSynchronize(st));
    h->ref2bit.release(); h->refN.release(); h->d_contig_off.release();
    h->contig_off.clear();
    if (n_contigs == 0) return CUDASW_OK;
    const int64_t total = contig_off[n_contigs];
    const size_t groups = (size_t)((total + 31) / 32);
    CK(h->ref2bit.ensure(2 * groups + 4));
    CK(h->refN.ensure(groups + 4));
    CK(h->d_contig_off.ensure((size_t)n_contigs + 1));
    CK(cudaMemcpyAsync(h->d_contig_off.p, contig_off, ((size_t)n_contigs + 1) * 8, cudaMemcpyHostToDevice, st));
    // ASCII goes up in slices of 256 MB (a multiple of 32 bases) through a staging buffer that is released afterwards
    const int64_t slice = (int64_t)256 << 20;
    DevBuf<uint8_t> stage;
    CK(stage.ensure((size_t)std::min(slice, total) + 32));
    for (int64_t g0 = 0; g0 < total; g0 += slice) {
        const int64_t len = std::min(slice, total - g0);
        CK(cudaMemcpyAsync(stage.p, bases + g0, (size_t)len, cudaMemcpyHostToDevice, st));
        const int64_t thr = (len + 31) / 32;
        ref_pack_kernel<<<(int)std::min<int64_t>((thr + 255) / 256, 148 * 16), 256, 0, st>>>(stage.p, len, g0, h->ref2bit.p, h->refN.p);
        CK(cudaStreamSynchronize(st));  // the staging buffer is reused by the next slice
    }
    CK(cudaGetLastError());
    stage.release();
    h->contig_off.assign(contig_off, contig_off + n_contigs + 1);
    return CUDASW_OK;
}

// Validates the descriptors and sizes every window; an invalid descriptor makes its window empty, which the
// pipeline reports as a per-pair CUDASW_EINVAL (like an empty `ref`, SswJniAligner.java:48-50).
static void window_offsets(const cudasw_t* h, int64_t n, const cudasw_window* win, const int64_t* win_off,
                           std::vector<int64_t>& ref_off) {
    // window lengths, then their prefix sums; several threads for large batches (one serial pass over a million descriptors
    // is milliseconds, on the critical path of every call)
    ref_off.assign((size_t)n + 1, 0);
    const int64_t ncontig = (int64_t)h->contig_off.size() - 1;
    const int nt = (int)std::min<int64_t>(8, std::max<int64_t>(1, n / 65536));
    std::vector<int64_t> part((size_t)nt + 1, 0);
    auto lengths = [&](int t) {
        const int64_t a = n * t / nt, b = n * (t + 1) / nt;
        int64_t sum = 0;
        for (int64_t p = a; p < b; p++) {
            int64_t len = 0;
            bool ok = win_off[p + 1] >= win_off[p];
            for (int64_t s = win_off[p]; ok && s < win_off[p + 1]; s++) {
                const cudasw_window& w = win[s];
                if (w.contig < 0 || w.contig >= ncontig || w.start < 0 || w.end < w.start ||
                    w.end > h->contig_off[w.contig + 1] - h->contig_off[w.contig]) ok = false;
                else len += w.end - w.start;
            }
            sum += ok ? len : 0;
            ref_off[p + 1] = sum;  // prefix inside the part
        }
        part[(size_t)t + 1] = sum;
    };
    auto shift = [&](int t) {
        const int64_t a = n * t / nt, b = n * (t + 1) / nt, base = part[(size_t)t];
        if (base != 0)
            for (int64_t p = a; p < b; p++) ref_off[p + 1] += base;
    };
    auto run_all = [&](auto&& fn) {
        if (nt == 1) { fn(0); return; }
        std::vector<std::thread> th;
        for (int t = 1; t < nt; t++) th.emplace_back(fn, t);
        fn(0);
        for (auto& x : th) x.join();
    };
    run_all(lengths);
    for (int t = 0; t < nt; t++) part[(size_t)t + 1] += part[(size_t)t];
    run_all(shift);
}

extern "C" int cudasw_align_batch_windows(cudasw_t* h, int64_t n, const uint8_t* seq_blob, const int64_t* seq_off,
                                          const cudasw_window* win, const int64_t* win_off, cudasw_result* out,
                                          uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used) {
    if (!h) return CUDASW_EINVAL;
    if (n > 0 && (!seq_blob || !seq_off || !win || !win_off)) { h->set_error("null input"); return CUDASW_EINVAL; }
    if (n > 0 && (seq_off[0] != 0 || win_off[0] != 0)) { h->set_error("offsets must start at 0"); return CUDASW_EINVAL; }
    std::shared_lock<std::shared_mutex> ref_lock(h->ref_mu);
    if (h->contig_off.empty()) { h->set_error("no reference loaded (cudasw_reference_load)"); return CUDASW_EINVAL; }
    std::vector<int64_t> ref_off;
    window_offsets(h, std::max<int64_t>(n, 0), win, win_off, ref_off);
    // a pair with a rejected descriptor has window length 0: the gather skips it and the pair reports EINVAL
    int64_t tq = n > 0 ? seq_off[n] : 0, tw = n > 0 ? ref_off[n] : 0;
    return h->run(n, nullptr, nullptr, nullptr, nullptr, tq, tw, out, cigar_blob, cigar_cap, cigar_used, true,
                  seq_blob, seq_off, nullptr, ref_off.data(), reinterpret_cast<const WindowSeg*>(win), win_off);
}

// ---- BreakpointHomology.calculate for a list of breakpoints (BreakpointHomology.java:38-103).
// Everything the reference does with host strings becomes segment arithmetic here: the breakend query
// (localBsSeq + inserted + remoteBsSeq) and both windows (localSeq + localRef, and the reverse complement of
// remoteRef + remoteSeq) are lists of reference segments, so 32 bytes per breakpoint go to the device, both
// sequences of both pairs are gathered from the 2-bit reference there, and only the score passes run: the homology
// lengths are functions of the two clip lengths (read_begin, q - read_end - 1).
namespace {
struct Seg { int32_t contig; int64_t start, end; bool rc; bool some; };
// BreakendSummary.getAnchorSequence (BreakendSummary.java:73-106) as a segment; dir 0 = forward, 1 = backward
inline Seg anchor_seg(const std::vector<int64_t>& coff, int32_t contig, int32_t dir, int64_t pos, int64_t length) {
    int64_t a_start, a_end;
    if (dir == 0) { a_end = pos; a_start = a_end - length + 1; }
    else { a_start = pos; a_end = a_start + length - 1; }
    a_start = std::max<int64_t>(1, a_start);
    a_end = std::min<int64_t>(coff[contig + 1] - coff[contig], a_end);
    return Seg{contig, a_start - 1, a_end, dir != 0, a_start <= a_end};
}
inline int64_t seg_len(const Seg& s) { return s.some ? s.end - s.start : 0; }
inline WindowSeg slot(const Seg& s, bool flip) {  // an empty anchor keeps its slot with start == end
    return s.some ? WindowSeg{s.contig, (s.rc != flip) ? 1 : 0, s.start, s.end} : WindowSeg{0, 0, 0, 0};
}
inline uint8_t comp_base(uint8_t c) {
    switch (c) {
        case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A';
        case 'a': return 't'; case 'c': return 'g'; case 'g': return 'c'; case 't': return 'a';
        default: return c;
    }
}
}  // namespace

extern "C" int cudasw_breakpoint_homology(cudasw_t* h, int64_t n, const cudasw_breakpoint* bp, const uint8_t* ins_blob,
                                          const int64_t* ins_off, int32_t max_breakend_length, int32_t margin,
                                          int32_t* local_hom, int32_t* remote_hom) {
    if (!h) return CUDASW_EINVAL;
    if (n < 0 || n > 0x3fffff00ll || (n > 0 && (!bp || !local_hom || !remote_hom)) || max_breakend_length < 0 || margin < 0) {
        h->set_error("bad argument");
        return CUDASW_EINVAL;
    }
    if (n == 0) return CUDASW_OK;
    if (h->tie_break != CUDASW_TIEBREAK_SSW) { h->set_error("cudasw_breakpoint_homology needs the SSW tie-break mode"); return CUDASW_EINVAL; }
    if (ins_off && (ins_off[0] != 0 || (!ins_blob && ins_off[n] > 0))) { h->set_error("bad inserted-sequence blob"); return CUDASW_EINVAL; }
    std::lock_guard<std::mutex> hom_lock(h->hom_mu);
    std::shared_lock<std::shared_mutex> ref_lock(h->ref_mu);
    if (h->contig_off.empty()) { h->set_error("no reference loaded (cudasw_reference_load)"); return CUDASW_EINVAL; }
    const std::vector<int64_t>& coff = h->contig_off;
    const int32_t ncontig = (int32_t)coff.size() - 1;
    // Fixed layout, so breakpoints can be planned independently (and in parallel): breakpoint k owns pairs 2k (local
    // window) and 2k+1 (remote window); a pair owns 3 query slots and 2 window slots; the literal blob holds, per
    // breakpoint, the inserted sequence as the breakend reads it followed by its reverse complement. A call the
    // reference skips (empty breakend or window) is a pair of length 0, which the engine rejects without work.
    const int64_t np = 2 * n;
    const int64_t lit_bytes = ins_off ? 2 * ins_off[n] : 0;
    auto alloc_fail = [&](cudaError_t e) { h->set_error(std::string("pinned staging: ") + cudaGetErrorString(e)); return CUDASW_ENOMEM; };
    cudaError_t ce;
    if ((ce = h->hom_qs.ensure((size_t)np * 3)) != cudaSuccess || (ce = h->hom_ws.ensure((size_t)np * 2)) != cudaSuccess ||
        (ce = h->hom_seq_off.ensure((size_t)np + 1)) != cudaSuccess || (ce = h->hom_ref_off.ensure((size_t)np + 1)) != cudaSuccess ||
        (ce = h->hom_lit.ensure((size_t)lit_bytes + 16)) != cudaSuccess || (ce = h->hom_res.ensure((size_t)np)) != cudaSuccess)
        return alloc_fail(ce);
    WindowSeg* qs = h->hom_qs.p;
    WindowSeg* ws = h->hom_ws.p;
    int64_t* seq_off = h->hom_seq_off.p;   // first lengths (slot p + 1), then their running sums
    int64_t* ref_off = h->hom_ref_off.p;
    uint8_t* lit = h->hom_lit.p;
    std::vector<int32_t> n_bs((size_t)np);  // [2k] = |localBsSeq|, [2k+1] = |remoteBsSeq|
    std::atomic<int64_t> bad{-1};
    auto plan = [&](int64_t k0, int64_t k1) {
        for (int64_t k = k0; k < k1; k++) {
            const cudasw_breakpoint& b = bp[k];
            const int64_t ins_len = ins_off ? ins_off[k + 1] - ins_off[k] : 0;
            if (b.contig1 < 0 || b.contig1 >= ncontig || b.contig2 < 0 || b.contig2 >= ncontig || ((b.dir1 | b.dir2) & ~1) || ins_len < 0) {
                bad.store(k);
                return;
            }
            int64_t seq_len = max_breakend_length, ref_len = (int64_t)max_breakend_length + ins_len + margin;
            if (b.contig1 == b.contig2) {  // BreakpointSummary.getEventSize (BreakpointSummary.java:226-236)
                const int64_t d = b.pos2 > b.pos1 ? b.pos2 - b.pos1 : b.pos1 - b.pos2;
                int64_t ev;
                if (b.dir1 == b.dir2) ev = d;
                else if ((b.pos1 < b.pos2 && b.dir1 == 0) || (b.pos2 < b.pos1 && b.dir1 == 1)) ev = d - 1;
                else ev = d + 1;
                seq_len = std::min(seq_len, ev);
                ref_len = std::min(ref_len, ev);
            }
            const int64_t adv1 = b.dir1 == 0 ? b.pos1 + ref_len : b.pos1 - ref_len;  // BreakendSummary.advance
            const int64_t adv2 = b.dir2 == 0 ? b.pos2 + ref_len : b.pos2 - ref_len;
            const Seg local_seq = anchor_seg(coff, b.contig1, b.dir1, b.pos1, ref_len);
            const Seg local_bs = anchor_seg(coff, b.contig1, b.dir1, b.pos1, seq_len);
            const Seg local_ref = anchor_seg(coff, b.contig1, b.dir1, adv1, ref_len);
            // the remote pieces are the reverse complements of these anchors (:64-66)
            const Seg remote_seq = anchor_seg(coff, b.contig2, b.dir2, b.pos2, ref_len);
            const Seg remote_bs = anchor_seg(coff, b.contig2, b.dir2, b.pos2, seq_len);
            const Seg remote_ref = anchor_seg(coff, b.contig2, b.dir2, adv2, ref_len);
            n_bs[(size_t)(2 * k)] = (int32_t)seg_len(local_bs);
            n_bs[(size_t)(2 * k + 1)] = (int32_t)seg_len(remote_bs);
            const int64_t q = seg_len(local_bs) + ins_len + seg_len(remote_bs);
            // inserted sequence as the breakend reads it (:45-47), then its reverse complement for the second pair
            const int64_t lit0 = ins_off ? 2 * ins_off[k] : 0;
            if (ins_len > 0) {
                const uint8_t* src = ins_blob + ins_off[k];
                uint8_t* fwd_ins = lit + lit0;
                uint8_t* rc_ins = fwd_ins + ins_len;
                if (b.dir1 == 1) for (int64_t x = 0; x < ins_len; x++) fwd_ins[x] = comp_base(src[ins_len - 1 - x]);
                else for (int64_t x = 0; x < ins_len; x++) fwd_ins[x] = src[x];
                for (int64_t x = 0; x < ins_len; x++) rc_ins[x] = comp_base(fwd_ins[ins_len - 1 - x]);
            }
            const int64_t w_local = seg_len(local_seq) + seg_len(local_ref);
            const int64_t w_remote = seg_len(remote_ref) + seg_len(remote_seq);
            WindowSeg* q0 = qs + 6 * k;
            WindowSeg* w0 = ws + 4 * k;
            // pair 2k: align(breakend, localSeq + localRef)  (:76-78)
            q0[0] = slot(local_bs, false);
            q0[1] = WindowSeg{-1, 0, lit0, lit0 + ins_len};
            q0[2] = slot(remote_bs, true);
            w0[0] = slot(local_seq, false);
            w0[1] = slot(local_ref, false);
            // pair 2k+1: align(rc(breakend), rc(remoteRef + remoteSeq))  (:86-94); rc(rc(anchor)) = the anchor itself
            q0[3] = slot(remote_bs, false);
            q0[4] = WindowSeg{-1, 0, lit0 + ins_len, lit0 + 2 * ins_len};
            q0[5] = slot(local_bs, true);
            w0[2] = slot(remote_seq, false);
            w0[3] = slot(remote_ref, false);
            const bool a = q > 0 && w_local > 0, c = q > 0 && w_remote > 0;
            seq_off[2 * k + 1] = a ? q : 0; ref_off[2 * k + 1] = a ? w_local : 0;
            seq_off[2 * k + 2] = c ? q : 0; ref_off[2 * k + 2] = c ? w_remote : 0;
        }
    };
    {
        const int nthreads = n < 32768 ? 1 : (int)std::min<int64_t>(std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u), n / 16384);
        if (nthreads <= 1) plan(0, n);
        else {
            std::vector<std::thread> th;
            for (int t = 0; t < nthreads; t++) th.emplace_back(plan, n * t / nthreads, n * (t + 1) / nthreads);
            for (auto& t : th) t.join();
        }
    }
    if (bad.load() >= 0) {
        h->set_error("breakpoint " + std::to_string(bad.load()) + ": contig, direction or inserted-sequence offsets out of range");
        return CUDASW_EINVAL;
    }
    seq_off[0] = ref_off[0] = 0;
    for (int64_t p = 0; p < np; p++) { seq_off[p + 1] += seq_off[p]; ref_off[p + 1] += ref_off[p]; }
    for (int64_t k = 0; k < n; k++) { local_hom[k] = 0; remote_hom[k] = 0; }
    if (seq_off[np] == 0 || ref_off[np] == 0) return CUDASW_OK;  // nothing to align anywhere
    cudasw_result* res = h->hom_res.p;
    const QuerySegs qsegs{qs, nullptr, 3, lit, lit_bytes, 2};
    const int rc = h->run(np, nullptr, nullptr, nullptr, nullptr, seq_off[np], ref_off[np], res, nullptr, 0, nullptr, true,
                          nullptr, seq_off, nullptr, ref_off, ws, nullptr, &qsegs, true);
    if (rc != CUDASW_OK) return rc;
    for (int64_t p = 0; p < np; p++) {
        const int64_t q = seq_off[p + 1] - seq_off[p];
        if (q == 0 || ref_off[p + 1] == ref_off[p]) continue;  // a call the reference skips (:75-76, :86)
        const cudasw_result& r = res[p];
        if (r.status != CUDASW_OK) { h->set_error("pair failed with status " + std::to_string(r.status)); return r.status; }
        const int64_t lead = r.read_begin, tail = q - r.read_end - 1;   // the clips of SswJniAligner.java:52-59
        // homology = bases of the OTHER side's breakend sequence that aligned; nothing if the anchor is clipped
        const int64_t other = n_bs[(size_t)(p ^ 1)];
        const int32_t v = lead > 0 ? 0 : (int32_t)std::max<int64_t>(0, other - tail);
        if (p & 1) local_hom[p >> 1] = v; else remote_hom[p >> 1] = v;
    }
    return CUDASW_OK;
}

extern "C" int cudasw_reference_fetch(cudasw_t* h, const cudasw_window* win, int32_t n_win, uint8_t* out, int64_t out_cap) {
    if (!h || !out || (n_win > 0 && !win) || n_win < 0) return CUDASW_EINVAL;
    std::shared_lock<std::shared_mutex> ref_lock(h->ref_mu);
    std::lock_guard<std::mutex> lock(h->mu);
    h->set_error("");
    auto set_error = [&](const std::string& m) { h->set_error(m); };
    if (h->contig_off.empty()) { h->set_error("no reference loaded (cudasw_reference_load)"); return CUDASW_EINVAL; }
    const int64_t woff[2] = {0, n_win};
    std::vector<int64_t> ref_off;
    window_offsets(h, 1, win, woff, ref_off);
    const int64_t len = ref_off[1];
    if (len == 0 && n_win > 0) {
        bool all_empty = true;
        for (int32_t s = 0; s < n_win; s++) all_empty &= win[s].end == win[s].start;
        if (!all_empty) { h->set_error("window outside the loaded reference"); return CUDASW_EINVAL; }
    }
    if (len > out_cap) { h->set_error("output buffer too small"); return CUDASW_ENOMEM; }
    if (len == 0) return 0;
    if (len > 0x7fffffffll) { h->set_error("window too long"); return CUDASW_ETOOLARGE; }
    CK(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    DevBuf<uint8_t> buf;
    DevBuf<int64_t> offs;
    DevBuf<WindowSeg> segs;
    CK(buf.ensure((size_t)len + 16));
    CK(offs.ensure(4));
    CK(segs.ensure((size_t)n_win));
    const int64_t hoffs[4] = {0, n_win, 0, len};
    CK(cudaMemcpyAsync(offs.p, hoffs, sizeof hoffs, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(segs.p, win, (size_t)n_win * sizeof(WindowSeg), cudaMemcpyHostToDevice, st));
    window_gather_kernel<<<1, 32, 0, st>>>(1, segs.p, offs.p, offs.p + 2, h->d_contig_off.p, h->ref2bit.p, h->refN.p, buf.p,
                                           nullptr, nullptr, 1);
    CK(cudaMemcpyAsync(out, buf.p, (size_t)len, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    buf.release(); offs.release(); segs.release();
    return (int)len;
}

extern "C" int cudasw_device_alloc(cudasw_t* h, size_t bytes, void** dptr) {
    if (!h || !dptr) return CUDASW_EINVAL;
    if (cudaSetDevice(h->device) != cudaSuccess) return CUDASW_ECUDA;
    cudaError_t e = cudaMalloc(dptr, bytes ? bytes : 1);
    if (e != cudaSuccess) { h->set_error(cudaGetErrorString(e)); return CUDASW_ENOMEM; }
    return CUDASW_OK;
}
extern "C" int cudasw_device_free(cudasw_t* h, void* dptr) {
    if (!h) return CUDASW_EINVAL;
    cudaSetDevice(h->device);
    return cudaFree(dptr) == cudaSuccess ? CUDASW_OK : CUDASW_ECUDA;
}
extern "C" int cudasw_upload(cudasw_t* h, void* dptr, const void* src, size_t bytes) {
    if (!h || !dptr || !src) return CUDASW_EINVAL;
    cudaSetDevice(h->device);
    cudaError_t e = cudaMemcpyAsync(dptr, src, bytes, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) { h->set_error(cudaGetErrorString(e)); return CUDASW_ECUDA; }
    return CUDASW_OK;
}
extern "C" int cudasw_host_alloc(size_t bytes, void** hptr) {
    if (!hptr) return CUDASW_EINVAL;
    return cudaMallocHost(hptr, bytes ? bytes : 1) == cudaSuccess ? CUDASW_OK : CUDASW_ENOMEM;
}
extern "C" int cudasw_host_free(void* hptr) { return cudaFreeHost(hptr) == cudaSuccess ? CUDASW_OK : CUDASW_ECUDA; }

extern "C" int cudasw_get_stats(cudasw_t* h, cudasw_stats* out) {
    if (!h || !out) return CUDASW_EINVAL;
    std::lock_guard<std::mutex> lock(h->mu);  // the counters of the last COMPLETED batch (waits for a running one)
    *out = h->stats;
    return CUDASW_OK;
}

extern "C" int cudasw_variant_for(cudasw_t* h, int32_t q, int32_t* lanes, int32_t* rpl, int32_t* strips) {
    if (!h || q <= 0) return CUDASW_EINVAL;
    int v = variant_of(h->vt, q);
    if (lanes) *lanes = g_var[v].G;
    if (rpl) *rpl = g_var[v].R;
    if (strips) *strips = (q + g_var[v].rows - 1) / g_var[v].rows;
    return CUDASW_OK;
}

// ---- async interface (StreamingAligner.java:10-44 contract) + dispatcher thread (SURVEY.md §8b "Threading")
// The queue has its own mutex (qmu): submit / poll / wait never block behind a running GPU batch (mu).
static int deliver_locked(cudasw_t* h, std::unordered_map<int64_t, cudasw::Done>::iterator it, cudasw_result* res,
                          uint32_t* cigar, int32_t cigar_cap, int32_t* cigar_len) {
    int32_t need = (int32_t)it->second.cigar.size();
    if (cigar_len) *cigar_len = need;
    if (need > cigar_cap || (need && !cigar)) return CUDASW_ENOMEM;
    *res = it->second.r;
    if (need) memcpy(cigar, it->second.cigar.data(), (size_t)need * 4);
    h->live.erase(it->first);
    h->done.erase(it);
    return 1;
}

extern "C" int cudasw_submit(cudasw_t* h, const uint8_t* seq, int64_t seq_len, const uint8_t* ref,
                             int64_t ref_len, int64_t* ticket) {
    if (!h || !ticket || seq_len < 0 || ref_len < 0 || (seq_len && !seq) || (ref_len && !ref)) return CUDASW_EINVAL;
    bool wake;
    {
        std::lock_guard<std::mutex> lock(h->qmu);
        if (h->pend_tickets.empty()) h->first_pending = std::chrono::steady_clock::now();
        h->pend_seq.insert(h->pend_seq.end(), seq, seq + seq_len);
        h->pend_ref.insert(h->pend_ref.end(), ref, ref + ref_len);
        h->pend_seq_off.push_back((int64_t)h->pend_seq.size());
        h->pend_ref_off.push_back((int64_t)h->pend_ref.size());
        *ticket = h->next_ticket++;
        h->pend_tickets.push_back(*ticket);
        h->live.insert(*ticket);
        // the dispatcher sleeps with a deadline once something is pending: wake it for the first pair and for a full batch
        wake = h->disp_run && (h->pend_tickets.size() == 1 || (int64_t)h->pend_tickets.size() >= h->disp_max_pairs);
    }
    if (wake) h->cv_work.notify_one();
    return CUDASW_OK;
}

extern "C" int cudasw_flush(cudasw_t* h) {
    if (!h) return CUDASW_EINVAL;
    std::vector<uint8_t> seq, ref;
    std::vector<int64_t> so, ro, tickets;
    {
        std::lock_guard<std::mutex> lock(h->qmu);
        if (h->pend_tickets.empty()) return CUDASW_OK;
        seq.swap(h->pend_seq); ref.swap(h->pend_ref);
        so.swap(h->pend_seq_off); ro.swap(h->pend_ref_off);
        tickets.swap(h->pend_tickets);
        h->pend_seq_off.assign(1, 0);
        h->pend_ref_off.assign(1, 0);
    }
    int64_t n = (int64_t)tickets.size();
    std::vector<cudasw_result> res(n);
    int64_t cap = (int64_t)seq.size() + 4 * n + 16, used = 0;
    std::vector<uint32_t> cig(cap);
    if (seq.empty()) seq.push_back(0);
    if (ref.empty()) ref.push_back(0);
    int rc = cudasw_align_batch(h, n, seq.data(), so.data(), ref.data(), ro.data(), res.data(), cig.data(), cap, &used);
    if (rc == CUDASW_ENOMEM && used > cap) {  // the estimate was short (a pair may need up to ~2q + 2 words): once more with the size asked for
        cap = used + 16;
        cig.assign((size_t)cap, 0u);
        rc = cudasw_align_batch(h, n, seq.data(), so.data(), ref.data(), ro.data(), res.data(), cig.data(), cap, &used);
    }
    {
        std::lock_guard<std::mutex> lock(h->qmu);
        for (int64_t i = 0; i < n; i++) {
            cudasw::Done d;
            d.r = res[i];
            if (rc != CUDASW_OK && d.r.status == CUDASW_OK) d.r.status = rc;
            if (rc == CUDASW_OK && d.r.status == CUDASW_OK)
                d.cigar.assign(cig.begin() + d.r.cigar_off, cig.begin() + d.r.cigar_off + d.r.cigar_len);
            d.r.cigar_off = 0;
            h->done.emplace(tickets[i], std::move(d));
        }
    }
    h->cv_done.notify_all();
    return rc;
}

extern "C" int64_t cudasw_outstanding(cudasw_t* h) {
    if (!h) return 0;
    std::lock_guard<std::mutex> lock(h->qmu);
    return (int64_t)h->pend_tickets.size();
}
extern "C" int64_t cudasw_processed(cudasw_t* h) {
    if (!h) return 0;
    std::lock_guard<std::mutex> lock(h->qmu);
    return (int64_t)h->done.size();
}

extern "C" int cudasw_poll(cudasw_t* h, int64_t ticket, cudasw_result* res, uint32_t* cigar,
                           int32_t cigar_cap, int32_t* cigar_len) {
    if (!h || !res) return CUDASW_EINVAL;
    std::lock_guard<std::mutex> lock(h->qmu);
    auto it = h->done.find(ticket);
    if (it == h->done.end()) return h->live.count(ticket) ? 0 : CUDASW_EINVAL;  // pending or on the GPU / unknown
    return deliver_locked(h, it, res, cigar, cigar_cap, cigar_len);
}

extern "C" int cudasw_wait(cudasw_t* h, int64_t ticket, int64_t timeout_ms, cudasw_result* res, uint32_t* cigar,
                           int32_t cigar_cap, int32_t* cigar_len) {
    if (!h || !res) return CUDASW_EINVAL;
    std::unique_lock<std::mutex> lock(h->qmu);
    if (!h->live.count(ticket)) return CUDASW_EINVAL;
    auto ready = [&] { return h->done.find(ticket) != h->done.end(); };
    if (timeout_ms < 0) h->cv_done.wait(lock, ready);
    else if (!h->cv_done.wait_for(lock, std::chrono::milliseconds(timeout_ms), ready)) return 0;
    return deliver_locked(h, h->done.find(ticket), res, cigar, cigar_cap, cigar_len);
}

// Dispatcher: one thread per engine that turns the submit stream into GPU batches, flushing when `max_pairs` are
// pending or the oldest pending pair has waited `max_delay_us` (BwaStreamingAligner.java:79-127: size-triggered flush
// by a single driver thread). Worker threads then only submit and wait.
static void dispatcher_main(cudasw_t* h) {
    std::unique_lock<std::mutex> lock(h->qmu);
    while (h->disp_run) {
        if (h->pend_tickets.empty()) { h->cv_work.wait(lock); continue; }
        const auto deadline = h->first_pending + std::chrono::microseconds(h->disp_delay_us);
        if ((int64_t)h->pend_tickets.size() < h->disp_max_pairs && std::chrono::steady_clock::now() < deadline) {
            h->cv_work.wait_until(lock, deadline);
            continue;
        }
        lock.unlock();
        cudasw_flush(h);  // errors reach the callers as per-pair status
        lock.lock();
    }
}

extern "C" int cudasw_dispatcher_start(cudasw_t* h, int64_t max_pairs, int64_t max_delay_us) {
    if (!h || max_pairs <= 0 || max_delay_us < 0) return CUDASW_EINVAL;
    std::lock_guard<std::mutex> lock(h->qmu);
    if (h->disp_run) return CUDASW_EINVAL;
    h->disp_max_pairs = max_pairs;
    h->disp_delay_us = max_delay_us;
    h->disp_run = true;
    h->first_pending = std::chrono::steady_clock::now();
    h->disp = std::thread(dispatcher_main, h);
    return CUDASW_OK;
}

extern "C" int cudasw_dispatcher_stop(cudasw_t* h) {
    if (!h) return CUDASW_EINVAL;
    {
        std::lock_guard<std::mutex> lock(h->qmu);
        if (!h->disp_run) return CUDASW_OK;
        h->disp_run = false;
    }
    h->cv_work.notify_all();
    if (h->disp.joinable()) h->disp.join();
    return cudasw_flush(h);  // nothing stays pending behind a stopped dispatcher
}

extern "C" int cudasw_cigar_string(const uint32_t* ops, int32_t n_ops, char* buf, int32_t cap) {
    static const char kOps[] = "MIDNSHP=X";
    if (!buf || cap <= 0 || (n_ops && !ops)) return CUDASW_EINVAL;
    int32_t o = 0;
    for (int32_t i = 0; i < n_ops; i++) {
        char tmp[16];
        uint32_t op = ops[i] & 15u;
        int k = snprintf(tmp, sizeof tmp, "%u%c", ops[i] >> 4, op < 9 ? kOps[op] : '?');
        if (o + k >= cap) return CUDASW_EINVAL;
        memcpy(buf + o, tmp, (size_t)k);
        o += k;
    }
    buf[o] = 0;
    return o;
}
