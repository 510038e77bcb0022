// noalloc.cu - cudasw_score_batch_noalloc: the forward score pass for callers that cannot allocate (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// It runs the strip-mined forward kernel sw_score_kernel<32, 20, REV = 0, MULTI = 1, PROF = 0> of
// sw_kernels.cuh - the variant whose strips hand their boundary rows (H, F per column) to each other through scratch
// memory - with every piece of scratch the engine normally cudaMallocs (encoded copies of the two blobs, the PRMT
// constant table, the job list, the per-pair int4 records, the strip boundary rows and their bump counter) laid out
// in ONE file-scope `static __device__` array. Every GPU that loads libcudasw.so holds that array until exit.
//
// The array's name is a host-side stand-in in host code: its device address is taken once per device, with that
// device current, by cudaGetSymbolAddress below, and only that pointer is handed to kernels.
//
// Every call shares the array and finds the last call's contents in it: the counters are cleared by the first kernel
// of a call, everything else is written before it is read. Two calls that overlap on one GPU (different streams, or
// two host threads without their own ordering) race on it.
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>
#include <string>

#include "../../include/cudasw.h"
#include "sw_kernels.cuh"

namespace cusw {
namespace {

constexpr int kNaRows = 20;  // R of the kernel variant: strips of 2 * 32 * 20 = 1280 query rows
constexpr int kNaStripRows = 2 * 32 * kNaRows;

// ---- layout of the scratch array (byte offsets, all multiples of 256)
constexpr size_t kNaMaxPairs = CUDASW_NOALLOC_MAX_PAIRS;
constexpr size_t kNaMaxSeq = CUDASW_NOALLOC_MAX_SEQ_BYTES;
constexpr size_t kNaMaxRef = CUDASW_NOALLOC_MAX_REF_BYTES;
constexpr size_t kOffPool = 0;                                      // 32 lanes x 16 words, + 8 statistics words
constexpr size_t kOffCtl = kOffPool + 32 * 16 * 4 + 256;            // [0] strip bytes used (u64), [2],[3] job list {start, count}
constexpr size_t kOffOrder = kOffCtl + 256;                         // job list: ordinals of the pairs that pass the checks
constexpr size_t kOffFwd = kOffOrder + kNaMaxPairs * 4;             // int4 {score, ref_end, read_end, failed} per pair
constexpr size_t kOffEncSeq = kOffFwd + kNaMaxPairs * 16;           // encoded queries (sw_kernels.cuh "Encoded sequence bytes")
constexpr size_t kOffEncRef = kOffEncSeq + kNaMaxSeq;               // encoded windows
constexpr size_t kOffStrip = kOffEncRef + kNaMaxRef;                // strip boundary rows
// Worst case of the strip memory: a pair takes it only when its query exceeds one strip, 2 * align16(4 w) bytes each,
// so all such pairs together need at most 8 * (window bytes) + 32 * (query bytes / strip rows): the bump allocation
// below cannot run out inside the stated maxima.
constexpr size_t kStripBytes = 8 * kNaMaxRef + 32 * (kNaMaxSeq / kNaStripRows + 1);
constexpr size_t kNaScratchBytes = (kOffStrip + kStripBytes + 255) & ~(size_t)255;
static_assert(kOffCtl % 256 == 0 && kOffOrder % 256 == 0 && kOffFwd % 256 == 0 && kOffEncSeq % 256 == 0 &&
              kOffEncRef % 256 == 0 && kOffStrip % 256 == 0, "scratch regions are 256-byte aligned");
static_assert(kNaScratchBytes <= (size_t)CUDASW_NOALLOC_SCRATCH_BYTES, "the header states the size of the array");

static __device__ __align__(256) unsigned char g_noalloc_scratch[CUDASW_NOALLOC_SCRATCH_BYTES];

// First kernel of a call: clears what must start at zero and writes the scoring constants (one copy per lane, the
// layout sw_score_kernel reads through ScoreConst::pool_dev).
__global__ void noalloc_init_kernel(unsigned char* scratch, ScoreConst sc, int match, int mismatch, int ambiguous) {
    uint32_t* pool = reinterpret_cast<uint32_t*>(scratch + kOffPool);
    uint32_t* ctl = reinterpret_cast<uint32_t*>(scratch + kOffCtl);
    const int t = threadIdx.x;  // 512 threads: lane t / 16, word t % 16
    const int wd = t & 15;
    uint32_t v = 0u;
    if (wd == 0) v = sc.poolA;
    else if (wd == 1) v = sc.poolB;
    else if (wd == 2) v = sc.BB;
    else if (wd == 3) v = sc.GO2;
    else if (wd == 4) v = sc.NGE2;
    else if (wd == 5) v = sc.GOBB;
    else if (wd >= 8) {  // query profile word of code wd - 8 (unused by the xor-selector flavour; kept for the layout)
        const int c = wd - 8;
        for (int x = 0; x < 4; x++) {
            const int s = c < 4 ? (c == x ? match : mismatch) : ambiguous;
            v |= (uint32_t)(uint8_t)(int8_t)s << (8 * x);
        }
    }
    pool[t] = v;
    if (t < 16) pool[512 + t] = 0u;  // step-statistics words of CUSW_STEP_STATS builds
    if (t < 64) ctl[t] = 0u;
}

// Per pair: the checks of fwd_keys_kernel (empty -> CUDASW_EINVAL, beyond the engine's limits -> CUDASW_ETOOLARGE,
// offsets outside the stated blob sizes -> CUDASW_EINVAL), the result record every pair starts from, and the job list.
__global__ void noalloc_prep_kernel(unsigned char* scratch, int n, const int64_t* __restrict__ seq_off, const int64_t* __restrict__ ref_off,
                                    int64_t seq_bytes, int64_t ref_bytes, int max_rows, int max_cols, int32_t* __restrict__ out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int* ctl = reinterpret_cast<int*>(scratch + kOffCtl);
    int* order = reinterpret_cast<int*>(scratch + kOffOrder);
    const int64_t q0 = seq_off[p], q1 = seq_off[p + 1], w0 = ref_off[p], w1 = ref_off[p + 1];
    int status = CUDASW_OK;
    if (q0 < 0 || w0 < 0 || q1 > seq_bytes || w1 > ref_bytes || q1 <= q0 || w1 <= w0) status = CUDASW_EINVAL;
    else if (q1 - q0 > max_rows || w1 - w0 > max_cols) status = CUDASW_ETOOLARGE;
    out[4 * p + 0] = 0; out[4 * p + 1] = -1; out[4 * p + 2] = 0; out[4 * p + 3] = status;
    if (status == CUDASW_OK) order[atomicAdd(&ctl[3], 1)] = p;  // ctl[2] = 0: the list starts at order[0]
}

// clean() + symbol codes (SswJniAligner.java:71-79; the byte layout of sw_kernels.cuh): query byte = code * 0x11,
// window byte = code * 0x11 ^ 0x80. Small inputs: one byte per thread, coalesced.
__global__ void noalloc_encode_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, uint32_t xor_mask) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = (uint8_t)((base_code(in[i]) * 0x11u) ^ xor_mask);
}

// Last kernel of a call: the forward records -> the caller's array, as ends_only_kernel reports them.
__global__ void noalloc_finish_kernel(const unsigned char* scratch, int n, int32_t* __restrict__ out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n || out[4 * p + 3] != CUDASW_OK) return;
    const int4 f = reinterpret_cast<const int4*>(scratch + kOffFwd)[p];
    if (f.w != 0) { out[4 * p + 3] = CUDASW_ENOMEM; return; }  // strip memory ran out: excluded by the sizing above
    out[4 * p + 0] = f.x;
    if (f.x > 0) { out[4 * p + 1] = f.y; out[4 * p + 2] = f.z; }  // score 0: ref_end -1, read_end 0
}

std::mutex g_na_mu;
unsigned char* g_na_ptr[64];  // device address of the scratch array, per device; nullptr = not set up yet

struct ErrBuf {
    char* buf;
    size_t cap;
    int fail(int code, const std::string& msg) const {
        if (buf && cap > 0) snprintf(buf, cap, "%s", msg.c_str());
        return code;
    }
};

}  // namespace
}  // namespace cusw

using namespace cusw;

extern "C" int cudasw_score_batch_noalloc(const cudasw_scoring* s, int device, int32_t n_pairs,
                                          const uint8_t* d_seq_blob, const int64_t* d_seq_off, int64_t seq_bytes,
                                          const uint8_t* d_ref_blob, const int64_t* d_ref_off, int64_t ref_bytes,
                                          int32_t* d_out, void* stream, char* errbuf, size_t errbuf_cap) {
    const ErrBuf err{errbuf, errbuf_cap};
    if (errbuf && errbuf_cap > 0) errbuf[0] = 0;
    cudasw_scoring sc = {1, -4, -4, 6, 1};  // AlignerFactory.java:90
    if (s) sc = *s;
    // the checks of cudasw_create
    if (sc.gap_open < 0 || sc.gap_open > 255 || sc.gap_extend < 0 || sc.gap_extend > 255)
        return err.fail(CUDASW_EINVAL, "Gap open and extension penalties must be in [0, 255]");
    if (std::abs(sc.match) > 127 || std::abs(sc.mismatch) > 127 || std::abs(sc.ambiguous) > 127)
        return err.fail(CUDASW_EINVAL, "Scoring matrix values must fit into signed 8-bit integer");
    if (sc.mismatch > 0 || sc.ambiguous > 0 || sc.match < 0)
        return err.fail(CUDASW_EINVAL, "cudasw requires match >= 0 and mismatch, ambiguous <= 0");
    if (n_pairs < 0 || seq_bytes < 0 || ref_bytes < 0) return err.fail(CUDASW_EINVAL, "negative size");
    if ((size_t)n_pairs > kNaMaxPairs || (size_t)seq_bytes > kNaMaxSeq || (size_t)ref_bytes > kNaMaxRef) {
        char buf[256];
        snprintf(buf, sizeof buf,
                 "cudasw_score_batch_noalloc: %d pairs / %lld query bytes / %lld window bytes exceed the static scratch "
                 "(at most %zu pairs, %zu query bytes, %zu window bytes per call): split the batch or use cudasw_align_batch",
                 n_pairs, (long long)seq_bytes, (long long)ref_bytes, kNaMaxPairs, kNaMaxSeq, kNaMaxRef);
        return err.fail(CUDASW_ETOOLARGE, buf);
    }
    if (n_pairs > 0 && (!d_seq_blob || !d_seq_off || !d_ref_blob || !d_ref_off || !d_out)) return err.fail(CUDASW_EINVAL, "null pointer");
    if (device < 0 || device >= 64) return err.fail(CUDASW_EINVAL, "device ordinal out of range");

    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return err.fail(CUDASW_ECUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) + " (libcudasw has no CPU fallback)");
    struct Restore {
        int prev, dev;
        ~Restore() { if (prev != dev) cudaSetDevice(prev); }
    } restore{prev, device};
    if (prev != device && (e = cudaSetDevice(device)) != cudaSuccess) {
        restore.dev = prev;
        return err.fail(CUDASW_ECUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    }
    // set-up, once per device, with that device current: the device address behind the array's name
    unsigned char* scratch = nullptr;
    {
        std::lock_guard<std::mutex> l(g_na_mu);
        if (!g_na_ptr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) return err.fail(CUDASW_ECUDA, std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e));
            g_na_ptr[device] = static_cast<unsigned char*>(p);
        }
        scratch = g_na_ptr[device];
    }
    if (n_pairs == 0) return CUDASW_OK;  // a set-up-only call (make one before a stream capture begins)

    auto pk2 = [](int v) { return ((uint32_t)v & 0xFFFFu) | (((uint32_t)v & 0xFFFFu) << 16); };
    auto b = [](int v) { return (uint32_t)(uint8_t)(int8_t)v; };
    ScoreConst k{};
    k.poolA = b(sc.match) | (b(sc.mismatch) << 8) | (b(sc.mismatch) << 16) | (b(sc.mismatch) << 24);
    k.poolB = b(sc.ambiguous) * 0x01010101u;
    k.B = sc.gap_open;
    k.BB = pk2(sc.gap_open);
    k.GO2 = pk2(sc.gap_open);
    k.GOBB = pk2(2 * sc.gap_open);
    k.NGE2 = pk2(-sc.gap_extend);
    k.ge = sc.gap_extend;
    k.match = sc.match;
    k.pool_dev = reinterpret_cast<const uint32_t*>(scratch + kOffPool);
    const int max_rows = std::min(32767, sc.match > 0 ? (32767 - sc.gap_open) / sc.match : 32767);
    const int max_cols = (1 << 24) - 2;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int n = n_pairs;
    uint8_t* enc_seq = scratch + kOffEncSeq;
    uint8_t* enc_ref = scratch + kOffEncRef;
    noalloc_init_kernel<<<1, 512, 0, st>>>(scratch, k, sc.match, sc.mismatch, sc.ambiguous);
    noalloc_prep_kernel<<<(n + 255) / 256, 256, 0, st>>>(scratch, n, d_seq_off, d_ref_off, seq_bytes, ref_bytes, max_rows, max_cols, d_out);
    if (seq_bytes > 0)
        noalloc_encode_kernel<<<(int)std::min<int64_t>((seq_bytes + 255) / 256, 148 * 8), 256, 0, st>>>(d_seq_blob, enc_seq, seq_bytes, 0u);
    if (ref_bytes > 0)
        noalloc_encode_kernel<<<(int)std::min<int64_t>((ref_bytes + 255) / 256, 148 * 8), 256, 0, st>>>(d_ref_blob, enc_ref, ref_bytes, 0x80u);
    PairIO io{enc_seq, enc_ref, d_seq_off, d_ref_off};
    StripPool sp{scratch + kOffStrip, (unsigned long long)kStripBytes, reinterpret_cast<unsigned long long*>(scratch + kOffCtl)};
    const int* slot = reinterpret_cast<const int*>(scratch + kOffCtl) + 2;
    // one warp per pair, four pairs per CTA; the grid covers n, the kernel reads the true list bounds from `slot`
    sw_score_kernel<32, kNaRows, false, true, false><<<(n + 3) / 4, 128, 0, st>>>(
        reinterpret_cast<const int*>(scratch + kOffOrder), n, io, nullptr, reinterpret_cast<int4*>(scratch + kOffFwd), k, sp, slot);
    noalloc_finish_kernel<<<(n + 255) / 256, 256, 0, st>>>(scratch, n, d_out);
    e = cudaPeekAtLastError();
    if (e != cudaSuccess) return err.fail(CUDASW_ECUDA, std::string("kernel launch: ") + cudaGetErrorString(e));
    return CUDASW_OK;
}
