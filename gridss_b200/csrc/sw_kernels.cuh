// sw_kernels.cuh - sm_100a device code of the batched Smith-Waterman engine.
//
// Replaces the native hot loops behind GRIDSS's Aligner.align_smith_waterman
// (alignment/SswJniAligner.java:42-61 -> libsswjni.so: sw_sse2_byte / sw_sse2_word / banded_sw,
// SURVEY.md §8a A4). Nothing here is translated from SSW: the score pass is an inter-task
// register-resident wavefront (one pair per G-lane group, two virtual lanes per thread packed as
// s16x2 DPX operands), not Farrar's striped layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/cudasw.h"

// -DCUSW_BOUNDS_CHECK (python -m gridss_b200.build --check -> libcudasw_check.so): every data-dependent global-memory
// index of the score, reverse and traceback kernels is tested against the allocation / the pair's own bytes it must stay
// inside, and the kernel traps (the CUDA call fails, the test fails) on the first violation. compute-sanitizer is not
// available on the GPU pool; this is the memcheck the GPU test-suite runs under instead (profiles/bounds_check_r02.log).
#ifdef CUSW_BOUNDS_CHECK
#include <cstdio>
#define CUSW_CHECK_PTR(p, lo, hi, what)                                                                              \
    do {                                                                                                             \
        if ((const char*)(p) < (const char*)(lo) || (const char*)(p) >= (const char*)(hi)) {                         \
            printf("CUSW_BOUNDS_CHECK: %s out of range at %s:%d (off %lld of %lld)\n", what, __FILE__, __LINE__,      \
                   (long long)((const char*)(p) - (const char*)(lo)), (long long)((const char*)(hi) - (const char*)(lo))); \
            __trap();                                                                                                \
        }                                                                                                            \
    } while (0)
#else
#define CUSW_CHECK_PTR(p, lo, hi, what) ((void)0)
#endif

namespace cusw {

// ------------------------------------------------------------------------------------------
// Encoded sequence bytes.
//   code: A0 C1 G2 T3 N4 (SswJniAligner.clean, SswJniAligner.java:71-79), window pad 5, query pad 6
//   query byte  = code * 0x11            (both nibbles = code)
//   window byte = code * 0x11 ^ 0x80     (high nibble additionally carries PRMT's sign-replicate bit)
// so that (query ^ window) is directly the PRMT selector pair {x | 8, x}, x = qcode ^ wcode, and
//   prmt(pool) yields the int8 substitution score pool[x] sign-extended to s16.
// x == 0 <=> equal symbols (match, N vs N included); x in 1..3 <=> both ACGT (mismatch);
// x in 4..7 <=> exactly one N (ambiguous)  [SswJniAligner.createMatrix, SswJniAligner.java:19-33].
// ------------------------------------------------------------------------------------------
constexpr uint32_t kQueryPad = 6u * 0x11u;
constexpr uint32_t kWindowPad = (5u * 0x11u) ^ 0x80u;

__device__ __forceinline__ uint32_t base_code(uint32_t c) {
    c &= 0xDFu;  // upper-case letters; non-letters never land on A/C/G/T
    return c == 'A' ? 0u : c == 'C' ? 1u : c == 'G' ? 2u : c == 'T' ? 3u : 4u;
}

#ifdef CUSW_MAIN_TU
// Elementwise encode of a whole blob (layout unchanged, so per-pair offsets still apply).
// HBM-bound: 1 byte in, 1 byte out; 16-byte vector accesses.
__device__ __forceinline__ uint32_t encode_word(uint32_t w, uint32_t xor_word, bool& seen_n) {
    // four bases per word, byte-parallel: upper-case, compare against A/C/G/T, everything else is N (4)
    const uint32_t u = w & 0xDFDFDFDFu;
    const uint32_t mC = __vcmpeq4(u, 0x43434343u), mG = __vcmpeq4(u, 0x47474747u), mT = __vcmpeq4(u, 0x54545454u);
    const uint32_t mN = ~(__vcmpeq4(u, 0x41414141u) | mC | mG | mT);
    const uint32_t cw = (mC & 0x01010101u) | (mG & 0x02020202u) | (mT & 0x03030303u) | (mN & 0x04040404u);
    seen_n |= mN != 0u;
    return (cw | (cw << 4)) ^ xor_word;
}

__global__ void __launch_bounds__(256) encode_kernel(const uint8_t* __restrict__ in,
                                                     uint8_t* __restrict__ out, int64_t n,
                                                     uint32_t xor_mask, int* __restrict__ any_n) {
    bool seen_n = false;
    const uint32_t xw = xor_mask * 0x01010101u;
    // a chunk starts at an arbitrary byte of the blob; input and output share the layout, hence the misalignment:
    // the first `head` bytes go one by one, the body in 16-byte vectors
    const bool same = ((reinterpret_cast<uintptr_t>(in) ^ reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    int64_t head = same ? (int64_t)((16 - (reinterpret_cast<uintptr_t>(in) & 15)) & 15) : n;
    if (head > n) head = n;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int64_t k = tid; k < head; k += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t code = base_code(in[k]);
        seen_n |= code == 4u;
        out[k] = (uint8_t)((code * 0x11u) ^ xor_mask);
    }
    in += head; out += head; n -= head;
    // 32 bytes per thread and iteration (two 16-byte loads in flight before the first use)
    int64_t i = tid * 32;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 32;
    for (; i < n; i += stride) {
        if (i + 32 <= n) {
            const uint4 v0 = __ldcs(reinterpret_cast<const uint4*>(in + i));
            const uint4 v1 = __ldcs(reinterpret_cast<const uint4*>(in + i + 16));
            uint4 o0, o1;
            o0.x = encode_word(v0.x, xw, seen_n); o0.y = encode_word(v0.y, xw, seen_n);
            o0.z = encode_word(v0.z, xw, seen_n); o0.w = encode_word(v0.w, xw, seen_n);
            o1.x = encode_word(v1.x, xw, seen_n); o1.y = encode_word(v1.y, xw, seen_n);
            o1.z = encode_word(v1.z, xw, seen_n); o1.w = encode_word(v1.w, xw, seen_n);
            *reinterpret_cast<uint4*>(out + i) = o0;
            *reinterpret_cast<uint4*>(out + i + 16) = o1;
        } else {
            for (int64_t k = i; k < n && k < i + 32; k++) {
                const uint32_t code = base_code(in[k]);
                seen_n |= code == 4u;
                out[k] = (uint8_t)((code * 0x11u) ^ xor_mask);
            }
        }
    }
    if (any_n && seen_n) *any_n = 1;  // some window of this batch holds an N: per-pair flags are needed
}

// Per-pair flag "the window contains N" (such pairs take the xor-selector flavour of the score kernel).
// One warp per pair over the encoded window; does nothing unless encode_kernel saw an N at all.
__global__ void __launch_bounds__(256) window_flags_kernel(int n, const uint8_t* __restrict__ enc_ref,
                                                           const int64_t* __restrict__ ref_off,
                                                           const int* __restrict__ any_n,
                                                           uint8_t* __restrict__ flags) {
    if (*any_n == 0) return;
    const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (p >= n) return;
    const int64_t b = ref_off[p], e = ref_off[p + 1];
    bool seen = false;
    for (int64_t i = b + (threadIdx.x & 31); i < e; i += 32) seen |= enc_ref[i] == ((4u * 0x11u) ^ 0x80u);
    seen = __any_sync(0xffffffffu, seen);
    if ((threadIdx.x & 31) == 0) flags[p] = seen ? 1 : 0;
}

// ------------------------------------------------------------------------------------------
// Device-resident reference genome. Mirrors PackedReferenceSequence
// (picard/TwoBitBufferedReferenceSequenceFile.java:97-125): 2 bits per base plus one "ambiguous" bit per base
// (every byte that clean() would turn into N, SswJniAligner.java:71-79). Contigs are concatenated in one linear
// coordinate g = contig_off[c] + pos; base g lives in bits 2*(g&15) of word g>>4, its N flag in bit g&31 of
// word g>>5. HBM-bound: 1 byte in, 3/8 byte out.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ref_pack_kernel(const uint8_t* __restrict__ in, int64_t n, int64_t g0,
                                                       uint32_t* __restrict__ two_bit, uint32_t* __restrict__ nmask) {
    // g0 is a multiple of 32; thread t packs bases [g0 + 32 t, g0 + 32 t + 32) of which `in` holds the first n
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; t * 32 < n; t += stride) {
        const uint8_t* p = in + t * 32;
        const int64_t left = n - t * 32;
        uint32_t lo = 0, hi = 0, nm = 0;
        uint32_t wds[8];
        if (left >= 32 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
            const uint4 a = reinterpret_cast<const uint4*>(p)[0], b = reinterpret_cast<const uint4*>(p)[1];
            wds[0] = a.x; wds[1] = a.y; wds[2] = a.z; wds[3] = a.w; wds[4] = b.x; wds[5] = b.y; wds[6] = b.z; wds[7] = b.w;
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                uint32_t v = 0;
                for (int b = 0; b < 4; b++) v |= (uint32_t)(4 * k + b < left ? p[4 * k + b] : 'A') << (8 * b);
                wds[k] = v;
            }
        }
#pragma unroll
        for (int k = 0; k < 32; k++) {
            const uint32_t code = base_code((wds[k >> 2] >> (8 * (k & 3))) & 0xFFu);
            const uint32_t two = code & 3u;  // an N stores 'A'; the mask bit overrides it
            if (k < 16) lo |= two << (2 * k); else hi |= two << (2 * (k - 16));
            nm |= (code >> 2) << k;
        }
        const int64_t g = (g0 >> 5) + t;
        two_bit[2 * g] = lo;
        two_bit[2 * g + 1] = hi;
        nmask[g] = nm;
    }
}

// Packed host inputs (cudasw_align_batch_packed): the same 2 bit + N mask layout as the resident reference, over
// the linear base coordinate of a whole blob. Expands bases [lo, hi) into the ENCODED byte layout the score
// kernels read (byte b of `out` = base b), i.e. what encode_kernel makes of the ASCII blob. HBM-bound: 3/8 byte
// in, 1 byte out per base. `two_bit` / `nmask` are indexed from word `word0` (in 16-base words; even), so a
// batch that starts deep inside a blob does not need the words before it. One thread per 16 bases.
__device__ __forceinline__ uint32_t spread4(uint32_t x) {  // 4 two-bit codes in bits 0..7 -> one code per byte
    uint32_t t = (x | (x << 12)) & 0x000F000Fu;
    return (t | (t << 6)) & 0x03030303u;
}
__global__ void __launch_bounds__(256) unpack_kernel(const uint32_t* __restrict__ two_bit,
                                                     const uint32_t* __restrict__ nmask, int64_t word0,
                                                     int64_t lo, int64_t hi, uint8_t* __restrict__ out,
                                                     uint32_t xor_mask, int* __restrict__ any_n) {
    const uint32_t xw = xor_mask * 0x01010101u;
    const int64_t g_lo = lo >> 4, g_hi = (hi + 15) >> 4;
    bool seen_n = false;
    for (int64_t g = g_lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < g_hi; g += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t w = two_bit[g - word0];
        uint32_t nm = nmask ? (nmask[(g - word0) >> 1] >> (16 * (int)(g & 1))) & 0xFFFFu : 0u;
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t nb = (((nm >> (4 * k)) & 0xFu) * 0x00204081u) & 0x01010101u;  // N flag of base k in byte k
            const uint32_t cw = (spread4((w >> (8 * k)) & 0xFFu) & ~(nb * 3u)) | (nb << 2);
            o[k] = (cw | (cw << 4)) ^ xw;
        }
        const int64_t b = g << 4;
        if (b >= lo && b + 16 <= hi) {
            *reinterpret_cast<uint4*>(out + b) = make_uint4(o[0], o[1], o[2], o[3]);
            seen_n |= nm != 0u;
        } else {  // first / last word of the range: only the bases inside [lo, hi) belong to this batch
#pragma unroll
            for (int k = 0; k < 16; k++)
                if (b + k >= lo && b + k < hi) {
                    out[b + k] = (uint8_t)(o[k >> 2] >> (8 * (k & 3)));
                    seen_n |= ((nm >> k) & 1u) != 0u;
                }
        }
    }
    if (any_n && seen_n) *any_n = 1;
}

// One window segment: bases [start, end) of contig `contig` (0-based, end exclusive), reverse-complemented when
// bit 0 of flags is set (BreakpointHomology.java:66-68 builds its windows from such pieces).
struct WindowSeg {
    int32_t contig, flags;
    int64_t start, end;
};

// Expands window segments straight into the ENCODED window layout the score kernels read (no ASCII round trip):
// one warp per pair, 4 bases per lane and iteration, 32-bit stores once the destination is aligned.
// ascii != 0 writes upper-case "ACGTN" instead (getSubsequenceAt, for cudasw_reference_fetch).
// The same kernel assembles QUERIES on the device (xor_mask 0 = the query encoding): a segment with contig < 0 takes
// bytes [start, end) of the literal blob `lit` (e.g. the untemplated sequence between the two breakend anchors,
// BreakpointHomology.java:67).
__global__ void __launch_bounds__(256)
window_gather_kernel(int n, const WindowSeg* __restrict__ segs, const int64_t* __restrict__ seg_off,
                     const int64_t* __restrict__ ref_off, const int64_t* __restrict__ contig_off,
                     const uint32_t* __restrict__ two_bit, const uint32_t* __restrict__ nmask,
                     uint8_t* __restrict__ out, uint8_t* __restrict__ flags, int* __restrict__ any_n, int ascii,
                     const uint8_t* __restrict__ lit = nullptr, uint32_t xor_mask = 0x80u, int seg_stride = 0) {
    const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= n) return;
    uint8_t* dst = out + ref_off[p];
    const int64_t wlen = ref_off[p + 1] - ref_off[p];
    int64_t done = 0;
    bool seen_n = false;
    bool literal = false;  // contig < 0: bytes [start, end) of `lit` (raw ASCII, never reverse-complemented here)
    auto fetch = [&](int64_t g, bool rc) -> uint32_t {
        uint32_t code;
        if (literal) {
            code = base_code(lit[g]);
            if (code == 4u) seen_n = true;
        } else {
            code = (two_bit[g >> 4] >> (2 * (int)(g & 15))) & 3u;
            if (rc) code = 3u - code;
            if ((nmask[g >> 5] >> (int)(g & 31)) & 1u) { code = 4u; seen_n = true; }
        }
        if (ascii) return code == 0 ? 'A' : code == 1 ? 'C' : code == 2 ? 'G' : code == 3 ? 'T' : 'N';
        return (code * 0x11u) ^ xor_mask;
    };
    // seg_stride > 0: every pair owns exactly that many descriptor slots (empty ones have start == end), no offsets
    const int64_t s_lo = seg_stride > 0 ? (int64_t)p * seg_stride : seg_off[p];
    const int64_t s_hi = seg_stride > 0 ? s_lo + seg_stride : seg_off[p + 1];
    for (int64_t s = s_lo; s < s_hi && done < wlen; s++) {
        const WindowSeg sg = segs[s];
        const int64_t len = min(sg.end - sg.start, wlen - done);  // the host sized ref_off from the same segments
        literal = sg.contig < 0;
        const bool rc = !literal && (sg.flags & 1) != 0;
        const int64_t g0 = literal ? sg.start : contig_off[sg.contig] + sg.start;
        uint8_t* d = dst + done;
        const int64_t head = min((int64_t)((4 - (reinterpret_cast<uintptr_t>(d) & 3)) & 3), len);
        if (lane < head) d[lane] = (uint8_t)fetch(rc ? g0 + len - 1 - lane : g0 + lane, rc);
        for (int64_t k = head + 4 * lane; k < len; k += 128) {
            if (k + 4 <= len) {
                uint32_t v = 0;
#pragma unroll
                for (int b = 0; b < 4; b++) v |= fetch(rc ? g0 + len - 1 - (k + b) : g0 + k + b, rc) << (8 * b);
                *reinterpret_cast<uint32_t*>(d + k) = v;
            } else {
                for (int64_t b = k; b < len; b++) d[b] = (uint8_t)fetch(rc ? g0 + len - 1 - b : g0 + b, rc);
            }
        }
        done += len;
    }
    seen_n = __any_sync(0xffffffffu, seen_n);
    if (lane == 0) {
        if (flags) flags[p] = seen_n ? 1 : 0;
        if (any_n && seen_n) *any_n = 1;
    }
}

#endif  // CUSW_MAIN_TU

// ------------------------------------------------------------------------------------------
// Score-pass variants: a pair occupies a group of G lanes; each thread owns two virtual lanes of
// R query rows each, so one group covers 2*G*R rows.
// ------------------------------------------------------------------------------------------
struct ScoreConst {
    uint32_t poolA, poolB;  // PRMT byte pool: pool[x] = substitution score of symbol-xor x
    uint32_t BB;            // bias B = gap_open in both halves (stored value = true value + B)
    uint32_t GO2;           // gap_open in both halves
    uint32_t GOBB;          // BB + GO2: the biased value of H == gap_open (quiet-step threshold)
    uint32_t NGE2;          // -gap_extend in both halves (two's complement per half)
    int32_t B;      // = gap_open
    int32_t ge, match;
    const uint32_t* pool_dev;  // 32 x {poolA, poolB, BB, GO2, NGE2, GOBB, 0, 0, profile[8]} in device memory (one copy per lane)
};

// What the k-mer filter (band_filter.cuh) decided for a pair. margin_bins == 0: no certificate sought, the forward pass
// sweeps the whole window.
struct FilterRec {
    int32_t c0, wlen;     // the forward pass sweeps window columns [c0, c0 + wlen)
    int32_t margin_bins;  // bins of margin either side of the core: m_band | (m_old << 8) [| kFltWholeWindow]
    int32_t noise;        // most hits in any run of m_old bins outside the band's core
    int32_t gap_bound;    // once certified: no alignment of the pair's score has more gap bases than this (-1: not known)
    int32_t dlo, dhi;     // kFltBand: the diagonals (column - row) the banded forward kernel sweeps, inclusive
};

constexpr int32_t kFltWholeWindow = 1 << 30;  // FilterRec::margin_bins: the range is the whole window, the certificate only yields gap_bound
constexpr int32_t kFltBand = 1 << 29;         // FilterRec::margin_bins: the pair's first forward pass is the banded kernel (rev_band.cuh, FWD)

struct PairIO {
    const uint8_t* enc_seq;
    const uint8_t* enc_ref;
    const int64_t* seq_off;
    const int64_t* ref_off;
    const FilterRec* sub = nullptr;  // forward pass only: per-pair column range (same indexing as the offsets)
};

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s));
    return d;
}

// Sort keys. Forward: (variant, window length); reverse: (variant, columns to walk).
// variant_of() is shared with the host dispatch table (cudasw.cu).
struct VariantTable {
    int n;          // all variants
    int n_single;   // the first n_single are single-strip, ascending by rows
    int rows[100];  // rows covered by variant v (rows per strip for the strip-mined ones)
    int lanes[100]; // G of variant v
};

__host__ __device__ __forceinline__ int variant_of(const VariantTable& vt, int q) {
    for (int v = 0; v < vt.n_single; v++)
        if (vt.rows[v] >= q) return v;
    // strip-mined: least padded rows, ties to the later (taller strips) variant
    int bestv = vt.n - 1;
    long long bestpad = -1;
    for (int v = vt.n_single; v < vt.n; v++) {
        long long padded = (long long)((q + vt.rows[v] - 1) / vt.rows[v]) * vt.rows[v];
        if (bestpad < 0 || padded <= bestpad) { bestpad = padded; bestv = v; }
    }
    return bestv;
}

#ifdef CUSW_MAIN_TU
// ------------------------------------------------------------------------------------------
// Scheduling hint (never affects a result): the window column where a pair's alignment probably starts,
// from 8-mer seeds. The four pairs that share a warp of the score kernel leave their quiet steps only
// while their own alignment crosses the wavefront; binning pairs with similar columns together makes
// those busy stretches coincide. One warp per pair: every 8-mer of the query goes into a small tagged hash
// table in shared memory, every 8th window 8-mer is looked up and votes for a diagonal bucket of 32 columns.
// ------------------------------------------------------------------------------------------
constexpr int kSeedK = 8, kSeedStride = 8, kSeedTable = 1024, kSeedBins = 512, kSeedWarps = 8;
__device__ __forceinline__ bool seed_kmer(const uint8_t* __restrict__ p, uint32_t& km) {
    km = 0;
    bool ok = true;
#pragma unroll
    for (int x = 0; x < kSeedK; x++) {
        const uint32_t c = p[x] & 7u;  // low nibble of an encoded byte = base code, >= 4: N / pad
        ok &= c < 4u;
        km = (km << 2) | (c & 3u);
    }
    return ok;
}
__global__ void __launch_bounds__(32 * kSeedWarps)
seed_anchor_kernel(int n, const uint8_t* __restrict__ enc_seq, const uint8_t* __restrict__ enc_ref,
                   const int64_t* __restrict__ seq_off, const int64_t* __restrict__ ref_off,
                   uint16_t* __restrict__ anchor, const FilterRec* __restrict__ sub = nullptr) {
    // sub: pairs the k-mer filter gave a column range (band_filter.cuh) need no anchor (fwd_keys_kernel ignores it)
    __shared__ uint32_t table[kSeedWarps][kSeedTable];
    __shared__ unsigned int bins[kSeedWarps][kSeedBins / 2];  // two 16-bit counters per word
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    uint32_t* tab = table[wib];
    unsigned int* bin = bins[wib];
    for (int i = lane; i < kSeedTable; i += 32) tab[i] = 0u;
    uint32_t tag = 0;
    for (int p = blockIdx.x * kSeedWarps + wib; p < n; p += gridDim.x * kSeedWarps) {
        const int64_t qo = seq_off[p], ro = ref_off[p];
        const int q = (int)(seq_off[p + 1] - qo), w = (int)(ref_off[p + 1] - ro);
        int col = 0;
        if (sub && sub[p].margin_bins != 0 && !(sub[p].margin_bins & kFltWholeWindow)) {
            if (lane == 0) anchor[p] = 0;
            continue;
        }
        if (q >= kSeedK && w >= kSeedK && q <= 4096 && w + q <= 32 * kSeedBins) {  // uniform per warp
            tag = tag == 0xFFFFFu ? 1u : tag + 1u;
            for (int i = lane; i < kSeedBins / 2; i += 32) bin[i] = 0u;
            const uint8_t* s = enc_seq + qo;
            for (int qp = lane; qp + kSeedK <= q; qp += 32) {
                uint32_t km;
                if (seed_kmer(s + qp, km)) tab[(km ^ (km >> 6)) & (kSeedTable - 1)] = (tag << 12) | (uint32_t)qp;
            }
            __syncwarp();
            const uint8_t* r = enc_ref + ro;
            for (int wp = kSeedStride * lane; wp + kSeedK <= w; wp += 32 * kSeedStride) {
                uint32_t km;
                const bool ok = seed_kmer(r + wp, km);
                const uint32_t ent = tab[(km ^ (km >> 6)) & (kSeedTable - 1)];
                if (ok && (ent >> 12) == tag) {
                    // a table slot is keyed by 10 hash bits only: confirm the 8-mer before voting
                    uint32_t k2;
                    const int qp = (int)(ent & 0xFFFu);
                    if (seed_kmer(s + qp, k2) && k2 == km) {
                        const int d = wp - qp + q;  // diagonal, shifted to be >= 0
                        atomicAdd(bin + (d >> 6), (d >> 5) & 1 ? 0x10000u : 1u);
                    }
                }
            }
            __syncwarp();
            int bestv = 0, bestb = 0;
            for (int i = lane; i < kSeedBins / 2; i += 32) {
                const unsigned int v2 = bin[i];
                const int lo = (int)(v2 & 0xFFFFu), hi = (int)(v2 >> 16);
                if (lo > bestv) { bestv = lo; bestb = 2 * i; }
                if (hi > bestv) { bestv = hi; bestb = 2 * i + 1; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const int ov = __shfl_xor_sync(0xffffffffu, bestv, o), ob = __shfl_xor_sync(0xffffffffu, bestb, o);
                if (ov > bestv || (ov == bestv && ob < bestb)) { bestv = ov; bestb = ob; }
            }
            col = bestv >= 2 ? max(bestb * 32 - q, 0) : 0;
            __syncwarp();
        }
        if (lane == 0) anchor[p] = (uint16_t)min(col >> 2, 4095);
    }
}

// keys for the forward pass; hist[slot] counts pairs per (flavour, variant); the last slot counts invalid pairs
__global__ void fwd_keys_kernel(int n, const int64_t* __restrict__ seq_off,
                                const int64_t* __restrict__ ref_off, VariantTable vt, int max_rows,
                                int max_cols, uint32_t* __restrict__ keys, int* __restrict__ vals,
                                int* __restrict__ hist, cudasw_result* __restrict__ res,
                                const int* __restrict__ any_n, const uint8_t* __restrict__ wflags,
                                const uint16_t* __restrict__ anchor, unsigned long long* __restrict__ cell_counts,
                                const FilterRec* __restrict__ sub = nullptr, const uint8_t* __restrict__ done = nullptr, int first_call = 1,
                                int skip_band = 0) {
    // sub: the k-mer filter's column ranges (band_filter.cuh): the forward pass sweeps ws = sub[p].wlen columns.
    // done != nullptr: pairs with done[p] != 0 have nothing left to do (their result is proven, band_filter.cuh).
    // first_call: the chunk's first binning initialises the result records and counts its q x w cells.
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    // cell_counts: [0] sum of q*w, [1] cells the forward kernels compute (padded rows, fill/drain columns)
    unsigned long long my_cells = 0, my_padded = 0;
    int64_t q = 0, w = 0, ws = 0;
    // skip_band: pairs the filter gave a diagonal band are swept by the banded forward kernel first (rev_band.cuh, FWD)
    const bool skip = p < n && ((done && done[p]) || (skip_band && sub && (sub[p].margin_bins & kFltBand)));
    bool banded = false;
    if (p < n) {
        q = seq_off[p + 1] - seq_off[p];
        w = ref_off[p + 1] - ref_off[p];
        ws = w;
        if (sub) { const FilterRec fr = sub[p]; ws = fr.wlen; banded = fr.margin_bins != 0 && !(fr.margin_bins & kFltWholeWindow); }
        if (q > 0 && w > 0) {
            if (first_call) my_cells = (unsigned long long)q * (unsigned long long)w;
            if (q <= max_rows && !skip) {
                const int v0 = variant_of(vt, (int)q);
                const unsigned long long strips = ((unsigned long long)q + vt.rows[v0] - 1) / vt.rows[v0];
                my_padded = strips * (unsigned long long)vt.rows[v0] * (unsigned long long)(ws + 2 * vt.lanes[v0] - 1);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o);
        my_padded += __shfl_xor_sync(0xffffffffu, my_padded, o);
    }
    if ((threadIdx.x & 31) == 0 && (my_cells | my_padded)) { atomicAdd(cell_counts, my_cells); atomicAdd(cell_counts + 1, my_padded); }
    if (p >= n) return;
    cudasw_result r;
    r.score = 0; r.ref_begin = -1; r.ref_end = -1; r.read_begin = 0; r.read_end = 0;
    r.cigar_off = 0; r.cigar_len = 0; r.status = CUDASW_OK;
    int v;
    // slots [0, n): query-profile flavour, [n, 2n): xor-selector flavour, 2n: nothing to do
    if (q <= 0 || w <= 0) { r.status = CUDASW_EINVAL; v = 2 * vt.n; }
    else if (q > max_rows || w > max_cols) { r.status = CUDASW_ETOOLARGE; v = 2 * vt.n; }
    else if (skip) v = 2 * vt.n;
    else v = variant_of(vt, (int)q) + ((*any_n != 0 && wflags[p]) ? vt.n : 0);
    if (first_call) res[p] = r;
    w = ws;
    // key = (slot : 7 | 0x1FFF - window length / 64 : 13 | seed anchor column / 4 : 12): inside a slot the LONGEST windows
    // come first, so the CTAs that run longest start first and a launch does not end in a tail of its slowest pairs
    // (mixed-length batches: C3's forward pass took twice as long with its variants launched one after the other)
    uint32_t wq = (uint32_t)(w >> 6);
    if (wq > 0x1FFFu) wq = 0x1FFFu;
    // (a banded pair's alignment sits in the middle of its column range: no anchor needed to line the warps up)
    keys[p] = ((uint32_t)v << 25) | ((0x1FFFu - wq) << 12) | (banded ? 0u : (uint32_t)anchor[p]);
    vals[p] = p;
    atomicAdd(&hist[v], 1);
}

// keys for the reverse pass (needs the forward result): rows = read_end+1, columns walked ~ rows
__global__ void rev_keys_kernel(int n, const int4* __restrict__ fwd,
                                cudasw_result* __restrict__ res, VariantTable vt,
                                uint32_t* __restrict__ keys, int* __restrict__ vals,
                                int* __restrict__ hist, const int* __restrict__ any_n,
                                const uint8_t* __restrict__ wflags) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int4 f = fwd[p];
    int v;
    uint32_t k2 = 0;
    if (res[p].status == CUDASW_OK && f.w != 0) res[p].status = CUDASW_ENOMEM;  // strip buffers exhausted
    if (res[p].status != CUDASW_OK || f.x <= 0) v = 2 * vt.n;  // nothing to walk back
    else {
        v = variant_of(vt, f.z + 1) + ((*any_n != 0 && wflags[p]) ? vt.n : 0);
        k2 = (uint32_t)min(f.y + 1, f.z + 1 + 64) >> 4;
        if (k2 > 0xFFFFFu) k2 = 0xFFFFFu;
    }
    keys[p] = ((uint32_t)v << 20) | k2;
    vals[p] = p;
    atomicAdd(&hist[v], 1);
}

#endif  // CUSW_MAIN_TU

// ------------------------------------------------------------------------------------------
// The score pass (forward: phase 1 of ssw_align; REV: phase 2, the reverse pass with early stop).
//
// Recurrences, in biased form v' = v + B with B = gap_open so that every
// stored half-word is >= 0 and `H - gap_open` is a plain 32-bit subtract:
//     H = max(0, Hdiag + s, E, F);  E' = max(E - ge, H - go);  F' = max(F - ge, H - go)
// Per packed cell pair: LOP3 (selector), PRMT (score), VIMNMX3, VIADDMNMX, IADD, 2x VIADDMNMX,
// + 1/2 VIMNMX3 for the column maximum.
//
// Wavefront: virtual lane v = 2*lig + half owns rows [v*R, (v+1)*R) and is at column t - v at
// step t. The low half hands its bottom row to the high half in-thread; the high half hands it
// to the next lane's low half by one shuffle per value.
//
// Tie-breaks (ssw.c:283-306 per SURVEY.md): result = first column that strictly raises the running
// maximum, then the smallest row holding it. Each virtual lane keeps (best, col, row) with strict
// improvement; the group reduces with the key (score, -col, -row).
// ------------------------------------------------------------------------------------------
// MULTI (G == 32 only): the query is strip-mined into blocks of 2*G*R rows processed one after the
// other by the same warp; the bottom row (H, F) of a strip is written per column to a boundary
// buffer in global memory (one 32-bit word per column) and read back by lane 0 of the next strip.
// This is the long-contig path (BASELINE.json C5: 10 kb contigs vs 20 kb windows).
struct StripPool {
    uint8_t* base;
    unsigned long long cap;
    unsigned long long* used;
};

// Most gap bases (insertions + deletions, over all gaps) an alignment over `rows` query rows can contain and still
// score `target`: its matches give at most rows * match, k gaps of L bases in total cost k*go + (L-k)*ge.
// -1 = unbounded (a gap can grow for free).
__host__ __device__ __forceinline__ long long max_gap_bases(long long rows, int match, int target, int go, int ge) {
    const long long budget = rows * match - target;
    if (budget < 0) return 0;
    if (go >= ge) {  // cheapest: one gap
        if (budget < go) return 0;
        if (ge == 0) return -1;
        return (budget - go) / ge + 1;
    }
    if (go == 0) return -1;  // cheapest: L gaps of one base
    return budget / go;
}

// Row (per half) of the first H equal to `target` = number of rows whose running prefix maximum is still below
// it. A half whose column is out of range may hold values above its target (they are never used): the prefix
// maximum is clamped so that the packed subtract cannot borrow across halves.
template <int R>
__device__ __forceinline__ uint32_t first_row_of(const uint32_t (&H)[R], uint32_t target, uint32_t floor2) {
    uint32_t pm = floor2, cnt = 0u;
#pragma unroll
    for (int k = 0; k < R; k++) {
        pm = __vmins2(__vmaxs2(pm, H[k]), target);
        cnt += __vminu2(target - pm, 0x00010001u);
    }
    return cnt;
}

// PROF (query profile flavour, the default): every row keeps its four substitution scores against A, C, G, T for
// both halves in two registers (QA: low virtual lane, QB: high one) and the window byte of the step is the PRMT
// selector, so a cell pair needs ONE PRMT and no LOP3. Query N and pad rows are profile rows of `ambiguous`.
// The window must be free of N (pairs whose window has one take the !PROF flavour, binned apart), and
// out-of-range columns read as 'A': a half's state is reset when it reaches column 0 and its column maximum is
// masked while it is outside [0, w), which is all a pad symbol was needed for.
// Resident CTAs per SM the register allocation is tuned for: with the query profile a thread holds 5 registers
// per row (HA, HB, E, QA, QB). Four CTAs (128 registers) measured 6.04 vs 5.63 TCUPS for three (R = 19, B200).
template <int R, bool PROF>
constexpr int score_min_blocks() { return (PROF ? 5 : 4) * R + 32 <= 128 ? 4 : (PROF ? 5 : 4) * R + 32 <= 168 ? 3 : 2; }

// CHN: the quiet-chain flavour of the forward pass (see qstep below). Both flavours are built: the host takes the
// chained one unless many different variants run side by side (cudasw.cu, launch_variants).
template <int G, int R, bool REV, bool MULTI, bool PROF, bool CHN = false>
__global__ void __launch_bounds__(128, score_min_blocks<R, PROF>())
sw_score_kernel(const int* __restrict__ order, int njobs, PairIO io, const int4* __restrict__ fwd,
                int4* __restrict__ out, ScoreConst sc, StripPool sp, const int* __restrict__ slot) {
    static_assert(!MULTI || G == 32, "strip mining uses one warp per pair");
    constexpr int GPB = 128 / G;
    constexpr int S = 2 * G * R;  // rows per strip
    const int lig = threadIdx.x % G;
    const int gidx = blockIdx.x * GPB + threadIdx.x / G;
    // slot != nullptr: the job list is order[slot[0] .. slot[0] + slot[1]), binned on the device; the grid was sized from
    // an upper bound the host knows (pairs per query-length class), so trailing warps may have nothing to do
    if (slot) { njobs = slot[1]; order += slot[0]; }
    const bool active = gidx < njobs;
    if (!__any_sync(0xffffffffu, active)) return;  // (warps never synchronise with each other in this kernel)

    int q = 0, w = 0, target = -1, pair = -1;
    const uint8_t* qptr = io.enc_seq;
    const uint8_t* rptr = io.enc_ref;
#ifdef CUSW_BOUNDS_CHECK
    const uint8_t *ck_q0 = nullptr, *ck_q1 = nullptr, *ck_r0 = nullptr, *ck_r1 = nullptr;  // this pair's own bytes
    const uint32_t *ck_b0 = nullptr, *ck_b1 = nullptr;                                      // its strip boundary rows
    if (active) {
        const int pr = order[gidx];
        ck_q0 = io.enc_seq + io.seq_off[pr]; ck_q1 = io.enc_seq + io.seq_off[pr + 1];
        ck_r0 = io.enc_ref + io.ref_off[pr]; ck_r1 = io.enc_ref + io.ref_off[pr + 1];
    }
#endif
    if (active) {
        pair = order[gidx];
        int64_t qo = io.seq_off[pair], ro = io.ref_off[pair];
        if (REV) {
            int4 f = fwd[pair];
            target = f.x;
            w = f.y + 1;
            q = f.z + 1;
            qptr += qo + f.z;
            rptr += ro + f.y;
            // An alignment of `target` inside q rows spans at most q + D columns, D = the deletions its
            // score budget can pay for; columns beyond cannot hold the first cell reaching `target`.
            const long long dmax = max_gap_bases(q, sc.match, target, sc.B, sc.ge);
            if (dmax >= 0) {
                const long long lim = (long long)q + dmax + 1;
                if (lim < w) w = (int)lim;
            }
        } else {
            q = (int)(io.seq_off[pair + 1] - qo);
            w = (int)(io.ref_off[pair + 1] - ro);
            qptr += qo;
            rptr += ro;
            if (io.sub) {  // the filter's column range: a window of its own as far as this kernel is concerned
                const FilterRec fr = io.sub[pair];
                rptr += fr.c0;
                w = fr.wlen;
            }
        }
    }
    constexpr int STEP = REV ? -1 : 1;
    // the PRMT pool lives in ordinary per-thread registers: read through a volatile load so that ptxas neither
    // keeps it in uniform registers nor re-materialises a copy in front of every PRMT of the row loops
    uint32_t cPA, cPB, cBB, cGO2, cNGE2, cGOBB;
    const uint32_t* cptr = sc.pool_dev + 16 * (threadIdx.x & 31);  // one copy per lane: ptxas cannot prove uniformity
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(cPA) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+4];" : "=r"(cPB) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+8];" : "=r"(cBB) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+12];" : "=r"(cGO2) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+16];" : "=r"(cNGE2) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+20];" : "=r"(cGOBB) : "l"(cptr));

    // strip-mining state
    int nstrips = 1;
    uint32_t* bnd_in = nullptr;
    uint32_t* bnd_out = nullptr;
    bool failed = false;
    if (MULTI) {
        nstrips = active ? (q + S - 1) / S : 0;
        if (nstrips > 1) {
            unsigned long long off = 0;
            const unsigned long long need = (((unsigned long long)w * 4ull + 15ull) & ~15ull) * 2ull;
            if (lig == 0) off = atomicAdd(sp.used, need);
            off = __shfl_sync(0xffffffffu, off, 0);
            if (off + need > sp.cap) { failed = true; nstrips = 0; }
            else {
                bnd_in = reinterpret_cast<uint32_t*>(sp.base + off);
                bnd_out = reinterpret_cast<uint32_t*>(sp.base + off + need / 2);
#ifdef CUSW_BOUNDS_CHECK
                ck_b0 = bnd_in; ck_b1 = reinterpret_cast<uint32_t*>(sp.base + off + need);
#endif
            }
        }
    }
    unsigned long long key = 0;  // running (score, -col, -row) maximum of this thread's virtual lanes

    for (int strip = 0; strip < (MULTI ? nstrips : 1); ++strip) {
    const int row0 = strip * S;

    // query rows -> PRMT selector halves (low byte: low virtual lane, high byte: high one), or (PROF) the
    // rows' profile words: byte x = substitution score of the row's base against window base x
    uint32_t Q[R], QB[PROF ? R : 1];
    {
        const int r0 = row0 + (2 * lig) * R;
#pragma unroll
        for (int k = 0; k < R; k++) {
            int a = r0 + k, b = r0 + R + k;
            if (a < q) CUSW_CHECK_PTR(qptr + (int64_t)a * STEP, ck_q0, ck_q1, "query row");
            if (b < q) CUSW_CHECK_PTR(qptr + (int64_t)b * STEP, ck_q0, ck_q1, "query row");
            uint32_t qa = a < q ? (uint32_t)qptr[(int64_t)a * STEP] : kQueryPad;
            uint32_t qb = b < q ? (uint32_t)qptr[(int64_t)b * STEP] : kQueryPad;
            if (PROF) {
                Q[k] = cptr[8 + (qa & 7u)];
                QB[k] = cptr[8 + (qb & 7u)];
            } else {
                Q[k] = qa | (qb << 8);
            }
        }
    }
    // H ping-pongs between two register arrays (HA -> HB -> HA ...) so that `diag = old H[k]` needs no
    // register moves; the step loop is therefore unrolled by two and the trip count rounded up to even
    // (an extra step only sees pad columns, which can never strictly improve a maximum).
    uint32_t HA[R], HB[R], E[R];
#pragma unroll
    for (int k = 0; k < R; k++) { HA[k] = sc.BB; HB[k] = sc.BB; E[k] = 0u; }

    uint32_t best = sc.BB;
    int col_lo = -1, col_hi = -1, row_lo = 0, row_hi = 0;
    uint32_t pend = 0u;  // `changed` word of the last step: a non-zero half = column recorded, row search still owed
    uint32_t Hout = sc.BB, Fout = 0u, Hdiag_top = sc.BB, Emax = 0u;
    constexpr uint32_t kOut = PROF ? 0x80u : kWindowPad;  // PROF: 'A'; state reset / maximum mask take care of it
    uint32_t Rsel = kOut | (kOut << 8);

    int t_end = active ? w + 2 * G - 1 : 0;
    int t_end_warp = (__reduce_max_sync(0xffffffffu, t_end) + 1) & ~1;

    // prefetch the window byte (and, when strip-mining, the boundary word) of step 0
    int c = -2 * lig;  // column of the low virtual lane at step t
    if ((unsigned)c < (unsigned)w) CUSW_CHECK_PTR(rptr + (int64_t)c * STEP, ck_r0, ck_r1, "window column");
    uint32_t rb_next = ((unsigned)c < (unsigned)w) ? (uint32_t)rptr[(int64_t)c * STEP] : kOut;
    uint32_t top_next = 0u;  // lane 0 only: (H | F << 16) of the row above this strip at column c
    const bool read_top = MULTI && strip > 0 && lig == 0;
    const bool write_bot = MULTI && strip + 1 < nstrips && lig == G - 1;
    if (MULTI) {
        top_next = sc.BB & 0xFFFFu;
        if (read_top && (unsigned)c < (unsigned)w) { CUSW_CHECK_PTR(bnd_in + c, ck_b0, ck_b1, "strip boundary read"); top_next = bnd_in[c]; }
    }

    // ---- quiet chains (CHAIN flavours only; see qstep below)
#ifndef CUSW_CHAIN_GATE  // A/B switch for tools/exp_build_run.sh
#define CUSW_CHAIN_GATE 2  // (3 before the k-mer filter: its column ranges are ~2.5 x the rows, and chains pay there: C2 forward 24.0 -> 23.05 ms per 400 k pairs)
#endif
    constexpr bool CHAIN = CHN && PROF && !REV && !MULTI;
    bool chain = false;  // warp-uniform: the previous step was a verified quiet step
    bool skipq = false;  // warp-uniform: the step about to be redone by `step` is known not to be quiet
    int tt = 0;          // step counter (warp-uniform), c == tt - 2 * lig
    // steps [2G, t_hi) are interior for every lane of the warp: both halves inside [2, w), columns up to c + 2 loadable
    // from a pair of steps that starts at tt + 2 < t_hi
    // Chains pay when most of the sweep is away from the alignment (measured, profiles/chain_sweep_r01j.txt): they are
    // switched off for warps whose shortest window is below CUSW_CHAIN_GATE times the rows of a group.
    int t_hi = 0;
    int t_hb = 0;           // steps [2G, t_hb) are interior in the same sense, whatever the gate says (busy steps, see bstep)
    bool was_full = false;  // warp-uniform: the last general step took the full recurrence
#ifndef CUSW_NO_BUSY  // A/B switch
    constexpr bool BUSY = PROF && !REV && !MULTI;
#else
    constexpr bool BUSY = false;
#endif
    if (CHAIN || BUSY) {
        const int w_min = __reduce_min_sync(0xffffffffu, active ? w : 0);
        t_hi = CHAIN && w_min >= CUSW_CHAIN_GATE * S ? w_min - 1 : 0;
        t_hb = w_min - 1;
    }

    auto step = [&](uint32_t (&Ho)[R], uint32_t (&Hn)[R]) {
        // ---- inputs from the virtual lane above
        uint32_t rH = __shfl_up_sync(0xffffffffu, Hout, 1, G);
        uint32_t rF = __shfl_up_sync(0xffffffffu, Fout, 1, G);
        if (lig == 0) {  // row above the strip: matrix row -1 (H = 0, F = -inf) or the previous strip
            rH = cBB; rF = 0u;
            if (MULTI) { rH = top_next << 16; rF = top_next & 0xFFFF0000u; }
        }
        const uint32_t Hin = prmt(rH, Hout, 0x5432u);  // lo <- upper lane's hi half, hi <- own lo half
        const uint32_t Fin = prmt(rF, Fout, 0x5432u);
        uint32_t F = 0u;  // stays dead through a quiet step

        // ---- window bytes: lo half column c, hi half column c-1
        const uint32_t rb = rb_next;
        {
            int cn = c + 1;
            rb_next = kOut;
            if ((unsigned)cn < (unsigned)w) { CUSW_CHECK_PTR(rptr + (int64_t)cn * STEP, ck_r0, ck_r1, "window column"); rb_next = (uint32_t)rptr[(int64_t)cn * STEP]; }
            if (MULTI) {
                top_next = cBB & 0xFFFFu;
                if (read_top && (unsigned)cn < (unsigned)w) { CUSW_CHECK_PTR(bnd_in + cn, ck_b0, ck_b1, "strip boundary read"); top_next = bnd_in[cn]; }
            }
        }
        Rsel = prmt(Rsel, rb, 0x3204u);  // byte1 <- old byte0, byte0 <- new
        const uint32_t Rs = Rsel + 0x4400u;  // PROF: the high half indexes pool bytes 4..7 (QB)

        if (PROF) {
            // a half that enters column 0 now starts from the matrix border: whatever it computed over the
            // out-of-range columns before (read as 'A') is discarded
            if (c == 0 || c == 1) {
                const uint32_t keep = c == 0 ? 0xFFFF0000u : 0x0000FFFFu;
#pragma unroll
                for (int k = 0; k < R; k++) {
                    Ho[k] = (Ho[k] & keep) | (cBB & ~keep);
                    E[k] &= keep;
                }
                Hdiag_top = (Hdiag_top & keep) | (cBB & ~keep);
                Emax &= keep;
            }
        }
        const uint32_t diag0 = Hdiag_top;
        Hdiag_top = Hin;
        uint32_t cm = cBB;

        // ---- quiet step (speculative): when every E of this warp and every incoming F is dead (<= 0) and no
        // cell of the step exceeds gap_open, no gap can be open or opened, so H = max(0, Hdiag + s) and E, F
        // keep being dead: 3.5 instead of 7.5 instructions per cell pair. Dead values never reach an H (H >= 0)
        // and every live E/F is re-derived from an H > gap_open, so skipping their updates is exact.
        // Emax is the exact maximum of this thread's E[] (the max distributes over the E recurrence).
        bool quiet = false;
#ifdef CUSW_STEP_STATS
        bool tried = false;
#endif
        {
            const uint32_t live = __vimax3_s16x2(Emax, Fin, cBB);
            // (skipq: a chained quiet step has just failed on this very step, see below: no second attempt)
            if (!(CHAIN && skipq) && __all_sync(0xffffffffu, live == cBB)) {
#ifdef CUSW_STEP_STATS
                tried = true;
#endif
                uint32_t diag = diag0;
#pragma unroll
                for (int k = 0; k < R; k++) {
                    const uint32_t s = PROF ? prmt(Q[k], QB[PROF ? k : 0], Rs) : prmt(cPA, cPB, Q[k] ^ Rsel);
                    const uint32_t h = __viaddmax_s16x2(diag, s, cBB);
                    diag = Ho[k];
                    Hn[k] = h;
                    if (k & 1) cm = __vimax3_s16x2(cm, Hn[k - 1], h);
                }
                if (R & 1) cm = __vmaxs2(cm, Hn[R - 1]);
                // verified after the fact: had any h exceeded gap_open, the step is redone in full from Ho
                quiet = __all_sync(0xffffffffu, __vmaxs2(cm, cGOBB) == cGOBB);
            }
        }
        if (!quiet) {
            uint32_t diag = diag0;
            cm = cBB;
            F = Fin;
#pragma unroll
            for (int k = 0; k < R; k++) {
                const uint32_t s = PROF ? prmt(Q[k], QB[PROF ? k : 0], Rs) : prmt(cPA, cPB, Q[k] ^ Rsel);
                const uint32_t m = __vimax3_s16x2(E[k], F, cBB);
                const uint32_t h = __viaddmax_s16x2(diag, s, m);
                diag = Ho[k];
                Hn[k] = h;
                const uint32_t hg = h - cGO2;
                E[k] = __viaddmax_s16x2(E[k], cNGE2, hg);
                F = __viaddmax_s16x2(F, cNGE2, hg);
                if (k & 1) cm = __vimax3_s16x2(cm, Hn[k - 1], h);
            }
            if (R & 1) cm = __vmaxs2(cm, Hn[R - 1]);
            Emax = __viaddmax_s16x2(Emax, cNGE2, cm - cGO2);
        }
#ifdef CUSW_STEP_STATS
        if ((threadIdx.x & 31) == 0) atomicAdd((unsigned long long*)(sc.pool_dev + 512) + (quiet ? 0 : tried ? 1 : 2), 1ull);
#endif
        Hout = Hn[R - 1];
        Fout = F;
        if (MULTI) {
            // bottom row of the strip (hi half = last virtual lane, column c-1) -> boundary buffer
            if (write_bot && (unsigned)(c - 1) < (unsigned)w) {
                CUSW_CHECK_PTR(bnd_out + c - 1, ck_b0, ck_b1, "strip boundary write");
                bnd_out[c - 1] = (Hout >> 16) | (Fout & 0xFFFF0000u);
            }
        }

        if (PROF) {
            // columns outside [0, w) must never raise the maximum
            const bool v_lo = (unsigned)c < (unsigned)w, v_hi = (unsigned)(c - 1) < (unsigned)w;
            if (!(v_lo && v_hi)) {
                const uint32_t keep = (v_lo ? 0x0000FFFFu : 0u) | (v_hi ? 0xFFFF0000u : 0u);
                cm = (cm & keep) | (cBB & ~keep);
            }
        }
        // ---- running best with strict improvement (ssw.c:283-306: first column that raises the maximum, then
        // the smallest row holding it). The column is recorded at once; the row search is deferred until the
        // half stops improving (along an alignment the maximum rises every step): the column that last raised
        // it is then still intact in Ho, the other ping-pong array.
        // (not __vibmax_s16x2: its header asm lets the output alias input `a`, which turns both
        //  predicates into constants when best is updated in place)
        const uint32_t nb = __vmaxs2(best, cm);
        const uint32_t changed = nb ^ best;
        const bool imp_lo = (changed & 0xFFFFu) != 0u, imp_hi = (changed >> 16) != 0u;
        const bool due_lo = (pend & 0xFFFFu) != 0u && !imp_lo, due_hi = (pend >> 16) != 0u && !imp_hi;
        if (due_lo || due_hi) {
            const uint32_t cnt = first_row_of(Ho, best, cBB);
            if (due_lo) row_lo = (int)(cnt & 0xFFFFu);
            if (due_hi) row_hi = (int)(cnt >> 16);
        }
        if (imp_lo) col_lo = c;
        if (imp_hi) col_hi = c - 1;
        pend = changed;
        best = nb;
        ++c;
        ++tt;
        if (CHAIN) { chain = quiet && tt >= 2 * G && tt + 2 < t_hi; skipq = false; }
        if (BUSY) was_full = !quiet;
    };

    // ---- chained quiet step. After a verified quiet step every F handed down is dead and every E of the warp
    // is dead, and a quiet step keeps both so: the next step needs neither the F shuffle nor the pre-vote, and on
    // interior columns neither the border reset nor the range masks. The step is speculative like the quiet path
    // of `step`: nothing is committed unless the whole warp stayed at or below gap_open; otherwise the caller
    // redoes it with `step` from the untouched Ho. E[] is not read here and is rewritten (any dead value is as
    // good as another, see above) by the caller when the chain ends, so its registers are free inside a chain.
    uint32_t rb2 = 0u;  // inside a chain: window byte of column c + 1 (rb_next holds column c)
    auto qstep = [&](uint32_t (&Ho)[R], uint32_t (&Hn)[R]) -> bool {
        uint32_t rH = __shfl_up_sync(0xffffffffu, Hout, 1, G);
        if (lig == 0) rH = cBB;
        const uint32_t Hin = prmt(rH, Hout, 0x5432u);
        const uint32_t Rnew = prmt(Rsel, rb_next, 0x3204u);
        const uint32_t Rs = Rnew + 0x4400u;
        uint32_t diag = Hdiag_top;
        uint32_t cm = cBB;
#pragma unroll
        for (int k = 0; k < R; k++) {
            const uint32_t s = prmt(Q[k], QB[PROF ? k : 0], Rs);
            const uint32_t h = __viaddmax_s16x2(diag, s, cBB);
            diag = Ho[k];
            Hn[k] = h;
            if (k & 1) cm = __vimax3_s16x2(cm, Hn[k - 1], h);
        }
        const uint32_t thr = (R & 1) ? __vimax3_s16x2(cm, Hn[R - 1], cGOBB) : __vmaxs2(cm, cGOBB);
        if (R & 1) cm = __vmaxs2(cm, Hn[R - 1]);
#ifdef CUSW_STEP_STATS
        {
            const bool okq = __all_sync(0xffffffffu, thr == cGOBB);
            if ((threadIdx.x & 31) == 0) atomicAdd((unsigned long long*)(sc.pool_dev + 512) + (okq ? 3 : 4), 1ull);
        }
#endif
        if (!__all_sync(0xffffffffu, thr == cGOBB)) return false;
        Hdiag_top = Hin;
        rb_next = rb2;  // loaded a whole step ago
        CUSW_CHECK_PTR(rptr + c + 2, ck_r0, ck_r1, "chained window column (c + 2)");  // unguarded by design: the chain gate keeps it inside
        rb2 = (uint32_t)rptr[c + 2];
        Rsel = Rnew;
        Hout = Hn[R - 1];
        const uint32_t nb = __vmaxs2(best, cm);
        const uint32_t changed = nb ^ best;
        if ((changed | pend) != 0u) {  // rare on quiet columns: an improvement now or a row search owed
            const bool imp_lo = (changed & 0xFFFFu) != 0u, imp_hi = (changed >> 16) != 0u;
            const bool due_lo = (pend & 0xFFFFu) != 0u && !imp_lo, due_hi = (pend >> 16) != 0u && !imp_hi;
            if (due_lo || due_hi) {
                const uint32_t cnt = first_row_of(Ho, best, cBB);
                if (due_lo) row_lo = (int)(cnt & 0xFFFFu);
                if (due_hi) row_hi = (int)(cnt >> 16);
            }
            if (imp_lo) col_lo = c;
            if (imp_hi) col_hi = c - 1;
            pend = changed;
            best = nb;
        }
        ++c;
        ++tt;
        return true;
    };
    // ---- busy step: the full recurrence of `step` on interior columns, for the stretches where step after step is full (the
    // alignment and the E/F tails it leaves behind: most steps of a filtered column range, every step of a squarish pair).
    // No quiet attempt (the caller looks every other step whether one has become possible), no border reset and no range
    // masks (interior), the window byte loaded unguarded, best / column tracking behind one test as in qstep.
    auto bstep = [&](uint32_t (&Ho)[R], uint32_t (&Hn)[R]) {
        uint32_t rH = __shfl_up_sync(0xffffffffu, Hout, 1, G);
        uint32_t rF = __shfl_up_sync(0xffffffffu, Fout, 1, G);
        if (lig == 0) { rH = cBB; rF = 0u; }
        const uint32_t Hin = prmt(rH, Hout, 0x5432u);
        uint32_t F = prmt(rF, Fout, 0x5432u);
        const uint32_t rb = rb_next;
        CUSW_CHECK_PTR(rptr + c + 1, ck_r0, ck_r1, "busy window column (c + 1)");  // unguarded by design: t_hb keeps it inside
        rb_next = (uint32_t)rptr[c + 1];
        Rsel = prmt(Rsel, rb, 0x3204u);
        const uint32_t Rs = Rsel + 0x4400u;
        uint32_t diag = Hdiag_top;
        Hdiag_top = Hin;
        uint32_t cm = cBB;
#pragma unroll
        for (int k = 0; k < R; k++) {
            const uint32_t s = prmt(Q[k], QB[PROF ? k : 0], Rs);
            const uint32_t m = __vimax3_s16x2(E[k], F, cBB);
            const uint32_t h = __viaddmax_s16x2(diag, s, m);
            diag = Ho[k];
            Hn[k] = h;
            const uint32_t hg = h - cGO2;
            E[k] = __viaddmax_s16x2(E[k], cNGE2, hg);
            F = __viaddmax_s16x2(F, cNGE2, hg);
            if (k & 1) cm = __vimax3_s16x2(cm, Hn[k - 1], h);
        }
        if (R & 1) cm = __vmaxs2(cm, Hn[R - 1]);
        Emax = __viaddmax_s16x2(Emax, cNGE2, cm - cGO2);
#ifdef CUSW_STEP_STATS
        if ((threadIdx.x & 31) == 0) atomicAdd((unsigned long long*)(sc.pool_dev + 512) + 2, 1ull);
#endif
        Hout = Hn[R - 1];
        Fout = F;
        const uint32_t nb = __vmaxs2(best, cm);
        const uint32_t changed = nb ^ best;
        if ((changed | pend) != 0u) {
            const bool imp_lo = (changed & 0xFFFFu) != 0u, imp_hi = (changed >> 16) != 0u;
            const bool due_lo = (pend & 0xFFFFu) != 0u && !imp_lo, due_hi = (pend >> 16) != 0u && !imp_hi;
            if (due_lo || due_hi) {
                const uint32_t cnt = first_row_of(Ho, best, cBB);
                if (due_lo) row_lo = (int)(cnt & 0xFFFFu);
                if (due_hi) row_hi = (int)(cnt >> 16);
            }
            if (imp_lo) col_lo = c;
            if (imp_hi) col_hi = c - 1;
            pend = changed;
            best = nb;
        }
        ++c;
        ++tt;
    };
    // full steps tend to come in long runs: stay on the lean twin while the columns are interior, and look every other step
    // whether every E of the warp and every F handed down has died (then `step` may try a quiet one)
    auto busy_run = [&]() {
        while (tt >= 2 * G && tt + 2 < t_hb) {
            bstep(HA, HB);
            bstep(HB, HA);
            if (__all_sync(0xffffffffu, __vimax3_s16x2(Emax, Fout, cBB) == cBB)) break;
        }
    };
    if (CHAIN) {
        while (tt < t_end_warp) {
            if (chain) {
                bool odd = false;
#ifdef CUSW_STEP_STATS
                if ((threadIdx.x & 31) == 0 && tt + 2 < t_hi) atomicAdd((unsigned long long*)(sc.pool_dev + 512) + 5, 1ull);
#endif
                if (tt + 2 < t_hi) { CUSW_CHECK_PTR(rptr + c + 1, ck_r0, ck_r1, "chained window column (c + 1)"); rb2 = (uint32_t)rptr[c + 1]; }
                while (tt + 2 < t_hi) {  // both steps of the pair are interior
#ifdef CUSW_NO_SKIPQ
                    if (!qstep(HA, HB)) break;
                    if (!qstep(HB, HA)) { odd = true; break; }
#else
                    if (!qstep(HA, HB)) { skipq = true; break; }
                    if (!qstep(HB, HA)) { skipq = true; odd = true; break; }
#endif
                }
#ifdef CUSW_CHAIN_REALIGN
                if (odd) {  // the state after an odd number of steps lives in HB: realign the ping-pong
#pragma unroll
                    for (int k = 0; k < R; k++) HA[k] = HB[k];
                }
#endif
                chain = false;
#pragma unroll
                for (int k = 0; k < R; k++) E[k] = 0u;
                Emax = 0u;
#ifndef CUSW_CHAIN_REALIGN
                if (odd) { step(HB, HA); continue; }  // a third copy of `step` instead of realigning the ping-pong (measured faster)
#endif
            }
            // (a realigned chain makes the step count odd: the extra step at the end only sees pad columns)
            step(HA, HB);
            step(HB, HA);
            if (BUSY && was_full) {
                busy_run();
                chain = false;
                skipq = false;
            }
        }
    } else if (BUSY) {
        while (tt < t_end_warp) {
            step(HA, HB);
            step(HB, HA);
            if (was_full) busy_run();
        }
    } else
    for (int t = 0; t < t_end_warp; t += 2) {
        step(HA, HB);
        step(HB, HA);
        if (REV) {
            // early stop (ssw.c:296): once a lane holds `target`, later lanes may still find it in an
            // earlier column, so run on until the last virtual lane has passed this column.
            if ((t & 2) == 2) {
                const int tg = target + sc.B;
                bool found = ((int)(best & 0xFFFFu) == tg) || ((int)(best >> 16) == tg);
                if (found) t_end = min(t_end, t + 1 + 2 * G + 1);
                // all lanes of a group must agree on t_end
#pragma unroll
                for (int o = G / 2; o > 0; o >>= 1)
                    t_end = min(t_end, __shfl_xor_sync(0xffffffffu, t_end, o));
                t_end_warp = (__reduce_max_sync(0xffffffffu, t_end) + 1) & ~1;
            }
        }
    }

    if (pend != 0u) {  // the last step wrote HA
        const uint32_t cnt = first_row_of(HA, best, cBB);
        if ((pend & 0xFFFFu) != 0u) row_lo = (int)(cnt & 0xFFFFu);
        if ((pend >> 16) != 0u) row_hi = (int)(cnt >> 16);
    }

    // ---- fold this strip's virtual lanes into the running key: max score, then min column, then min row
    {
        const int rbase = row0 + (2 * lig) * R;
        unsigned long long klo = ((unsigned long long)(best & 0xFFFFu) << 40) |
                                 ((unsigned long long)(0xFFFFFFu - (uint32_t)(col_lo + 1)) << 16) |
                                 (unsigned long long)(0xFFFFu - (uint32_t)(rbase + row_lo));
        unsigned long long khi = ((unsigned long long)(best >> 16) << 40) |
                                 ((unsigned long long)(0xFFFFFFu - (uint32_t)(col_hi + 1)) << 16) |
                                 (unsigned long long)(0xFFFFu - (uint32_t)(rbase + R + row_hi));
        const unsigned long long k2 = klo > khi ? klo : khi;
        key = k2 > key ? k2 : key;
    }
    if (MULTI) {
        // later strips read what this one wrote (same warp, different lanes)
        __syncwarp();
        uint32_t* tmp = bnd_in; bnd_in = bnd_out; bnd_out = tmp;
        if (REV) {
            // columns past the first one that reached `target` cannot matter any more
            unsigned long long kk = key;
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) {
                unsigned long long other = __shfl_xor_sync(0xffffffffu, kk, o);
                kk = other > kk ? other : kk;
            }
            if ((int)(kk >> 40) == target + sc.B) {
                int colf = (int)(0xFFFFFFu - (uint32_t)((kk >> 16) & 0xFFFFFFu)) - 1;
                w = min(w, colf + 1);
            }
        }
    }
    }  // strips

    // ---- group reduction
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    if (active && lig == 0) {
        int score = (int)(key >> 40) - sc.B;
        int col = (int)(0xFFFFFFu - (uint32_t)((key >> 16) & 0xFFFFFFu)) - 1;
        int row = (int)(0xFFFFu - (uint32_t)(key & 0xFFFFu));
        if (!REV && io.sub) col += io.sub[pair].c0;
        out[pair] = make_int4(score, col, row, failed ? 1 : 0);
    }
}

#ifdef CUSW_MAIN_TU
// ------------------------------------------------------------------------------------------
// Phase 3: banded traceback (banded_sw, ssw.c:557-734). One thread per pair.
// The band rule, the zero borders (including the `edge` quirk, ssw.c:596), every tie-break and the
// `while (i > 0)` walk are reproduced; state is laid out differently: one in-place rolling row of
// (H, E) pairs and ONE direction byte per cell (bits 0-2: H source 1..5, bit 3: E opened,
// bit 4: F opened) addressed with the reference's flat cell index, so even an out-of-band step of
// the walk lands on the same neighbouring cell. Only the last band iteration's directions exist
// here; a walk that leaves the cells written by it is reported as CUDASW_EUNDEF.
// ------------------------------------------------------------------------------------------
struct TbConst {
    int32_t match, mismatch, ambiguous, go, ge;
    uint32_t poolA, poolB;  // the PRMT byte pool of ScoreConst: byte x = substitution score of symbol-xor x
};

struct TbPool {
    uint8_t* base;
    unsigned long long cap;
    unsigned long long* used;
};

__device__ __forceinline__ int tb_sub(const TbConst& tc, uint32_t qb, uint32_t rb) {
    // x == 0: equal symbols (match); 1..3: both ACGT (mismatch); 4..7: exactly one N (ambiguous). One PRMT picks byte x of the
    // pool and replicates its sign into the upper three bytes (selector nibbles x, x|8, x|8, x|8).
    const uint32_t x = (qb ^ rb) & 7u;
    return (int)prmt(tc.poolA, tc.poolB, x * 0x1111u + 0x8880u);
}

// Walks the direction matrix. When `outp` != nullptr writes ops (reversed order resolved by index).
// Returns number of M/I/D ops, or a negative status.
// center < 0: the reference's flat layout (row i starts at its first in-band column). center >= 0: band-frame layout
// of sw_traceback_kernel, cell (i, j) at dir[width_d * i + (j - i + center)].
// Ops come out last-to-first. With `outp` they are written to their final place (needs the total, i.e. a second walk);
// with `buf` the first kTbOpsBuf of them are kept in walk order, so that the usual short CIGAR needs one walk only.
constexpr int kTbOpsBuf = 12;
__device__ int tb_walk(const uint8_t* dir, int width_d, int band, int readLen,
                       int refLen, uint32_t* outp, int total_ops, int center = -1, uint32_t* buf = nullptr) {
#ifdef CUSW_BOUNDS_CHECK
    const uint8_t* ck_d1 = dir + (long long)width_d * readLen;
#endif
    int i = readLen - 1, j = refLen - 1;
    int e = 0, l = 0, state = 2;
    int op = 0, prev_op = 0;  // 0 = M, 1 = I, 2 = D
    auto put = [&](int idx, uint32_t v) {
        if (outp) { CUSW_CHECK_PTR(outp + total_ops - idx, outp, outp + total_ops, "CIGAR word"); outp[total_ops - idx] = v; }
        if (buf && idx <= kTbOpsBuf) buf[idx - 1] = v;
    };
    while (i > 0) {
        const int xi = max(i - band, 0);
        const int jj = j - xi;
        const int lo = max(i - band, 0) - xi, hi = min(i + band, refLen - 1) - xi;
        if (jj < lo || jj > hi) return CUDASW_EUNDEF;
        CUSW_CHECK_PTR(dir + (int64_t)width_d * i + (center >= 0 ? j - i + center : jj), dir, ck_d1, "direction byte (walk)");
        const uint32_t b = dir[(int64_t)width_d * i + (center >= 0 ? j - i + center : jj)];
        int dv = state == 2 ? (int)(b & 7u) : state == 0 ? (2 + (int)((b >> 3) & 1u)) : (4 + (int)((b >> 4) & 1u));
        switch (dv) {
            case 1: --i; --j; state = 2; op = 0; break;
            case 2: --i; state = 0; op = 1; break;
            case 3: --i; state = 2; op = 1; break;
            case 4: --j; state = 1; op = 2; break;
            case 5: --j; state = 2; op = 2; break;
            default: return CUDASW_EUNDEF;
        }
        if (op == prev_op) ++e;
        else {
            ++l;
            put(l, ((uint32_t)e << 4) | (uint32_t)prev_op);
            prev_op = op;
            e = 1;
        }
    }
    if (op == 0) {
        ++l;
        put(l, ((uint32_t)(e + 1) << 4) | 0u);
    } else {
        l += 2;
        put(l - 1, ((uint32_t)e << 4) | (uint32_t)op);
        put(l, (1u << 4) | 0u);
    }
    return l;
}

// Sort keys for the traceback: estimated banded cells readLen * (2*band + 1), descending, so that the threads of a
// warp get pairs of similar cost and the long ones start first. A pair whose (starting) band exceeds kTbThreadBand
// goes to the warp-cooperative kernel; `tb_band` remembers the band a pair reached in an earlier round.
constexpr uint32_t kTbKeyMax = (1u << 22) - 1u;
constexpr int kTbThreadBand = 16;  // widest band sw_traceback_kernel keeps in registers (33 diagonals)
constexpr int kTbClasses = 5;      // register frames of half-width 1, 2, 4, 8, 16
constexpr int kTbWarpSmallD = 8;   // sw_traceback_warp_kernel<8>: bands of up to 64*8 diagonals; <32>: everything else
__host__ __device__ __forceinline__ int tb_class_of(int band) { return band <= 1 ? 0 : band <= 2 ? 1 : band <= 4 ? 2 : band <= 8 ? 3 : 4; }
__host__ __device__ __forceinline__ int tb_frame_bytes(int cls) { return ((2 << cls) + 1 + 3) & ~3; }  // direction bytes per row
// Keys: warp-kernel pairs first (by descending cost), then the one-thread pairs grouped by frame class and, inside a
// class, by descending read length: the 32 pairs of a warp then run the same instantiation for about as many rows.
// cls_count[c] = pairs of class c; est_bytes = scratch the first band pass of every pair will ask for.
__global__ void tb_keys_kernel(const int* __restrict__ list, int n, const int4* __restrict__ fwd,
                               const int4* __restrict__ rev, const cudasw_result* __restrict__ res,
                               long long wide_cells, int all_big, int thread_band, const int* __restrict__ tb_band, uint32_t* __restrict__ keys,
                               int* __restrict__ vals, int* __restrict__ wide_count, int* __restrict__ cls_count,
                               unsigned long long* __restrict__ est_bytes) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= n) return;
    const int pair = list ? list[tid] : tid;
    uint32_t key = ((uint32_t)kTbClasses << 19) | 0x7FFFu | (1u << 22);  // nothing to trace: sorts last (class slot 5)
    int cls = kTbClasses;
    if (res[pair].status == CUDASW_OK && fwd[pair].x > 0) {
        const int4 rv = rev[pair];
        const int refLen = rv.y + 1, readLen = rv.z + 1;
        const int band = max(abs(refLen - readLen) + 1, tb_band[pair]);
        long long cost = (long long)readLen * (2 * band + 1);
        // wide_cells < 0: the default rule; >= 0: additionally every pair above that many cells (0 = all of them)
        const bool wide = band > thread_band || (wide_cells >= 0 && cost > wide_cells);
        if (wide) {
            // warp-kernel pairs: those that start beyond 16 diagonals per lane first (their kernel holds more state
            // per thread), each group by descending cost; wide_count = {all, big}
            const bool big = all_big || 2 * band + 1 > 64 * kTbWarpSmallD;
            atomicAdd(wide_count, 1);
            if (big) atomicAdd(wide_count + 1, 1);
            if (cost > (long long)(kTbKeyMax >> 1) - 1) cost = (kTbKeyMax >> 1) - 1;
            key = (big ? 0u : (1u << 21)) | ((kTbKeyMax >> 1) - (uint32_t)cost);
            cls = -1;
            atomicAdd(est_bytes, (unsigned long long)readLen * (unsigned long long)((2 * band + 1 + 3) & ~3) + 16ull);
        } else {
            cls = tb_class_of(band);
            key = (1u << 22) | ((uint32_t)cls << 19) | (0x7FFFu - (uint32_t)min(readLen, 0x7FFF));
            atomicAdd(est_bytes, (unsigned long long)readLen * (unsigned long long)tb_frame_bytes(cls) + 16ull);
        }
    }
    if (cls >= 0) atomicAdd(&cls_count[cls], 1);
    keys[tid] = key;
    vals[tid] = pair;
}

// (start, count) of the traceback lists inside the sorted order, from the counters tb_keys_kernel left behind, so that
// the kernels can be launched without the host knowing the counts. Slots: 0 = big pairs for the CTA-per-pair kernel,
// 1 = big pairs for the one-warp kernel (one of the two is empty: a CTA per pair pays while they fit one wave),
// 2 = other wide pairs (small warp kernel), 3 + k = one-thread class k (k = kTbClasses: nothing to trace).
constexpr int kTbSlots = 3 + kTbClasses + 1;
__global__ void tb_slots_kernel(const int* __restrict__ wide_count, const int* __restrict__ cls_count, int cta_wave, int* __restrict__ slots) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int nwide = wide_count[0], nbig = wide_count[1];
    const bool cta = nbig <= cta_wave;
    slots[0] = 0; slots[1] = cta ? nbig : 0;
    slots[2] = 0; slots[3] = cta ? 0 : nbig;
    slots[4] = nbig; slots[5] = nwide - nbig;
    int start = nwide;
    for (int k = 0; k <= kTbClasses; k++) {
        slots[6 + 2 * k] = start;
        slots[7 + 2 * k] = cls_count[k];
        start += cls_count[k];
    }
    slots[2 * kTbSlots] = 0; slots[2 * kTbSlots + 1] = 0;  // an empty list (launches that only drain an escalation list)
}

// Reserves CIGAR words for one pair and writes soft clips + body (SswJniAligner.java:52-59).
__device__ __forceinline__ void tb_emit(int pair, int q, int score, int ref_begin, int ref_end, int read_begin,
                                        int read_end, int nbody, const uint8_t* dirp, int width_d, int band,
                                        int readLen, int refLen, cudasw_result r, cudasw_result* __restrict__ res,
                                        uint32_t* __restrict__ cig, unsigned long long cig_cap,
                                        unsigned long long* __restrict__ cig_used, int center = -1,
                                        const uint32_t* buf = nullptr) {
    const int lead = read_begin != 0;
    const int tail = (q - read_end - 1) != 0;
    const int nops = nbody + lead + tail;
    const unsigned long long coff = atomicAdd(cig_used, (unsigned long long)nops);
    r.score = score;
    r.ref_begin = ref_begin;
    r.ref_end = ref_end;
    r.read_begin = read_begin;
    r.read_end = read_end;
    r.cigar_len = nops;
    r.cigar_off = (int32_t)coff;
    if (coff + (unsigned long long)nops <= cig_cap && coff < 0x7fffffffull) {
        uint32_t* o = cig + coff;
        if (lead) *o++ = ((uint32_t)read_begin << 4) | CUDASW_CIGAR_S;
        if (score <= 0) *o = (1u << 4) | CUDASW_CIGAR_M;
        else if (buf && nbody <= kTbOpsBuf) {
            for (int k = 0; k < nbody; k++) o[k] = buf[nbody - 1 - k];  // kept by the counting walk, last op first
        } else tb_walk(dirp, width_d, band, readLen, refLen, o, nbody, center);
        if (tail) o[nbody] = ((uint32_t)(q - read_end - 1) << 4) | CUDASW_CIGAR_S;
    } else {
        r.status = CUDASW_ENOMEM;
    }
    res[pair] = r;
}

// Results without a CIGAR (callers that only read the clip lengths, BreakpointHomology.java:82-84,96-98): the ends of
// the forward pass and the begins of the reverse pass, exactly as tb_emit reports them; cigar_len = 0.
__global__ void ends_only_kernel(int n, const int4* __restrict__ fwd, const int4* __restrict__ rev,
                                 cudasw_result* __restrict__ res) {
    const int pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= n) return;
    cudasw_result r = res[pair];
    if (r.status != CUDASW_OK) return;
    const int4 f = fwd[pair];
    r.score = f.x;
    r.cigar_off = 0; r.cigar_len = 0;
    if (f.x <= 0) {
        r.ref_begin = -1; r.ref_end = -1; r.read_begin = 0; r.read_end = 0;
    } else {
        const int4 rv = rev[pair];
        if (rv.w != 0) r.status = CUDASW_ENOMEM;
        else if (rv.x != f.x) r.status = CUDASW_ECUDA;
        r.ref_end = f.y; r.read_end = f.z;
        r.ref_begin = f.y - rv.y; r.read_begin = f.z - rv.z;
    }
    res[pair] = r;
}

}  // namespace cusw
#include "tb_warp.cuh"
namespace cusw {

// Banded DP of one pair by one thread with the whole band in registers. The frame is WMAX = 2*BMAX + 1 diagonals
// centred on the main diagonal (x = j - i + BMAX); a narrower band simply leaves the outer slots idle. Same cell
// arithmetic and border rules as tb_dp_warp (tb_warp.cuh header): out-of-band neighbours read as 0, row 0 uses
// -gapOpen / -gapExtend, plus the reference's `edge` quirk. The window bytes of the frame slide along in registers
// (one new byte per row) and the direction bytes of a row leave as packed 32-bit words.
template <int BMAX>
__device__ int tb_dp_thread(const TbConst& tc, const uint8_t* __restrict__ ref, const uint8_t* __restrict__ read,
                            int refLen, int readLen, int band, uint8_t* __restrict__ dir) {
    constexpr int WMAX = 2 * BMAX + 1;
    constexpr int WP = (WMAX + 3) & ~3;
    int H[WMAX], E[WMAX];
    uint32_t rw[WMAX];
#pragma unroll
    for (int x = 0; x < WMAX; x++) {
        H[x] = 0; E[x] = 0;
        const int j = x - BMAX;
        rw[x] = (unsigned)j < (unsigned)refLen ? (uint32_t)ref[j] : 0u;
    }
    int mx = 0;
    // the two bytes a row needs (its query base, the window byte entering the frame after it) are loaded one row ahead
    uint32_t qb_next = read[0];
    uint32_t rw_next = BMAX + 1 < refLen ? (uint32_t)ref[BMAX + 1] : 0u;
    for (int i = 0; i < readLen; i++) {
        const int beg = max(i - band, 0);
        const int end = min(i + band, refLen - 1);
        const int endp = min(i - 1 + band, refLen - 1);  // last column of the previous row
        const bool quirk = i >= 1 && i <= band + 1 && end == refLen - 1 && endp == end && end <= 2 * band + 1;
        const int quirk_col = quirk ? end : -1;  // the reference's `edge` quirk: that cell's up neighbour reads as (0, 0)
        const uint32_t qb = qb_next;
        const uint32_t rw_in = rw_next;
        if (i + 1 < readLen) qb_next = read[i + 1];
        rw_next = i + 2 + BMAX < refLen ? (uint32_t)ref[i + 2 + BMAX] : 0u;
        const int j0 = i - BMAX;
        int hleft = 0, fgap = 0;
        int cur_h = H[0];
        uint32_t* drow = reinterpret_cast<uint32_t*>(dir + (long long)WP * i);
        uint32_t dw = 0;
#pragma unroll
        for (int x = 0; x < WMAX; x++) {
            const int j = j0 + x;
            // row i-1 at the next diagonal = the cell above; a slot outside the band was never written in this pass: (0, 0)
            const int nxt_h = x + 1 < WMAX ? H[x + 1 < WMAX ? x + 1 : 0] : 0;
            const int nxt_e = x + 1 < WMAX ? E[x + 1 < WMAX ? x + 1 : 0] : 0;
            if (j >= beg && j <= end) {
                const bool qk = j == quirk_col;
                int h, e;
                uint8_t dv;
                tb_cell(tc, qk ? 0 : nxt_h, qk ? 0 : nxt_e, cur_h, tb_sub(tc, qb, rw[x]), hleft, fgap, mx, h, e, dv);
                H[x] = h;
                E[x] = e;
                dw |= (uint32_t)dv << (8 * (x & 3));
            }
            cur_h = nxt_h;
            if ((x & 3) == 3 || x == WMAX - 1) {
                CUSW_CHECK_PTR(drow + (x >> 2), dir, dir + (long long)WP * readLen, "direction word (one-thread DP)");
                drow[x >> 2] = dw; dw = 0;
            }
        }
        // slide the window bytes: the frame of row i+1 starts one column further right
#pragma unroll
        for (int x = 0; x + 1 < WMAX; x++) rw[x] = rw[x + 1];
        rw[WMAX - 1] = rw_in;
    }
    return mx;
}

// Phase 3 for pairs whose band fits a register frame: one thread per pair, one instantiation (and launch) per frame
// class so that a warp never mixes frame widths. Runs the band-doubling loop of banded_sw (ssw.c:630-631) while the
// band fits THIS frame. A pair that needs a wider band is appended to `esc_out`, the input of the next wider class
// (launched behind this one on the same stream; the widest class hands over to the warp-cooperative kernel), so the
// whole escalation chain needs no host round trip. A pair that finds the scratch pool exhausted goes to the retry
// list of the host loop. tb_band remembers the band reached.
template <int BMAX>
__global__ void __launch_bounds__(128)
sw_traceback_kernel(const int* __restrict__ list, int n_own, const int* __restrict__ esc_in,
                    const int* __restrict__ esc_in_count, int* __restrict__ esc_out, int* __restrict__ esc_out_count,
                    PairIO io, const int4* __restrict__ fwd, const int4* __restrict__ rev, TbConst tc, TbPool pool,
                    cudasw_result* __restrict__ res, uint32_t* __restrict__ cig,
                    unsigned long long cig_cap, unsigned long long* __restrict__ cig_used,
                    int* __restrict__ retry_list, int* __restrict__ retry_count, int* __restrict__ tb_band,
                    const int* __restrict__ slot = nullptr) {
    if (slot) { list += slot[0]; n_own = slot[1]; }  // own list bounds from device memory (tb_slots_kernel)
    const int n_all = n_own + (esc_in_count ? *esc_in_count : 0);
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n_all; idx += gridDim.x * blockDim.x) {
        const int pair = idx < n_own ? list[idx] : esc_in[idx - n_own];
        cudasw_result r = res[pair];
        if (r.status != CUDASW_OK) continue;
        const int q = (int)(io.seq_off[pair + 1] - io.seq_off[pair]);
        const int4 f = fwd[pair];
        const int score = f.x;

        int nbody = 0;
        uint32_t opsbuf[kTbOpsBuf];
        int ref_begin = -1, read_begin = 0, ref_end = f.y, read_end = f.z;
        const uint8_t* dirp = nullptr;
        constexpr int width_d = (2 * BMAX + 1 + 3) & ~3;
        int band = 0, readLen = 0, refLen = 0;

        if (score <= 0) {
            // no positive cell -> start -1, "1M" + (q-1)S
            ref_end = -1; read_end = 0;
            nbody = 1;
        } else {
            const int4 rv = rev[pair];
            if (rv.w != 0) { res[pair].status = CUDASW_ENOMEM; continue; }     // strip buffers exhausted
            if (rv.x != score) { res[pair].status = CUDASW_ECUDA; continue; }  // internal inconsistency
            ref_begin = ref_end - rv.y;
            read_begin = read_end - rv.z;
            refLen = ref_end - ref_begin + 1;
            readLen = read_end - read_begin + 1;
            const uint8_t* ref = io.enc_ref + io.ref_off[pair] + ref_begin;
            const uint8_t* read = io.enc_seq + io.seq_off[pair] + read_begin;
            band = max(abs(refLen - readLen) + 1, tb_band[pair]);
            int mx = 0;
            bool done = false;
            for (;;) {
                if (band > BMAX) {  // a wider frame class, or the warp-cooperative kernel, takes over at this band
                    tb_band[pair] = band;
                    esc_out[atomicAdd(esc_out_count, 1)] = pair;
                    break;
                }
                // the reference gives up (exit(1)) when its own direction matrix would not fit an int (ssw.c:580-585)
                if ((long long)(2 * band + 1) * readLen * 3 >= (1ll << 31)) { res[pair].status = CUDASW_ETOOLARGE; break; }
                const unsigned long long need = (((unsigned long long)width_d * (unsigned long long)readLen) + 15ull) & ~15ull;
                if (need > pool.cap) { res[pair].status = CUDASW_ETOOLARGE; break; }
                const unsigned long long off = dirp ? (unsigned long long)(dirp - pool.base) : atomicAdd(pool.used, need);
                if (off + need > pool.cap) {
                    tb_band[pair] = band;
                    retry_list[atomicAdd(retry_count, 1)] = pair;
                    break;
                }
                uint8_t* dir = pool.base + off;  // a wider band inside the same frame overwrites the failed pass
                dirp = dir;
                mx = max(mx, tb_dp_thread<BMAX>(tc, ref, read, refLen, readLen, band, dir));  // the reference never resets its maximum
                if (mx >= score) { done = true; break; }
                band *= 2;
            }
            if (!done) continue;
            nbody = tb_walk(dirp, width_d, band, readLen, refLen, nullptr, 0, BMAX, opsbuf);
            if (nbody < 0) { res[pair].status = nbody; continue; }
        }
        tb_emit(pair, q, score, ref_begin, ref_end, read_begin, read_end, nbody, dirp, width_d, band, readLen, refLen, r,
                res, cig, cig_cap, cig_used, BMAX, opsbuf);
    }
}

#endif  // CUSW_MAIN_TU

}  // namespace cusw

#ifdef CUSW_MAIN_TU
namespace cusw {

// One warp per pair of `list` (the pairs sw_traceback_kernel handed over): banded DP by tb_dp_warp2 (every lane busy in
// every half-step, tb_warp.cuh), then the warp walks the direction matrix together (tb_walk_warp) and lane 0 emits the
// CIGAR exactly like the one-thread kernel.
// DMAX = 8: chunks of up to 8 diagonals (bands up to 255) with a small register footprint; a pair that outgrows it is
// appended to esc_out, the input of the DMAX = 32 instantiation launched behind it (bands up to 1023 in registers, and
// the memory-backed path for even wider ones).
template <int DMAX>
__global__ void __launch_bounds__(128)
sw_traceback_warp_kernel(const int* __restrict__ list, int n, const int* __restrict__ n_dev, int* __restrict__ esc_out,
                         int* __restrict__ esc_out_count, PairIO io, const int4* __restrict__ fwd,
                         const int4* __restrict__ rev, TbConst tc, TbPool pool, cudasw_result* __restrict__ res,
                         uint32_t* __restrict__ cig, unsigned long long cig_cap,
                         unsigned long long* __restrict__ cig_used, int* __restrict__ retry_list,
                         int* __restrict__ retry_count, int* __restrict__ tb_band, const int* __restrict__ slot = nullptr) {
    const int lane = threadIdx.x & 31;
    if (n_dev) n = *n_dev;  // the list was filled on the device (pairs escalated by sw_traceback_kernel<16>)
    if (slot) { list += slot[0]; n = slot[1]; }  // list bounds from device memory (tb_slots_kernel)
    for (int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; wid < n; wid += (gridDim.x * blockDim.x) >> 5) {
    const int pair = list[wid];
    cudasw_result r = res[pair];
    if (r.status != CUDASW_OK) continue;
    const int q = (int)(io.seq_off[pair + 1] - io.seq_off[pair]);
    const int4 f = fwd[pair];
    const int4 rv = rev[pair];
    const int score = f.x;
    const int ref_end = f.y, read_end = f.z;
    const int ref_begin = ref_end - rv.y, read_begin = read_end - rv.z;
    const int refLen = ref_end - ref_begin + 1, readLen = read_end - read_begin + 1;
    const uint8_t* ref = io.enc_ref + io.ref_off[pair] + ref_begin;
    const uint8_t* read = io.enc_seq + io.seq_off[pair] + read_begin;
    int band = max(abs(refLen - readLen) + 1, tb_band[pair]);  // tb_band: where an earlier round / kernel stopped
    int mx = 0, width_d = 0;
    const uint8_t* dirp = nullptr;
    bool ok = true;
    for (;;) {
        const int W = 2 * band + 1;
        width_d = tb_frame_width(band);
        const long long ndir = (long long)width_d * readLen;
        int status = CUDASW_OK;
        const int dper = (W + 63) / 64;  // diagonals per chunk
        if (DMAX < 32 && dper > DMAX) {  // the wide-state instantiation takes over at this band
            if (lane == 0) { tb_band[pair] = band; esc_out[atomicAdd(esc_out_count, 1)] = pair; }
            ok = false;
            break;
        }
        const unsigned long long row_bytes = dper > 32 ? (((unsigned long long)W * 8ull + 15ull) & ~15ull) : 0ull;
        const unsigned long long need = row_bytes + (((unsigned long long)ndir + 15ull) & ~15ull);
        unsigned long long off = 0;
        // the reference gives up (exit(1)) when its own direction matrix would not fit an int (ssw.c:580-585)
        if ((long long)W * readLen * 3 >= (1ll << 31) || need > pool.cap) status = CUDASW_ETOOLARGE;
        else {
            if (lane == 0) off = atomicAdd(pool.used, need);
            off = __shfl_sync(0xffffffffu, off, 0);
            if (off + need > pool.cap) status = 1;  // scratch exhausted: try again in the next round
        }
        if (status != CUDASW_OK) {
            if (lane == 0) {
                if (status == 1) { tb_band[pair] = band; retry_list[atomicAdd(retry_count, 1)] = pair; }
                else res[pair].status = status;
            }
            ok = false;
            break;
        }
        int2* row = reinterpret_cast<int2*>(pool.base + off);
        uint8_t* dir = pool.base + off + row_bytes;
        dirp = dir;
        int m;
        if (dper <= 1) m = tb_dp_warp2<1>(tc, ref, read, refLen, readLen, band, dir);
        else if (dper <= 2) m = tb_dp_warp2<2>(tc, ref, read, refLen, readLen, band, dir);
        else if (dper <= 4) m = tb_dp_warp2<4>(tc, ref, read, refLen, readLen, band, dir);
        else if (dper <= 8) m = tb_dp_warp2<8>(tc, ref, read, refLen, readLen, band, dir);
        else if (DMAX >= 32 && dper <= 16) m = tb_dp_warp2<DMAX >= 32 ? 16 : 1>(tc, ref, read, refLen, readLen, band, dir);
        else if (DMAX >= 32 && dper <= 32) m = tb_dp_warp2<DMAX >= 32 ? 32 : 1>(tc, ref, read, refLen, readLen, band, dir);
        else m = tb_dp_warp_mem(tc, ref, read, refLen, readLen, band, dir, row, (W + 31) / 32);
        mx = max(mx, m);  // the reference never resets its running maximum between band widths
        if (mx >= score) break;
        band *= 2;
    }
    __syncwarp();  // every lane's direction bytes are visible to the whole warp
    if (ok) {
        __shared__ uint32_t ops_sh[4][kTbOpsBuf];
        uint32_t* opsbuf = ops_sh[threadIdx.x >> 5];
        const int nbody = tb_walk_warp(dirp, width_d, band, readLen, refLen, nullptr, 0, opsbuf);
        if (nbody < 0) { if (lane == 0) res[pair].status = nbody; }
        else {
            // tb_emit, with the second walk (CIGARs longer than the ops buffer) done by the whole warp
            const int lead = read_begin != 0, tail = (q - read_end - 1) != 0;
            const int nops = nbody + lead + tail;
            unsigned long long coff = 0;
            if (lane == 0) coff = atomicAdd(cig_used, (unsigned long long)nops);
            coff = __shfl_sync(0xffffffffu, coff, 0);
            const bool fits = coff + (unsigned long long)nops <= cig_cap && coff < 0x7fffffffull;
            uint32_t* o = cig + coff + lead;
            if (fits && nbody > kTbOpsBuf) tb_walk_warp(dirp, width_d, band, readLen, refLen, o, nbody, nullptr);
            if (lane == 0) {
                r.score = score; r.ref_begin = ref_begin; r.ref_end = ref_end; r.read_begin = read_begin; r.read_end = read_end;
                r.cigar_len = nops; r.cigar_off = (int32_t)coff;
                if (fits) {
                    if (lead) cig[coff] = ((uint32_t)read_begin << 4) | CUDASW_CIGAR_S;
                    if (nbody <= kTbOpsBuf)
                        for (int k = 0; k < nbody; k++) o[k] = opsbuf[nbody - 1 - k];  // kept by the counting walk, last op first
                    if (tail) o[nbody] = ((uint32_t)(q - read_end - 1) << 4) | CUDASW_CIGAR_S;
                } else r.status = CUDASW_ENOMEM;
                res[pair] = r;
            }
        }
    }
    __syncwarp();
    }  // pairs of this warp
}

// One CTA (4 warps) per pair: the latency-oriented flavour of the kernel above for batches with too few wide pairs to
// fill the machine with one warp each (BASELINE.json C5: 1000 contigs of 10 kb, where a lone warp needs tens of
// milliseconds per pair). 256 chunks instead of 64, so a chunk is a quarter as wide and an iteration a quarter as long;
// the warps hand over through shared memory (tb_dp_warp2<D, 4>). Bands beyond 256*4 diagonals go to esc_out (the
// one-warp kernel); keeping D <= 4 holds the kernel at 8 CTAs per SM.
__global__ void __launch_bounds__(128, 8)
sw_traceback_block_kernel(const int* __restrict__ list, int n, int* __restrict__ esc_out, int* __restrict__ esc_out_count,
                          PairIO io, const int4* __restrict__ fwd, const int4* __restrict__ rev, TbConst tc, TbPool pool,
                          cudasw_result* __restrict__ res, uint32_t* __restrict__ cig, unsigned long long cig_cap,
                          unsigned long long* __restrict__ cig_used, int* __restrict__ retry_list,
                          int* __restrict__ retry_count, int* __restrict__ tb_band, const int* __restrict__ slot = nullptr) {
    if (slot) { list += slot[0]; n = slot[1]; }
    __shared__ TbXchg xc;
    __shared__ unsigned long long off_sh;
    __shared__ uint32_t ops_sh[kTbOpsBuf];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    for (int wid = blockIdx.x; wid < n; wid += gridDim.x) {
    const int pair = list[wid];
    cudasw_result r = res[pair];
    if (r.status != CUDASW_OK) continue;  // (uniform: every thread reads the same record)
    const int q = (int)(io.seq_off[pair + 1] - io.seq_off[pair]);
    const int4 f = fwd[pair];
    const int4 rv = rev[pair];
    const int score = f.x;
    const int ref_end = f.y, read_end = f.z;
    const int ref_begin = ref_end - rv.y, read_begin = read_end - rv.z;
    const int refLen = ref_end - ref_begin + 1, readLen = read_end - read_begin + 1;
    const uint8_t* ref = io.enc_ref + io.ref_off[pair] + ref_begin;
    const uint8_t* read = io.enc_seq + io.seq_off[pair] + read_begin;
    int band = max(abs(refLen - readLen) + 1, tb_band[pair]);
    __syncthreads();  // tb_band[pair] is rewritten below by thread 0
    int mx = 0, width_d = 0;
    const uint8_t* dirp = nullptr;
    bool ok = true;
    for (;;) {
        const int W = 2 * band + 1;
        width_d = tb_frame_width(band);
        const long long ndir = (long long)width_d * readLen;
        int status = CUDASW_OK;
        const int dper = (W + 255) / 256;  // diagonals per chunk
        if (dper > 4) {  // wider than this kernel keeps in registers (bands above 511)
            if (threadIdx.x == 0) { tb_band[pair] = band; esc_out[atomicAdd(esc_out_count, 1)] = pair; }
            ok = false;
            break;
        }
        const unsigned long long need = ((unsigned long long)ndir + 15ull) & ~15ull;
        if ((long long)W * readLen * 3 >= (1ll << 31) || need > pool.cap) status = CUDASW_ETOOLARGE;  // ssw.c:580-585
        else {
            __syncthreads();
            if (threadIdx.x == 0) off_sh = atomicAdd(pool.used, need);
            __syncthreads();
            if (off_sh + need > pool.cap) status = 1;  // scratch exhausted: try again in the next round
        }
        if (status != CUDASW_OK) {
            if (threadIdx.x == 0) {
                if (status == 1) { tb_band[pair] = band; retry_list[atomicAdd(retry_count, 1)] = pair; }
                else res[pair].status = status;
            }
            ok = false;
            break;
        }
        uint8_t* dir = pool.base + off_sh;
        dirp = dir;
        int m;
        if (dper <= 1) m = tb_dp_warp2<1, 4>(tc, ref, read, refLen, readLen, band, dir, &xc);
        else if (dper <= 2) m = tb_dp_warp2<2, 4>(tc, ref, read, refLen, readLen, band, dir, &xc);
        else m = tb_dp_warp2<4, 4>(tc, ref, read, refLen, readLen, band, dir, &xc);
        mx = max(mx, m);  // the reference never resets its running maximum between band widths
        if (mx >= score) break;
        band *= 2;
    }
    __syncthreads();  // every warp's direction bytes are visible to warp 0
    if (ok && wib == 0) {
        const int nbody = tb_walk_warp(dirp, width_d, band, readLen, refLen, nullptr, 0, ops_sh);
        if (nbody < 0) { if (lane == 0) res[pair].status = nbody; }
        else {
            const int lead = read_begin != 0, tail = (q - read_end - 1) != 0;
            const int nops = nbody + lead + tail;
            unsigned long long coff = 0;
            if (lane == 0) coff = atomicAdd(cig_used, (unsigned long long)nops);
            coff = __shfl_sync(0xffffffffu, coff, 0);
            const bool fits = coff + (unsigned long long)nops <= cig_cap && coff < 0x7fffffffull;
            uint32_t* o = cig + coff + lead;
            if (fits && nbody > kTbOpsBuf) tb_walk_warp(dirp, width_d, band, readLen, refLen, o, nbody, nullptr);
            __syncwarp();
            if (lane == 0) {
                r.score = score; r.ref_begin = ref_begin; r.ref_end = ref_end; r.read_begin = read_begin; r.read_end = read_end;
                r.cigar_len = nops; r.cigar_off = (int32_t)coff;
                if (fits) {
                    if (lead) cig[coff] = ((uint32_t)read_begin << 4) | CUDASW_CIGAR_S;
                    if (nbody <= kTbOpsBuf)
                        for (int k = 0; k < nbody; k++) o[k] = ops_sh[nbody - 1 - k];
                    if (tail) o[nbody] = ((uint32_t)(q - read_end - 1) << 4) | CUDASW_CIGAR_S;
                } else r.status = CUDASW_ENOMEM;
                res[pair] = r;
            }
        }
    }
    __syncthreads();
    }  // pairs of this CTA
}

}  // namespace cusw
#endif  // CUSW_MAIN_TU

