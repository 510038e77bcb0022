// band_filter.cuh - a LOSSLESS k-mer filter in front of the forward score pass: most pairs only need the columns that a
// diagonal band around their alignment can reach, and the filter PROVES after the fact that nothing outside mattered.
// Pairs without such a proof run the forward pass over their whole window as before; results are identical either way.
// (No counterpart in the reference - ssw.c sweeps the full matrix - the results are what must match, and do: see below.)
//
// Why it is exact. Let a = match > 0, p = min(|mismatch|, |ambiguous|), go / ge the gap costs (a gap of L bases costs
// go + (L-1) ge >= go), n = min(q, w). Take ANY local alignment path P of score S(P) >= S0 with M match steps, X other
// diagonal steps and g gaps with Lg gap bases in total:
//   (1) S(P) <= a M - p X - go g   and   M + X <= n;
//   (2) its match steps form at most X + g + 1 maximal runs, and a run of r matches contains r - k + 1 positions (i, j)
//       with query[i .. i+k) == window[j .. j+k) ("hits"; the window holds no N here, so a match is an equal A/C/G/T
//       symbol), hence P passes through at least M - (k-1)(X + g + 1) hits. Minimised over (1) - a fractional knapsack
//       with one budget, solved in closed form by flt_hits_floor() - that is T(S0), non-decreasing in S0;
//   (3) the diagonals j - i it visits span at most Lg <= L(S0) = max_gap_bases(n, a, S0, go, ge) (the bound the banded
//       reverse pass rests on, rev_band.cuh), i.e. they touch at most m(S0) = ceil(L / W) + 1 consecutive bins of W = 16 diagonals.
// The filter counts ALL hits of the pair per diagonal bin (never fewer: query k-mers are chained, not hashed away),
// picks the densest bin pair [b, b+1], calls it and c = kFltCore bins either side of it the CORE [b-c, b+1+c], adds a margin of
// m_r - 1 bins either side, and runs the forward pass on the columns [c0, c1] that the diagonals of those bins can reach, which gives a true score S0 (a DP over a
// sub-rectangle never exceeds the full DP). If m(S0) <= m_r and every run of m_r bins that lies outside
// the core holds fewer than T(S0) hits ("noise"), then a path of score >= S0 that leaves the band would have to
// collect >= T(S0) hits inside a window of m(S0) bins that is not inside the band - there is none. So every cell whose
// true H is >= S0 is reached by paths inside the computed rectangle only: its value there is exact, the maximum is the
// true maximum and the set of cells holding it - all the reference's tie-break (first column, then smallest row,
// ssw.c:283-306) looks at - is the same. Otherwise the pair is handed to the full-window pass.
//
// A tighter margin from the hits themselves ("evidence"). A run of r matches holds at most max(0, r - (k-1)) + (k-1)
// bases, so M <= hits(P) + (k-1)(X + g + 1); with (1): S0 <= a hits(P) + a(k-1) + X (a(k-1) - p) + g (a(k-1) - go)
// - ge (Lg - g), and X <= Xmax(S0) = (a n - S0) / (p + a) (from M + X <= n and a M >= S0 + p X). When go > a(k-1) and
// ge > 0 that bounds the gap bases of P by its hits: Lg <= R / min(go - a(k-1), ge), R = a hits(P) + a(k-1) - S0 +
// max(0, a(k-1) - p) Xmax. A path whose bins all lie outside the core is excluded by the noise test above; any
// other path spanning m bins lies inside bins [b-c-(m-1), b+1+c+(m-1)], whose hit total hc(m-1) bounds hits(P). Starting
// from m = m(L(S0)) and iterating m <- min(m, m(Lg bound from hc(m-1))) only ever applies sound implications, so its
// fixed point m_t bounds the bins of EVERY path of score >= S0, and the band only needs a margin of m_t bins. The same
// bound on Lg (gap_bound) narrows the band of the reverse pass (rev_band.cuh), whose paths are alignment paths of
// score S0 ending in the forward pass's cell.
#pragma once
#include "sw_kernels.cuh"

namespace cusw {

constexpr int kFltK = 6;                      // k-mer length: 4^6 direct-mapped heads
constexpr int kFltHeads = 1 << (2 * kFltK);
#ifndef CUSW_FLT_BIN_SHIFT  // A/B switch
#define CUSW_FLT_BIN_SHIFT 4
#endif
constexpr int kFltBinShift = CUSW_FLT_BIN_SHIFT;  // 16 diagonals per bin (32 at first: C2's column ranges were 12 % longer)
constexpr int kFltBinW = 1 << kFltBinShift;
// The core: the densest pair of adjacent bins [b, b + 1] and kFltCore bins either side of it, 96 diagonals in all (an
// alignment with a 50-base indel still has its hits inside). Everything outside it is "noise" to the certificate.
constexpr int kFltCore = kFltBinShift == 5 ? 1 : 2;
constexpr int kFltMaxBins = 32768 >> kFltBinShift;  // q + w - 1 <= 32 k diagonals
constexpr int kFltMaxQ = 2040;                // chain links are 16 bit; a bin then holds at most 16 * 2040 hits
constexpr int kFltWarps = 4;                  // pairs per CTA of filter_kernel
constexpr size_t kFltSmemPerWarp = kFltHeads * 2 + 2048 * 2 + kFltMaxBins * 2;  // heads, links, bins: 16 bit each (14 KB at most)
// The launch sizes the links and bins for the batch's longest query / most diagonals (the host knows both), so that short
// pairs leave room for more warps per SM: C2 (300 x 4000) needs 9.5 KB per warp instead of 14.
struct FilterSmem { int link_cap, bins_cap; };  // entries; both multiples of 64
__host__ __device__ __forceinline__ size_t flt_smem_per_warp(FilterSmem m) { return (size_t)kFltHeads * 2 + (size_t)m.link_cap * 2 + (size_t)m.bins_cap * 2; }
inline FilterSmem make_filter_smem(long long max_q, long long max_w) {
    long long lc = std::min<long long>(std::max<long long>(max_q, 64), 2048), bc = std::min<long long>((max_q + max_w) / kFltBinW + 4, kFltMaxBins);
    FilterSmem m;
    m.link_cap = (int)((lc + 63) & ~63ll);
    m.bins_cap = (int)std::min<long long>((bc + 63) & ~63ll, kFltMaxBins);
    return m;
}

struct FilterConst {
    int ok;              // 0: this scoring gives no usable bound (match <= 0, free mismatches or free gaps)
    int a, go, ge;       // match, gap open, gap extend
    long long cn, cd;    // rho = cn / cd = min(0, (p - (k-1) a) / (p + a), (go - (k-1) a) / go), cd > 0
    int est_pct;         // the margin is sized for a score of est_pct % of (a * hits of the densest bin pair)
    int ev_div;          // min(go - a(k-1), ge) when both are positive, else 0: no evidence bound for this scoring
    int ev_alpha, ev_pa; // max(0, a(k-1) - p); p + a
    int band;            // 1: certifiable pairs take the banded forward kernel (a parallelogram of diagonals); 0: the column
                         // range [c0, c1] through the wavefront kernels (a rectangle)
};
constexpr int kFltBandMaxWd = 1024;  // widest band the banded forward kernel takes (its rings are sized for it)
constexpr int kFltBandMaxBins = (kFltBandMaxWd / kFltBinW - 2 - 2 * kFltCore) / 2 + 1;  // the margin that gives that width
// Cells the banded forward kernel computes for a pair: its strips of S rows each sweep S + Wd - 1 columns.
constexpr int kBandVariants = 5;  // strips of 32, 64, 128, 256, 1024 rows (rev_band.cuh)
__host__ __device__ __forceinline__ int band_variant_rows(int v) { return v == 0 ? 32 : v == 1 ? 64 : v == 2 ? 128 : v == 3 ? 256 : 1024; }
__host__ __device__ __forceinline__ long long band_cells(int q, int w, long long wd, int v) {
    const long long S = band_variant_rows(v);
    const long long cols = S + wd - 1 < w ? S + wd - 1 : w;
    return ((q + S - 1) / S) * S * cols;
}
__host__ __device__ __forceinline__ int band_best_variant(int q, int w, long long wd, int few_pairs) {
    if (few_pairs && q > 256) return 4;  // a whole warp per pair when there are few of them (latency)
    int best = 0;
    long long bc = band_cells(q, w, wd, 0);
    for (int v = 1; v < 4; v++) {
        // a taller strip wastes triangle cells but crosses the ring less often: take it when it costs at most 10 % more
        const long long c = band_cells(q, w, wd, v);
        if (c * 10 <= bc * 11) { best = v; if (c < bc) bc = c; }
    }
    return best;
}
constexpr int kFltHc = 16;  // hc(j), j < kFltHc, is kept per pair for the certificate

inline FilterConst make_filter_const(int match, int mismatch, int ambiguous, int go, int ge) {
    FilterConst f{};
    f.a = match; f.go = go; f.ge = ge; f.cn = 0; f.cd = 1; f.est_pct = 80;
    const long long a = match, p = std::min(-mismatch, -ambiguous), k1 = kFltK - 1;
    f.ok = match > 0 && p > 0 && go > 0 && max_gap_bases(1000, match, 1, go, ge) >= 0;
    if (!f.ok) return f;
    auto take = [&](long long n, long long d) {  // d > 0
        if (n * f.cd < f.cn * d) { f.cn = n; f.cd = d; }
    };
    take(p - k1 * a, p + a);
    take(go - k1 * a, go);
    f.ev_div = (go - k1 * a > 0 && ge > 0) ? (int)std::min<long long>(go - k1 * a, ge) : 0;
    f.ev_alpha = (int)std::max<long long>(0, k1 * a - p);
    f.ev_pa = (int)(p + a);
    return f;
}

// T(S0): floor of the least number of hits on a path of score >= S0 over at most n diagonal steps (see (2) above)
__host__ __device__ __forceinline__ long long flt_hits_floor(const FilterConst& f, long long S0, long long n) {
    long long bud = (long long)f.a * n - S0;
    if (bud < 0) bud = 0;
    const long long num = (S0 - (long long)(kFltK - 1) * f.a) * f.cd + bud * f.cn;
    const long long den = (long long)f.a * f.cd;
    long long t = num / den;
    if (num % den != 0 && num < 0) --t;
    return t;
}
// m(L): bins that a window of L + 1 consecutive diagonals can touch
__host__ __device__ __forceinline__ int flt_bins_of(long long L) { return (int)((L + kFltBinW - 1) >> kFltBinShift) + 1; }

// The evidence iteration (header): hc(j) = hits in the core bins extended by j bins either side. Starts from m0 bins (from the
// worst-case gap bound L0), returns the fixed point m_t and leaves the gap-base bound that goes with it in *gap_bound.
template <typename HC>
__host__ __device__ __forceinline__ int flt_tight_bins(const FilterConst& f, long long S, long long n, int m0, long long L0, HC&& hc, long long* gap_bound) {
    *gap_bound = L0;
    if (f.ev_div == 0 || m0 - 1 >= kFltHc) return m0;
    const long long xmax = ((long long)f.a * n - S) / f.ev_pa;
    int m = m0;
    for (int it = 0; it < 16; it++) {
        long long R = (long long)f.a * (long long)hc(m - 1) + (long long)f.a * (kFltK - 1) - S + (long long)f.ev_alpha * (xmax > 0 ? xmax : 0);
        long long Lt = R > 0 ? R / f.ev_div : 0;
        if (Lt < *gap_bound) *gap_bound = Lt;
        const int m2 = flt_bins_of(*gap_bound);
        if (m2 >= m) break;
        m = m2;
    }
    return m;
}

#ifdef CUSW_MAIN_TU
// One warp per pair. Shared memory per warp: heads[4^k] (first query row + 1 of every k-mer), link[row] (next row with
// the same k-mer), bins[diagonal / kFltBinW]; all 16 bit.
__global__ void __launch_bounds__(32 * kFltWarps)
filter_kernel(int n, PairIO io, const int* __restrict__ any_n, const uint8_t* __restrict__ wflags, FilterConst fc,
              FilterRec* __restrict__ rec, uint32_t* __restrict__ hc_out, unsigned long long* __restrict__ counters, FilterSmem sm) {
    extern __shared__ __align__(16) unsigned char flt_smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    unsigned short* heads = reinterpret_cast<unsigned short*>(flt_smem + (size_t)wib * flt_smem_per_warp(sm));
    unsigned short* link = heads + kFltHeads;
    uint32_t* bins2 = reinterpret_cast<uint32_t*>(link + sm.link_cap);  // two 16-bit counters per word (a bin holds <= kFltBinW * kFltMaxQ hits)
    uint32_t* psum = reinterpret_cast<uint32_t*>(heads);         // prefix sums of the bins: overlays the heads once the scan is over
    for (int p = blockIdx.x * kFltWarps + wib; p < n; p += gridDim.x * kFltWarps) {
        const int64_t qo = io.seq_off[p], ro = io.ref_off[p];
        const int q = (int)(io.seq_off[p + 1] - qo), w = (int)(io.ref_off[p + 1] - ro);
        FilterRec r{0, w, 0, 0, -1, 0, 0};
        const bool eligible = fc.ok && q >= kFltK && w >= kFltK && q <= kFltMaxQ && (long long)q + w <= (long long)kFltBinW * kFltMaxBins - 64 &&
                              q <= sm.link_cap && ((q + w - 2) >> kFltBinShift) + 3 <= sm.bins_cap && !(*any_n != 0 && wflags[p]);
        if (!eligible) {  // warp-uniform
            if (lane == 0) rec[p] = r;
            continue;
        }
        const uint8_t* qs = io.enc_seq + qo;
        const uint8_t* ws = io.enc_ref + ro;
        const int nb = ((q + w - 2) >> kFltBinShift) + 1;
        {   // clear
            uint4* h4 = reinterpret_cast<uint4*>(heads);
            for (int i = lane; i < kFltHeads / 8; i += 32) h4[i] = make_uint4(0u, 0u, 0u, 0u);
            for (int i = lane; i < (nb + 3) / 2; i += 32) bins2[i] = 0u;
        }
        __syncwarp();
        // Both sequences are read 512 bases per step, one aligned 16-byte vector per lane (prefetched two steps ahead); a lane
        // owns the k-mers that END in its vector and gets the five bases before it from the lane below (lane 0: from the
        // previous step). body(vc, vn, pos0, checked): k-mer u of the lane (u = 0 .. 15) starts at base pos0 + u, its value is
        // bits [2u, 2u + 12) of vc (sum of code[pos + t] << 2t), bits [2u, 2u + 12) of vn are non-zero when it holds an N
        // (check_n only); !checked: every pos0 + u of the step is inside [0, len - k] for every lane.
        auto scan = [&](const uint8_t* sq, int len, bool check_n, auto&& body) {
            const int mis = (int)(reinterpret_cast<uintptr_t>(sq) & 15u);
            const uint4* al = reinterpret_cast<const uint4*>(sq - mis);
            const int total = len + mis;  // bytes of the aligned stream that are ours (the first `mis` are not)
            const int ntiles = (total + 511) >> 9;
            auto load = [&](int vi) {
                if (16 * vi >= total) return make_uint4(0u, 0u, 0u, 0u);
                // (a vector may start up to 15 bytes before the sequence and end up to 15 bytes after it: inside the blob, or its 16 bytes of slack)
                CUSW_CHECK_PTR(reinterpret_cast<const uint8_t*>(al + vi) + 15, sq - 15, sq + len + 16, "filter: sequence vector");
                return al[vi];
            };
            auto c8 = [](uint32_t x) { return ((x & 0x03030303u) * 0x01041040u) >> 24; };  // four bytes -> four 2-bit codes
            uint4 nxt = load(lane), nxt2 = load(32 + lane);  // two steps in flight: the kernel lives on HBM latency
            uint32_t carry_c = 0u, carry_n = 0u;
            for (int t = 0; t < ntiles; t++) {
                const uint4 cur = nxt;
                nxt = nxt2;
                nxt2 = load(32 * (t + 2) + lane);
                const uint32_t c32 = c8(cur.x) | (c8(cur.y) << 8) | (c8(cur.z) << 16) | (c8(cur.w) << 24);
                uint32_t n32 = 0u;
                if (check_n) n32 = c8(cur.x >> 2) | (c8(cur.y >> 2) << 8) | (c8(cur.z >> 2) << 16) | (c8(cur.w >> 2) << 24);
                uint32_t pc = __shfl_up_sync(0xffffffffu, c32, 1), pn = check_n ? __shfl_up_sync(0xffffffffu, n32, 1) : 0u;
                if (lane == 0) { pc = carry_c; pn = carry_n; }
                carry_c = __shfl_sync(0xffffffffu, c32, 31);
                if (check_n) carry_n = __shfl_sync(0xffffffffu, n32, 31);
                const unsigned long long vc = ((unsigned long long)c32 << 10) | (pc >> 22);
                const unsigned long long vn = ((unsigned long long)n32 << 10) | (pn >> 22);
                body(vc, vn, 512 * t + 16 * lane - mis - (kFltK - 1), t == 0 || t + 2 > ntiles);
            }
        };
        // query k-mers -> chains (rows holding an N have none). Plain stores, verified: of the lanes that push onto the same
        // head in the same step one wins, the others see it and try again on top of it.
        scan(qs, q, true, [&](unsigned long long vc, unsigned long long vn, int pos0, bool) {
            for (int u = 0; u < 16; u++) {
                const int i = pos0 + u;
                const uint32_t km = (uint32_t)(vc >> (2 * u)) & (kFltHeads - 1);
                bool pending = i >= 0 && i <= q - kFltK && ((uint32_t)(vn >> (2 * u)) & (kFltHeads - 1)) == 0u;
                while (__any_sync(0xffffffffu, pending)) {
                    unsigned short old = 0;
                    if (pending) { old = heads[km]; }
                    __syncwarp();
                    if (pending) heads[km] = (unsigned short)(i + 1);
                    __syncwarp();
                    if (pending && heads[km] == (unsigned short)(i + 1)) {
                        CUSW_CHECK_PTR(link + i, link, link + sm.link_cap, "filter: chain link");
                        link[i] = old;
                        pending = false;
                    }
                    __syncwarp();
                }
            }
        });
        __syncwarp();
        // window k-mers -> hits per diagonal bin (the window holds A/C/G/T only): sixteen independent look-ups per lane and
        // step, then the (rare) chains of the ones that found a head
        scan(ws, w, false, [&](unsigned long long vc, unsigned long long, int pos0, bool checked) {
            uint32_t found = 0u;
            if (!checked) {
#pragma unroll
                for (int u = 0; u < 16; u++)
                    if (heads[(uint32_t)(vc >> (2 * u)) & (kFltHeads - 1)] != 0) found |= 1u << u;
            } else {
#pragma unroll
                for (int u = 0; u < 16; u++)
                    if (pos0 + u >= 0 && pos0 + u <= w - kFltK && heads[(uint32_t)(vc >> (2 * u)) & (kFltHeads - 1)] != 0) found |= 1u << u;
            }
            while (found != 0u) {
                const int u = __ffs((int)found) - 1;
                found &= found - 1u;
                const int j = pos0 + u;
                uint32_t i1p = heads[(uint32_t)(vc >> (2 * u)) & (kFltHeads - 1)];
                while (i1p != 0u) {
                    const int i = (int)i1p - 1;
                    const int b = (j - i + q - 1) >> kFltBinShift;
                    CUSW_CHECK_PTR(bins2 + (b >> 1), bins2, bins2 + sm.bins_cap / 2, "filter: diagonal bin");
                    CUSW_CHECK_PTR(link + i, link, link + sm.link_cap, "filter: chain link (walk)");
                    atomicAdd(&bins2[b >> 1], 1u << (16 * (b & 1)));
                    i1p = link[i];
                }
            }
        });
        __syncwarp();
        auto bin_at = [&](int b) { return (bins2[b >> 1] >> (16 * (b & 1))) & 0xFFFFu; };
        // densest pair of adjacent bins (ties: the first)
        uint32_t bh = 0u;
        int bb = 0;
        for (int b = lane; b < nb; b += 32) {
            const uint32_t s = bin_at(b) + bin_at(b + 1);  // bin nb is zero
            if (s > bh) { bh = s; bb = b; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint32_t oh = __shfl_xor_sync(0xffffffffu, bh, o);
            const int ob = __shfl_xor_sync(0xffffffffu, bb, o);
            if (oh > bh || (oh == bh && ob < bb)) { bh = oh; bb = ob; }
        }
        const long long nmin = min(q, w);
        const long long s_est = min((long long)fc.a * bh * fc.est_pct / 100, 32767ll);
        const long long L = s_est > 0 ? max_gap_bases(nmin, fc.a, (int)s_est, fc.go, fc.ge) : -1;
        bool go_on = L >= 0, evidence_only = false, use_band = false;
        long long band_lo = 0, band_hi = 0;
        int m_old = 0, m_band = 0;
        long long c0 = 0, c1 = w - 1;
        uint32_t noise = 0u;
        if (go_on) {  // warp-uniform
            m_old = flt_bins_of(L);
            // inclusive prefix sums of the bins
            uint32_t carry = 0u;
            for (int base = 0; base < nb; base += 32) {
                const int b = base + lane;
                uint32_t v = b < nb ? bin_at(b) : 0u;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
                    if (lane >= o) v += u;
                }
                v += carry;
                if (b < nb) psum[b] = v;
                carry = __shfl_sync(0xffffffffu, v, 31);
            }
            __syncwarp();
            // runs [bs, be) of m_old bins that lie entirely left or entirely right of the core [bb - kFltCore, bb + 1 + kFltCore]
            for (int bs = lane; bs < nb; bs += 32) {
                const int be = min(bs + m_old, nb);
                if (be - 1 <= bb - kFltCore - 1 || bs >= bb + kFltCore + 2) {
                    const uint32_t s = psum[be - 1] - (bs > 0 ? psum[bs - 1] : 0u);
                    noise = max(noise, s);
                }
            }
            noise = __reduce_max_sync(0xffffffffu, noise);
            // hc(lane): hits of the core bins extended by `lane` bins either side
            uint32_t my_hc;
            {
                const int lo = max(0, bb - kFltCore - lane), hi = min(nb - 1, bb + 1 + kFltCore + lane);
                my_hc = psum[hi] - (lo > 0 ? psum[lo - 1] : 0u);
            }
            if (lane < kFltHc) hc_out[(size_t)p * kFltHc + lane] = my_hc;
            long long gb;
            m_band = flt_tight_bins(fc, s_est, nmin, m_old, L, [&](int j) { return __shfl_sync(0xffffffffu, my_hc, j); }, &gb);
            // a path through the core spans at most m_band bins: bins [bb - kFltCore - (m_band - 1), bb + 1 + kFltCore + (m_band - 1)]
            const long long Dlo = (long long)kFltBinW * (bb - kFltCore - m_band + 1), Dhi = (long long)kFltBinW * (bb + 1 + kFltCore + m_band) - 1;  // diagonal index j - i + q - 1
            c0 = max(0ll, Dlo - (q - 1));
            c1 = min((long long)w - 1, Dhi);
            // nothing to win when the range is most of the window; not worth trying when even the estimated score could not
            // be told from the noise
            const bool told_apart = m_old <= 255 && flt_hits_floor(fc, (long long)fc.a * bh, nmin) > (long long)noise;
            go_on = told_apart && (c1 - c0 + 1) * 10 <= 7ll * w;
            if (fc.band && told_apart) {
                // the parallelogram pays when its strips compute clearly fewer cells than the whole window holds
                const long long Dl = max(Dlo, 0ll), Dh = min(Dhi, (long long)q + w - 2), wd = Dh - Dl + 1;
                band_lo = Dl - (q - 1);
                band_hi = Dh - (q - 1);
                const int v = band_best_variant(q, w, wd, 0);
                use_band = m_band <= kFltBandMaxBins && band_cells(q, w, wd, v) * 10 <= 6ll * q * w;
                go_on = use_band;
            }
            // A pair that sweeps its whole window anyway (long query in a short window) can still have the gap bases of its
            // alignments bounded by the hits, which is what the reverse pass's band is made of: keep the certificate data.
            evidence_only = told_apart && !go_on;
        }
        if (lane == 0) {
            if (go_on && use_band) {  // (c0, wlen stay the whole window: what an unproven pair falls back to)
                r.margin_bins = kFltBand | m_band | (m_old << 8); r.noise = (int32_t)noise; r.dlo = (int32_t)band_lo; r.dhi = (int32_t)band_hi;
            } else if (go_on) { r.c0 = (int32_t)c0; r.wlen = (int32_t)(c1 - c0 + 1); r.margin_bins = m_band | (m_old << 8); r.noise = (int32_t)noise; }
            else if (evidence_only) { r.margin_bins = kFltWholeWindow | m_band | (m_old << 8); r.noise = (int32_t)noise; }
            rec[p] = r;
            if (counters) atomicAdd(counters + (go_on ? 0 : 1), 1ull);
        }
        __syncwarp();
    }
}

// Bins the kFltBand pairs of a chunk for the banded forward kernel: slot = strip-height variant, kBandVariants = not banded;
// inside a slot the costly pairs come first. Also counts the cells those kernels will compute (cell_counts[1]).
__global__ void band_keys_kernel(int n, const int64_t* __restrict__ seq_off, const int64_t* __restrict__ ref_off,
                                 const FilterRec* __restrict__ rec, int few_pairs, uint32_t* __restrict__ keys, int* __restrict__ vals,
                                 int* __restrict__ hist, unsigned long long* __restrict__ cell_counts) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long cells = 0;
    if (p < n) {
        const FilterRec r = rec[p];
        int slot = kBandVariants;
        uint32_t k2 = 0xFFFFFu;
        if (r.margin_bins & kFltBand) {
            const int q = (int)(seq_off[p + 1] - seq_off[p]), w = (int)(ref_off[p + 1] - ref_off[p]);
            const long long wd = (long long)r.dhi - r.dlo + 1;
            slot = band_best_variant(q, w, wd, few_pairs);
            cells = (unsigned long long)band_cells(q, w, wd, slot);
            const unsigned long long c = cells >> 10;
            k2 = 0xFFFFFu - (uint32_t)(c > 0xFFFFFull ? 0xFFFFFull : c);
        }
        keys[p] = ((uint32_t)slot << 20) | k2;
        vals[p] = p;
        atomicAdd(&hist[slot], 1);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cell_counts + 1, cells);
}

// After the sub-window forward pass: is the score it found a proof (see the header)? done[p] = 1 keeps the result,
// 0 sends the pair to the full-window pass (its record is then reset to the whole window).
__global__ void filter_certify_kernel(int n, const int64_t* __restrict__ seq_off, const int64_t* __restrict__ ref_off,
                                      const int4* __restrict__ fwd, FilterConst fc, FilterRec* __restrict__ rec,
                                      const uint32_t* __restrict__ hc_in, uint8_t* __restrict__ done,
                                      unsigned long long* __restrict__ counters) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    FilterRec r = rec[p];
    if (r.margin_bins == 0) { done[p] = 1; return; }  // the first pass already swept the whole window, nothing to prove
    const bool whole = (r.margin_bins & kFltWholeWindow) != 0;  // swept the whole window; only the gap bound is sought
    const int m_band = whole ? 1 << 20 : r.margin_bins & 0xFF, m_old = (r.margin_bins >> 8) & 0xFF;
    // (a kFltBand pair was swept by the banded forward kernel over bins [b - m_band, b + 1 + m_band]: the same test)
    const long long q = seq_off[p + 1] - seq_off[p], w = ref_off[p + 1] - ref_off[p];
    const long long nmin = q < w ? q : w;
    const int4 f = fwd[p];
    bool ok = f.w == 0 && f.x > 0;
    long long gb = -1;
    if (ok) {
        const long long L = max_gap_bases(nmin, fc.a, f.x, fc.go, fc.ge);
        // paths outside the core: too few hits in any window they could span
        ok = L >= 0 && flt_bins_of(L) <= m_old && flt_hits_floor(fc, f.x, nmin) > (long long)r.noise;
        if (ok) {  // paths through the core: inside the band
            const uint32_t* hc = hc_in + (size_t)p * kFltHc;
            ok = flt_tight_bins(fc, f.x, nmin, flt_bins_of(L), L, [&](int j) { return hc[j]; }, &gb) <= m_band;
        }
    }
    done[p] = (ok || whole) ? 1 : 0;
    if (ok) {
        r.gap_bound = (int32_t)gb;
        rec[p] = r;
    } else if (whole) {
        // (the forward result is the full sweep's; the record stays as it is, without a gap bound)
    } else {
        r.c0 = 0; r.wlen = (int32_t)w; r.margin_bins = 0; r.noise = 0; r.gap_bound = -1; r.dlo = r.dhi = 0;
        rec[p] = r;
        if (counters) atomicAdd(counters + 2, 1ull);
    }
}
#endif  // CUSW_MAIN_TU

}  // namespace cusw
