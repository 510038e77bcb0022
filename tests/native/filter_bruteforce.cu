// Test-only host program logic: the k-mer filter's THEOREM on small instances, by brute force. For every generated pair it
// (1) fills the full affine Smith-Waterman matrix (the forward pass's semantics: score, first column holding it, smallest row
// in that column), (2) runs the filter's decision exactly as filter_kernel does (hits per diagonal bin, densest bin pair,
// margins, noise, hc[]), (3) fills the matrix again restricted to the column range, (4) evaluates the certificate with the
// kernels' own functions (band_filter.cuh) and, when it holds, demands the restricted result to equal the full one.
// Built and driven by tests/test_filter_math.py; not part of libcudasw.so.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../gridss_b200/csrc/band_filter.cuh"

using namespace cusw;

namespace {

struct Rng {
    uint64_t s;
    uint64_t next() { s += 0x9e3779b97f4a7c15ULL; uint64_t z = s; z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL; return z ^ (z >> 31); }
    int below(int n) { return (int)((next() >> 33) % (uint64_t)n); }
    double unif() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
};

struct Res { int score, col, row; };

// codes 0..3 = ACGT, 4 = N (query only)
inline int sub(int a, int b, int match, int mismatch, int ambiguous) { return a == b ? match : (a < 4 && b < 4) ? mismatch : ambiguous; }

// Full-matrix semantics of the forward pass over window columns [c0, c0 + wl): H = max(0, diag + s, E, F), a gap of L bases
// costs go + (L - 1) ge; result = maximum, first column holding it, smallest row in that column (ssw.c:283-306).
Res sw(const std::vector<int>& q, const std::vector<int>& w, int c0, int wl, int match, int mismatch, int ambiguous, int go, int ge) {
    const int n = (int)q.size();
    const int NEG = -(1 << 28);
    std::vector<int> H(n + 1, 0), E(n + 1, NEG);  // previous column: H[i + 1] = H(i, j - 1)
    Res best{0, -1, 0};
    for (int jj = 0; jj < wl; jj++) {
        const int j = c0 + jj;
        int diag = 0, F = NEG, up = 0;  // row -1
        int colmax = 0, colrow = 0;
        for (int i = 0; i < n; i++) {
            const int e = std::max(E[i + 1] - ge, H[i + 1] - go);  // gap in the query direction along the row (from column j - 1)
            F = std::max(F - ge, up - go);                          // from the row above in this column
            int h = std::max(0, diag + sub(q[i], w[j], match, mismatch, ambiguous));
            h = std::max(h, std::max(e, F));
            diag = H[i + 1];
            H[i + 1] = h;
            E[i + 1] = e;
            up = h;
            if (h > colmax) { colmax = h; colrow = i; }
        }
        if (colmax > best.score) best = Res{colmax, j, colrow};
    }
    return best;
}

std::vector<int> rnd(Rng& r, int n, int alphabet) {
    std::vector<int> v(n);
    for (auto& x : v) x = r.below(alphabet);
    return v;
}

void mutate_into(Rng& r, const std::vector<int>& src, int s0, std::vector<int>& dst, int at, int len, double p_sub, double p_indel) {
    int i = s0;
    for (int o = 0; o < len && i < (int)src.size() && at + o < (int)dst.size();) {
        const double u = r.unif();
        if (u < p_sub) { dst[at + o++] = (src[i++] + 1 + r.below(3)) & 3; }
        else if (u < p_sub + p_indel / 2) { dst[at + o++] = r.below(4); }   // insertion
        else if (u < p_sub + p_indel) { i += 1 + r.below(4); }              // deletion
        else dst[at + o++] = src[i++];
    }
}

}  // namespace

// out[0] pairs the filter gave a range, out[1] certified, out[2] certified AND different from the full sweep (must be 0),
// out[3] pairs where the restricted score exceeded the full one (must be 0), out[4] evidence bound violated by a certified
// pair's optimal... (not computable here; 0)
extern "C" int fb_run(uint64_t seed, int n_pairs, int match, int mismatch, int ambiguous, int go, int ge, int qmax, int wmax, long long* out) {
    Rng r{seed};
    FilterConst fc = make_filter_const(match, mismatch, ambiguous, go, ge);
    for (int k = 0; k < 5; k++) out[k] = 0;
    if (!fc.ok) return 0;
    for (int t = 0; t < n_pairs; t++) {
        const int kind = t % 8;
        const int q = kFltK + 4 + r.below(qmax - kFltK - 4), wl = std::max(q + 8, kFltK + r.below(wmax));
        const int alphabet = kind == 6 ? 2 : 4;
        std::vector<int> win = rnd(r, wl, alphabet), qry = rnd(r, q, alphabet);
        auto place = [&](int at, double ps, double pi) { at = std::max(0, std::min(wl - q, at)); std::vector<int> tmp(q); mutate_into(r, qry, 0, tmp, 0, q, ps, pi); for (int x = 0; x < q; x++) win[at + x] = tmp[x]; };
        switch (kind) {
            case 0: place(r.below(wl), 0.0, 0.0); break;                                  // one exact copy
            case 1: place(r.below(wl), 0.03, 0.01); break;                                // one mutated copy
            case 2: place(r.below(wl / 2), 0.0, 0.0); place(wl / 2 + r.below(wl / 2), 0.0, 0.0); break;   // two exact copies (ties)
            case 3: place(r.below(wl / 2), 0.04, 0.0); place(wl / 2 + r.below(wl / 2), 0.02, 0.0); break; // decoy first
            case 4: { const int u = 2 + r.below(9); for (int x = 0; x < q; x++) qry[x] = qry[x % u]; for (int x = 0; x < wl; x++) if (r.below(3)) win[x] = qry[x % u]; break; }  // tandem
            case 5: place(r.below(wl), 0.01, 0.04); break;                                // indel-heavy copy
            case 6: break;                                                               // two-letter noise
            default: { place(r.below(wl), 0.02, 0.0); for (int x = q / 2; x < q; x++) qry[x] = r.below(4); break; }  // half of the query is noise
        }
        if (r.below(10) == 0) qry[r.below(q)] = 4;  // an N in the query
        // ---- the filter, as filter_kernel decides
        const int nb = ((q + wl - 2) >> kFltBinShift) + 1;
        std::vector<long long> bins(nb + 2, 0);
        for (int j = 0; j + kFltK <= wl; j++)
            for (int i = 0; i + kFltK <= q; i++) {
                bool eq = true;
                for (int x = 0; x < kFltK && eq; x++) eq = qry[i + x] < 4 && qry[i + x] == win[j + x];
                if (eq) bins[(j - i + q - 1) >> kFltBinShift]++;
            }
        long long bh = 0; int bb = 0;
        for (int b = 0; b < nb; b++) { const long long s = bins[b] + bins[b + 1]; if (s > bh) { bh = s; bb = b; } }
        const long long nmin = std::min(q, wl);
        const long long s_est = std::min<long long>((long long)fc.a * bh * fc.est_pct / 100, 32767);
        const long long L = s_est > 0 ? max_gap_bases(nmin, fc.a, (int)s_est, fc.go, fc.ge) : -1;
        if (L < 0) continue;
        const int m_old = flt_bins_of(L);
        std::vector<long long> ps(nb + 1, 0);
        for (int b = 0; b < nb; b++) ps[b + 1] = ps[b] + bins[b];
        long long noise = 0;
        for (int bs = 0; bs < nb; bs++) {
            const int be = std::min(bs + m_old, nb);
            if (be - 1 <= bb - kFltCore - 1 || bs >= bb + kFltCore + 2) noise = std::max(noise, ps[be] - ps[bs]);
        }
        uint32_t hc[kFltHc];
        for (int j = 0; j < kFltHc; j++) { const int lo = std::max(0, bb - kFltCore - j), hi = std::min(nb - 1, bb + 1 + kFltCore + j); hc[j] = (uint32_t)(ps[hi + 1] - ps[lo]); }
        long long gb;
        const int m_band = flt_tight_bins(fc, s_est, nmin, m_old, L, [&](int j) { return hc[j]; }, &gb);
        const long long Dlo = (long long)kFltBinW * (bb - kFltCore - m_band + 1), Dhi = (long long)kFltBinW * (bb + 1 + kFltCore + m_band) - 1;
        const long long c0 = std::max(0ll, Dlo - (q - 1)), c1 = std::min<long long>(wl - 1, Dhi);
        // (the kernel also requires told_apart and a range below 70 % of the window before it bothers; the theorem does not)
        if (m_old > 255 || c1 < c0) continue;
        out[0]++;
        // ---- restricted and full sweeps
        const Res full = sw(qry, win, 0, wl, match, mismatch, ambiguous, go, ge);
        const Res part = sw(qry, win, (int)c0, (int)(c1 - c0 + 1), match, mismatch, ambiguous, go, ge);
        if (part.score > full.score) out[3]++;
        // ---- the certificate, as filter_certify_kernel evaluates it
        bool ok = part.score > 0;
        if (ok) {
            const long long L0 = max_gap_bases(nmin, fc.a, part.score, fc.go, fc.ge);
            ok = L0 >= 0 && flt_bins_of(L0) <= m_old && flt_hits_floor(fc, part.score, nmin) > noise;
            if (ok) ok = flt_tight_bins(fc, part.score, nmin, flt_bins_of(L0), L0, [&](int j) { return hc[j]; }, &gb) <= m_band;
        }
        if (!ok) continue;
        out[1]++;
        if (part.score != full.score || part.col != full.col || part.row != full.row) out[2]++;
    }
    return 0;
}
