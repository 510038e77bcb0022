// rev_band.cuh - phase 2 of ssw_align (the reverse pass that finds the begin coordinates; ssw.c:786-868 per
// SURVEY.md §8a A4) as a BANDED, strip-mined, persistent kernel.
//
// Why a band is exact here. The reverse pass runs Smith-Waterman over the reversed prefixes read[read_end..0] /
// ref[ref_end..0] and reports the first column holding `target` (= the forward score) and the smallest row in it.
// The forward pass picked (read_end, ref_end) as the FIRST cell of its tie-break order that holds the matrix maximum,
// so every alignment of score `target` inside the prefixes ends exactly there, i.e. starts at the reverse origin; and a
// path of score `target` over at most q' rows has at most q' matches, hence at most
// L = max_gap_bases(q', match, target, gapOpen, gapExtend) gap bases in total, so it never leaves |row - col| <= L.
// A DP restricted to that band (everything outside read as 0 / no gap) computes values <= the unrestricted ones
// (the recurrence is monotone) and == on every in-band path, and the unrestricted values never exceed `target`.
// So the set of cells holding `target` - all the reference's result depends on - is reproduced exactly.
//
// Layout. A pair belongs to a group of G lanes (128/G pairs per CTA, two virtual lanes per thread, R rows each, as in
// sw_score_kernel); the query rows are processed in strips of S = 2*G*R rows, and strip s only sweeps the columns
// [s*S - L, s*S + S - 1 + L] its band rows can reach: q' * (S + 2L) cells instead of q' * (q' + L). The bottom row of a
// strip goes through a small per-group ring in global memory (two halves, ping-pong; sized by the host from the batch's
// longest query / window, and small enough to stay in L2) to lane 0 of the next strip. The kernel is persistent: a
// fixed grid strides over the job list, whose length it reads from device memory, so the host never needs the per-
// variant pair counts (no synchronisation between the forward pass and this one).
#pragma once
#include "sw_kernels.cuh"

namespace cusw {

constexpr int kRevVariants = 5;  // G = 1, 2, 4, 8, 32 with R = 16: strips of 32, 64, 128, 256, 1024 rows
constexpr int kRevR = 16;
__host__ __device__ __forceinline__ int rev_variant_lanes(int v) { return v == 0 ? 1 : v == 1 ? 2 : v == 2 ? 4 : v == 3 ? 8 : 32; }
// Which strip height a pair takes: the band is 2L+1 wide, a strip sweeps S + 2L columns, so S should not be much
// larger than L (wasted triangle cells) nor than the rows there are; small batches of long queries take a whole warp
// per pair (latency).
__host__ __device__ __forceinline__ int rev_variant_of(int rows, long long gap_bases, int few_pairs) {
    if (few_pairs && rows > 256) return 4;
    const long long want = rows < 32 ? rows : (gap_bases < 32 ? 32 : gap_bases);
    const long long need = rows < want ? rows : want;
    return need <= 32 ? 0 : need <= 64 ? 1 : need <= 128 ? 2 : 3;
}

// Columns a multi-strip pair of variant v can sweep in one strip (what its ring must hold), given the batch's longest
// query and window. Variants 0-2 only take multi-strip pairs whose gap budget is at most their strip height.
__host__ __device__ __forceinline__ long long rev_ring_half(int v, long long max_q, long long max_w, int match, int go, int ge) {
    const long long S = 32ll * rev_variant_lanes(v);
    long long L = v < 3 ? S : max_gap_bases(max_q, match, 1, go, ge);
    if (L < 0 || L > max_w) L = max_w;
    long long need = S + 2 * L + 8;
    if (need > max_w + 8) need = max_w + 8;
    return need;
}

#ifdef CUSW_MAIN_TU
// Bins the pairs of a batch for the banded reverse pass: slot = strip-height variant (+ kRevVariants for the flavour
// that tolerates N in the window), 2 * kRevVariants = nothing to walk back; inside a slot the long jobs come first.
__global__ void rev_band_keys_kernel(int n, const int4* __restrict__ fwd, cudasw_result* __restrict__ res, int match, int go,
                                     int ge, int few_pairs, uint32_t* __restrict__ keys, int* __restrict__ vals,
                                     int* __restrict__ hist, const int* __restrict__ any_n, const uint8_t* __restrict__ wflags,
                                     const FilterRec* __restrict__ sub = nullptr) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int4 f = fwd[p];
    int slot = 2 * kRevVariants;
    uint32_t k2 = 0xFFFFFu;
    if (res[p].status == CUDASW_OK && f.w != 0) res[p].status = CUDASW_ENOMEM;  // strip buffers exhausted in the forward pass
    if (res[p].status == CUDASW_OK && f.x > 0) {
        const int rows = f.z + 1;
        long long L = max_gap_bases(rows, match, f.x, go, ge);
        if (L < 0 || L > f.y + 1) L = f.y + 1;
        if (sub) { const int gb = sub[p].gap_bound; if (gb >= 0 && gb < L) L = gb; }
        const int v = rev_variant_of(rows, L, few_pairs);
        slot = v + ((*any_n != 0 && wflags[p]) ? kRevVariants : 0);
        const long long S = 32ll * rev_variant_lanes(v);
        long long steps = ((rows + S - 1) / S) * (S + 2 * L < (long long)f.y + 1 ? S + 2 * L : (long long)f.y + 1);
        steps >>= 3;
        k2 = 0xFFFFFu - (uint32_t)(steps > 0xFFFFF ? 0xFFFFF : steps);
    }
    keys[p] = ((uint32_t)slot << 20) | k2;
    vals[p] = p;
    atomicAdd(&hist[slot], 1);
}

// (start, count) of every slot of a histogram, for kernels that read their job list bounds from device memory
__global__ void slot_scan_kernel(const int* __restrict__ hist, int nslots, int* __restrict__ slots) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int start = 0;
    for (int s = 0; s < nslots; s++) {
        slots[2 * s] = start;
        slots[2 * s + 1] = hist[s];
        start += hist[s];
    }
}
#endif  // CUSW_MAIN_TU

// FWD = true: the same strips run FORWARDS over the whole query inside the diagonal band [dlo, dhi] (d = column - row) that
// the k-mer filter chose for the pair (io.sub, band_filter.cuh): the forward score pass of a banded pair, with the forward
// pass's result convention (score, first column holding it, smallest row in that column). No target, no early stop; a DP
// over a region that contains the band never exceeds the full DP and equals it on every path inside the band, which is all
// the filter's certificate needs.
template <int G, int R, bool PROF, bool FWD = false>
__global__ void __launch_bounds__(128, score_min_blocks<R, PROF>())
sw_rev_band_kernel(const int* __restrict__ order, const int* __restrict__ slot, PairIO io, const int4* __restrict__ fwd,
                   int4* __restrict__ out, ScoreConst sc, uint32_t* __restrict__ ring, int ring_half) {
    constexpr int DIR = FWD ? 1 : -1;  // sequence bytes are read at base + DIR * index
    constexpr int GPB = 128 / G;
    constexpr int S = 2 * G * R;
    const int lig = threadIdx.x % G;
    const int grp = threadIdx.x / G;
    const int njobs = slot[1];
    order += slot[0];
    uint32_t* ring_a = ring + (size_t)(blockIdx.x * GPB + grp) * (size_t)(2 * ring_half);
    uint32_t* ring_b = ring_a + ring_half;

    // the packed constants live in ordinary registers (see sw_score_kernel)
    uint32_t cPA, cPB, cBB, cGO2, cNGE2;
    const uint32_t* cptr = sc.pool_dev + 16 * (threadIdx.x & 31);
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(cPA) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+4];" : "=r"(cPB) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+8];" : "=r"(cBB) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+12];" : "=r"(cGO2) : "l"(cptr));
    asm volatile("ld.volatile.global.u32 %0, [%1+16];" : "=r"(cNGE2) : "l"(cptr));
    constexpr uint32_t kOut = PROF ? 0x80u : kWindowPad;

    for (int job0 = blockIdx.x * GPB; job0 < njobs; job0 += gridDim.x * GPB) {
    const int gidx = job0 + grp;
    const bool active = gidx < njobs;
    int q = 0, w = 0, target = -1, pair = -1;
    long long dmax = 0;
    long long dlo = 0, dhi = 0;  // the band: diagonals column - row in [dlo, dhi] (reverse pass: [-dmax, dmax])
    const uint8_t* qptr = io.enc_seq;
    const uint8_t* rptr = io.enc_ref;
#ifdef CUSW_BOUNDS_CHECK
    const uint8_t *ck_q0 = nullptr, *ck_q1 = nullptr, *ck_r0 = nullptr, *ck_r1 = nullptr;  // this pair's own bytes
    const uint32_t* ck_b1 = ring_a + 2 * (size_t)ring_half;                                // end of this group's ring
    if (active) {
        const int pr = order[gidx];
        ck_q0 = io.enc_seq + io.seq_off[pr]; ck_q1 = io.enc_seq + io.seq_off[pr + 1];
        ck_r0 = io.enc_ref + io.ref_off[pr]; ck_r1 = io.enc_ref + io.ref_off[pr + 1];
    }
#endif
    if (active && FWD) {
        pair = order[gidx];
        const FilterRec fr = io.sub[pair];
        q = (int)(io.seq_off[pair + 1] - io.seq_off[pair]);
        w = (int)(io.ref_off[pair + 1] - io.ref_off[pair]);
        qptr += io.seq_off[pair];
        rptr += io.ref_off[pair];
        dlo = fr.dlo;
        dhi = fr.dhi;
    }
    if (active && !FWD) {
        pair = order[gidx];
        const int4 f = fwd[pair];
        target = f.x;
        w = f.y + 1;
        q = f.z + 1;
        qptr += io.seq_off[pair] + f.z;
        rptr += io.ref_off[pair] + f.y;
        dmax = max_gap_bases(q, sc.match, target, sc.B, sc.ge);
        if (dmax < 0 || dmax > (long long)q + w) dmax = (long long)q + w;  // unbounded: no cell is outside the band
        if (io.sub) {  // the k-mer filter's proven bound on the gap bases of any alignment of this score (band_filter.cuh)
            const int gb = io.sub[pair].gap_bound;
            if (gb >= 0 && gb < dmax) dmax = gb;
        }
        // an alignment of `target` inside q rows spans at most q + dmax columns
        const long long lim = (long long)q + dmax + 1;
        if (lim < w) w = (int)lim;
        dlo = -dmax;
        dhi = dmax;
    }
    unsigned long long key = 0;  // running (score, -col, -row) maximum of this thread's virtual lanes
    uint32_t* bnd_in = ring_a;
    uint32_t* bnd_out = ring_b;
    int p0 = 0, plen = 0;  // the previous strip's first column and number of columns
    bool failed = false;

    for (int strip = 0;; ++strip) {
    const int row0 = strip * S;
    int c0 = 0, wlen = 0;
    const bool alive = active && row0 < q;
    bool has = alive;
    if (has) {
        const long long c0l = max(0ll, (long long)row0 + dlo);
        const long long c1 = min((long long)w - 1, (long long)row0 + S - 1 + dhi);
        c0 = (int)min(c0l, (long long)w);
        wlen = (int)(c1 - c0l + 1);
        if (wlen <= 0) { has = false; wlen = 0; }
        else if (q > S && wlen > ring_half) { has = false; wlen = 0; failed = true; }  // (multi-strip pairs only) the host sized the ring from the batch maxima
    }
    // (reverse pass: the first strip without columns is the last one that matters; forward: the band may start below row 0 of
    //  the window's reach or end above the last rows, so only the rows decide)
    if (!__any_sync(0xffffffffu, FWD ? alive : has)) break;

    uint32_t Q[R], QB[PROF ? R : 1];
    {
        const int r0 = row0 + (2 * lig) * R;
#pragma unroll
        for (int k = 0; k < R; k++) {
            const int a = r0 + k, b = r0 + R + k;
            if (has && a < q) CUSW_CHECK_PTR(qptr + (int64_t)a * DIR, ck_q0, ck_q1, "query row (band)");
            if (has && b < q) CUSW_CHECK_PTR(qptr + (int64_t)b * DIR, ck_q0, ck_q1, "query row (band)");
            const uint32_t qa = (has && a < q) ? (uint32_t)qptr[(int64_t)a * DIR] : kQueryPad;
            const uint32_t qb = (has && b < q) ? (uint32_t)qptr[(int64_t)b * DIR] : kQueryPad;
            if (PROF) {
                Q[k] = cptr[8 + (qa & 7u)];
                QB[k] = cptr[8 + (qb & 7u)];
            } else {
                Q[k] = qa | (qb << 8);
            }
        }
    }
    uint32_t HA[R], HB[R], E[R];
#pragma unroll
    for (int k = 0; k < R; k++) { HA[k] = cBB; HB[k] = cBB; E[k] = 0u; }
    uint32_t best = cBB;
    int col_lo = -1, col_hi = -1, row_lo = 0, row_hi = 0;
    uint32_t pend = 0u;
    uint32_t Hout = cBB, Fout = 0u, Hdiag_top = cBB;
    uint32_t Rsel = kOut | (kOut << 8);
    const uint8_t* rp = rptr + (int64_t)c0 * DIR;  // local column cl of this strip = byte rp[DIR * cl]
    int t_end = has ? wlen + 2 * G - 1 : 0;
    int t_end_warp = (__reduce_max_sync(0xffffffffu, t_end) + 1) & ~1;
    int c = -2 * lig;  // local column of the low virtual lane
    if ((unsigned)c < (unsigned)wlen) CUSW_CHECK_PTR(rp + (int64_t)c * DIR, ck_r0, ck_r1, "window column (band)");
    uint32_t rb_next = ((unsigned)c < (unsigned)wlen) ? (uint32_t)rp[(int64_t)c * DIR] : kOut;
    const bool read_top = has && strip > 0 && lig == 0;
    const bool write_bot = has && lig == G - 1 && row0 + S < q;
    const int shift = c0 - p0;  // local column -> index into the previous strip's bottom row
    uint32_t top_next = cBB & 0xFFFFu;  // lane 0: (H | F << 16) of the row above the strip at column c
    uint32_t topdiag = cBB & 0xFFFFu;   // lane 0: H of the row above the strip at column c0 - 1 (diagonal of the first cell)
    if (read_top) {
        if ((unsigned)c < (unsigned)wlen && c + shift < plen) { CUSW_CHECK_PTR(bnd_in + c + shift, ring_a, ck_b1, "ring read"); top_next = bnd_in[c + shift]; }
        if (shift >= 1 && shift - 1 < plen) { CUSW_CHECK_PTR(bnd_in + shift - 1, ring_a, ck_b1, "ring read (diagonal)"); topdiag = bnd_in[shift - 1] & 0xFFFFu; }
    }
    if (lig == 0) Hdiag_top = (Hdiag_top & 0xFFFF0000u) | topdiag;

    auto step = [&](uint32_t (&Ho)[R], uint32_t (&Hn)[R]) {
        uint32_t rH = __shfl_up_sync(0xffffffffu, Hout, 1, G);
        uint32_t rF = __shfl_up_sync(0xffffffffu, Fout, 1, G);
        if (lig == 0) { rH = top_next << 16; rF = top_next & 0xFFFF0000u; }
        const uint32_t Hin = prmt(rH, Hout, 0x5432u);  // lo <- upper lane's hi half, hi <- own lo half
        const uint32_t Fin = prmt(rF, Fout, 0x5432u);
        const uint32_t rb = rb_next;
        {
            const int cn = c + 1;
            rb_next = kOut;
            top_next = cBB & 0xFFFFu;
            if ((unsigned)cn < (unsigned)wlen) {
                CUSW_CHECK_PTR(rp + (int64_t)cn * DIR, ck_r0, ck_r1, "window column (band)");
                rb_next = (uint32_t)rp[(int64_t)cn * DIR];
                if (read_top && cn + shift < plen) { CUSW_CHECK_PTR(bnd_in + cn + shift, ring_a, ck_b1, "ring read"); top_next = bnd_in[cn + shift]; }
            }
        }
        Rsel = prmt(Rsel, rb, 0x3204u);  // byte1 <- old byte0, byte0 <- new
        const uint32_t Rs = Rsel + 0x4400u;
        if (PROF) {
            // a half that enters local column 0 starts from the strip's left border (outside the band: zeros), except the
            // very first cell of the strip, whose diagonal neighbour belongs to the previous strip
            if (c == 0 || c == 1) {
                const uint32_t keep = c == 0 ? 0xFFFF0000u : 0x0000FFFFu;
#pragma unroll
                for (int k = 0; k < R; k++) {
                    Ho[k] = (Ho[k] & keep) | (cBB & ~keep);
                    E[k] &= keep;
                }
                const uint32_t border = (lig == 0 && c == 0) ? topdiag : cBB;
                Hdiag_top = (Hdiag_top & keep) | (border & ~keep);
            }
        }
        uint32_t diag = Hdiag_top;
        Hdiag_top = Hin;
        uint32_t cm = cBB;
        uint32_t F = Fin;
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
        Hout = Hn[R - 1];
        Fout = F;
        if (write_bot && (unsigned)(c - 1) < (unsigned)wlen) {
            CUSW_CHECK_PTR(bnd_out + c - 1, ring_a, ck_b1, "ring write");
            // the two halves of the ring must not run into each other either
            CUSW_CHECK_PTR(bnd_out + c - 1, bnd_out, bnd_out + ring_half, "ring write (half)");
            bnd_out[c - 1] = (Hout >> 16) | (Fout & 0xFFFF0000u);
        }
        {
            // columns outside the strip's range must never raise the maximum
            const bool v_lo = (unsigned)c < (unsigned)wlen, v_hi = (unsigned)(c - 1) < (unsigned)wlen;
            if (!(v_lo && v_hi)) {
                const uint32_t keep = (v_lo ? 0x0000FFFFu : 0u) | (v_hi ? 0xFFFF0000u : 0u);
                cm = (cm & keep) | (cBB & ~keep);
            }
        }
        // running best with strict improvement; the row search is deferred until the half stops improving
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
    };

    for (int t = 0; t < t_end_warp; t += 2) {
        step(HA, HB);
        step(HB, HA);
        // early stop (ssw.c:296): once a lane holds `target`, later lanes may still find it in an earlier column, so
        // run on until the last virtual lane has passed this column
        if (!FWD && (t & 2) == 2) {
            const int tg = target + sc.B;
            const bool found = has && (((int)(best & 0xFFFFu) == tg) || ((int)(best >> 16) == tg));
            if (found) t_end = min(t_end, t + 1 + 2 * G + 1);
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) t_end = min(t_end, __shfl_xor_sync(0xffffffffu, t_end, o));
            t_end_warp = (__reduce_max_sync(0xffffffffu, t_end) + 1) & ~1;
        }
    }
    if (pend != 0u) {  // the last step wrote HA
        const uint32_t cnt = first_row_of(HA, best, cBB);
        if ((pend & 0xFFFFu) != 0u) row_lo = (int)(cnt & 0xFFFFu);
        if ((pend >> 16) != 0u) row_hi = (int)(cnt >> 16);
    }
    if (has) {
        const int rbase = row0 + (2 * lig) * R;
        const int a_lo = col_lo < 0 ? -1 : col_lo + c0, a_hi = col_hi < 0 ? -1 : col_hi + c0;  // absolute columns
        const unsigned long long klo = ((unsigned long long)(best & 0xFFFFu) << 40) |
                                       ((unsigned long long)(0xFFFFFFu - (uint32_t)(a_lo + 1)) << 16) |
                                       (unsigned long long)(0xFFFFu - (uint32_t)(rbase + row_lo));
        const unsigned long long khi = ((unsigned long long)(best >> 16) << 40) |
                                       ((unsigned long long)(0xFFFFFFu - (uint32_t)(a_hi + 1)) << 16) |
                                       (unsigned long long)(0xFFFFu - (uint32_t)(rbase + R + row_hi));
        const unsigned long long k2 = klo > khi ? klo : khi;
        key = k2 > key ? k2 : key;
    }
    // later strips read what this one wrote (same warp, other lanes)
    __syncwarp();
    if (has) {
        uint32_t* tmp = bnd_in; bnd_in = bnd_out; bnd_out = tmp;
        p0 = c0;
        plen = wlen;
    } else if (FWD) {
        plen = 0;  // a strip without columns leaves nothing above the next one
    }
    if (!FWD) {
        // columns from the first one that reached `target` on cannot matter any more (later strips hold larger rows)
        unsigned long long kk = key;
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, kk, o);
            kk = other > kk ? other : kk;
        }
        if (has && (int)(kk >> 40) == target + sc.B) {
            const int colf = (int)(0xFFFFFFu - (uint32_t)((kk >> 16) & 0xFFFFFFu)) - 1;
            w = min(w, colf + 1);
        }
    }
    }  // strips

#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    if (G > 1) failed = __shfl_sync(0xffffffffu, failed ? 1 : 0, 0, G) != 0;
    if (active && lig == 0) {
        const int score = (int)(key >> 40) - sc.B;
        const int col = (int)(0xFFFFFFu - (uint32_t)((key >> 16) & 0xFFFFFFu)) - 1;
        const int row = (int)(0xFFFFu - (uint32_t)(key & 0xFFFFu));
        out[pair] = make_int4(score, col, row, failed ? 1 : 0);
    }
    __syncwarp();  // the rings are reused by this group's next job
    }  // jobs
}

}  // namespace cusw
