// tb_warp.cuh - warp-cooperative banded traceback for long / wide-band pairs (BASELINE.json C3, C5).
//
// Same arithmetic, tie-breaks and direction bytes as sw_traceback_kernel (banded_sw, ssw.c:557-734;
// SURVEY.md §8a A4-tb); only the schedule differs. In band coordinates k = j - i + band a cell (i, k)
// needs (i, k-1) [left: H and F], (i-1, k) [diagonal H] and (i-1, k+1) [up: H and E].
//
// Schedule (round 2): the band is cut into 64 chunks of D consecutive diagonals; chunk c processes row i at
// half-step 2*i + c, so its left neighbour (chunk c-1, row i) and its up neighbour (chunk c+1, row i-1) were both
// produced one half-step earlier. Lane L hosts chunks 2L and 2L+1: in loop iteration u it processes row u - L,
// first its even chunk, then its odd one, so EVERY lane computes D cells in EVERY half-step (the round-1 schedule
// gave a lane one chunk and left it idle every other step). Per iteration a lane exchanges four values: the left
// (H, F) of its even chunk comes from lane L-1's odd chunk of the previous iteration, the up (H, E) of its odd chunk
// from lane L+1's even chunk of this iteration; the two hand-overs inside a lane stay in registers. (H, E) of the
// previous row and the window bytes of the lane's 2*D diagonals live in registers; the window bytes slide by one
// column per row, so a row costs one global byte load per lane, issued a row ahead.
//
// Directions are stored in the band-frame layout of the one-thread kernel: cell (i, k) at dir[Wp*i + k],
// Wp = (2*band+1) rounded up to 4, so the D cells of a chunk leave as packed 32-bit words.
//
// What "outside the band" reads as (derived from the reference's rolling h_b/e_b/h_c rows,
// ssw.c:586-628): 0 for H, E above or left of the band and left of column 0;
// row 0 uses -gapOpen / -gapExtend; F and the left H start every row at 0. One quirk of the reference's
// `edge` zeroing (ssw.c:596) is reproduced: while i <= band + 1, if the band is clipped by the end of
// the reference in this row and the previous one and refLen <= 2*band + 2, the up neighbour of the
// last column reads as (0, 0) although it is inside the band.
#pragma once
#include <type_traits>

namespace cusw {

// One cell of banded_sw (ssw.c:607-626). The values are the reference's; the evaluation order is arranged so that the
// chain from one cell of a row to the next (left H -> F -> H) is two dependent instructions (VIADDMNMX, VIMNMX3):
// everything that only depends on the previous row (E, the diagonal candidate, their maximum) and the direction bits
// sit beside that chain.
// Borders need no special cases here because of how the callers keep their state: every (H, E) slot starts a band pass
// at (0, 0) and is only ever written for an in-band cell, so a neighbour that lies outside the band or the matrix reads
// as (0, 0) - which is exactly what the reference's rolling rows hold there (h_b = e_b = 0, ssw.c:596), and what its
// first-row rule (-gapOpen / -gapExtend, ssw.c:607-608) amounts to. The one exception is the `edge` quirk, which the
// callers apply by passing (0, 0) for the up neighbour.
__device__ __forceinline__ void tb_cell(const TbConst& tc, int up_h, int up_e, int hdiag, int s,
                                        int& hleft, int& fgap, int& mx, int& h_out, int& e_out,
                                        uint8_t& dir_out) {
    // E (gap along the read, `I`): from the previous row only
    const int te1 = up_h - tc.go;
    const int te2 = up_e - tc.ge;
    const int E = max(te1, te2);
    const uint32_t dE = te1 > te2 ? 1u : 0u;
    const int e1 = max(E, 0);
    const int d = hdiag + s;
    // F (gap along the reference, `D`) and H: the chain along the row
    const int tf2 = fgap - tc.ge;
    const int tf1 = hleft - tc.go;
    const int F = __viaddmax_s32(hleft, -tc.go, tf2);  // max(hleft - go, fgap - ge)
    const int h = __vimax3_s32(max(e1, d), F, 0);      // max(max(e1, f1), d), f1 = max(F, 0)
    fgap = F;
    hleft = h;
    // direction bits (ties: extend beats open, diagonal beats gaps, F beats E); h == d <=> max(e1, f1) <= d
    const uint32_t dF = tf1 > tf2 ? 1u : 0u;
    const int f1 = max(F, 0);
    const uint32_t dH = (h == d) ? 1u : (e1 > f1 ? 2u + dE : 4u + dF);
    dir_out = (uint8_t)(dH | (dE << 3) | (dF << 4));
    mx = max(mx, h);
    h_out = h;
    e_out = E;
}

__host__ __device__ __forceinline__ int tb_frame_width(int band) { return (2 * band + 1 + 3) & ~3; }

// Hand-over slots between the warps of a CTA that works on one pair together (tb_dp_warp2 with NW > 1).
struct TbXchg {
    int2 out[4];    // (H, F) behind the last diagonal of a warp's last lane
    int2 first[4];  // (H, E) of the first diagonal of a warp's first lane
    int mx[4];
};

// Banded DP of one pair by NW cooperating warps (NW = 1: one warp, no shared memory; NW = 4: the 128 threads of a CTA,
// for batches with too few pairs to fill the machine with one warp each), state in registers: 2*D diagonals per lane
// (band width 2*band+1 <= 64*NW*D). "Lane" below is the thread's index among the 32*NW cooperating lanes; neighbouring
// lanes exchange by shuffles inside a warp and through `xc` (shared memory, two CTA barriers per iteration) across
// warps. Returns the banded maximum (uniform). `dir` is the band frame (Wp bytes per row, 4-byte aligned).
template <int D, int NW = 1>
__device__ int tb_dp_warp2(const TbConst& tc, const uint8_t* __restrict__ ref, const uint8_t* __restrict__ read,
                           int refLen, int readLen, int band, uint8_t* __restrict__ dir, TbXchg* xc = nullptr) {
    const int wlane = threadIdx.x & 31;
    const int wib = NW > 1 ? (threadIdx.x >> 5) : 0;
    const int lane = NW > 1 ? (int)threadIdx.x : wlane;
    const int W = 2 * band + 1;
    const int Wp = tb_frame_width(band);
    const int kb = lane * 2 * D;  // first diagonal of this lane's even chunk
    int sh[2 * D], se[2 * D];     // (H, E) of the previous row at the lane's diagonals
    uint32_t rwin[2 * D];         // window bytes at the lane's diagonals for its next row
#pragma unroll
    for (int x = 0; x < 2 * D; x++) {
        sh[x] = 0; se[x] = 0;
        const int j = kb - band + x;  // row 0 of this lane
        rwin[x] = (unsigned)j < (unsigned)refLen ? (uint32_t)ref[j] : 0u;
    }
    int mx = 0;
    uint32_t qb_next = read[0];
    int outA_h = 0, outA_f = 0, outB_h = 0, outB_f = 0;          // (H, F) after the last diagonal of a chunk, latest row
    int firstA_h = 0, firstA_e = 0, firstB_h = 0, firstB_e = 0;  // (H, E) of the first diagonal of a chunk, latest row
    const int iters = readLen + 32 * NW - 1;
    for (int u = 0; u < iters; ++u) {
        const int i = u - lane;
        const bool act = i >= 0 && i < readLen && kb < W;
        const int beg = max(i - band, 0);
        const int end = min(i + band, refLen - 1);
        const int endp = min(i - 1 + band, refLen - 1);  // last column of the previous row
        const bool quirk = i >= 1 && i <= band + 1 && end == refLen - 1 && endp == end && end <= 2 * band + 1;
        const int quirk_col = quirk ? end : -1;
        const uint32_t qb = qb_next;
        uint32_t rw_in = 0u;
        if (act) {  // what the next row of this lane needs, in flight while this row is computed
            if (i + 1 < readLen) qb_next = read[i + 1];
            const int jn = kb + i - band + 2 * D;
            if ((unsigned)jn < (unsigned)refLen) rw_in = ref[jn];
        }
        uint8_t* drow = dir + (long long)Wp * i;

        // one chunk of D diagonals starting at register slot X0; (lh, lf) = left neighbour, (uh, ue) = up neighbour
        // of its last diagonal
        auto half = [&](auto X0c, int lh, int lf, int uh, int ue, int& out_h, int& out_f, int& first_h, int& first_e) {
            constexpr int X0 = decltype(X0c)::value;
            const int j0 = kb + X0 + i - band;  // column of the chunk's first diagonal
            int hleft = 0, fgap = 0;
            if (j0 > beg) { hleft = lh; fgap = lf; }  // else: row border (h_c[0] = 0, f = 0)
            int cur_h = sh[X0];
            uint32_t dw = 0;
            bool any = false;
#pragma unroll
            for (int x = 0; x < D; x++) {
                const int j = j0 + x;
                const int nh = x + 1 < D ? sh[X0 + (x + 1 < D ? x + 1 : 0)] : uh;  // row i-1 at diagonal k+1
                const int ne = x + 1 < D ? se[X0 + (x + 1 < D ? x + 1 : 0)] : ue;
                if (j >= beg && j <= end) {
                    const bool qk = j == quirk_col;  // the reference's `edge` quirk: this cell's up neighbour reads as (0, 0)
                    int h, e;
                    uint8_t dv;
                    tb_cell(tc, qk ? 0 : nh, qk ? 0 : ne, cur_h, tb_sub(tc, qb, rwin[X0 + x]), hleft, fgap, mx, h, e, dv);
                    sh[X0 + x] = h; se[X0 + x] = e;
                    if (x == 0) { first_h = h; first_e = e; }
                    out_h = h; out_f = fgap;
                    if (D >= 4) { dw |= (uint32_t)dv << (8 * (x & 3)); any = true; }
                    else { CUSW_CHECK_PTR(drow + kb + X0 + x, dir, dir + (long long)Wp * readLen, "direction byte (cooperative DP)"); drow[kb + X0 + x] = dv; }
                }
                cur_h = nh;
                if (D >= 4 && (x & 3) == 3) {
                    if (any) {
                        CUSW_CHECK_PTR(drow + kb + X0 + x, dir, dir + (long long)Wp * readLen, "direction word (cooperative DP)");
                        *reinterpret_cast<uint32_t*>(drow + kb + X0 + x - 3) = dw;
                    }
                    dw = 0; any = false;
                }
            }
        };
        // ---- even chunk: left from lane L-1's odd chunk (previous iteration), up from the own odd chunk (previous iteration)
        if (NW > 1) {
            if (wlane == 31) xc->out[wib] = make_int2(outB_h, outB_f);
            __syncthreads();
        }
        int lh = __shfl_up_sync(0xffffffffu, outB_h, 1);
        int lf = __shfl_up_sync(0xffffffffu, outB_f, 1);
        if (NW > 1 && wlane == 0 && wib > 0) { const int2 v = xc->out[wib - 1]; lh = v.x; lf = v.y; }
        if (act) half(std::integral_constant<int, 0>{}, lh, lf, firstB_h, firstB_e, outA_h, outA_f, firstA_h, firstA_e);
        // ---- odd chunk: left from the own even chunk (just computed), up from lane L+1's even chunk (this iteration)
        if (NW > 1) {
            if (wlane == 0) xc->first[wib] = make_int2(firstA_h, firstA_e);
            __syncthreads();
        }
        int uh = __shfl_down_sync(0xffffffffu, firstA_h, 1);
        int ue = __shfl_down_sync(0xffffffffu, firstA_e, 1);
        if (NW > 1 && wlane == 31 && wib + 1 < NW) { const int2 v = xc->first[wib + 1]; uh = v.x; ue = v.y; }
        if (act) {
            half(std::integral_constant<int, D>{}, outA_h, outA_f, uh, ue, outB_h, outB_f, firstB_h, firstB_e);
#pragma unroll
            for (int x = 0; x + 1 < 2 * D; x++) rwin[x] = rwin[x + 1];
            rwin[2 * D - 1] = rw_in;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (NW > 1) {
        __syncthreads();  // (also: every warp's direction bytes are written before anyone walks them)
        if (wlane == 0) xc->mx[wib] = mx;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < NW; k++) mx = max(mx, xc->mx[k]);
    }
    return mx;
}

// Bands too wide for registers (2*band+1 > 2048): state in `row` (global scratch, one int2 per diagonal), Dn diagonals
// per lane, the round-1 schedule (lane L processes row i at step 2*i + L). Same band-frame direction layout.
__device__ int tb_dp_warp_mem(const TbConst& tc, const uint8_t* __restrict__ ref, const uint8_t* __restrict__ read,
                              int refLen, int readLen, int band, uint8_t* __restrict__ dir, int2* __restrict__ row,
                              int nd) {
    const int lane = threadIdx.x & 31;
    const int W = 2 * band + 1;
    const int Wp = tb_frame_width(band);
    const int kb = lane * nd;
    for (int x = 0; x < nd; x++)
        if (kb + x < W) row[kb + x] = make_int2(0, 0);
    int mx = 0;
    int out_h = 0, out_f = 0;      // (H, F) after this lane's last diagonal in its latest row
    int first_h = 0, first_e = 0;  // (H, E) of this lane's first diagonal in its latest row
    const int steps = 2 * readLen + 31;
    for (int t = 0; t < steps; ++t) {
        const int left_h = __shfl_up_sync(0xffffffffu, out_h, 1);
        const int left_f = __shfl_up_sync(0xffffffffu, out_f, 1);
        const int upn_h = __shfl_down_sync(0xffffffffu, first_h, 1);
        const int upn_e = __shfl_down_sync(0xffffffffu, first_e, 1);
        const int ti = t - lane;
        const int i = ti >> 1;
        if (ti >= 0 && !(ti & 1) && i < readLen && kb < W) {
            const int beg = max(i - band, 0);
            const int end = min(i + band, refLen - 1);
            const int endp = min(i - 1 + band, refLen - 1);
            const bool quirk = i >= 1 && i <= band + 1 && end == refLen - 1 && endp == end && end <= 2 * band + 1;
            const uint32_t qb = (uint32_t)read[i];
            uint8_t* drow = dir + (long long)Wp * i + (band - i);  // drow[j] = cell (i, k = j - i + band)
            const int j0 = kb + i - band;
            int hleft = 0, fgap = 0;
            if (j0 > beg) { hleft = left_h; fgap = left_f; }
            int2 cur = row[kb];
            for (int x = 0; x < nd; x++) {
                const int j = j0 + x;
                int2 nxt;
                if (x + 1 < nd) nxt = row[min(kb + x + 1, W - 1)];
                else nxt = make_int2(upn_h, upn_e);
                if (j >= beg && j <= end) {
                    // (the scratch row is shared by all band widths of this pass only: zeroed above, written for in-band cells)
                    int up_h = 0, up_e = 0;
                    if (j <= endp && !(quirk && j == end)) { up_h = nxt.x; up_e = nxt.y; }
                    const int hdiag = (i >= 1 && j >= 1) ? cur.x : 0;
                    int h, e;
                    uint8_t dv;
                    tb_cell(tc, up_h, up_e, hdiag, tb_sub(tc, qb, (uint32_t)ref[j]), hleft, fgap, mx, h, e, dv);
                    CUSW_CHECK_PTR(drow + j, dir, dir + (long long)Wp * readLen, "direction byte (memory-backed DP)");
                    drow[j] = dv;
                    row[kb + x] = make_int2(h, e);
                    if (x == 0) { first_h = h; first_e = e; }
                    out_h = h; out_f = fgap;
                }
                cur = nxt;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    return mx;
}

// Warp-cooperative walk of the direction matrix (band-frame layout, center = band): same moves and the same run-length
// bookkeeping as tb_walk, but while the walk is in state H the 32 lanes read the next 32 cells down the diagonal at once
// and every leading "diagonal" cell is consumed in one go - along an alignment most cells are M, so the dependent
// global loads of the walk drop by up to 32x. All lanes hold the same walk state; only lane 0 writes.
// (`dir` was written by other lanes of this warp in the same kernel: no __restrict__, so it is never read through the
// non-coherent path.)
__device__ int tb_walk_warp(const uint8_t* dir, int width_d, int band, int readLen, int refLen,
                            uint32_t* outp, int total_ops, uint32_t* buf) {
    const int lane = threadIdx.x & 31;
    int i = readLen - 1, j = refLen - 1;
    int e = 0, l = 0, state = 2;
    int op = 0, prev_op = 0;  // 0 = M, 1 = I, 2 = D
    auto put = [&](int idx, uint32_t v) {
        if (lane == 0) {
            if (outp) { CUSW_CHECK_PTR(outp + total_ops - idx, outp, outp + total_ops, "CIGAR word"); outp[total_ops - idx] = v; }
            if (buf && idx <= kTbOpsBuf) buf[idx - 1] = v;
        }
    };
#ifdef CUSW_BOUNDS_CHECK
    const uint8_t* ck_d1 = dir + (long long)width_d * readLen;
#endif
    auto in_band = [&](int ii, int jj) { return jj >= max(ii - band, 0) && jj <= min(ii + band, refLen - 1); };
    while (i > 0) {
        if (state == 2) {
            // lane x looks at cell (i - x, j - x)
            const int ii = i - lane, jj = j - lane;
            bool diag = false;
            if (ii > 0 && in_band(ii, jj)) {
                CUSW_CHECK_PTR(dir + (long long)width_d * ii + (jj - ii + band), dir, ck_d1, "direction byte (cooperative walk)");
                diag = (dir[(long long)width_d * ii + (jj - ii + band)] & 7u) == 1u;
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, diag);
            const uint32_t run = bal == 0xffffffffu ? 32u : (uint32_t)(__ffs((int)~bal) - 1);  // leading diagonal cells (0..32)
            if (run > 0) {
                i -= (int)run; j -= (int)run;
                op = 0;
                if (prev_op == 0) e += (int)run;
                else { ++l; put(l, ((uint32_t)e << 4) | (uint32_t)prev_op); prev_op = 0; e = (int)run; }
                continue;
            }
        }
        if (!in_band(i, j)) return CUDASW_EUNDEF;
        CUSW_CHECK_PTR(dir + (long long)width_d * i + (j - i + band), dir, ck_d1, "direction byte (cooperative walk)");
        const uint32_t b = dir[(long long)width_d * i + (j - i + band)];
        const int dv = state == 2 ? (int)(b & 7u) : state == 0 ? (2 + (int)((b >> 3) & 1u)) : (4 + (int)((b >> 4) & 1u));
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

}  // namespace cusw
