// tb_warp.cuh - warp-cooperative banded traceback for long / wide-band pairs (BASELINE.json C3, C5).
//
// Same arithmetic, tie-breaks and direction bytes as sw_traceback_kernel (banded_sw, ssw.c:557-734;
// SURVEY.md §8a A4-tb); only the schedule differs. In band coordinates k = j - i + band a cell (i, k)
// needs (i, k-1) [left: H and F], (i-1, k) [diagonal H] and (i-1, k+1) [up: H and E]. Lane L owns the
// D consecutive diagonals [L*D, (L+1)*D) with their (H, E) of the previous row in registers (or in a
// per-pair scratch row when the band is too wide) and processes row i at step t = 2*i + L, so the left
// neighbour (lane L-1, row i) and the up neighbour (lane L+1, row i-1) were both produced at step t-1
// and arrive by one shuffle each. A pair therefore takes 2*readLen + 32 steps of D cells per lane.
//
// What "outside the band" reads as (derived from the reference's rolling h_b/e_b/h_c rows,
// ssw.c:586-628): 0 for H, E above or left of the band and left of column 0;
// row 0 uses -gapOpen / -gapExtend; F and the left H start every row at 0. One quirk of the reference's
// `edge` zeroing (ssw.c:596) is reproduced: while i <= band + 1, if the band is clipped by the end of
// the reference in this row and the previous one and refLen <= 2*band + 2, the up neighbour of the
// last column reads as (0, 0) although it is inside the band.
#pragma once

namespace cusw {

struct TbCell {
    int h, e;
};

// One cell, identical to the loop body of sw_traceback_kernel.
__device__ __forceinline__ void tb_cell(const TbConst& tc, int i, int up_h, int up_e, int hdiag, int s,
                                        int& hleft, int& fgap, int& mx, int& h_out, int& e_out,
                                        uint8_t& dir_out) {
    int t1 = i == 0 ? -tc.go : up_h - tc.go;
    int t2 = i == 0 ? -tc.ge : up_e - tc.ge;
    const int E = max(t1, t2);
    const uint32_t dE = t1 > t2 ? 1u : 0u;
    t1 = hleft - tc.go;
    t2 = fgap - tc.ge;
    fgap = max(t1, t2);
    const uint32_t dF = t1 > t2 ? 1u : 0u;
    const int e1 = max(E, 0), f1 = max(fgap, 0);
    const int g = max(e1, f1);
    const int d = hdiag + s;
    const int h = max(g, d);
    mx = max(mx, h);
    const uint32_t dH = (g <= d) ? 1u : (e1 > f1 ? 2u + dE : 4u + dF);
    dir_out = (uint8_t)(dH | (dE << 3) | (dF << 4));
    h_out = h;
    e_out = E;
    hleft = h;
}

// Banded DP of one pair by one warp. D > 0: state in registers (band width 2*band+1 <= 32*D);
// D == 0: state in `row` (global scratch, one int2 per diagonal), Dn diagonals per lane.
// Returns the banded maximum (warp-uniform).
template <int D>
__device__ int tb_dp_warp(const TbConst& tc, const uint8_t* __restrict__ ref, const uint8_t* __restrict__ read,
                          int refLen, int readLen, int band, uint8_t* __restrict__ dir, int2* __restrict__ row,
                          int Dn) {
    const int lane = threadIdx.x & 31;
    const int W = 2 * band + 1;
    const int nd = D > 0 ? D : Dn;
    const int kb = lane * nd;
    int2 st[D > 0 ? D : 1];
#pragma unroll
    for (int x = 0; x < (D > 0 ? D : 1); x++) st[x] = make_int2(0, 0);
    if (D == 0)
        for (int x = 0; x < nd; x++)
            if (kb + x < W) row[kb + x] = make_int2(0, 0);
    int mx = 0;
    // register-resident instantiations keep the window bytes of the lane's diagonals in registers too: they slide by
    // one column per row, so a row costs one global byte load (issued a row ahead) instead of D
    uint32_t rwin[D > 0 ? D : 1];
    uint32_t qb_next = 0;
    if (D > 0) {
#pragma unroll
        for (int x = 0; x < (D > 0 ? D : 1); x++) {
            const int j = kb - band + x;
            rwin[x] = (unsigned)j < (unsigned)refLen ? (uint32_t)ref[j] : 0u;
        }
        qb_next = read[0];
    }
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
            const int endp = min(i - 1 + band, refLen - 1);  // last column of the previous row
            const bool quirk = i >= 1 && i <= band + 1 && end == refLen - 1 && endp == end && end <= 2 * band + 1;
            const uint32_t qb = D > 0 ? qb_next : (uint32_t)read[i];
            uint8_t* drow = dir + (long long)W * i - beg;
            const int j0 = kb + i - band;  // column of this lane's first diagonal
            uint32_t rw_in = 0u;
            if (D > 0) {  // what the next row of this lane needs, in flight while this row is computed
                if (i + 1 < readLen) qb_next = read[i + 1];
                const int jn = j0 + D;
                if ((unsigned)jn < (unsigned)refLen) rw_in = ref[jn];
            }
            int hleft = 0, fgap = 0;
            if (j0 > beg) { hleft = left_h; fgap = left_f; }  // else: row border (h_c[0] = 0, f = 0)
            // (H, E) of row i-1 at this lane's first diagonal is the diagonal neighbour of the first cell
            int2 cur = D > 0 ? st[0] : row[kb];
            const int xn = D > 0 ? D : nd;
#pragma unroll
            for (int x = 0; x < xn; x++) {
                const int j = j0 + x;
                int2 nxt;  // row i-1 at diagonal k+1 = up neighbour of this cell, diagonal neighbour of the next
                if (x + 1 < nd) nxt = D > 0 ? st[x + 1 < D ? x + 1 : 0] : row[min(kb + x + 1, W - 1)];
                else nxt = make_int2(upn_h, upn_e);
                if (j >= beg && j <= end) {
                    int up_h = 0, up_e = 0;
                    if (j <= endp && !(quirk && j == end)) { up_h = nxt.x; up_e = nxt.y; }
                    const int hdiag = (i >= 1 && j >= 1) ? cur.x : 0;
                    int h, e;
                    uint8_t dv;
                    tb_cell(tc, i, up_h, up_e, hdiag, tb_sub(tc, qb, D > 0 ? rwin[x < (D > 0 ? D : 1) ? x : 0] : (uint32_t)ref[j]), hleft, fgap, mx, h, e, dv);
                    drow[j] = dv;
                    if (D > 0) st[x < D ? x : 0] = make_int2(h, e); else row[kb + x] = make_int2(h, e);
                    if (x == 0) { first_h = h; first_e = e; }
                    out_h = h; out_f = fgap;
                }
                cur = nxt;
            }
            if (D > 0) {
#pragma unroll
                for (int x = 0; x + 1 < (D > 0 ? D : 1); x++) rwin[x] = rwin[x + 1];
                rwin[(D > 0 ? D : 1) - 1] = rw_in;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    return mx;
}

}  // namespace cusw
