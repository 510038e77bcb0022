// ja_mode.cuh - the JAligner-compatible tie-break mode (CUDASW_TIEBREAK_JALIGNER).
//
// The reference falls back to JAlignerAligner when sswjni is disabled (AlignerFactory.java:108-117, `--nojni`).
// JAlignerAligner (alignment/JAlignerAligner.java:30-37) calls jaligner.SmithWatermanGotoh.align(ref, seq, matrix, open,
// extend): one Gotoh pass that stores a pointer per cell (STOP when H == 0, else DIAGONAL, else UP = gap in the read = `D`,
// else LEFT = `I`, in that priority), the first strictly greater cell in (ref, read) order as the end, and a walk from
// the end along the pointers - whole gaps at a time, a gap being extended only when extending is STRICTLY better than
// opening - until a STOP pointer. So: same score and end cell as SSW, but the LONGEST zero-free start instead of SSW's
// nearest one, open-beats-extend on ties, and no band rule. PARITY UNPINNED: the arithmetic ships as class files only
// (jaligner-1.0.jar) and this image has no JVM; the kernel is tested (tests/test_gpu_parity.py) against the CPU restatement
// of that published algorithm, and the three contract vectors the reference runs against every backend. Symbols go
// through the same clean() encoding as the SSW mode, which is what JAlignerAligner's raw-byte matrix gives for
// A, C, G, T, N in either case (other IUPAC letters score differently there; documented limit).
//
// No reverse pass. A banded pointer DP over rows [0, read_end] x the diagonals within L of the end cell's diagonal is
// exact for every cell the walk can visit: all alignments of the maximal score that end in the end cell spend at most
// L = max_gap_bases(...) gap bases, so the walk (which only ever moves between cells of such alignments) and every
// candidate predecessor that TIES with the chosen one stay inside the band, where banded and full values agree; a
// candidate that does not tie is smaller in the full matrix and can only be smaller still in the band.
//
// One CTA per pair (tb_dp-style chunks of diagonals, hand-over by shuffles and shared memory), then warp 0 walks.
#pragma once

namespace cusw {

constexpr int kJaNeg = -(1 << 29);  // "-infinity" of the gap scores (jaligner uses Float.NEGATIVE_INFINITY)

struct JaGeom {
    int rows;      // read_end + 1
    int d0;        // ref_end - read_end: the end cell's diagonal; virtual column j = c - d0, so the end cell is (rows-1, rows-1)
    int band;      // L: half-width of the band around j == i
    int jmin;      // smallest virtual column inside the window (-d0 clipped)
};

__device__ __forceinline__ void ja_cell(const TbConst& tc, int up_v, int up_gi, int diag_v, int s, int& vleft, int& gd,
                                        int& v_out, int& gi_out, uint32_t& dir_out) {
    const int gi1 = up_gi - tc.ge, gi2 = up_v - tc.go;   // insertion (consumes the read): extend / open
    const int gi = max(gi1, gi2);
    const int gd1 = gd - tc.ge, gd2 = vleft - tc.go;     // deletion (consumes the reference): extend / open
    const int g = max(gd1, gd2);
    const int f = diag_v + s;
    const int x = max(max(f, g), max(gi, 0));
    const uint32_t ptr = x == 0 ? 0u : x == f ? 1u : x == g ? 2u : 3u;
    dir_out = ptr | (gd1 > gd2 ? 4u : 0u) | (gi1 > gi2 ? 8u : 0u);
    vleft = x;
    gd = g;
    v_out = x;
    gi_out = gi;
}

// Pointer DP of one pair by the 128 threads of a CTA: 256 chunks of D diagonals (band width 2*band+1 <= 256*D), the
// schedule of tb_dp_warp2<D, 4>. dir: band frame, Wp bytes per row. Returns V at the end cell (uniform).
template <int D>
__device__ int ja_dp_block(const TbConst& tc, const uint8_t* __restrict__ ref, const uint8_t* __restrict__ read, JaGeom gm,
                           uint8_t* __restrict__ dir, TbXchg* xc) {
    const int wlane = threadIdx.x & 31, wib = threadIdx.x >> 5, lane = threadIdx.x;
    const int band = gm.band, W = 2 * band + 1, Wp = tb_frame_width(band);
    const int kb = lane * 2 * D;
    int sv[2 * D], sg[2 * D];  // (V, insertion score) of the previous row at the lane's diagonals
    uint32_t rwin[2 * D];
#pragma unroll
    for (int x = 0; x < 2 * D; x++) {
        sv[x] = 0; sg[x] = kJaNeg;
        const int j = kb - band + x;  // row 0 of this lane
        rwin[x] = (j >= gm.jmin && j < gm.rows) ? (uint32_t)ref[j + gm.d0] : 0u;
    }
    uint32_t qb_next = read[0];
    int outA_v = 0, outA_g = kJaNeg, outB_v = 0, outB_g = kJaNeg;        // (V, deletion score) behind a chunk's last diagonal
    int firstA_v = 0, firstA_g = kJaNeg, firstB_v = 0, firstB_g = kJaNeg;  // (V, insertion score) of a chunk's first diagonal
    int v_end = 0;
    const int iters = gm.rows + 32 * 4 - 1;
    for (int u = 0; u < iters; ++u) {
        const int i = u - lane;
        const bool act = i >= 0 && i < gm.rows && kb < W;
        const int beg = max(i - band, gm.jmin);
        const int end = min(i + band, gm.rows - 1);
        const int endp = min(i - 1 + band, gm.rows - 1);
        const uint32_t qb = qb_next;
        uint32_t rw_in = 0u;
        if (act) {
            if (i + 1 < gm.rows) qb_next = read[i + 1];
            const int jn = kb + i - band + 2 * D;
            if (jn >= gm.jmin && jn < gm.rows) rw_in = ref[jn + gm.d0];
        }
        uint8_t* drow = dir + (long long)Wp * i;
        auto half = [&](auto X0c, int lv, int lg, int uv, int ug, int& out_v, int& out_g, int& first_v, int& first_g) {
            constexpr int X0 = decltype(X0c)::value;
            const int j0 = kb + X0 + i - band;
            int vleft = 0, gd = kJaNeg;
            if (j0 > beg) { vleft = lv; gd = lg; }  // else: left border of the band / the matrix
            int cur_v = sv[X0];
            uint32_t dw = 0;
            bool any = false;
#pragma unroll
            for (int x = 0; x < D; x++) {
                const int j = j0 + x;
                const int nv = x + 1 < D ? sv[X0 + (x + 1 < D ? x + 1 : 0)] : uv;  // row i-1 at diagonal k+1 (the cell above)
                const int ng = x + 1 < D ? sg[X0 + (x + 1 < D ? x + 1 : 0)] : ug;
                if (j >= beg && j <= end) {
                    int up_v = 0, up_g = kJaNeg;
                    if (i >= 1 && j <= endp) { up_v = nv; up_g = ng; }
                    const int dg = (i >= 1 && j - 1 >= gm.jmin) ? cur_v : 0;
                    int v, gi;
                    uint32_t dv;
                    ja_cell(tc, up_v, up_g, dg, tb_sub(tc, qb, rwin[X0 + x]), vleft, gd, v, gi, dv);
                    sv[X0 + x] = v; sg[X0 + x] = gi;
                    if (x == 0) { first_v = v; first_g = gi; }
                    out_v = v; out_g = gd;
                    if (i == gm.rows - 1 && j == gm.rows - 1) v_end = v;
                    if (D >= 4) { dw |= dv << (8 * (x & 3)); any = true; }
                    else drow[kb + X0 + x] = (uint8_t)dv;
                }
                cur_v = nv;
                if (D >= 4 && (x & 3) == 3) {
                    if (any) *reinterpret_cast<uint32_t*>(drow + kb + X0 + x - 3) = dw;
                    dw = 0; any = false;
                }
            }
        };
        if (wlane == 31) xc->out[wib] = make_int2(outB_v, outB_g);
        __syncthreads();
        int lv = __shfl_up_sync(0xffffffffu, outB_v, 1);
        int lg = __shfl_up_sync(0xffffffffu, outB_g, 1);
        if (wlane == 0 && wib > 0) { const int2 t = xc->out[wib - 1]; lv = t.x; lg = t.y; }
        if (act) half(std::integral_constant<int, 0>{}, lv, lg, firstB_v, firstB_g, outA_v, outA_g, firstA_v, firstA_g);
        if (wlane == 0) xc->first[wib] = make_int2(firstA_v, firstA_g);
        __syncthreads();
        int uv = __shfl_down_sync(0xffffffffu, firstA_v, 1);
        int ug = __shfl_down_sync(0xffffffffu, firstA_g, 1);
        if (wlane == 31 && wib + 1 < 4) { const int2 t = xc->first[wib + 1]; uv = t.x; ug = t.y; }
        if (act) {
            half(std::integral_constant<int, D>{}, outA_v, outA_g, uv, ug, outB_v, outB_g, firstB_v, firstB_g);
#pragma unroll
            for (int x = 0; x + 1 < 2 * D; x++) rwin[x] = rwin[x + 1];
            rwin[2 * D - 1] = rw_in;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v_end = max(v_end, __shfl_xor_sync(0xffffffffu, v_end, o));
    __syncthreads();
    if (wlane == 0) xc->mx[wib] = v_end;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; k++) v_end = max(v_end, xc->mx[k]);
    return v_end;
}

// The walk of jaligner's traceback (whole gaps by their stored extension flags, until a STOP pointer or the matrix
// border) by one warp; all lanes hold the same state, runs of DIAGONAL pointers are consumed 32 cells at a time.
// Ops come out last-to-first: with `outp` they are written to their final place (total known), else only counted.
// Returns the number of M/I/D runs of JAlignerAligner.alignmentToCigar (at least 1: an empty walk is "0M"); the stop cell
// is returned in (stop_i, stop_j) (virtual column).
__device__ int ja_walk_warp(const uint8_t* dir, JaGeom gm, uint32_t* outp, int total_ops, int& stop_i, int& stop_j) {
    const int lane = threadIdx.x & 31;
    const int band = gm.band, Wp = tb_frame_width(band);
    int i = gm.rows - 1, j = gm.rows - 1;
    int e = 0, l = 0, prev_op = -1;  // prev_op: 0 = M, 1 = I, 2 = D; -1 = no op yet
    auto in_band = [&](int ii, int jj) { return ii >= 0 && jj >= gm.jmin && jj >= ii - band && jj <= min(ii + band, gm.rows - 1); };
    auto cell = [&](int ii, int jj) -> uint32_t { return dir[(long long)Wp * ii + (jj - ii + band)]; };
    auto put = [&](int idx, uint32_t v) { if (lane == 0 && outp) outp[total_ops - idx] = v; };
    auto add = [&](int op, int n) {
        if (op == prev_op) e += n;
        else {
            if (prev_op >= 0) { ++l; put(l, ((uint32_t)e << 4) | (uint32_t)prev_op); }
            prev_op = op;
            e = n;
        }
    };
    for (;;) {
        {   // runs of diagonal pointers
            const int ii = i - lane, jj = j - lane;
            bool dg = false;
            if (in_band(ii, jj)) dg = (cell(ii, jj) & 3u) == 1u;
            const uint32_t bal = __ballot_sync(0xffffffffu, dg);
            const int run = bal == 0xffffffffu ? 32 : (__ffs((int)~bal) - 1);
            if (run > 0) { add(0, run); i -= run; j -= run; continue; }
        }
        if (!in_band(i, j)) break;  // matrix border (pointers there are STOP) - or outside the band, which an exact walk never is
        const uint32_t b = cell(i, j);
        const uint32_t p = b & 3u;
        if (p == 0u) break;
        if (p == 2u) {  // UP in jaligner's matrix: the whole deletion, along the reference
            int n = 0;
            bool ext;
            do {
                ext = in_band(i, j) && (cell(i, j) & 4u) != 0u;
                --j; ++n;
            } while (ext);
            add(2, n);
        } else {        // LEFT: the whole insertion, along the read
            int n = 0;
            bool ext;
            do {
                ext = in_band(i, j) && (cell(i, j) & 8u) != 0u;
                --i; ++n;
            } while (ext);
            add(1, n);
        }
    }
    // alignmentToCigar (JAlignerAligner.java:59-93): the last run is always appended, even when it is empty ("0M")
    ++l;
    put(l, ((uint32_t)(prev_op >= 0 ? e : 0) << 4) | (uint32_t)(prev_op >= 0 ? prev_op : 0));
    stop_i = i;
    stop_j = j;
    return l;
}

// Sort keys of the JAligner-mode traceback: slot 0 = pairs whose band fits the small frames (up to 256*4 diagonals),
// slot 1 = up to 256*32, slot 2 = nothing to do (failed pairs); inside a slot by descending cost. A band wider than that
// is reported per pair as CUDASW_ETOOLARGE.
constexpr int kJaSmallD = 4, kJaBigD = 32;
__device__ __forceinline__ JaGeom ja_geometry(const int4 f, const TbConst& tc, int q, int w) {
    JaGeom g;
    g.rows = f.z + 1;
    g.d0 = f.y - f.z;
    long long L = max_gap_bases(g.rows, tc.match, f.x, tc.go, tc.ge);
    const long long full = (long long)q + w;
    if (L < 0 || L > full) L = full;
    g.band = (int)L;
    g.jmin = max(-g.d0, -(1 << 28));
    return g;
}

__global__ void ja_keys_kernel(const int* __restrict__ list, int n, PairIO io, const int4* __restrict__ fwd, cudasw_result* __restrict__ res,
                               TbConst tc, uint32_t* __restrict__ keys, int* __restrict__ vals, int* __restrict__ hist,
                               unsigned long long* __restrict__ est_bytes) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= n) return;
    const int pair = list ? list[tid] : tid;
    int slot = 2;
    uint32_t k2 = 0xFFFFFu;
    if (res[pair].status == CUDASW_OK && fwd[pair].w != 0) res[pair].status = CUDASW_ENOMEM;  // strip buffers exhausted
    if (res[pair].status == CUDASW_OK) {
        const int4 f = fwd[pair];
        const int q = (int)(io.seq_off[pair + 1] - io.seq_off[pair]), w = (int)(io.ref_off[pair + 1] - io.ref_off[pair]);
        if (f.x <= 0) slot = 0;  // "0M<q>S": no DP, any kernel emits it
        else {
            const JaGeom g = ja_geometry(f, tc, q, w);
            // the band never needs to reach past the window's first column or the end cell's column
            const long long W = 2ll * g.band + 1;
            const long long bytes = (long long)tb_frame_width(g.band) * g.rows;
            if (W > 256ll * kJaBigD || bytes >= (1ll << 31)) res[pair].status = CUDASW_ETOOLARGE;
            else {
                slot = W <= 256 * kJaSmallD ? 0 : 1;
                const long long cost = bytes >> 6;
                k2 = 0xFFFFFu - (uint32_t)(cost > 0xFFFFF ? 0xFFFFF : cost);
                atomicAdd(est_bytes, (unsigned long long)bytes + 16ull);
            }
        }
    }
    keys[tid] = ((uint32_t)slot << 20) | k2;
    vals[tid] = pair;
    atomicAdd(&hist[slot], 1);
}

template <int DMAX>
__global__ void __launch_bounds__(128, DMAX <= 4 ? 8 : 2)
sw_ja_traceback_kernel(const int* __restrict__ list, const int* __restrict__ slot, PairIO io, const int4* __restrict__ fwd,
                       TbConst tc, TbPool pool, cudasw_result* __restrict__ res, uint32_t* __restrict__ cig,
                       unsigned long long cig_cap, unsigned long long* __restrict__ cig_used, int* __restrict__ retry_list,
                       int* __restrict__ retry_count) {
    __shared__ TbXchg xc;
    __shared__ unsigned long long off_sh;
    const int n = slot[1];
    list += slot[0];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    for (int wid = blockIdx.x; wid < n; wid += gridDim.x) {
        const int pair = list[wid];
        cudasw_result r = res[pair];
        if (r.status != CUDASW_OK) continue;  // (uniform)
        const int q = (int)(io.seq_off[pair + 1] - io.seq_off[pair]), w = (int)(io.ref_off[pair + 1] - io.ref_off[pair]);
        const int4 f = fwd[pair];
        const int score = f.x;
        JaGeom gm{};
        int nbody = 1, stop_i = -1, stop_j = -1;
        const uint8_t* dirp = nullptr;
        bool ok = true;
        if (score > 0) {
            gm = ja_geometry(f, tc, q, w);
            const int W = 2 * gm.band + 1;
            const int dper = (W + 255) / 256;
            const unsigned long long need = (((unsigned long long)tb_frame_width(gm.band) * (unsigned long long)gm.rows) + 15ull) & ~15ull;
            int status = CUDASW_OK;
            if (need > pool.cap) status = CUDASW_ETOOLARGE;
            else {
                __syncthreads();
                if (threadIdx.x == 0) off_sh = atomicAdd(pool.used, need);
                __syncthreads();
                if (off_sh + need > pool.cap) status = 1;  // scratch exhausted: try again in the next round
            }
            if (status != CUDASW_OK) {
                if (threadIdx.x == 0) {
                    if (status == 1) retry_list[atomicAdd(retry_count, 1)] = pair;
                    else res[pair].status = status;
                }
                ok = false;
            } else {
                uint8_t* dir = pool.base + off_sh;
                dirp = dir;
                const uint8_t* ref = io.enc_ref + io.ref_off[pair];
                const uint8_t* read = io.enc_seq + io.seq_off[pair];
                int v;
                if (dper <= 1) v = ja_dp_block<1>(tc, ref, read, gm, dir, &xc);
                else if (dper <= 2) v = ja_dp_block<2>(tc, ref, read, gm, dir, &xc);
                else if (dper <= 4) v = ja_dp_block<4>(tc, ref, read, gm, dir, &xc);
                else if (DMAX >= 32 && dper <= 8) v = ja_dp_block<DMAX >= 32 ? 8 : 1>(tc, ref, read, gm, dir, &xc);
                else if (DMAX >= 32 && dper <= 16) v = ja_dp_block<DMAX >= 32 ? 16 : 1>(tc, ref, read, gm, dir, &xc);
                else v = ja_dp_block<DMAX >= 32 ? 32 : 1>(tc, ref, read, gm, dir, &xc);
                if (v != score) {  // the band argument (or the kernel) is broken: never report a guess
                    if (threadIdx.x == 0) res[pair].status = CUDASW_ECUDA;
                    ok = false;
                }
            }
        }
        __syncthreads();  // every warp's pointer bytes are visible to warp 0
        if (ok && wib == 0) {
            if (score > 0) nbody = ja_walk_warp(dirp, gm, nullptr, 0, stop_i, stop_j);
            // start2 = first aligned read base, start1 = first aligned reference base (jaligner.Alignment.getStart1/2)
            const int read_begin = score > 0 ? stop_i + 1 : 0;
            const int ref_begin = score > 0 ? stop_j + 1 + gm.d0 : 0;
            // bases of the read the M / I runs consume: everything between the start and the end cell
            const int consumed = score > 0 ? f.z - read_begin + 1 : 0;
            const int lead = read_begin > 0, tail = (read_begin + consumed) != q;
            const int nops = nbody + lead + tail;
            unsigned long long coff = 0;
            if (lane == 0) coff = atomicAdd(cig_used, (unsigned long long)nops);
            coff = __shfl_sync(0xffffffffu, coff, 0);
            const bool fits = coff + (unsigned long long)nops <= cig_cap && coff < 0x7fffffffull;
            uint32_t* o = cig + coff + lead;
            if (fits) {
                if (score > 0) { int a, b; ja_walk_warp(dirp, gm, o, nbody, a, b); }
                else if (lane == 0) o[0] = 0u;  // "0M"
            }
            __syncwarp();
            if (lane == 0) {
                r.score = score; r.ref_begin = ref_begin; r.ref_end = score > 0 ? f.y : -1; r.read_begin = read_begin;
                r.read_end = score > 0 ? f.z : 0;
                r.cigar_len = nops; r.cigar_off = (int32_t)coff;
                if (fits) {
                    if (lead) cig[coff] = ((uint32_t)read_begin << 4) | CUDASW_CIGAR_S;
                    if (tail) o[nbody] = ((uint32_t)(q - read_begin - consumed) << 4) | CUDASW_CIGAR_S;
                } else r.status = CUDASW_ENOMEM;
                res[pair] = r;
            }
        }
        __syncthreads();
    }
}

}  // namespace cusw
