// rev_variants.cu - instantiates sw_rev_band_kernel (rev_band.cuh) for its five strip heights and both flavours.
#include "rev_band.cuh"
#include "variants.h"

namespace cusw {

template <int G, bool PROF>
static void launch_one(int blocks, const int* order, const int* slot, PairIO io, const int4* fwd, int4* out, ScoreConst sc,
                       uint32_t* ring, int ring_half, cudaStream_t st) {
    sw_rev_band_kernel<G, kRevR, PROF><<<blocks, 128, 0, st>>>(order, slot, io, fwd, out, sc, ring, ring_half);
}

void launch_rev_band(int v, bool prof, int blocks, const int* order, const int* slot, PairIO io, const int4* fwd, int4* out,
                     ScoreConst sc, uint32_t* ring, int ring_half, cudaStream_t st) {
#define RB(G)                                                                                        \
    if (prof) launch_one<G, true>(blocks, order, slot, io, fwd, out, sc, ring, ring_half, st);      \
    else launch_one<G, false>(blocks, order, slot, io, fwd, out, sc, ring, ring_half, st)
    switch (v) {
        case 0: RB(1); break;
        case 1: RB(2); break;
        case 2: RB(4); break;
        case 3: RB(8); break;
        default: RB(32); break;
    }
#undef RB
}

// The forward flavour of the same strips (rev_band.cuh, FWD): the banded forward pass of the k-mer filter's pairs
// (band_filter.cuh). Query-profile flavour only: the filter never takes a window that holds an N.
template <int G>
static void launch_fwd_one(int blocks, const int* order, const int* slot, PairIO io, int4* out, ScoreConst sc, uint32_t* ring, int ring_half,
                           cudaStream_t st) {
    sw_rev_band_kernel<G, kRevR, true, true><<<blocks, 128, 0, st>>>(order, slot, io, nullptr, out, sc, ring, ring_half);
}

void launch_fwd_band(int v, int blocks, const int* order, const int* slot, PairIO io, int4* out, ScoreConst sc, uint32_t* ring,
                     int ring_half, cudaStream_t st) {
    switch (v) {
        case 0: launch_fwd_one<1>(blocks, order, slot, io, out, sc, ring, ring_half, st); break;
        case 1: launch_fwd_one<2>(blocks, order, slot, io, out, sc, ring, ring_half, st); break;
        case 2: launch_fwd_one<4>(blocks, order, slot, io, out, sc, ring, ring_half, st); break;
        case 3: launch_fwd_one<8>(blocks, order, slot, io, out, sc, ring, ring_half, st); break;
        default: launch_fwd_one<32>(blocks, order, slot, io, out, sc, ring, ring_half, st); break;
    }
}

}  // namespace cusw
