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

}  // namespace cusw
