// variants.h - the (G, R) instantiation table of sw_score_kernel, split over several translation
// units (score_variants.cu compiled once per G) so the build parallelises.
#pragma once
#include <cuda_runtime.h>

#include <vector>

#include "sw_kernels.cuh"

namespace cusw {

typedef void (*score_launch_fn)(const int*, int, PairIO, const int4*, int4*, ScoreConst, StripPool, cudaStream_t, const int* slot);

struct Variant {
    int G, R, rows;  // rows = 2*G*R; for multi (strip-mined) variants rows per strip
    bool multi;
    score_launch_fn fwd, rev;    // query-profile flavour (window free of N)
    score_launch_fn fwdx, revx;  // xor-selector flavour (any symbols)
    score_launch_fn fwdc;        // query-profile forward pass with quiet chains (strip-mined variants: same as fwd)
};

// defined in score_variants.cu, one per -DSCORE_G
void variants_g1(std::vector<Variant>&);
void variants_g2(std::vector<Variant>&);
void variants_g4(std::vector<Variant>&);
void variants_g8(std::vector<Variant>&);
void variants_g16(std::vector<Variant>&);
void variants_g32(std::vector<Variant>&);

// single-strip variants ascending by rows, then the strip-mined ones; for equal rows the larger R (fewer lanes) wins; CUDASW_VARIANTS="8x19,4x38" restricts
std::vector<Variant> collect_variants();

// rev_variants.cu: the banded reverse pass (rev_band.cuh), variant v in [0, kRevVariants), persistent grid of `blocks`
void launch_rev_band(int v, bool prof, int blocks, const int* order, const int* slot, PairIO io, const int4* fwd, int4* out,
                     ScoreConst sc, uint32_t* ring, int ring_half, cudaStream_t st);

// ... and their forward flavour: the banded forward pass of the pairs the k-mer filter gave a diagonal band (band_filter.cuh)
void launch_fwd_band(int v, int blocks, const int* order, const int* slot, PairIO io, int4* out, ScoreConst sc, uint32_t* ring,
                     int ring_half, cudaStream_t st);

}  // namespace cusw
