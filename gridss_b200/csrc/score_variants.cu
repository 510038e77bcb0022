// score_variants.cu - instantiates sw_score_kernel<SCORE_G, R, REV> for one group width.
#include "variants.h"

#ifndef SCORE_G
#error "compile with -DSCORE_G=<1|2|4|8|16|32>"
#endif

namespace cusw {

template <int G, int R, bool REV, bool MULTI, bool PROF, bool CHN = false>
static void launch_score(const int* order, int njobs, PairIO io, const int4* fwd, int4* out, ScoreConst sc,
                         StripPool sp, cudaStream_t st, const int* slot) {
    // njobs sizes the grid (with `slot` it is an upper bound; the kernel reads the true list bounds from device memory)
    constexpr int GPB = 128 / G;
    int blocks = (njobs + GPB - 1) / GPB;
    sw_score_kernel<G, R, REV, MULTI, PROF, CHN><<<blocks, 128, 0, st>>>(order, njobs, io, fwd, out, sc, sp, slot);
}

#define V(R) v.push_back(Variant{SCORE_G, R, 2 * SCORE_G * R, false, launch_score<SCORE_G, R, false, false, true>, launch_score<SCORE_G, R, true, false, true>, launch_score<SCORE_G, R, false, false, false>, launch_score<SCORE_G, R, true, false, false>, launch_score<SCORE_G, R, false, false, true, true>})
#define VM(R) v.push_back(Variant{SCORE_G, R, 2 * SCORE_G * R, true, launch_score<SCORE_G, R, false, true, true>, launch_score<SCORE_G, R, true, true, true>, launch_score<SCORE_G, R, false, true, false>, launch_score<SCORE_G, R, true, true, false>, launch_score<SCORE_G, R, false, true, true>})
#define CAT2(a, b) a##b
#define CAT(a, b) CAT2(a, b)

void CAT(variants_g, SCORE_G)(std::vector<Variant>& v) {
#if SCORE_G == 1
    V(8); V(12); V(16); V(20); V(24);
#else
    V(16); V(17); V(18); V(19); V(20); V(21); V(22); V(23); V(24);
#endif
#if SCORE_G == 4 || SCORE_G == 8
    V(28); V(32); V(38);
#endif
#if SCORE_G == 32
    V(26); V(28); V(30); V(32);
    VM(20); VM(32);  // strip-mined: any query length
#endif
}

}  // namespace cusw
