// Test-only host library: the certificate arithmetic of gridss_b200/csrc/band_filter.cuh behind a C ABI, so that
// tests/test_filter_math.py can check the very functions the kernels use (closed forms against brute force, against the
// Python model of tools/filter_probe.py, and against optimal paths of the oracle). Built by the test with nvcc; not part
// of libcudasw.so.
#include <algorithm>
#include <cstdint>

#include "../../gridss_b200/csrc/band_filter.cuh"

using namespace cusw;

extern "C" int fm_ok(int match, int mismatch, int ambiguous, int go, int ge) { return make_filter_const(match, mismatch, ambiguous, go, ge).ok; }
extern "C" int fm_ev_div(int match, int mismatch, int ambiguous, int go, int ge) { return make_filter_const(match, mismatch, ambiguous, go, ge).ev_div; }
extern "C" long long fm_hits_floor(int match, int mismatch, int ambiguous, int go, int ge, long long S0, long long n) {
    return flt_hits_floor(make_filter_const(match, mismatch, ambiguous, go, ge), S0, n);
}
extern "C" long long fm_max_gap_bases(long long rows, int match, int target, int go, int ge) { return max_gap_bases(rows, match, target, go, ge); }
extern "C" int fm_bins_of(long long L) { return flt_bins_of(L); }
// hc: kFltHc values; returns m_t and writes the gap bound
extern "C" int fm_tight_bins(int match, int mismatch, int ambiguous, int go, int ge, long long S, long long n, int m0, long long L0,
                             const uint32_t* hc, long long* gap_bound) {
    return flt_tight_bins(make_filter_const(match, mismatch, ambiguous, go, ge), S, n, m0, L0, [&](int j) { return hc[j]; }, gap_bound);
}
extern "C" int fm_k(void) { return kFltK; }
