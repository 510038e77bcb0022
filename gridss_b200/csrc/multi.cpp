// multi.cpp - one host process driving several B200s: the batch is cut into contiguous,
// cost-balanced (sum of query_len * window_len) slices, one per GPU; every GPU has its own engine handle
// (own streams, pinned staging, device buffers) and its own host thread; results land directly in the
// caller's arrays at the pairs' original ordinals. Pairs are independent, so there is no exchange step and
// no collective (no NCCL) on this path.
//
// This is what a single JVM needs (the reference's worker pool lives in one process,
// AnnotateInexactHomology.java:17-21); bench.py's torchrun arm uses one process per GPU instead.
// Built only on the public C ABI of include/cudasw.h.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/cudasw.h"

struct cudasw_multi {
    std::vector<cudasw_t*> eng;
    std::vector<int> devices;
    std::vector<int64_t> last_first;  // slice boundaries of the last batch (n_gpus + 1)
    std::mutex mu;
    std::string error;
};

namespace {
thread_local std::string g_multi_create_error;

// contiguous slices with balanced cost; cost_i = q_i * w_i + a constant for the per-pair passes.
// Two levels so that the O(n) part runs on `parts` threads: per-block sums in parallel, a prefix over the (few)
// blocks, then each cut is refined inside its block.
std::vector<int64_t> balance(int64_t n, int parts, const int64_t* seq_off, const int64_t* ref_off) {
    std::vector<int64_t> cut((size_t)parts + 1, n);
    cut[0] = 0;
    if (parts <= 1 || n <= 0) return cut;
    auto cost = [&](int64_t i) {
        const double q = (double)std::max<int64_t>(seq_off[i + 1] - seq_off[i], 0);
        const double w = (double)std::max<int64_t>(ref_off[i + 1] - ref_off[i], 0);
        return q * w + 20000.0;
    };
    const int64_t B = 4096;
    const int64_t nb = (n + B - 1) / B;
    std::vector<double> blk((size_t)nb, 0.0);
    {
        std::vector<std::thread> th;
        const int nt = (int)std::min<int64_t>(parts, nb);
        for (int t = 0; t < nt; t++)
            th.emplace_back([&, t] {
                for (int64_t b = t; b < nb; b += nt) {
                    double s = 0.0;
                    const int64_t e = std::min(n, (b + 1) * B);
                    for (int64_t i = b * B; i < e; i++) s += cost(i);
                    blk[(size_t)b] = s;
                }
            });
        for (auto& x : th) x.join();
    }
    std::vector<double> pre((size_t)nb + 1, 0.0);
    for (int64_t b = 0; b < nb; b++) pre[(size_t)b + 1] = pre[(size_t)b] + blk[(size_t)b];
    for (int d = 1; d < parts; d++) {
        const double want = pre[(size_t)nb] * d / parts;
        int64_t b = (std::upper_bound(pre.begin(), pre.end(), want) - pre.begin()) - 1;
        b = std::min<int64_t>(std::max<int64_t>(b, 0), nb - 1);
        double acc = pre[(size_t)b];
        int64_t i = b * B;
        const int64_t e = std::min(n, (b + 1) * B);
        while (i < e && acc < want) acc += cost(i++);
        cut[(size_t)d] = std::max(i, cut[(size_t)d - 1]);
    }
    return cut;
}

struct SliceOut {
    int rc = CUDASW_OK;
    int64_t used = 0;
    std::string err;
};
}  // namespace

extern "C" int cudasw_multi_create(const cudasw_scoring* scoring, const int* devices, int n_devices, cudasw_multi_t** out) {
    if (!out) return CUDASW_EINVAL;
    *out = nullptr;
    std::vector<int> devs;
    if (n_devices <= 0 || !devices) {  // all visible devices
        const int nd = cudasw_device_count();
        for (int d = 0; d < nd; d++) devs.push_back(d);
    } else {
        devs.assign(devices, devices + n_devices);
    }
    if (devs.empty()) {
        g_multi_create_error = "no CUDA device (libcudasw has no CPU fallback)";
        return CUDASW_ECUDA;
    }
    cudasw_multi_t* m = new cudasw_multi();
    for (int d : devs) {
        cudasw_t* e = nullptr;
        const int rc = cudasw_create(scoring, d, &e);
        if (rc != CUDASW_OK) {
            g_multi_create_error = std::string("device ") + std::to_string(d) + ": " + cudasw_last_error(nullptr);
            for (auto* x : m->eng) cudasw_destroy(x);
            delete m;
            return rc;
        }
        m->eng.push_back(e);
        m->devices.push_back(d);
    }
    *out = m;
    return CUDASW_OK;
}

extern "C" void cudasw_multi_destroy(cudasw_multi_t* m) {
    if (!m) return;
    for (auto* e : m->eng) cudasw_destroy(e);
    delete m;
}

extern "C" int cudasw_multi_device_count(cudasw_multi_t* m) { return m ? (int)m->eng.size() : 0; }

extern "C" const char* cudasw_multi_last_error(cudasw_multi_t* m) {
    return m ? m->error.c_str() : g_multi_create_error.c_str();
}

extern "C" int cudasw_multi_get_stats(cudasw_multi_t* m, int idx, cudasw_stats* out, int64_t* first_pair, int64_t* n_pairs) {
    if (!m || idx < 0 || idx >= (int)m->eng.size()) return CUDASW_EINVAL;
    if (first_pair) *first_pair = m->last_first.size() > (size_t)idx + 1 ? m->last_first[(size_t)idx] : 0;
    if (n_pairs) *n_pairs = m->last_first.size() > (size_t)idx + 1 ? m->last_first[(size_t)idx + 1] - m->last_first[(size_t)idx] : 0;
    return out ? cudasw_get_stats(m->eng[(size_t)idx], out) : CUDASW_OK;
}

// Shared driver of both batch entry points: `call(d, first, cnt, rebased seq_off, out slice, cigar region, cap, &used)`.
template <typename Call>
static int run_slices(cudasw_multi_t* m, int64_t n, const int64_t* seq_off, const std::vector<int64_t>& cut,
                      cudasw_result* out, uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used, Call call) {
    const int parts = (int)m->eng.size();
    std::vector<SliceOut> so((size_t)parts);
    std::vector<int64_t> region((size_t)parts + 1, 0);
    // CIGAR words of GPU d go to its own region of the caller's blob (sized by its share of the query bases, which
    // bounds the words a pair can need), so no GPU waits for another and nothing is copied twice
    const int64_t tq = std::max<int64_t>(seq_off[n] - seq_off[0], 1);
    for (int d = 1; d <= parts; d++)
        region[(size_t)d] = d == parts ? cigar_cap : (int64_t)((double)cigar_cap * (double)(seq_off[cut[(size_t)d]] - seq_off[0]) / (double)tq);
    std::vector<std::thread> th;
    for (int d = 0; d < parts; d++) {
        th.emplace_back([&, d] {
            const int64_t first = cut[(size_t)d], cnt = cut[(size_t)d + 1] - first;
            if (cnt <= 0) return;
            const int64_t base = region[(size_t)d], cap = region[(size_t)d + 1] - base;
            so[(size_t)d].rc = call(d, first, cnt, out + first, cigar_blob ? cigar_blob + base : nullptr, cap, &so[(size_t)d].used);
            if (so[(size_t)d].rc != CUDASW_OK) so[(size_t)d].err = cudasw_last_error(m->eng[(size_t)d]);
            if (so[(size_t)d].rc == CUDASW_OK && base != 0) {
                cudasw_result* r = out + first;
                for (int64_t i = 0; i < cnt; i++)
                    if (r[i].cigar_len > 0) r[i].cigar_off += (int32_t)base;
            }
        });
    }
    for (auto& t : th) t.join();
    int rc = CUDASW_OK;
    int64_t used = 0, need = 0;
    for (int d = 0; d < parts; d++) {
        used += so[(size_t)d].used;
        const int64_t cnt = cut[(size_t)d + 1] - cut[(size_t)d];
        if (so[(size_t)d].rc == CUDASW_ENOMEM && so[(size_t)d].used > region[(size_t)d + 1] - region[(size_t)d] && cnt > 0) {
            // scale this GPU's need up to the whole blob so that a retry with *cigar_used words succeeds
            const double share = std::max(1e-9, (double)(seq_off[cut[(size_t)d + 1]] - seq_off[cut[(size_t)d]]) / (double)tq);
            need = std::max(need, (int64_t)((double)so[(size_t)d].used / share) + 64);
            if (rc == CUDASW_OK) rc = CUDASW_ENOMEM;
            m->error = "cigar buffer too small";
        } else if (so[(size_t)d].rc != CUDASW_OK && (rc == CUDASW_OK || rc == CUDASW_ENOMEM)) {
            rc = so[(size_t)d].rc;
            m->error = "device " + std::to_string(m->devices[(size_t)d]) + ": " + so[(size_t)d].err;
        }
    }
    if (cigar_used) *cigar_used = rc == CUDASW_ENOMEM && need > 0 ? need : used;
    return rc;
}

extern "C" int cudasw_multi_align_batch(cudasw_multi_t* m, int64_t n, const uint8_t* seq_blob, const int64_t* seq_off,
                                        const uint8_t* ref_blob, const int64_t* ref_off, cudasw_result* out,
                                        uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used) {
    if (!m) return CUDASW_EINVAL;
    std::lock_guard<std::mutex> lock(m->mu);
    m->error.clear();
    if (cigar_used) *cigar_used = 0;
    if (n < 0) { m->error = "n_pairs out of range"; return CUDASW_EINVAL; }
    if (n == 0) return CUDASW_OK;
    if (!seq_blob || !seq_off || !ref_blob || !ref_off || !out) { m->error = "null input"; return CUDASW_EINVAL; }
    if (seq_off[0] != 0 || ref_off[0] != 0) { m->error = "offsets must start at 0"; return CUDASW_EINVAL; }
    const std::vector<int64_t> cut = balance(n, (int)m->eng.size(), seq_off, ref_off);
    m->last_first = cut;
    return run_slices(m, n, seq_off, cut, out, cigar_blob, cigar_cap, cigar_used,
                      [&](int d, int64_t first, int64_t cnt, cudasw_result* o, uint32_t* cg, int64_t cap, int64_t* used) {
                          // the single-GPU entry point wants offsets from 0: rebase this slice's view
                          std::vector<int64_t> so((size_t)cnt + 1), ro((size_t)cnt + 1);
                          const int64_t sb = seq_off[first], rb = ref_off[first];
                          for (int64_t i = 0; i <= cnt; i++) { so[(size_t)i] = seq_off[first + i] - sb; ro[(size_t)i] = ref_off[first + i] - rb; }
                          return cudasw_align_batch(m->eng[(size_t)d], cnt, seq_blob + sb, so.data(), ref_blob + rb, ro.data(), o, cg, cap, used);
                      });
}

// Packed host inputs (cudasw_align_batch_packed) over all GPUs: a slice is a view of the same word arrays. Its offsets
// are rebased to the 32-base group its first base lives in, so each engine copies and expands only its own groups.
extern "C" int cudasw_multi_align_batch_packed(cudasw_multi_t* m, int64_t n, const uint32_t* seq_two_bit, const uint32_t* seq_nmask,
                                               const int64_t* seq_off, const uint32_t* ref_two_bit, const uint32_t* ref_nmask,
                                               const int64_t* ref_off, cudasw_result* out, uint32_t* cigar_blob,
                                               int64_t cigar_cap, int64_t* cigar_used) {
    if (!m) return CUDASW_EINVAL;
    std::lock_guard<std::mutex> lock(m->mu);
    m->error.clear();
    if (cigar_used) *cigar_used = 0;
    if (n < 0) { m->error = "n_pairs out of range"; return CUDASW_EINVAL; }
    if (n == 0) return CUDASW_OK;
    if (!seq_two_bit || !seq_off || !ref_two_bit || !ref_off || !out) { m->error = "null input"; return CUDASW_EINVAL; }
    if (seq_off[0] < 0 || ref_off[0] < 0) { m->error = "offsets must be non-negative"; return CUDASW_EINVAL; }
    const std::vector<int64_t> cut = balance(n, (int)m->eng.size(), seq_off, ref_off);
    m->last_first = cut;
    return run_slices(m, n, seq_off, cut, out, cigar_blob, cigar_cap, cigar_used,
                      [&](int d, int64_t first, int64_t cnt, cudasw_result* o, uint32_t* cg, int64_t cap, int64_t* used) {
                          std::vector<int64_t> so((size_t)cnt + 1), ro((size_t)cnt + 1);
                          const int64_t sg = seq_off[first] >> 5, rg = ref_off[first] >> 5;  // 32-base groups to skip
                          for (int64_t i = 0; i <= cnt; i++) { so[(size_t)i] = seq_off[first + i] - 32 * sg; ro[(size_t)i] = ref_off[first + i] - 32 * rg; }
                          return cudasw_align_batch_packed(m->eng[(size_t)d], cnt, seq_two_bit + 2 * sg, seq_nmask ? seq_nmask + sg : nullptr, so.data(),
                                                           ref_two_bit + 2 * rg, ref_nmask ? ref_nmask + rg : nullptr, ro.data(), o, cg, cap, used);
                      });
}

extern "C" int cudasw_multi_reference_load(cudasw_multi_t* m, int32_t n_contigs, const uint8_t* bases, const int64_t* contig_off) {
    if (!m) return CUDASW_EINVAL;
    std::lock_guard<std::mutex> lock(m->mu);
    m->error.clear();
    // every GPU keeps its own replica (a one-time broadcast, not on the timed path: SURVEY.md §8e)
    std::vector<int> rcs(m->eng.size(), CUDASW_OK);
    std::vector<std::thread> th;
    for (size_t d = 0; d < m->eng.size(); d++)
        th.emplace_back([&, d] { rcs[d] = cudasw_reference_load(m->eng[d], n_contigs, bases, contig_off); });
    for (auto& t : th) t.join();
    for (size_t d = 0; d < m->eng.size(); d++)
        if (rcs[d] != CUDASW_OK) {
            m->error = "device " + std::to_string(m->devices[d]) + ": " + cudasw_last_error(m->eng[d]);
            return rcs[d];
        }
    return CUDASW_OK;
}

extern "C" int cudasw_multi_align_batch_windows(cudasw_multi_t* m, int64_t n, const uint8_t* seq_blob, const int64_t* seq_off,
                                                const cudasw_window* win, const int64_t* win_off, cudasw_result* out,
                                                uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used) {
    if (!m) return CUDASW_EINVAL;
    std::lock_guard<std::mutex> lock(m->mu);
    m->error.clear();
    if (cigar_used) *cigar_used = 0;
    if (n < 0) { m->error = "n_pairs out of range"; return CUDASW_EINVAL; }
    if (n == 0) return CUDASW_OK;
    if (!seq_blob || !seq_off || !win || !win_off || !out) { m->error = "null input"; return CUDASW_EINVAL; }
    if (seq_off[0] != 0 || win_off[0] != 0) { m->error = "offsets must start at 0"; return CUDASW_EINVAL; }
    // window lengths from the descriptors (bounds are checked per pair by the engine)
    std::vector<int64_t> ro((size_t)n + 1, 0);
    for (int64_t p = 0; p < n; p++) {
        int64_t len = 0;
        for (int64_t s = win_off[p]; s < win_off[p + 1]; s++) len += std::max<int64_t>(win[s].end - win[s].start, 0);
        ro[(size_t)p + 1] = ro[(size_t)p] + len;
    }
    const std::vector<int64_t> cut = balance(n, (int)m->eng.size(), seq_off, ro.data());
    m->last_first = cut;
    return run_slices(m, n, seq_off, cut, out, cigar_blob, cigar_cap, cigar_used,
                      [&](int d, int64_t first, int64_t cnt, cudasw_result* o, uint32_t* cg, int64_t cap, int64_t* used) {
                          std::vector<int64_t> so((size_t)cnt + 1), wo((size_t)cnt + 1);
                          const int64_t sb = seq_off[first], wb = win_off[first];
                          for (int64_t i = 0; i <= cnt; i++) { so[(size_t)i] = seq_off[first + i] - sb; wo[(size_t)i] = win_off[first + i] - wb; }
                          return cudasw_align_batch_windows(m->eng[(size_t)d], cnt, seq_blob + sb, so.data(), win + wb, wo.data(), o, cg, cap, used);
                      });
}

// BreakpointHomology.calculate across all GPUs: equal contiguous slices of the breakpoint list (every breakpoint costs
// about the same two alignments), each against that GPU's replica of the reference; outputs land at the original
// ordinals. No exchange step.
extern "C" int cudasw_multi_breakpoint_homology(cudasw_multi_t* m, int64_t n, const cudasw_breakpoint* bp, const uint8_t* ins_blob,
                                                const int64_t* ins_off, int32_t max_breakend_length, int32_t margin,
                                                int32_t* local_hom, int32_t* remote_hom) {
    if (!m) return CUDASW_EINVAL;
    std::lock_guard<std::mutex> lock(m->mu);
    m->error.clear();
    if (n < 0 || (n > 0 && (!bp || !local_hom || !remote_hom))) { m->error = "bad argument"; return CUDASW_EINVAL; }
    if (n == 0) return CUDASW_OK;
    const int parts = (int)m->eng.size();
    std::vector<int64_t> cut((size_t)parts + 1);
    for (int d = 0; d <= parts; d++) cut[(size_t)d] = n * d / parts;
    m->last_first = cut;
    std::vector<int> rcs((size_t)parts, CUDASW_OK);
    std::vector<std::thread> th;
    for (int d = 0; d < parts; d++) {
        th.emplace_back([&, d] {
            const int64_t first = cut[(size_t)d], cnt = cut[(size_t)d + 1] - first;
            if (cnt == 0) return;
            std::vector<int64_t> io;
            if (ins_off) {
                io.resize((size_t)cnt + 1);
                for (int64_t i = 0; i <= cnt; i++) io[(size_t)i] = ins_off[first + i] - ins_off[first];
            }
            rcs[(size_t)d] = cudasw_breakpoint_homology(m->eng[(size_t)d], cnt, bp + first, ins_off ? ins_blob + ins_off[first] : nullptr,
                                                        ins_off ? io.data() : nullptr, max_breakend_length, margin, local_hom + first,
                                                        remote_hom + first);
        });
    }
    for (auto& t : th) t.join();
    for (int d = 0; d < parts; d++)
        if (rcs[(size_t)d] != CUDASW_OK) {
            m->error = "device " + std::to_string(m->devices[(size_t)d]) + ": " + cudasw_last_error(m->eng[(size_t)d]);
            return rcs[(size_t)d];
        }
    return CUDASW_OK;
}
