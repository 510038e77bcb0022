/*
 * synth.c - synthetic workload generator for the parity tests and bench.py.
 *
 * Counter-based: every pair is generated from splitmix64 keyed by (seed, global pair index), so any
 * shard [first, first+n) of any config is reproducible without materialising the rest (this is how
 * ranks of a multi-GPU run generate their own slice).
 *
 * Configs (BASELINE.json `configs`):
 *   1  C1  150-bp soft-clip queries vs 2 kb windows
 *   2  C2  300-bp breakend assembly contigs vs 4 kb windows
 *   3  C3  mixed 50 bp-2 kb contigs vs 1-10 kb windows (log-uniform)
 *   4  C4  WGS-scale mix: 70 % C1 shape / 30 % C2 shape
 *   5  C5  10 kb contigs vs 20 kb windows, 5 % divergence, indels up to 50 bp
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { uint64_t s; } rng_t;
static inline uint64_t mix(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
static inline uint64_t next(rng_t* r) { r->s += 0x9e3779b97f4a7c15ULL; return mix(r->s); }
static inline rng_t seed_pair(uint64_t seed, uint64_t pair, uint64_t stream) {
    rng_t r; r.s = mix(seed * 0x100000001b3ULL + stream) ^ mix(pair + 0x632be59bd9b4e019ULL); return r;
}
static inline uint32_t below(rng_t* r, uint32_t n) { return (uint32_t)((next(r) >> 32) * (uint64_t)n >> 32); }
static inline double unif(rng_t* r) { return (double)(next(r) >> 11) * (1.0 / 9007199254740992.0); }
static const char ACGT[4] = {'A', 'C', 'G', 'T'};

typedef struct { int cfg; int q, w; } shape_t;

static shape_t shape_of(int cfg, uint64_t seed, uint64_t pair) {
    shape_t s; s.cfg = cfg;
    rng_t r = seed_pair(seed, pair, 1);
    switch (cfg) {
        case 1: s.q = 150; s.w = 2000; break;
        case 2: s.q = 300; s.w = 4000; break;
        case 3: {
            double lq = log(50.0) + unif(&r) * (log(2000.0) - log(50.0));
            double lw = log(1000.0) + unif(&r) * (log(10000.0) - log(1000.0));
            s.q = (int)exp(lq); s.w = (int)exp(lw);
            if (s.q < 50) s.q = 50; if (s.q > 2000) s.q = 2000;
            if (s.w < 1000) s.w = 1000; if (s.w > 10000) s.w = 10000;
            if (s.w < s.q) s.w = s.q;
            break;
        }
        case 4:
            if (below(&r, 100) < 70) { s.cfg = 1; s.q = 150; s.w = 2000; }
            else { s.cfg = 2; s.q = 300; s.w = 4000; }
            break;
        case 5: s.q = 10000; s.w = 20000; break;
        default: s.q = 150; s.w = 2000; break;
    }
    return s;
}

/* copy window[s..] into out[0..len) applying substitutions / short indels; returns bases written */
static int mutate_copy(rng_t* r, const uint8_t* win, int w, int start, uint8_t* out, int len,
                       double p_sub, double p_ins, double p_del, int max_indel) {
    int o = 0, i = start;
    while (o < len && i < w) {
        double u = unif(r);
        if (u < p_sub) {
            out[o++] = ACGT[(below(r, 3) + 1 + (win[i] == 'A' ? 0 : win[i] == 'C' ? 1 : win[i] == 'G' ? 2 : 3)) & 3];
            i++;
        } else if (u < p_sub + p_ins) {
            int k = 1 + below(r, max_indel);
            while (k-- > 0 && o < len) out[o++] = ACGT[below(r, 4)];
        } else if (u < p_sub + p_ins + p_del) {
            i += 1 + below(r, max_indel);
        } else {
            out[o++] = win[i++];
        }
    }
    while (o < len) out[o++] = ACGT[below(r, 4)];
    return o;
}

static void gen_pair(int cfg, uint64_t seed, uint64_t pair, uint8_t* seq, uint8_t* ref) {
    shape_t sh = shape_of(cfg, seed, pair);
    rng_t r = seed_pair(seed, pair, 2);
    int q = sh.q, w = sh.w;
    /* window: iid uniform ACGT, 32 bases per draw */
    for (int i = 0; i < w; i += 32) {
        uint64_t v = next(&r);
        int m = w - i < 32 ? w - i : 32;
        for (int k = 0; k < m; k++) { ref[i + k] = ACGT[v & 3]; v >>= 2; }
    }
    switch (sh.cfg) {
        case 1: case 3: {
            int s = below(&r, (uint32_t)(w - q + 1));
            mutate_copy(&r, ref, w, s, seq, q, 0.02, 0.0025, 0.0025, 6);
            if (below(&r, 2)) { /* soft clip: random tail */
                int k = 20 + below(&r, 81);
                if (k > q - 10) k = q / 2;
                for (int i = q - k; i < q; i++) seq[i] = ACGT[below(&r, 4)];
            }
            break;
        }
        case 2: {
            int anchor = 100 + below(&r, 101);
            int s = below(&r, (uint32_t)(w - anchor - 60));
            if (below(&r, 10) == 0) { /* one 10-50 bp indel inside the anchor */
                int k = 10 + below(&r, 41);
                int pos = 20 + below(&r, (uint32_t)(anchor - 40));
                int o = mutate_copy(&r, ref, w, s, seq, pos, 0.01, 0, 0, 1);
                if (below(&r, 2)) { /* insertion in the contig */
                    for (int i = 0; i < k && o < anchor; i++) seq[o++] = ACGT[below(&r, 4)];
                    if (o < anchor) mutate_copy(&r, ref, w, s + pos, seq + o, anchor - o, 0.01, 0, 0, 1);
                } else {            /* deletion from the contig */
                    mutate_copy(&r, ref, w, s + pos + k, seq + o, anchor - o, 0.01, 0, 0, 1);
                }
            } else {
                mutate_copy(&r, ref, w, s, seq, anchor, 0.01, 0, 0, 1);
            }
            for (int i = anchor; i < q; i++) seq[i] = ACGT[below(&r, 4)];
            if (below(&r, 2)) { /* breakend on the other side: anchor at the contig end */
                for (int i = 0; i < q / 2; i++) { uint8_t t = seq[i]; seq[i] = seq[q - 1 - i]; seq[q - 1 - i] = t; }
                for (int i = 0; i < w / 2; i++) { uint8_t t = ref[i]; ref[i] = ref[w - 1 - i]; ref[w - 1 - i] = t; }
            }
            break;
        }
        case 5: {
            int s = below(&r, (uint32_t)(w - q - 2000));
            mutate_copy(&r, ref, w, s, seq, q, 0.04, 0.0005, 0.0005, 50);
            break;
        }
    }
    /* 0.1 % of query bases -> N ; 1 % of windows lower-cased */
    int nn = (int)(q * 0.001 + unif(&r));
    for (int k = 0; k < nn; k++) seq[below(&r, (uint32_t)q)] = 'N';
    if (below(&r, 100) == 0)
        for (int i = 0; i < w; i++) ref[i] |= 0x20;
}

void synth_lengths(int cfg, uint64_t seed, int64_t first, int64_t n, int64_t* seq_off, int64_t* ref_off) {
    seq_off[0] = ref_off[0] = 0;
    for (int64_t i = 0; i < n; i++) {
        shape_t s = shape_of(cfg, seed, (uint64_t)(first + i));
        seq_off[i + 1] = seq_off[i] + s.q;
        ref_off[i + 1] = ref_off[i] + s.w;
    }
}

typedef struct {
    int cfg; uint64_t seed; int64_t first, lo, hi;
    uint8_t *seq_blob, *ref_blob; const int64_t *seq_off, *ref_off;
} job_t;
static void* worker(void* p) {
    job_t* j = (job_t*)p;
    for (int64_t i = j->lo; i < j->hi; i++)
        gen_pair(j->cfg, j->seed, (uint64_t)(j->first + i), j->seq_blob + j->seq_off[i], j->ref_blob + j->ref_off[i]);
    return NULL;
}
void synth_fill(int cfg, uint64_t seed, int64_t first, int64_t n, uint8_t* seq_blob,
                const int64_t* seq_off, uint8_t* ref_blob, const int64_t* ref_off, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 128) threads = 128;
    pthread_t th[128]; job_t jobs[128];
    for (int t = 0; t < threads; t++) {
        job_t* j = &jobs[t];
        j->cfg = cfg; j->seed = seed; j->first = first;
        j->lo = n * t / threads; j->hi = n * (t + 1) / threads;
        j->seq_blob = seq_blob; j->ref_blob = ref_blob; j->seq_off = seq_off; j->ref_off = ref_off;
        if (threads == 1) worker(j); else pthread_create(&th[t], NULL, worker, j);
    }
    if (threads > 1) for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
}
