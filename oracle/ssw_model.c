/*
 * ssw_model.c - CPU ORACLE (restatement). TEST / BENCH INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this; nothing under gridss_b200/ does. It is the checker, never the product.
 *
 * What it restates: the semantics of one reference call
 *   Aligner.align_smith_waterman(seq, ref)            alignment/Aligner.java:10
 *   -> SswJniAligner.do_align_smith_waterman           alignment/SswJniAligner.java:42-61
 *   -> ssw.Aligner.align -> Java_ssw_Aligner_align     repo/ssw/ssw/1.0/ssw-1.0.jar, libsswjni.so 0xcac
 *   -> ssw_init / ssw_align                            libsswjni.so 0x4748 / 0x48cb
 * The arithmetic lives in a third-party dependency that is NOT in /root/reference in source form:
 * ssw:ssw:1.0 (pom.xml:124-128), vendored as a class-only jar + prebuilt libsswjni.so built from a
 * 2013-era ssw.c of mengyao/Complete-Striped-Smith-Waterman-Library. This file restates the
 * published SSW algorithm (Zhao et al. 2013: striped SW score pass, reverse pass, banded traceback)
 * as plain scalar loops, following SURVEY.md §8a rows A2-A5 (`ssw.c:NNN` = DWARF line numbers of
 * the file the binary was built from).
 *
 * Parity status: PINNED. tests/test_oracle.py checks this model against (a) the reference's own
 * golden vectors (SmithWatermanAlignerTest.java:13-34, AlignmentHelperTest.java:11-41 and the
 * indirect pins) and (b) the reference's shipped binary itself, driven by oracle/ref_driver.c, on
 * large randomized/adversarial sets.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int32_t score, ref_begin, ref_end, read_begin, read_end, cigar_off, cigar_len, status;
} sw_result;

typedef struct {
    int match, mismatch, ambiguous, go, ge;
} scoring_t;

#define ST_OK 0
#define ST_EINVAL (-1)
#define ST_ENOMEM (-2)
#define ST_ETOOLARGE (-4)
#define ST_EUNDEF (-5)

/* SswJniAligner.clean (SswJniAligner.java:71-79): upper-case; everything but ACGT becomes N. */
static inline uint8_t code_of(uint8_t c) {
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return 4;
    }
}

/* SswJniAligner.createMatrix (SswJniAligner.java:19-33) restricted to the cleaned alphabet. */
static inline int sub(const scoring_t* s, int a, int b) {
    if (a == b) return s->match;
    if (a < 4 && b < 4) return s->mismatch;
    return s->ambiguous;
}

/*
 * Score pass (phases 1 and 2 of ssw_align; sw_sse2_byte / sw_sse2_word, ssw.c:135-546).
 * Reference base is the outer loop. Recurrences of SURVEY.md §8a A4-rec with E/F floored at 0
 * exactly as the unsigned-saturating SSE2 code does. Tie-breaks:
 *   end_ref  = first column whose column maximum is strictly greater than all earlier columns
 *              (ssw.c:283-289);   end_read = smallest row holding that maximum in that column
 *              (ssw.c:300-306).
 * `terminate` > 0 stops after the first column whose running maximum equals it (ssw.c:296).
 * rstep = +1 walks ref[0..], -1 walks ref[rstart], ref[rstart-1], ... (the reverse pass);
 * qstep likewise for the read, so no reversed copies are needed.
 */
static void score_pass(const uint8_t* read, int q, int qstart, int qstep, const uint8_t* ref, int w,
                       int rstart, int rstep, const scoring_t* sc, int terminate, int32_t* H,
                       int32_t* E, int* best_out, int* end_col, int* end_row) {
    int best = 0, bcol = -1, brow = 0;
    for (int i = 0; i < q; i++) H[i] = E[i] = 0;
    for (int j = 0; j < w; j++) {
        int rb = ref[rstart + j * rstep];
        int f = 0, diag = 0, colmax = 0, colrow = 0;
        for (int i = 0; i < q; i++) {
            int h = diag + sub(sc, read[qstart + i * qstep], rb);
            int e = E[i];
            if (h < e) h = e;
            if (h < f) h = f;
            if (h < 0) h = 0;
            diag = H[i];
            H[i] = h;
            int t = h - sc->go;
            e -= sc->ge;
            if (e < t) e = t;
            if (e < 0) e = 0;
            E[i] = e;
            f -= sc->ge;
            if (f < t) f = t;
            if (f < 0) f = 0;
            if (h > colmax) { colmax = h; colrow = i; }
        }
        if (colmax > best) {
            best = colmax;
            bcol = j;
            brow = colrow;
            if (best == terminate) break;
        }
    }
    *best_out = best;
    *end_col = bcol;
    *end_row = brow;
}

#define POISON 0x7f

typedef struct {
    int32_t *h_b, *e_b, *h_c;
    uint8_t* hb_init; /* 1 = entry has been written at least once (e_b/h_b/h_c share index space) */
    uint8_t *eb_init, *hc_init;
    int64_t s1;
    int8_t* dir;
    int64_t dcap;
    int32_t *H, *E;
    int64_t hcap;
} scratch_t;

static int grow_rows(scratch_t* s, int64_t need) {
    if (need < s->s1) return 0;
    int64_t n = need + 64;
    s->h_b = (int32_t*)realloc(s->h_b, n * sizeof(int32_t));
    s->e_b = (int32_t*)realloc(s->e_b, n * sizeof(int32_t));
    s->h_c = (int32_t*)realloc(s->h_c, n * sizeof(int32_t));
    s->hb_init = (uint8_t*)realloc(s->hb_init, n);
    s->eb_init = (uint8_t*)realloc(s->eb_init, n);
    s->hc_init = (uint8_t*)realloc(s->hc_init, n);
    if (!s->h_b || !s->e_b || !s->h_c || !s->hb_init || !s->eb_init || !s->hc_init) return -1;
    s->s1 = n;
    return 0;
}

/*
 * Phase 3: banded_sw (ssw.c:557-734), emulated buffer for buffer because the
 * emitted CIGAR depends on its quirks: rolling h_b/e_b rows addressed in band coordinates with
 * e_b updated in place, the `edge` zeroing that uses an absolute column as a band index
 * (ssw.c:596), the never-reset running `max`, band doubling while max < score (ssw.c:630-631), the
 * flat 3-bytes-per-cell direction array whose stale contents from narrower bands stay visible, and
 * a traceback that runs `while (i > 0)` from the bottom-right corner and assumes the final cell is
 * an M (ssw.c:641-712). Reads of memory the reference never initialised are reported as
 * ST_EUNDEF (the reference's answer then depends on heap garbage).
 * Returns number of ops written (M/I/D only), or <0 status.
 */
static int banded_traceback(const uint8_t* ref, const uint8_t* read, int refLen, int readLen,
                            int score, const scoring_t* sc, int band, scratch_t* s, uint32_t* out,
                            int64_t cap) {
    int32_t max = 0;
    int width, width_d;
    int undef = 0;
    /* per-call fresh buffers in the reference (malloc): nothing is initialised yet */
    memset(s->hb_init, 0, (size_t)s->s1);
    memset(s->eb_init, 0, (size_t)s->s1);
    memset(s->hc_init, 0, (size_t)s->s1);
    int64_t dir_written_hi = 0; /* bytes [0, hi) may hold data; beyond is never written */
    do {
        width = band * 2 + 3;
        width_d = band * 2 + 1;
        int64_t old = s->s1;
        if (grow_rows(s, (int64_t)width + 1)) return ST_ENOMEM;
        if (s->s1 > old) {
            memset(s->hb_init + old, 0, (size_t)(s->s1 - old));
            memset(s->eb_init + old, 0, (size_t)(s->s1 - old));
            memset(s->hc_init + old, 0, (size_t)(s->s1 - old));
        }
        int64_t need = (int64_t)width_d * readLen * 3;
        if (need >= (int64_t)1 << 31) return ST_ETOOLARGE; /* reference: int32 overflow / exit(1) */
        if (need + 8 > s->dcap) {
            int64_t n = need * 2 + 1024;
            int8_t* nd = (int8_t*)realloc(s->dir, (size_t)n);
            if (!nd) return ST_ENOMEM;
            s->dir = nd;
            s->dcap = n;
        }
        if (need > dir_written_hi) {
            memset(s->dir + dir_written_hi, POISON, (size_t)(need - dir_written_hi));
            dir_written_hi = need;
        }
        for (int j = 1; j < width - 1; j++) { s->h_b[j] = 0; s->hb_init[j] = 1; }
        for (int i = 0; i < readLen; i++) {
            int beg = i - band > 0 ? i - band : 0;
            int end = i + band < refLen - 1 ? i + band : refLen - 1;
            int edge = end + 1 < width - 1 ? end + 1 : width - 1;
            int32_t f = 0;
            s->h_b[0] = s->e_b[0] = s->h_b[edge] = s->e_b[edge] = s->h_c[0] = 0;
            s->hb_init[0] = s->eb_init[0] = s->hb_init[edge] = s->eb_init[edge] = s->hc_init[0] = 1;
            int8_t* dl = s->dir + (int64_t)width_d * i * 3;
            int xi = i - band > 0 ? i - band : 0;
            int xp = i - 1 - band > 0 ? i - 1 - band : 0;
            int u = 0;
            int rdb = read[i];
            for (int j = beg; j <= end; j++) {
                u = j - xi + 1;
                int e = j - xp + 1, b = u - 1, d = e - 1;
                int de = (j - xi) * 3, df = de + 1, dh = de + 2;
                int32_t t1, t2;
                if (i == 0) {
                    t1 = -sc->go;
                    t2 = -sc->ge;
                } else {
                    if (!s->hb_init[e] || !s->eb_init[e]) undef = 1;
                    t1 = s->h_b[e] - sc->go;
                    t2 = s->e_b[e] - sc->ge;
                }
                s->e_b[u] = t1 > t2 ? t1 : t2;
                s->eb_init[u] = 1;
                dl[de] = t1 > t2 ? 3 : 2;
                if (!s->hc_init[b]) undef = 1;
                t1 = s->h_c[b] - sc->go;
                t2 = f - sc->ge;
                f = t1 > t2 ? t1 : t2;
                dl[df] = t1 > t2 ? 5 : 4;
                int32_t e1 = s->e_b[u] > 0 ? s->e_b[u] : 0;
                int32_t f1 = f > 0 ? f : 0;
                t1 = e1 > f1 ? e1 : f1;
                if (!s->hb_init[d]) undef = 1;
                t2 = s->h_b[d] + sub(sc, ref[j], rdb);
                s->h_c[u] = t1 > t2 ? t1 : t2;
                s->hc_init[u] = 1;
                if (s->h_c[u] > max) max = s->h_c[u];
                if (t1 <= t2) dl[dh] = 1;
                else dl[dh] = e1 > f1 ? dl[de] : dl[df];
            }
            for (int j = 1; j <= u; j++) { s->h_b[j] = s->h_c[j]; s->hb_init[j] = s->hc_init[j]; }
        }
        band *= 2;
    } while (max < score);
    band /= 2;
    if (undef) return ST_EUNDEF;

    /* traceback */
    int i = readLen - 1, j = refLen - 1;
    int e = 0, l = 0, state = 2;
    char op = 'M', prev_op = 'M';
    int64_t total = (int64_t)width_d * readLen * 3;
    while (i > 0) {
        int xi = i - band > 0 ? i - band : 0;
        int64_t addr = (int64_t)width_d * i * 3 + (int64_t)(j - xi) * 3 + state;
        if (addr < 0 || addr >= total) return ST_EUNDEF;
        int dv = s->dir[addr];
        switch (dv) {
            case 1: --i; --j; state = 2; op = 'M'; break;
            case 2: --i; state = 0; op = 'I'; break;
            case 3: --i; state = 2; op = 'I'; break;
            case 4: --j; state = 1; op = 'D'; break;
            case 5: --j; state = 2; op = 'D'; break;
            default: return ST_EUNDEF; /* poison, or the reference's "Trace back error" */
        }
        if (op == prev_op) ++e;
        else {
            ++l;
            if (l > cap) return ST_ENOMEM;
            out[l - 1] = ((uint32_t)e << 4) | (prev_op == 'M' ? 0u : prev_op == 'I' ? 1u : 2u);
            prev_op = op;
            e = 1;
        }
    }
    if (op == 'M') {
        ++l;
        if (l > cap) return ST_ENOMEM;
        out[l - 1] = ((uint32_t)(e + 1) << 4) | 0u;
    } else {
        l += 2;
        if (l > cap) return ST_ENOMEM;
        out[l - 2] = ((uint32_t)e << 4) | (op == 'I' ? 1u : 2u);
        out[l - 1] = (1u << 4) | 0u;
    }
    for (int a = 0, b = l - 1; a < b; a++, b--) {
        uint32_t t = out[a];
        out[a] = out[b];
        out[b] = t;
    }
    return l;
}

/* One Aligner.align_smith_waterman call. Returns ops written or -(status). */
static int model_align_one(const uint8_t* seq, int64_t q64, const uint8_t* refa, int64_t w64,
                           const scoring_t* sc, scratch_t* s, sw_result* r, uint32_t* cig,
                           int64_t cap) {
    memset(r, 0, sizeof(*r));
    if (q64 <= 0 || w64 <= 0) { r->status = ST_EINVAL; return 0; }
    if (q64 > 32000 || w64 > (1 << 30)) { r->status = ST_ETOOLARGE; return 0; }
    int q = (int)q64, w = (int)w64;
    uint8_t* read = (uint8_t*)malloc((size_t)q);
    uint8_t* ref = (uint8_t*)malloc((size_t)w);
    if (q > s->hcap) {
        s->H = (int32_t*)realloc(s->H, (size_t)(q + 64) * sizeof(int32_t));
        s->E = (int32_t*)realloc(s->E, (size_t)(q + 64) * sizeof(int32_t));
        s->hcap = q + 64;
    }
    for (int i = 0; i < q; i++) read[i] = code_of(seq[i]);
    for (int i = 0; i < w; i++) ref[i] = code_of(refa[i]);

    int score, ref_end, read_end;
    score_pass(read, q, 0, 1, ref, w, 0, 1, sc, -1, s->H, s->E, &score, &ref_end, &read_end);
    int ref_begin = -1, read_begin = 0;
    int nops = 0;
    if (score > 0) {
        int sc2, rcol, rrow;
        score_pass(read, read_end + 1, read_end, -1, ref, ref_end + 1, ref_end, -1, sc, score,
                   s->H, s->E, &sc2, &rcol, &rrow);
        ref_begin = ref_end - rcol;
        read_begin = read_end - rrow;
        int refSpan = ref_end - ref_begin + 1, readSpan = read_end - read_begin + 1;
        int band = abs(refSpan - readSpan) + 1;
        uint32_t* body = cig + (read_begin != 0);
        int n = banded_traceback(ref + ref_begin, read + read_begin, refSpan, readSpan, score, sc,
                                 band, s, body, cap - 2);
        if (n < 0) {
            r->status = n;
        } else {
            nops = n;
        }
    } else {
        /* no positive cell -> ref_begin1 = -1, cigar "1M" (+ (q-1)S) */
        if (cap >= 2) cig[0] = (1u << 4) | 0u;
        nops = 1;
        ref_end = -1;
        read_end = 0;
    }
    r->score = score;
    r->ref_begin = ref_begin;
    r->ref_end = ref_end;
    r->read_begin = read_begin;
    r->read_end = read_end;
    if (r->status == ST_OK) {
        /* SswJniAligner.java:52-59 */
        if (read_begin != 0) { cig[0] = ((uint32_t)read_begin << 4) | 4u; nops++; }
        int tail = q - read_end - 1;
        if (tail != 0) { cig[nops++] = ((uint32_t)tail << 4) | 4u; }
        r->cigar_len = nops;
    }
    free(read);
    free(ref);
    return nops;
}

typedef struct {
    int64_t lo, hi;
    const uint8_t *seq_blob, *ref_blob;
    const int64_t *seq_off, *ref_off;
    scoring_t sc;
    sw_result* out;
    uint32_t* cig;
    int64_t cig_stride;
} job_t;

static void* worker(void* p) {
    job_t* j = (job_t*)p;
    scratch_t s;
    memset(&s, 0, sizeof(s));
    grow_rows(&s, 64);
    for (int64_t i = j->lo; i < j->hi; i++) {
        model_align_one(j->seq_blob + j->seq_off[i], j->seq_off[i + 1] - j->seq_off[i],
                        j->ref_blob + j->ref_off[i], j->ref_off[i + 1] - j->ref_off[i], &j->sc, &s,
                        &j->out[i], j->cig + i * j->cig_stride, j->cig_stride);
        j->out[i].cigar_off = (int32_t)(i * j->cig_stride);
    }
    free(s.h_b); free(s.e_b); free(s.h_c);
    free(s.hb_init); free(s.eb_init); free(s.hc_init);
    free(s.dir); free(s.H); free(s.E);
    return NULL;
}

/* Same batch interface as ref_driver.c:ref_align_batch. */
int model_align_batch(int64_t n, const uint8_t* seq_blob, const int64_t* seq_off,
                      const uint8_t* ref_blob, const int64_t* ref_off, int match, int mismatch,
                      int ambiguous, int go, int ge, sw_result* out, uint32_t* cig,
                      int64_t cig_stride, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    job_t jobs[256];
    for (int t = 0; t < threads; t++) {
        job_t* j = &jobs[t];
        j->lo = n * t / threads;
        j->hi = n * (t + 1) / threads;
        j->seq_blob = seq_blob; j->seq_off = seq_off;
        j->ref_blob = ref_blob; j->ref_off = ref_off;
        j->sc.match = match; j->sc.mismatch = mismatch; j->sc.ambiguous = ambiguous;
        j->sc.go = go; j->sc.ge = ge;
        j->out = out; j->cig = cig; j->cig_stride = cig_stride;
        if (threads == 1) worker(j);
        else pthread_create(&th[t], NULL, worker, j);
    }
    if (threads > 1)
        for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
    return 0;
}
