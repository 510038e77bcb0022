/*
 * ref_driver.c - TEST / BENCH INFRASTRUCTURE ONLY (never linked or loaded by gridss_b200/).
 *
 * Drives the reference's own shipped native aligner (oracle/_ref/libsswjni.so, staged unmodified by
 * oracle/make_ref.sh from /root/reference/src/main/resources/libsswjni.so) with exactly the call
 * sequence the reference's Java side performs for one Aligner.align_smith_waterman call:
 *
 *   SswJniAligner.clean()                (SswJniAligner.java:71-79)  upper-case, non-ACGT -> 'N'
 *   ssw.Aligner.align()                  (repo/ssw/ssw/1.0/ssw-1.0.jar)
 *        symbols renumbered by first appearance (read, then ref); n x n int8 matrix
 *   Java_ssw_Aligner_align               (libsswjni.so 0xcac)
 *        ssw_init(read, q, mat, n, 2); ssw_align(p, ref, w, go, ge, flag=1, 0, 0, max(15, q/2))
 *   SswJniAligner.do_align_smith_waterman (SswJniAligner.java:52-60)  soft-clip wrap, start = ref_begin1
 *
 * i.e. SswJniAligner minus the JVM. Results use the same packed layout as include/cudasw.h.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int32_t score, ref_begin, ref_end, read_begin, read_end, cigar_off, cigar_len, status;
} sw_result;

/* s_align as laid out in libsswjni.so (ctor sig (SSIIIIILjava/lang/String;)V) */
typedef struct {
    uint16_t score1, score2;
    int32_t ref_begin1, ref_end1, read_begin1, read_end1, ref_end2;
    uint32_t* cigar;
    int32_t cigarLen;
} s_align;

typedef void* (*ssw_init_fn)(const int8_t*, int32_t, const int8_t*, int32_t, int8_t);
typedef s_align* (*ssw_align_fn)(const void*, const int8_t*, int32_t, uint8_t, uint8_t, uint8_t,
                                 uint16_t, int32_t, int32_t);
typedef void (*init_destroy_fn)(void*);
typedef void (*align_destroy_fn)(s_align*);

static void* g_lib;
static ssw_init_fn p_init;
static ssw_align_fn p_align;
static init_destroy_fn p_init_destroy;
static align_destroy_fn p_align_destroy;

int ref_load(const char* path) {
    if (g_lib) return 0;
    g_lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!g_lib) {
        fprintf(stderr, "ref_driver: dlopen(%s): %s\n", path, dlerror());
        return -1;
    }
    p_init = (ssw_init_fn)dlsym(g_lib, "ssw_init");
    p_align = (ssw_align_fn)dlsym(g_lib, "ssw_align");
    p_init_destroy = (init_destroy_fn)dlsym(g_lib, "init_destroy");
    p_align_destroy = (align_destroy_fn)dlsym(g_lib, "align_destroy");
    if (!p_init || !p_align || !p_init_destroy || !p_align_destroy) return -2;
    return 0;
}

static inline uint8_t clean_base(uint8_t c) {
    if (c >= 'a' && c <= 'z') c -= 32;
    return (c == 'A' || c == 'C' || c == 'G' || c == 'T') ? c : 'N';
}

/* One pair. cig must hold cap words; returns number of ops or -(needed) if too small. */
static int align_one(const uint8_t* seq, int64_t q, const uint8_t* ref, int64_t w, int match,
                     int mismatch, int ambiguous, int go, int ge, sw_result* r, uint32_t* cig,
                     int64_t cap) {
    memset(r, 0, sizeof(*r));
    if (q <= 0 || w <= 0) {
        r->status = -1;
        return 0;
    }
    int8_t* rd = (int8_t*)malloc((size_t)q);
    int8_t* rf = (int8_t*)malloc((size_t)w);
    int table[256];
    uint8_t sym[8];
    int n = 0;
    for (int i = 0; i < 256; i++) table[i] = -1;
    for (int64_t i = 0; i < q; i++) {
        uint8_t c = clean_base(seq[i]);
        if (table[c] < 0) { table[c] = n; sym[n++] = c; }
        rd[i] = (int8_t)table[c];
    }
    for (int64_t i = 0; i < w; i++) {
        uint8_t c = clean_base(ref[i]);
        if (table[c] < 0) { table[c] = n; sym[n++] = c; }
        rf[i] = (int8_t)table[c];
    }
    int8_t mat[64];
    for (int a = 0; a < n; a++)
        for (int b = 0; b < n; b++) {
            int s;
            if (sym[a] == sym[b]) s = match;
            else if (sym[a] != 'N' && sym[b] != 'N') s = mismatch;
            else s = ambiguous;
            mat[a * n + b] = (int8_t)s;
        }
    int32_t mask = (int32_t)(q / 2);
    if (mask < 15) mask = 15;
    void* prof = p_init(rd, (int32_t)q, mat, n, 2);
    s_align* a = p_align(prof, rf, (int32_t)w, (uint8_t)go, (uint8_t)ge, 1, 0, 0, mask);
    int nops = 0;
    if (!a) {
        r->status = -3;
    } else {
        r->score = a->score1;
        r->ref_begin = a->ref_begin1;
        r->ref_end = a->ref_end1;
        r->read_begin = a->read_begin1;
        r->read_end = a->read_end1;
        int lead = a->read_begin1 != 0;
        int64_t tail = q - a->read_end1 - 1;
        nops = a->cigarLen + lead + (tail != 0);
        if (nops <= cap) {
            int k = 0;
            if (lead) cig[k++] = ((uint32_t)a->read_begin1 << 4) | 4u;
            for (int i = 0; i < a->cigarLen; i++) cig[k++] = a->cigar[i];
            if (tail != 0) cig[k++] = ((uint32_t)tail << 4) | 4u;
            r->cigar_len = nops;
        } else {
            r->status = -2;
            nops = -nops;
        }
        p_align_destroy(a);
    }
    p_init_destroy(prof);
    free(rd);
    free(rf);
    return nops;
}

typedef struct {
    int64_t lo, hi;
    const uint8_t *seq_blob, *ref_blob;
    const int64_t *seq_off, *ref_off;
    int match, mismatch, ambiguous, go, ge;
    sw_result* out;
    uint32_t* cig;       /* per-pair slot of cig_stride words at cig + i*cig_stride */
    int64_t cig_stride;
} job_t;

static void* worker(void* p) {
    job_t* j = (job_t*)p;
    for (int64_t i = j->lo; i < j->hi; i++) {
        align_one(j->seq_blob + j->seq_off[i], j->seq_off[i + 1] - j->seq_off[i],
                  j->ref_blob + j->ref_off[i], j->ref_off[i + 1] - j->ref_off[i], j->match,
                  j->mismatch, j->ambiguous, j->go, j->ge, &j->out[i], j->cig + i * j->cig_stride,
                  j->cig_stride);
        j->out[i].cigar_off = (int32_t)(i * j->cig_stride);
    }
    return NULL;
}

/* Batch over `threads` host threads (contiguous slices). CIGARs go to fixed-stride slots:
 * pair i's ops are cig[i*cig_stride ...]. */
int ref_align_batch(int64_t n, const uint8_t* seq_blob, const int64_t* seq_off,
                    const uint8_t* ref_blob, const int64_t* ref_off, int match, int mismatch,
                    int ambiguous, int go, int ge, sw_result* out, uint32_t* cig,
                    int64_t cig_stride, int threads) {
    if (!g_lib) return -1;
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
        j->match = match; j->mismatch = mismatch; j->ambiguous = ambiguous;
        j->go = go; j->ge = ge;
        j->out = out; j->cig = cig; j->cig_stride = cig_stride;
        if (threads == 1) worker(j);
        else pthread_create(&th[t], NULL, worker, j);
    }
    if (threads > 1)
        for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
    return 0;
}
