/*
 * cudasw.h - C ABI of libcudasw.so, the B200-native batched Smith-Waterman engine that drops in
 * behind GRIDSS's `Aligner.align_smith_waterman` (reference: src/main/java/au/edu/wehi/idsv/
 * alignment/Aligner.java:10, SswJniAligner.java:42-61).
 *
 * Plain pointers and sizes only; no torch / C++ types. Every entry point names the reference
 * interface it replaces. All functions return 0 (CUDASW_OK) or a negative CUDASW_E* code and never
 * abort the process (the reference's native code calls exit(1) on oversize direction matrices,
 * SURVEY.md §5 "Failure detection").
 *
 * There is NO CPU fallback in this library: if no CUDA device is usable, cudasw_create fails with
 * CUDASW_ECUDA.
 */
#ifndef CUDASW_H
#define CUDASW_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CUDASW_OK 0
#define CUDASW_EINVAL (-1)    /* bad argument (maps to java.lang.IllegalArgumentException) */
#define CUDASW_ENOMEM (-2)    /* host or device allocation failed */
#define CUDASW_ECUDA (-3)     /* CUDA runtime error; see cudasw_last_error */
#define CUDASW_ETOOLARGE (-4) /* pair exceeds the engine's limits (per-pair status only) */
#define CUDASW_EUNDEF (-5)    /* per-pair: the reference's own traceback is undefined here */

/* CIGAR op codes of the packed `(len << 4) | op` words (BAM numbering; the reference's native
 * to_cigar_int uses M=0,I=1,D=2 and SswJniAligner.java:52-59 adds the S runs in Java). */
#define CUDASW_CIGAR_M 0u
#define CUDASW_CIGAR_I 1u
#define CUDASW_CIGAR_D 2u
#define CUDASW_CIGAR_S 4u

/* One result per pair. Replaces ssw.Alignment (ctor sig (SSIIIIILjava/lang/String;)V in
 * repo/ssw/ssw/1.0/ssw-1.0.jar) + the soft-clip wrapping of SswJniAligner.java:52-60.
 * `ref_begin` is Alignment.getStartPosition() (Alignment.java:23); the CIGAR words
 * cigar_blob[cigar_off .. cigar_off+cigar_len) render to Alignment.getCigar() (Alignment.java:30). */
typedef struct cudasw_result {
    int32_t score;      /* score1 */
    int32_t ref_begin;  /* 0-based start on ref; -1 when score == 0 (reference behaviour) */
    int32_t ref_end;    /* 0-based inclusive end on ref */
    int32_t read_begin; /* 0-based first aligned query base  (= leading soft clip length) */
    int32_t read_end;   /* 0-based last aligned query base */
    int32_t cigar_off;  /* offset into the cigar blob, in uint32 words */
    int32_t cigar_len;  /* number of packed ops, soft clips included */
    int32_t status;     /* CUDASW_OK or a per-pair CUDASW_E* */
} cudasw_result;

/* Scoring. Replaces SswJniAligner(int match, int mismatch, int ambiguous, int gapOpen,
 * int gapExtend) (SswJniAligner.java:13-17) as built by AlignerFactory.create(...)
 * (AlignerFactory.java:108-120; defaults 1,-4,-4,6,1 at :90). Symbol classes after clean():
 * equal symbols (N vs N included) -> match; both in ACGT and different -> mismatch; otherwise
 * ambiguous (SswJniAligner.java:19-33). A k-base gap costs gapOpen + (k-1)*gapExtend. */
typedef struct cudasw_scoring {
    int32_t match, mismatch, ambiguous, gap_open, gap_extend;
} cudasw_scoring;

typedef struct cudasw cudasw_t; /* opaque engine handle: owns device memory, pinned staging, streams */

/* Counters of the last cudasw_align_batch call on a handle (device time from CUDA events). */
typedef struct cudasw_stats {
    uint64_t pairs;
    uint64_t cells;            /* sum of query_len * window_len (forward matrix cells) */
    uint64_t kernel_launches;  /* kernels of this library launched for the batch */
    float ms_total;            /* H2D + kernels + D2H, device-timed */
    float ms_h2d, ms_pack, ms_forward, ms_reverse, ms_traceback, ms_d2h;
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t forward_cells_padded; /* cells actually computed by the forward kernels */
    float ms_encode;               /* part of ms_pack: the HBM-bound encode (or window gather) kernels alone */
    uint64_t encode_bytes;         /* bytes those kernels read + wrote */
    uint64_t filter_banded;        /* pairs whose forward pass first swept a k-mer-filtered column range (band_filter.cuh) */
    uint64_t filter_failed;        /* ... of which the certificate failed: swept their whole window in a second pass */
} cudasw_stats;

/* Create an engine on CUDA device `device`. Replaces the AlignerFactory static initialiser's
 * System.load + self-test (AlignerFactory.java:55-84): fails (no fallback) if the device or the
 * kernels are unusable. Validation mirrors ssw.Aligner.align: 0 <= gap < 256 and |score| <= 127. */
int cudasw_create(const cudasw_scoring* scoring, int device, cudasw_t** out);
void cudasw_destroy(cudasw_t* h);

/* Which of the reference's CPU backends the results are bit-compatible with.
 * CUDASW_TIEBREAK_SSW (what cudasw_create gives): SswJniAligner, the reference's default (AlignerFactory.java:108-117).
 * CUDASW_TIEBREAK_JALIGNER: JAlignerAligner (alignment/JAlignerAligner.java:30-93), the fallback under --nojni /
 *   -Dsswjni.disable=true: same score and end cell, but jaligner's pointer walk - DIAGONAL, then the gap in the read
 *   (`D`), then `I`; a gap is extended only when that is strictly better than opening; the walk runs until H == 0, i.e.
 *   the longest zero-free start - and its CIGAR assembly ("0M<q>S" for score 0, start 0). UNPINNED: jaligner ships as
 *   class files only and no JVM is available; tested against the restatement oracle/jaligner_model.py. Sequences are
 *   clean()ed as in the SSW mode, which equals JAlignerAligner's raw-byte matrix for A/C/G/T/N in either case only.
 *   Limits: bands above 8192 diagonals -> per-pair CUDASW_ETOOLARGE; cudasw_breakpoint_homology is SSW-mode only. */
#define CUDASW_TIEBREAK_SSW 0
#define CUDASW_TIEBREAK_JALIGNER 1
int cudasw_create_ex(const cudasw_scoring* scoring, int device, int tie_break, cudasw_t** out);

/* Batched Aligner.align_smith_waterman (Aligner.java:10): pair i aligns
 * seq_blob[seq_off[i]..seq_off[i+1]) inside ref_blob[ref_off[i]..ref_off[i+1]).
 * Bytes are raw ASCII in any case; the clean() map of SswJniAligner.java:71-79 (upper-case,
 * non-ACGT -> N) is applied on the device. Offsets have n_pairs+1 entries. Empty seq or ref ->
 * per-pair CUDASW_EINVAL (SswJniAligner.java:45-50 throws IllegalArgumentException).
 * `out` has n_pairs entries; CIGAR words are appended to cigar_blob (capacity cigar_cap words);
 * if capacity is exceeded the call returns CUDASW_ENOMEM and *cigar_used holds the needed size.
 * Host buffers may be pageable or pinned; results are complete on return. Thread-safe per handle
 * (calls serialise on the handle, like -Dsswjni.sync=true, SswJniAligner.java:36-40,62-64). */
int cudasw_align_batch(cudasw_t* h, int64_t n_pairs,
                       const uint8_t* seq_blob, const int64_t* seq_off,
                       const uint8_t* ref_blob, const int64_t* ref_off,
                       cudasw_result* out, uint32_t* cigar_blob, int64_t cigar_cap,
                       int64_t* cigar_used);

/* ---- Packed host inputs (BASELINE.json north_star (1): the shim "packs 2-bit or byte sequences and reference windows
 * into pinned host buffers"; SURVEY.md 7 "Feeding the GPU"). Same result as cudasw_align_batch on the ASCII bytes, with
 * 3 bits instead of 8 per base crossing PCIe. Layout = the reference's own packed genome
 * (picard/TwoBitBufferedReferenceSequenceFile.java:97-125: 2 bits per base + an "ambiguous" bit set): all bases of a blob
 * share one linear coordinate b (pair i = bases [off[i], off[i+1])), base b sits in bits 2*(b&15) of two_bit[b>>4]
 * (A0 C1 G2 T3; an ambiguous base stores 0) and its N flag in bit b&31 of nmask[b>>5]. A NULL nmask says that no base of
 * that blob is ambiguous. off[0] may be > 0 (a slice of a larger blob; the pointers still name the blob's word 0).
 * Both word arrays must be readable up to the end of the last 32-base group they touch. */
int cudasw_align_batch_packed(cudasw_t* h, int64_t n_pairs,
                              const uint32_t* seq_two_bit, const uint32_t* seq_nmask, const int64_t* seq_off,
                              const uint32_t* ref_two_bit, const uint32_t* ref_nmask, const int64_t* ref_off,
                              cudasw_result* out, uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used);
/* Host-side packer for that layout (what the JNI shim calls while it collects a batch; replaces the two array copies +
 * clean() of SswJniAligner.java:42-44,71-79): packs ascii[0..n) as bases [base0, base0+n) of the blob. Upper/lower case
 * A,C,G,T -> 0..3, everything else -> N (mask bit; nmask may be NULL if the caller does not keep one). Words that also
 * hold earlier bases keep them, later bits of the last word are cleared, so a blob is appended to pair by pair.
 * Returns the number of ambiguous bases packed (>= 0) or CUDASW_EINVAL. Pure host code; needs no GPU. */
int64_t cudasw_pack_bases(const uint8_t* ascii, int64_t n, int64_t base0, uint32_t* two_bit, uint32_t* nmask);

/* Same hot path with inputs already resident in device memory (what bench.py's `value` times).
 * d_* are device pointers on the handle's device; out/cigar are HOST buffers. */
int cudasw_align_batch_device(cudasw_t* h, int64_t n_pairs,
                              const uint8_t* d_seq_blob, const int64_t* d_seq_off,
                              const uint8_t* d_ref_blob, const int64_t* d_ref_off,
                              const int64_t* h_seq_off, const int64_t* h_ref_off,
                              cudasw_result* out, uint32_t* cigar_blob, int64_t cigar_cap,
                              int64_t* cigar_used);

/* ---- Device-resident reference genome + window descriptors.
 * Replaces, for the aligner's callers, TwoBitBufferedReferenceSequenceFile.getSubsequenceAt
 * (picard/TwoBitBufferedReferenceSequenceFile.java:97-125,210-216: 2 bits per base + an "ambiguous" BitSet) and the
 * host-side reverse complements of BreakpointHomology.calculate (BreakpointHomology.java:62-72): the genome is packed
 * once into HBM (3 bits per base) and a batch then names its windows instead of shipping their bytes.
 * A window is the concatenation of one or more segments; a segment is bases [start, end) (0-based, end exclusive) of
 * contig `contig`, reverse-complemented when (flags & CUDASW_WINDOW_REVCOMP). Every byte that clean()
 * (SswJniAligner.java:71-79) would turn into N is stored as N; case is not kept (the aligner ignores it). */
#define CUDASW_WINDOW_REVCOMP 1
typedef struct cudasw_window {
    int32_t contig;
    int32_t flags;
    int64_t start, end;
} cudasw_window;

/* `bases` holds the contigs back to back (raw ASCII, any case), contig c = bases[contig_off[c] .. contig_off[c+1]).
 * Replaces any previously loaded reference; n_contigs == 0 unloads. */
int cudasw_reference_load(cudasw_t* h, int32_t n_contigs, const uint8_t* bases, const int64_t* contig_off);
/* getSubsequenceAt for one window of n_win segments: writes upper-case ACGT/N to `out`; returns the length or <0. */
int cudasw_reference_fetch(cudasw_t* h, const cudasw_window* win, int32_t n_win, uint8_t* out, int64_t out_cap);
/* cudasw_align_batch with windows named instead of shipped: pair i aligns seq_blob[seq_off[i]..seq_off[i+1]) inside
 * the concatenation of win[win_off[i] .. win_off[i+1]). A descriptor outside the loaded reference, or an empty
 * window, gives that pair status CUDASW_EINVAL. Results are identical to cudasw_align_batch on the fetched bytes. */
int cudasw_align_batch_windows(cudasw_t* h, int64_t n_pairs, const uint8_t* seq_blob, const int64_t* seq_off,
                               const cudasw_window* win, const int64_t* win_off, cudasw_result* out,
                               uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used);

/* ---- BreakpointHomology.calculate for a list of breakpoints (SURVEY.md 8a A7, 8f N1+N3).
 * Replaces the loop body of AnnotateInexactHomology / AnnotateInexactHomologyBedpe.java:68 around
 * BreakpointHomology.calculate(lookup, bs, insertedSequence, maxBreakendLength, margin)
 * (alignment/BreakpointHomology.java:38-103): for every breakpoint two alignments of the breakend sequence
 * (local anchor + inserted sequence + remote anchor) against the local and the remote reference context, and from their
 * clip lengths the two homology lengths. Needs a loaded reference (cudasw_reference_load): 32 bytes per breakpoint
 * (+ its inserted bases) go to the device, queries and windows are gathered there, only the score passes run.
 * Positions are 1-based and exact (start == end), as the reference requires (:39-41); dir 0 = forward, 1 = backward
 * (BreakendDirection). ins_off (n + 1 offsets into ins_blob) may be NULL = no inserted sequence anywhere.
 * Returns CUDASW_EINVAL for a contig / direction out of range (the reference throws). */
typedef struct cudasw_breakpoint {
    int32_t contig1, dir1;
    int64_t pos1;
    int32_t contig2, dir2;
    int64_t pos2;
} cudasw_breakpoint;
int cudasw_breakpoint_homology(cudasw_t* h, int64_t n, const cudasw_breakpoint* bp, const uint8_t* ins_blob,
                               const int64_t* ins_off, int32_t max_breakend_length, int32_t margin,
                               int32_t* local_hom, int32_t* remote_hom);

/* Device-memory helpers so a torch-free host (the JNI shim, the ctypes mirror) can stage inputs. */
int cudasw_device_alloc(cudasw_t* h, size_t bytes, void** dptr);
int cudasw_device_free(cudasw_t* h, void* dptr);
int cudasw_upload(cudasw_t* h, void* dptr, const void* src, size_t bytes);
/* Pinned host staging (the JNI shim hands these out as direct ByteBuffers). */
int cudasw_host_alloc(size_t bytes, void** hptr);
int cudasw_host_free(void* hptr);

/* Async trio modelled on StreamingAligner.asyncAlign/flush/getAlignment
 * (StreamingAligner.java:10-44) and BwaStreamingAligner's size-triggered flush
 * (BwaStreamingAligner.java:79-127): submit copies the pair into the handle's pending batch and
 * returns a ticket; flush runs every pending pair as one batch; poll fetches a finished ticket. */
int cudasw_submit(cudasw_t* h, const uint8_t* seq, int64_t seq_len, const uint8_t* ref,
                  int64_t ref_len, int64_t* ticket);
int cudasw_flush(cudasw_t* h);
int64_t cudasw_outstanding(cudasw_t* h); /* submitted, not yet flushed */
int64_t cudasw_processed(cudasw_t* h);   /* flushed, not yet polled */
/* Returns 1 and fills res/cigar (up to cigar_cap words, *cigar_len = words needed) if the ticket
 * is done, 0 if it is still pending, <0 on error (unknown ticket -> CUDASW_EINVAL). */
int cudasw_poll(cudasw_t* h, int64_t ticket, cudasw_result* res, uint32_t* cigar,
                int32_t cigar_cap, int32_t* cigar_len);

/* ---- Several GPUs behind one handle, for a single host process such as the JVM (the reference's
 * own worker pool is in-process: AnnotateInexactHomology.java:17-21). A batch is cut into contiguous slices of
 * balanced cost (sum of query_len * window_len), one per GPU, each driven by its own host thread and streams; results
 * land at the pairs' original ordinals. No exchange step exists between pairs, hence no collective.
 * devices == NULL or n_devices <= 0 selects every visible device. CIGAR words of GPU d are written to its own
 * region of cigar_blob (the blob is not compact; cigar_off is absolute), *cigar_used = words used in total, or on
 * CUDASW_ENOMEM a capacity with which a retry succeeds. Results are identical to cudasw_align_batch. */
typedef struct cudasw_multi cudasw_multi_t;
int cudasw_multi_create(const cudasw_scoring* scoring, const int* devices, int n_devices, cudasw_multi_t** out);
void cudasw_multi_destroy(cudasw_multi_t* m);
int cudasw_multi_device_count(cudasw_multi_t* m);
int cudasw_multi_align_batch(cudasw_multi_t* m, int64_t n_pairs,
                             const uint8_t* seq_blob, const int64_t* seq_off,
                             const uint8_t* ref_blob, const int64_t* ref_off,
                             cudasw_result* out, uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used);
int cudasw_multi_align_batch_packed(cudasw_multi_t* m, int64_t n_pairs,
                                    const uint32_t* seq_two_bit, const uint32_t* seq_nmask, const int64_t* seq_off,
                                    const uint32_t* ref_two_bit, const uint32_t* ref_nmask, const int64_t* ref_off,
                                    cudasw_result* out, uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used);
/* The reference is replicated on every GPU of the handle (one-time, outside any timed path). */
int cudasw_multi_reference_load(cudasw_multi_t* m, int32_t n_contigs, const uint8_t* bases, const int64_t* contig_off);
int cudasw_multi_align_batch_windows(cudasw_multi_t* m, int64_t n_pairs, const uint8_t* seq_blob, const int64_t* seq_off,
                                     const cudasw_window* win, const int64_t* win_off, cudasw_result* out,
                                     uint32_t* cigar_blob, int64_t cigar_cap, int64_t* cigar_used);
/* cudasw_breakpoint_homology over all GPUs of the handle (equal contiguous slices of the list, reference replicated by
 * cudasw_multi_reference_load). */
int cudasw_multi_breakpoint_homology(cudasw_multi_t* m, int64_t n, const cudasw_breakpoint* bp, const uint8_t* ins_blob,
                                     const int64_t* ins_off, int32_t max_breakend_length, int32_t margin,
                                     int32_t* local_hom, int32_t* remote_hom);
/* Stats of GPU `idx` for the last batch and the slice [first_pair, first_pair + n_pairs) it was given. */
int cudasw_multi_get_stats(cudasw_multi_t* m, int idx, cudasw_stats* out, int64_t* first_pair, int64_t* n_pairs);
const char* cudasw_multi_last_error(cudasw_multi_t* m); /* m may be NULL: last create() error */

/* Blocking poll: waits up to timeout_ms (< 0: no limit) for the ticket; returns like cudasw_poll (0 = timed out). */
int cudasw_wait(cudasw_t* h, int64_t ticket, int64_t timeout_ms, cudasw_result* res, uint32_t* cigar,
                int32_t cigar_cap, int32_t* cigar_len);
/* Dispatcher thread of the batch-collect layer (SURVEY.md 8b "Threading", 8f N1): flushes when max_pairs are pending or
 * the oldest pending pair has waited max_delay_us, so that GRIDSS's worker threads (ParallelTransformIterator,
 * AnnotateInexactHomology.java:17-21) only cudasw_submit and cudasw_wait. Modelled on BwaStreamingAligner's single
 * driver thread with a size-triggered flush (BwaStreamingAligner.java:79-127). submit / poll / wait are thread-safe and
 * do not block behind a running GPU batch. stop() flushes what is pending; destroy() stops the dispatcher. */
int cudasw_dispatcher_start(cudasw_t* h, int64_t max_pairs, int64_t max_delay_us);
int cudasw_dispatcher_stop(cudasw_t* h);

int cudasw_get_stats(cudasw_t* h, cudasw_stats* out);
/* Which score-pass kernel variant a query of `query_len` rows is dispatched to: lanes per pair (G),
 * rows per virtual lane (R), strips (1 unless strip-mined). For reporting (bench.py roofline line). */
int cudasw_variant_for(cudasw_t* h, int32_t query_len, int32_t* lanes, int32_t* rows_per_lane, int32_t* strips);
const char* cudasw_last_error(cudasw_t* h); /* h may be NULL: last create() error */
int cudasw_device_count(void);
const char* cudasw_version(void);

/* Render packed ops as the reference's CIGAR string ("4S78M45S"); returns bytes written
 * (excluding NUL) or CUDASW_EINVAL if `cap` is too small. Replaces s_align_cigar_to_jstring
 * (libsswjni.so 0xa6c) + the string concat in SswJniAligner.java:52-59. */
int cudasw_cigar_string(const uint32_t* ops, int32_t n_ops, char* buf, int32_t cap);

/* ---- Forward score pass for callers that cannot allocate. NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * Needs no handle and calls no cudaMalloc / cudaFree / allocating library routine: every piece of scratch the engine
 * normally allocates (encoded blobs, scoring constants, job list, strip boundary rows of the strip-mined forward kernel
 * sw_score_kernel<32, 20, MULTI>) lives in one file-scope `static __device__` array of CUDASW_NOALLOC_SCRATCH_BYTES
 * (gridss_b200/csrc/noalloc.cu) that every GPU loading libcudasw.so holds until exit. Hence the limits per call, which are
 * rejected with CUDASW_ETOOLARGE and a message in errbuf: n_pairs <= CUDASW_NOALLOC_MAX_PAIRS, seq_bytes <=
 * CUDASW_NOALLOC_MAX_SEQ_BYTES, ref_bytes <= CUDASW_NOALLOC_MAX_REF_BYTES.
 * ALL data pointers are DEVICE pointers on `device`, provided by the caller: raw ASCII blobs and n_pairs + 1 offsets each
 * as in cudasw_align_batch (clean() is applied on the device); seq_bytes / ref_bytes = size of the blobs (>= the last
 * offset). d_out receives 4 int32 per pair: {score, ref_end, read_end, status}, equal to the same fields of the
 * cudasw_result that cudasw_align_batch gives for the pair (score 0 -> ref_end -1, read_end 0; per-pair CUDASW_EINVAL /
 * CUDASW_ETOOLARGE as there). scoring == NULL -> 1,-4,-4,6,1. The kernels are launched on `stream` (a cudaStream_t; NULL
 * = the default stream) and the call returns WITHOUT synchronising: d_out is complete once the stream reaches that point.
 * The first call on a device takes the array's device address (cudaGetSymbolAddress) and must be made outside a stream
 * capture; n_pairs == 0 does only that. The current device is restored before returning.
 * Every call shares the array: calls that overlap on one GPU (different streams, concurrent host threads) RACE; order
 * them on one stream or with events. errbuf (may be NULL) receives a NUL-terminated message when the result is not 0. */
#define CUDASW_NOALLOC_MAX_PAIRS 4096
#define CUDASW_NOALLOC_MAX_SEQ_BYTES (1 << 20)
#define CUDASW_NOALLOC_MAX_REF_BYTES (1 << 20)
#define CUDASW_NOALLOC_SCRATCH_BYTES (10 * 1024 * 1024 + 128 * 1024) /* 10.125 MiB per GPU */
int cudasw_score_batch_noalloc(const cudasw_scoring* scoring, int device, int32_t n_pairs,
                               const uint8_t* d_seq_blob, const int64_t* d_seq_off, int64_t seq_bytes,
                               const uint8_t* d_ref_blob, const int64_t* d_ref_off, int64_t ref_bytes,
                               int32_t* d_out, void* stream, char* errbuf, size_t errbuf_cap);

#ifdef __cplusplus
}
#endif
#endif /* CUDASW_H */
