/* JNI glue: libcudaswjni.so = this file + -lcudasw. Replaces Java_ssw_Aligner_align (libsswjni.so 0xcac).
 * gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -I../../include cudaswjni.c -L../../gridss_b200 -lcudasw -o libcudaswjni.so
 *
 * Rules kept throughout (JNI spec, "GetPrimitiveArrayCritical"): no call that can block - CUDA synchronisation, the
 * engine's mutex - is ever made inside a critical region; the only critical regions here wrap the pure-host packer.
 * Heap arrays that a CUDA call needs are copied out first (Get<Type>ArrayRegion); the batch path does not need any,
 * because Java packs into page-locked direct buffers. */
#include <jni.h>
#include <stdlib.h>
#include <string.h>
#include "cudasw.h"

JNIEXPORT jlong JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_create(JNIEnv* env, jclass c, jint match, jint mismatch,
        jint ambiguous, jint gapOpen, jint gapExtend, jint device) {
    cudasw_scoring sc = {match, mismatch, ambiguous, gapOpen, gapExtend};
    cudasw_t* h = NULL;
    int rc = cudasw_create(&sc, device, &h);
    if (rc == CUDASW_EINVAL) {
        (*env)->ThrowNew(env, (*env)->FindClass(env, "java/lang/IllegalArgumentException"), cudasw_last_error(NULL));
        return 0;
    }
    return rc == CUDASW_OK ? (jlong)(intptr_t)h : 0;
}
JNIEXPORT void JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_destroy(JNIEnv* env, jclass c, jlong h) {
    cudasw_destroy((cudasw_t*)(intptr_t)h);
}
JNIEXPORT jstring JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_lastError(JNIEnv* env, jclass c, jlong h) {
    return (*env)->NewStringUTF(env, cudasw_last_error((cudasw_t*)(intptr_t)h));
}

/* Results -> int[] start, String[] cigar, int[] status. A CIGAR that cannot be rendered leaves a non-zero status behind
 * (never a null string with status 0); local references are dropped as the loop goes (n may be millions). */
static void cudaswjni_deliver(JNIEnv* env, jint n, const cudasw_result* res, const uint32_t* cig, jintArray startOut,
                              jobjectArray cigarOut, jintArray statusOut) {
    jint* st = (*env)->GetIntArrayElements(env, startOut, NULL);
    jint* status = (*env)->GetIntArrayElements(env, statusOut, NULL);
    size_t cap = 1 << 12;
    char* buf = (char*)malloc(cap);
    for (jint i = 0; i < n; i++) {
        st[i] = res[i].ref_begin;
        status[i] = res[i].status;
        if (res[i].status != 0) continue;
        const size_t need = 12 * (size_t)res[i].cigar_len + 16;  /* "4294967295M" per op at most */
        if (need > cap) {
            char* nb = (char*)realloc(buf, need);
            if (nb) { buf = nb; cap = need; }
        }
        jstring s = NULL;
        if (buf && need <= cap && cudasw_cigar_string(cig + res[i].cigar_off, res[i].cigar_len, buf, (int32_t)cap) >= 0)
            s = (*env)->NewStringUTF(env, buf);
        if (s) {
            (*env)->SetObjectArrayElement(env, cigarOut, i, s);
            (*env)->DeleteLocalRef(env, s);
        } else status[i] = CUDASW_ENOMEM;
    }
    free(buf);
    (*env)->ReleaseIntArrayElements(env, startOut, st, 0);
    (*env)->ReleaseIntArrayElements(env, statusOut, status, 0);
}

/* Device-resident reference (replaces TwoBitBufferedReferenceSequenceFile's per-JVM cache for the aligner's callers,
 * picard/TwoBitBufferedReferenceSequenceFile.java:97-125): called once per contig set. The bases are copied out of the
 * heap first: the upload synchronises with the device slice by slice. */
JNIEXPORT jint JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_loadReference(JNIEnv* env, jclass c, jlong h,
        jbyteArray bases, jlongArray contigOff) {
    const jsize nc = (*env)->GetArrayLength(env, contigOff) - 1;
    const jsize nb = (*env)->GetArrayLength(env, bases);
    if (nc < 0) return CUDASW_EINVAL;
    jbyte* b = (jbyte*)malloc((size_t)(nb > 0 ? nb : 1));
    jlong* o = (jlong*)malloc(sizeof(jlong) * ((size_t)nc + 1));
    if (!b || !o) { free(b); free(o); return CUDASW_ENOMEM; }
    (*env)->GetByteArrayRegion(env, bases, 0, nb, b);
    (*env)->GetLongArrayRegion(env, contigOff, 0, nc + 1, o);
    int rc = cudasw_reference_load((cudasw_t*)(intptr_t)h, (int32_t)nc, (const uint8_t*)b, (const int64_t*)o);
    free(b); free(o);
    return rc;
}

/* BreakpointHomology.calculate for n breakpoints (alignment/BreakpointHomology.java:38-103): parallel arrays in,
 * the two homology lengths out. */
JNIEXPORT jint JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_breakpointHomology(JNIEnv* env, jclass c, jlong h, jint n,
        jintArray contig1, jintArray dir1, jlongArray pos1, jintArray contig2, jintArray dir2, jlongArray pos2,
        jbyteArray insBlob, jlongArray insOff, jint maxBreakendLength, jint margin, jintArray localOut, jintArray remoteOut) {
    cudasw_breakpoint* bp = (cudasw_breakpoint*)malloc(sizeof(cudasw_breakpoint) * (size_t)(n > 0 ? n : 1));
    int32_t* out = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)(n > 0 ? n : 1));
    if (!bp || !out) { free(bp); free(out); return CUDASW_ENOMEM; }
    jint* c1 = (*env)->GetIntArrayElements(env, contig1, NULL);
    jint* d1 = (*env)->GetIntArrayElements(env, dir1, NULL);
    jlong* p1 = (*env)->GetLongArrayElements(env, pos1, NULL);
    jint* c2 = (*env)->GetIntArrayElements(env, contig2, NULL);
    jint* d2 = (*env)->GetIntArrayElements(env, dir2, NULL);
    jlong* p2 = (*env)->GetLongArrayElements(env, pos2, NULL);
    for (jint i = 0; i < n; i++) {
        bp[i].contig1 = c1[i]; bp[i].dir1 = d1[i]; bp[i].pos1 = p1[i];
        bp[i].contig2 = c2[i]; bp[i].dir2 = d2[i]; bp[i].pos2 = p2[i];
    }
    (*env)->ReleaseIntArrayElements(env, contig1, c1, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, dir1, d1, JNI_ABORT);
    (*env)->ReleaseLongArrayElements(env, pos1, p1, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, contig2, c2, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, dir2, d2, JNI_ABORT);
    (*env)->ReleaseLongArrayElements(env, pos2, p2, JNI_ABORT);
    jbyte* ib = (*env)->GetByteArrayElements(env, insBlob, NULL);   /* copies or pins; not a critical region */
    jlong* io = (*env)->GetLongArrayElements(env, insOff, NULL);
    int rc = cudasw_breakpoint_homology((cudasw_t*)(intptr_t)h, n, bp, (const uint8_t*)ib, (const int64_t*)io, maxBreakendLength,
                                        margin, out, out + (n > 0 ? n : 1));
    (*env)->ReleaseLongArrayElements(env, insOff, io, JNI_ABORT);
    (*env)->ReleaseByteArrayElements(env, insBlob, ib, JNI_ABORT);
    if (rc == CUDASW_OK) {
        (*env)->SetIntArrayRegion(env, localOut, 0, n, (const jint*)out);
        (*env)->SetIntArrayRegion(env, remoteOut, 0, n, (const jint*)(out + (n > 0 ? n : 1)));
    }
    free(bp); free(out);
    return rc;
}

/* Windows named as segments: parallel arrays (contig, flags, start, end) of nSeg entries + winOff[n+1]. The query bytes and
 * offsets are copied out of the heap before the CUDA call. */
JNIEXPORT jint JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_alignBatchWindows(JNIEnv* env, jclass c, jlong h, jint n,
        jbyteArray seqBlob, jlongArray seqOff, jintArray segContig, jintArray segFlags, jlongArray segStart,
        jlongArray segEnd, jlongArray winOff, jintArray startOut, jobjectArray cigarOut, jintArray statusOut) {
    if (n < 0) return CUDASW_EINVAL;
    const jsize nseg = (*env)->GetArrayLength(env, segContig);
    const jsize qbytes = (*env)->GetArrayLength(env, seqBlob);
    int64_t cap = 2 * (int64_t)qbytes + 4 * (int64_t)n + 16, used = 0;
    cudasw_window* win = (cudasw_window*)malloc(sizeof(cudasw_window) * (size_t)(nseg > 0 ? nseg : 1));
    cudasw_result* res = (cudasw_result*)malloc(sizeof(cudasw_result) * (size_t)(n > 0 ? n : 1));
    uint32_t* cig = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)cap);
    jbyte* s = (jbyte*)malloc((size_t)(qbytes > 0 ? qbytes : 1));
    jlong* so = (jlong*)malloc(sizeof(jlong) * ((size_t)n + 1));
    jlong* wo = (jlong*)malloc(sizeof(jlong) * ((size_t)n + 1));
    int rc = CUDASW_ENOMEM;
    if (win && res && cig && s && so && wo) {
        jint* sc = (*env)->GetIntArrayElements(env, segContig, NULL);
        jint* sf = (*env)->GetIntArrayElements(env, segFlags, NULL);
        jlong* ss = (*env)->GetLongArrayElements(env, segStart, NULL);
        jlong* se = (*env)->GetLongArrayElements(env, segEnd, NULL);
        for (jsize k = 0; k < nseg; k++) {
            win[k].contig = sc[k]; win[k].flags = sf[k]; win[k].start = ss[k]; win[k].end = se[k];
        }
        (*env)->ReleaseIntArrayElements(env, segContig, sc, JNI_ABORT);
        (*env)->ReleaseIntArrayElements(env, segFlags, sf, JNI_ABORT);
        (*env)->ReleaseLongArrayElements(env, segStart, ss, JNI_ABORT);
        (*env)->ReleaseLongArrayElements(env, segEnd, se, JNI_ABORT);
        (*env)->GetByteArrayRegion(env, seqBlob, 0, qbytes, s);
        (*env)->GetLongArrayRegion(env, seqOff, 0, n + 1, so);
        (*env)->GetLongArrayRegion(env, winOff, 0, n + 1, wo);
        rc = cudasw_align_batch_windows((cudasw_t*)(intptr_t)h, n, (const uint8_t*)s, (const int64_t*)so, win,
                                        (const int64_t*)wo, res, cig, cap, &used);
        if (rc == CUDASW_ENOMEM && used > cap) {  /* the CIGAR estimate was short: once more with the size asked for */
            uint32_t* nc = (uint32_t*)realloc(cig, sizeof(uint32_t) * (size_t)(used + 16));
            if (nc) {
                cig = nc; cap = used + 16;
                rc = cudasw_align_batch_windows((cudasw_t*)(intptr_t)h, n, (const uint8_t*)s, (const int64_t*)so, win,
                                                (const int64_t*)wo, res, cig, cap, &used);
            }
        }
        if (rc == CUDASW_OK) cudaswjni_deliver(env, n, res, cig, startOut, cigarOut, statusOut);
    }
    free(wo); free(so); free(s); free(cig); free(res); free(win);
    return rc;
}

/* Pinned staging as direct ByteBuffers (north_star: "packs 2-bit or byte sequences and reference windows into pinned host
 * buffers"): the Java side packs a batch straight into page-locked memory, so the H2D copies run at PCIe speed and
 * overlap the kernels; no JVM-heap array is touched while CUDA runs. */
JNIEXPORT jobject JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_allocPinned(JNIEnv* env, jclass c, jlong bytes) {
    void* p = NULL;
    if (bytes <= 0 || cudasw_host_alloc((size_t)bytes, &p) != CUDASW_OK) return NULL;
    memset(p, 0, (size_t)bytes);
    return (*env)->NewDirectByteBuffer(env, p, bytes);
}
JNIEXPORT void JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_freePinned(JNIEnv* env, jclass c, jobject buf) {
    void* p = buf ? (*env)->GetDirectBufferAddress(env, buf) : NULL;
    if (p) cudasw_host_free(p);
}
/* cudasw_pack_bases on one sequence: the only critical region of this file, around pure host code (no CUDA, no locks). */
JNIEXPORT jlong JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_packBases(JNIEnv* env, jclass c, jbyteArray ascii,
        jlong base0, jobject twoBit, jobject nmask) {
    uint32_t* t2 = (uint32_t*)(*env)->GetDirectBufferAddress(env, twoBit);
    uint32_t* nm = (uint32_t*)(*env)->GetDirectBufferAddress(env, nmask);
    const jsize n = (*env)->GetArrayLength(env, ascii);
    if (!t2 || !nm || base0 < 0) return CUDASW_EINVAL;
    /* the words of whole 32-base groups must fit the buffers */
    const jlong groups = (base0 + n + 31) / 32;
    if ((*env)->GetDirectBufferCapacity(env, twoBit) < 8 * groups || (*env)->GetDirectBufferCapacity(env, nmask) < 4 * groups)
        return CUDASW_EINVAL;
    jbyte* a = (jbyte*)(*env)->GetPrimitiveArrayCritical(env, ascii, NULL);
    if (!a) return CUDASW_ENOMEM;
    const int64_t amb = cudasw_pack_bases((const uint8_t*)a, n, base0, t2, nm);
    (*env)->ReleasePrimitiveArrayCritical(env, ascii, a, JNI_ABORT);
    return amb;
}
/* All inputs are direct buffers from allocPinned (offsets as little-endian int64, n+1 entries each); a null mask buffer
 * says that blob holds no ambiguous base. */
JNIEXPORT jint JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_alignBatchPacked(JNIEnv* env, jclass c, jlong h, jint n,
        jobject seqTwoBit, jobject seqMask, jobject seqOff, jobject refTwoBit, jobject refMask, jobject refOff,
        jintArray startOut, jobjectArray cigarOut, jintArray statusOut) {
    const uint32_t* s2 = (const uint32_t*)(*env)->GetDirectBufferAddress(env, seqTwoBit);
    const uint32_t* sn = seqMask ? (const uint32_t*)(*env)->GetDirectBufferAddress(env, seqMask) : NULL;
    const int64_t* so = (const int64_t*)(*env)->GetDirectBufferAddress(env, seqOff);
    const uint32_t* r2 = (const uint32_t*)(*env)->GetDirectBufferAddress(env, refTwoBit);
    const uint32_t* rn = refMask ? (const uint32_t*)(*env)->GetDirectBufferAddress(env, refMask) : NULL;
    const int64_t* ro = (const int64_t*)(*env)->GetDirectBufferAddress(env, refOff);
    if (!s2 || !so || !r2 || !ro || n < 0) return CUDASW_EINVAL;
    cudasw_result* res = (cudasw_result*)malloc(sizeof(cudasw_result) * (size_t)(n > 0 ? n : 1));
    int64_t cap = (n > 0 ? so[n] : 0) + 4 * (int64_t)n + 16, used = 0;
    uint32_t* cig = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)cap);
    int rc = CUDASW_ENOMEM;
    if (res && cig) {
        rc = cudasw_align_batch_packed((cudasw_t*)(intptr_t)h, n, s2, sn, so, r2, rn, ro, res, cig, cap, &used);
        if (rc == CUDASW_ENOMEM && used > cap) {  /* the CIGAR estimate was short: once more with the size asked for */
            uint32_t* nc = (uint32_t*)realloc(cig, sizeof(uint32_t) * (size_t)(used + 16));
            if (nc) {
                cig = nc; cap = used + 16;
                rc = cudasw_align_batch_packed((cudasw_t*)(intptr_t)h, n, s2, sn, so, r2, rn, ro, res, cig, cap, &used);
            }
        }
        if (rc == CUDASW_OK) cudaswjni_deliver(env, n, res, cig, startOut, cigarOut, statusOut);
    }
    free(cig);
    free(res);
    return rc;
}
