/* JNI glue: libcudaswjni.so = this file + -lcudasw. Replaces Java_ssw_Aligner_align (libsswjni.so 0xcac).
 * gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -I../../include cudaswjni.c -L../../gridss_b200 -lcudasw -o libcudaswjni.so */
#include <jni.h>
#include <stdlib.h>
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
JNIEXPORT jint JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_alignBatch(JNIEnv* env, jclass c, jlong h, jint n,
        jbyteArray seqBlob, jlongArray seqOff, jbyteArray refBlob, jlongArray refOff, jintArray startOut,
        jobjectArray cigarOut, jintArray statusOut) {
    cudasw_result* res = (cudasw_result*)malloc(sizeof(cudasw_result) * (size_t)(n > 0 ? n : 1));
    jsize qbytes = (*env)->GetArrayLength(env, seqBlob);
    int64_t cap = (int64_t)qbytes + 4 * (int64_t)n + 16, used = 0;
    uint32_t* cig = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)cap);
    /* critical sections: no copies of the blobs (the reference's GetByteArrayElements copies, .so 0xcde-0xe94) */
    jbyte* s = (*env)->GetPrimitiveArrayCritical(env, seqBlob, NULL);
    jbyte* r = (*env)->GetPrimitiveArrayCritical(env, refBlob, NULL);
    jlong* so = (*env)->GetPrimitiveArrayCritical(env, seqOff, NULL);
    jlong* ro = (*env)->GetPrimitiveArrayCritical(env, refOff, NULL);
    int rc = cudasw_align_batch((cudasw_t*)(intptr_t)h, n, (const uint8_t*)s, (const int64_t*)so, (const uint8_t*)r,
                                (const int64_t*)ro, res, cig, cap, &used);
    (*env)->ReleasePrimitiveArrayCritical(env, refOff, ro, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, seqOff, so, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, refBlob, r, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, seqBlob, s, JNI_ABORT);
    if (rc == CUDASW_OK) {
        jint* st = (*env)->GetIntArrayElements(env, startOut, NULL);
        jint* status = (*env)->GetIntArrayElements(env, statusOut, NULL);
        char buf[1 << 16];
        for (jint i = 0; i < n; i++) {
            st[i] = res[i].ref_begin;
            status[i] = res[i].status;
            if (res[i].status == 0 && cudasw_cigar_string(cig + res[i].cigar_off, res[i].cigar_len, buf, sizeof buf) >= 0)
                (*env)->SetObjectArrayElement(env, cigarOut, i, (*env)->NewStringUTF(env, buf));
        }
        (*env)->ReleaseIntArrayElements(env, startOut, st, 0);
        (*env)->ReleaseIntArrayElements(env, statusOut, status, 0);
    }
    free(cig);
    free(res);
    return rc;
}

/* Shared tail of both batch entry points: results -> int[] start, String[] cigar, int[] status. */
static void cudaswjni_deliver(JNIEnv* env, jint n, const cudasw_result* res, const uint32_t* cig, jintArray startOut,
                              jobjectArray cigarOut, jintArray statusOut) {
    jint* st = (*env)->GetIntArrayElements(env, startOut, NULL);
    jint* status = (*env)->GetIntArrayElements(env, statusOut, NULL);
    char buf[1 << 16];
    for (jint i = 0; i < n; i++) {
        st[i] = res[i].ref_begin;
        status[i] = res[i].status;
        if (res[i].status == 0 && cudasw_cigar_string(cig + res[i].cigar_off, res[i].cigar_len, buf, sizeof buf) >= 0)
            (*env)->SetObjectArrayElement(env, cigarOut, i, (*env)->NewStringUTF(env, buf));
    }
    (*env)->ReleaseIntArrayElements(env, startOut, st, 0);
    (*env)->ReleaseIntArrayElements(env, statusOut, status, 0);
}

/* Device-resident reference (replaces TwoBitBufferedReferenceSequenceFile's per-JVM cache for the aligner's callers,
 * picard/TwoBitBufferedReferenceSequenceFile.java:97-125): called once per contig set. */
JNIEXPORT jint JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_loadReference(JNIEnv* env, jclass c, jlong h,
        jbyteArray bases, jlongArray contigOff) {
    jsize nc = (*env)->GetArrayLength(env, contigOff) - 1;
    jbyte* b = (*env)->GetPrimitiveArrayCritical(env, bases, NULL);
    jlong* o = (*env)->GetPrimitiveArrayCritical(env, contigOff, NULL);
    int rc = cudasw_reference_load((cudasw_t*)(intptr_t)h, (int32_t)nc, (const uint8_t*)b, (const int64_t*)o);
    (*env)->ReleasePrimitiveArrayCritical(env, contigOff, o, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, bases, b, JNI_ABORT);
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
    jbyte* ib = (*env)->GetByteArrayElements(env, insBlob, NULL);
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

/* Windows named as segments: parallel arrays (contig, flags, start, end) of nSeg entries + winOff[n+1]. */
JNIEXPORT jint JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_alignBatchWindows(JNIEnv* env, jclass c, jlong h, jint n,
        jbyteArray seqBlob, jlongArray seqOff, jintArray segContig, jintArray segFlags, jlongArray segStart,
        jlongArray segEnd, jlongArray winOff, jintArray startOut, jobjectArray cigarOut, jintArray statusOut) {
    jsize nseg = (*env)->GetArrayLength(env, segContig);
    cudasw_window* win = (cudasw_window*)malloc(sizeof(cudasw_window) * (size_t)(nseg > 0 ? nseg : 1));
    {
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
    }
    cudasw_result* res = (cudasw_result*)malloc(sizeof(cudasw_result) * (size_t)(n > 0 ? n : 1));
    jsize qbytes = (*env)->GetArrayLength(env, seqBlob);
    int64_t cap = (int64_t)qbytes + 4 * (int64_t)n + 16, used = 0;
    uint32_t* cig = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)cap);
    jbyte* s = (*env)->GetPrimitiveArrayCritical(env, seqBlob, NULL);
    jlong* so = (*env)->GetPrimitiveArrayCritical(env, seqOff, NULL);
    jlong* wo = (*env)->GetPrimitiveArrayCritical(env, winOff, NULL);
    int rc = cudasw_align_batch_windows((cudasw_t*)(intptr_t)h, n, (const uint8_t*)s, (const int64_t*)so, win,
                                        (const int64_t*)wo, res, cig, cap, &used);
    (*env)->ReleasePrimitiveArrayCritical(env, winOff, wo, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, seqOff, so, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, seqBlob, s, JNI_ABORT);
    if (rc == CUDASW_OK) cudaswjni_deliver(env, n, res, cig, startOut, cigarOut, statusOut);
    free(cig);
    free(res);
    free(win);
    return rc;
}

/* Pinned staging as direct ByteBuffers (north_star: "packs ... sequences and reference windows into pinned host
 * buffers"): the Java side packs a batch straight into page-locked memory, so cudasw_align_batch's H2D copies run at
 * PCIe speed and overlap the kernels; no JVM-heap array is touched by the native side. */
JNIEXPORT jobject JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_allocPinned(JNIEnv* env, jclass c, jlong bytes) {
    void* p = NULL;
    if (bytes <= 0 || cudasw_host_alloc((size_t)bytes, &p) != CUDASW_OK) return NULL;
    return (*env)->NewDirectByteBuffer(env, p, bytes);
}
JNIEXPORT void JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_freePinned(JNIEnv* env, jclass c, jobject buf) {
    void* p = buf ? (*env)->GetDirectBufferAddress(env, buf) : NULL;
    if (p) cudasw_host_free(p);
}
/* All four inputs are direct buffers from allocPinned (offsets as little-endian int64, n+1 entries each). */
JNIEXPORT jint JNICALL Java_au_edu_wehi_idsv_alignment_CudaAligner_alignBatchPinned(JNIEnv* env, jclass c, jlong h, jint n,
        jobject seqBlob, jobject seqOff, jobject refBlob, jobject refOff, jintArray startOut, jobjectArray cigarOut,
        jintArray statusOut) {
    const uint8_t* s = (const uint8_t*)(*env)->GetDirectBufferAddress(env, seqBlob);
    const int64_t* so = (const int64_t*)(*env)->GetDirectBufferAddress(env, seqOff);
    const uint8_t* r = (const uint8_t*)(*env)->GetDirectBufferAddress(env, refBlob);
    const int64_t* ro = (const int64_t*)(*env)->GetDirectBufferAddress(env, refOff);
    if (!s || !so || !r || !ro || n < 0) return CUDASW_EINVAL;
    cudasw_result* res = (cudasw_result*)malloc(sizeof(cudasw_result) * (size_t)(n > 0 ? n : 1));
    int64_t cap = (n > 0 ? so[n] : 0) + 4 * (int64_t)n + 16, used = 0;
    uint32_t* cig = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)cap);
    int rc = cudasw_align_batch((cudasw_t*)(intptr_t)h, n, s, so, r, ro, res, cig, cap, &used);
    if (rc == CUDASW_OK) cudaswjni_deliver(env, n, res, cig, startOut, cigarOut, statusOut);
    free(cig);
    free(res);
    return rc;
}
