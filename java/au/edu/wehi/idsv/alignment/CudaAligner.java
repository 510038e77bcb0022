package au.edu.wehi.idsv.alignment;

import java.nio.charset.StandardCharsets;

/**
 * Drop-in Aligner backed by libcudasw.so (B200 batched Smith-Waterman).
 * Bit-exact with SswJniAligner (start position, CIGAR) by construction of the native library.
 * Written against include/cudasw.h; cannot be compiled in the build container (no JDK) - see INTEGRATION.md.
 */
public class CudaAligner implements Aligner, AutoCloseable {
	static { System.loadLibrary("cudaswjni"); }
	private long handle;
	public CudaAligner(int match, int mismatch, int ambiguous, int gapOpen, int gapExtend, int device) {
		handle = create(match, mismatch, ambiguous, gapOpen, gapExtend, device);
		if (handle == 0) throw new IllegalStateException("cudasw_create failed: " + lastError(0));
	}
	@Override
	public Alignment align_smith_waterman(byte[] seq, byte[] ref) {
		if (seq == null || seq.length == 0) throw new IllegalArgumentException("seq must be non-zero size");
		if (ref == null || ref.length == 0) throw new IllegalArgumentException("ref must be non-zero size");
		return alignBatch(new byte[][] { seq }, new byte[][] { ref })[0];
	}
	/** One JNI round trip for the whole batch. */
	public Alignment[] alignBatch(byte[][] seqs, byte[][] refs) {
		int n = seqs.length;
		long[] seqOff = new long[n + 1], refOff = new long[n + 1];
		for (int i = 0; i < n; i++) {
			seqOff[i + 1] = seqOff[i] + seqs[i].length;
			refOff[i + 1] = refOff[i] + refs[i].length;
		}
		byte[] seqBlob = new byte[(int) seqOff[n]], refBlob = new byte[(int) refOff[n]];
		for (int i = 0; i < n; i++) {
			System.arraycopy(seqs[i], 0, seqBlob, (int) seqOff[i], seqs[i].length);
			System.arraycopy(refs[i], 0, refBlob, (int) refOff[i], refs[i].length);
		}
		int[] start = new int[n], status = new int[n];
		String[] cigar = new String[n];
		int rc = alignBatch(handle, n, seqBlob, seqOff, refBlob, refOff, start, cigar, status);
		if (rc != 0) throw new IllegalStateException("cudasw_align_batch failed (" + rc + "): " + lastError(handle));
		Alignment[] out = new Alignment[n];
		for (int i = 0; i < n; i++) {
			if (status[i] == -1) throw new IllegalArgumentException("seq and ref must be non-zero size");
			if (status[i] != 0) throw new IllegalStateException("cudasw pair status " + status[i]);
			out[i] = new Alignment(start[i], cigar[i]);
		}
		return out;
	}
	@Override
	public void close() { if (handle != 0) { destroy(handle); handle = 0; } }
	private static native long create(int match, int mismatch, int ambiguous, int gapOpen, int gapExtend, int device);
	private static native void destroy(long handle);
	private static native String lastError(long handle);
	private static native int alignBatch(long handle, int n, byte[] seqBlob, long[] seqOff, byte[] refBlob, long[] refOff,
			int[] startOut, String[] cigarOut, int[] statusOut);
}
