package au.edu.wehi.idsv.alignment;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;

/**
 * Drop-in Aligner backed by libcudasw.so (B200 batched Smith-Waterman).
 * Bit-exact with SswJniAligner (start position, CIGAR) by construction of the native library.
 * Written against include/cudasw.h; cannot be compiled in the build container (no JDK) - see INTEGRATION.md.
 *
 * A batch crosses JNI once. The pairs are packed 2 bits per base (+ one "ambiguous" bit, the layout of
 * picard/TwoBitBufferedReferenceSequenceFile.java:97-125) straight into page-locked direct buffers by the native host
 * packer (cudasw_pack_bases), so 3/8 of the sequence bytes cross PCIe and no JVM-heap array is touched while CUDA runs.
 */
public class CudaAligner implements Aligner, AutoCloseable {
	static { System.loadLibrary("cudaswjni"); }
	public static final int STATUS_OK = 0, STATUS_EINVAL = -1;
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
	// pinned staging (page-locked direct buffers from the native side), grown on demand and reused across batches
	private ByteBuffer pinSeq2, pinSeqN, pinRef2, pinRefN, pinSeqOff, pinRefOff;
	private ByteBuffer pinned(ByteBuffer cur, long bytes) {
		if (cur != null && cur.capacity() >= bytes) { cur.clear(); return cur; }
		if (cur != null) freePinned(cur);
		ByteBuffer b = allocPinned(Math.max(4096, bytes + bytes / 4));
		if (b == null) throw new OutOfMemoryError("cudasw_host_alloc failed");
		return b.order(ByteOrder.LITTLE_ENDIAN);
	}
	/** Throws for the first failed pair (IllegalArgumentException for an empty seq/ref, as SswJniAligner.java:45-50). */
	public Alignment[] alignBatch(byte[][] seqs, byte[][] refs) {
		int[] status = new int[seqs.length];
		Alignment[] out = alignBatch(seqs, refs, status);
		for (int i = 0; i < out.length; i++) {
			if (status[i] != STATUS_OK) throw failure(status[i]);
		}
		return out;
	}
	/** The exception the reference would have raised for a pair with this status. */
	public static RuntimeException failure(int status) {
		if (status == STATUS_EINVAL) return new IllegalArgumentException("seq and ref must be non-zero size");
		return new IllegalStateException("cudasw pair status " + status);
	}
	/**
	 * One JNI round trip for the whole batch. Pairs are independent, as calls to the reference's aligner are: a pair that
	 * cannot be aligned gets its status in status[i] and a null Alignment, and does not affect its neighbours.
	 */
	public synchronized Alignment[] alignBatch(byte[][] seqs, byte[][] refs, int[] status) {
		int n = seqs.length;
		long tq = 0, tw = 0;
		for (int i = 0; i < n; i++) { tq += seqs[i] == null ? 0 : seqs[i].length; tw += refs[i] == null ? 0 : refs[i].length; }
		// whole 32-base groups: 8 bytes of 2-bit words + 4 bytes of mask per group
		long gq = (tq + 31) / 32 + 1, gw = (tw + 31) / 32 + 1;
		pinSeq2 = pinned(pinSeq2, 8 * gq); pinSeqN = pinned(pinSeqN, 4 * gq);
		pinRef2 = pinned(pinRef2, 8 * gw); pinRefN = pinned(pinRefN, 4 * gw);
		pinSeqOff = pinned(pinSeqOff, 8L * (n + 1)); pinRefOff = pinned(pinRefOff, 8L * (n + 1));
		long so = 0, ro = 0, ambQ = 0, ambW = 0;
		pinSeqOff.putLong(0); pinRefOff.putLong(0);
		for (int i = 0; i < n; i++) {
			if (seqs[i] != null && seqs[i].length > 0) { ambQ += packBases(seqs[i], so, pinSeq2, pinSeqN); so += seqs[i].length; }
			if (refs[i] != null && refs[i].length > 0) { ambW += packBases(refs[i], ro, pinRef2, pinRefN); ro += refs[i].length; }
			pinSeqOff.putLong(so); pinRefOff.putLong(ro);
		}
		int[] start = new int[n];
		String[] cigar = new String[n];
		// a blob without a single ambiguous base travels without its mask
		int rc = alignBatchPacked(handle, n, pinSeq2, ambQ > 0 ? pinSeqN : null, pinSeqOff, pinRef2, ambW > 0 ? pinRefN : null, pinRefOff,
				start, cigar, status);
		if (rc != 0) throw new IllegalStateException("cudasw_align_batch_packed failed (" + rc + "): " + lastError(handle));
		Alignment[] out = new Alignment[n];
		for (int i = 0; i < n; i++) {
			if (status[i] == STATUS_OK && cigar[i] != null) out[i] = new Alignment(start[i], cigar[i]);
			else if (status[i] == STATUS_OK) status[i] = -3;
		}
		return out;
	}
	/**
	 * Packs the reference into GPU memory once (2 bits per base + ambiguity bit); replaces the per-window
	 * ReferenceLookup.getSubsequenceAt + reverseComplement work of SAMRecordUtil.realign and BreakpointHomology.calculate.
	 * contigOff has one entry per contig plus the total length.
	 */
	public void loadReference(byte[] bases, long[] contigOff) {
		int rc = loadReference(handle, bases, contigOff);
		if (rc != 0) throw new IllegalStateException("cudasw_reference_load failed (" + rc + "): " + lastError(handle));
	}
	/** A window made of reference segments [start, end) (0-based) of contig referenceIndex, optionally reverse-complemented. */
	public static final class Window {
		public final int[] contig; public final boolean[] revcomp; public final long[] start, end;
		public Window(int[] contig, boolean[] revcomp, long[] start, long[] end) { this.contig = contig; this.revcomp = revcomp; this.start = start; this.end = end; }
		public static Window of(int contig, long start, long end, boolean revcomp) {
			return new Window(new int[] { contig }, new boolean[] { revcomp }, new long[] { start }, new long[] { end });
		}
	}
	/** alignBatch with the windows named instead of shipped: one JNI round trip, ~24 bytes per window segment. */
	public Alignment[] alignBatch(byte[][] seqs, Window[] windows) {
		int n = seqs.length, nseg = 0;
		long[] seqOff = new long[n + 1], winOff = new long[n + 1];
		for (int i = 0; i < n; i++) {
			seqOff[i + 1] = seqOff[i] + seqs[i].length;
			nseg += windows[i].contig.length;
			winOff[i + 1] = nseg;
		}
		byte[] seqBlob = new byte[(int) seqOff[n]];
		int[] segContig = new int[nseg], segFlags = new int[nseg];
		long[] segStart = new long[nseg], segEnd = new long[nseg];
		for (int i = 0, k = 0; i < n; i++) {
			System.arraycopy(seqs[i], 0, seqBlob, (int) seqOff[i], seqs[i].length);
			for (int j = 0; j < windows[i].contig.length; j++, k++) {
				segContig[k] = windows[i].contig[j]; segFlags[k] = windows[i].revcomp[j] ? 1 : 0;
				segStart[k] = windows[i].start[j]; segEnd[k] = windows[i].end[j];
			}
		}
		int[] start = new int[n], status = new int[n];
		String[] cigar = new String[n];
		int rc = alignBatchWindows(handle, n, seqBlob, seqOff, segContig, segFlags, segStart, segEnd, winOff, start, cigar, status);
		if (rc != 0) throw new IllegalStateException("cudasw_align_batch_windows failed (" + rc + "): " + lastError(handle));
		Alignment[] out = new Alignment[n];
		for (int i = 0; i < n; i++) {
			if (status[i] == -1) throw new IllegalArgumentException("seq and ref must be non-zero size and inside the reference");
			if (status[i] != 0 || cigar[i] == null) throw new IllegalStateException("cudasw pair status " + status[i]);
			out[i] = new Alignment(start[i], cigar[i]);
		}
		return out;
	}
	/**
	 * BreakpointHomology.calculate(lookup, bs, insertedSequence, maxBreakendLength, margin) for a whole list of exact
	 * breakpoints in one native call (replaces the per-variant loop of AnnotateInexactHomology /
	 * AnnotateInexactHomologyBedpe.java:68). Arrays are parallel; dir 0 = BreakendDirection.Forward, 1 = Backward;
	 * positions are 1-based; insertedSequence[i] may be null. Returns {localHomologyLength[], remoteHomologyLength[]}.
	 */
	public int[][] breakpointHomology(int[] referenceIndex, int[] direction, long[] position, int[] referenceIndex2, int[] direction2,
			long[] position2, String[] insertedSequence, int maxBreakendLength, int margin) {
		int n = referenceIndex.length;
		long[] insOff = new long[n + 1];
		for (int i = 0; i < n; i++) insOff[i + 1] = insOff[i] + (insertedSequence == null || insertedSequence[i] == null ? 0 : insertedSequence[i].length());
		byte[] insBlob = new byte[(int) Math.max(1, insOff[n])];
		for (int i = 0; i < n; i++) {
			if (insOff[i + 1] > insOff[i]) {
				byte[] b = insertedSequence[i].getBytes(java.nio.charset.StandardCharsets.US_ASCII);
				System.arraycopy(b, 0, insBlob, (int) insOff[i], b.length);
			}
		}
		int[] local = new int[n], remote = new int[n];
		int rc = breakpointHomology(handle, n, referenceIndex, direction, position, referenceIndex2, direction2, position2, insBlob, insOff,
				maxBreakendLength, margin, local, remote);
		if (rc == -1) throw new IllegalArgumentException("Breakpoint position must be exact and inside the reference: " + lastError(handle));
		if (rc != 0) throw new IllegalStateException("cudasw_breakpoint_homology failed (" + rc + "): " + lastError(handle));
		return new int[][] { local, remote };
	}
	@Override
	public synchronized void close() {
		for (ByteBuffer b : new ByteBuffer[] { pinSeq2, pinSeqN, pinRef2, pinRefN, pinSeqOff, pinRefOff }) if (b != null) freePinned(b);
		pinSeq2 = pinSeqN = pinRef2 = pinRefN = pinSeqOff = pinRefOff = null;
		if (handle != 0) { destroy(handle); handle = 0; }
	}
	private static native long create(int match, int mismatch, int ambiguous, int gapOpen, int gapExtend, int device);
	private static native void destroy(long handle);
	private static native String lastError(long handle);
	private static native ByteBuffer allocPinned(long bytes);
	private static native void freePinned(ByteBuffer buf);
	/** cudasw_pack_bases: ascii as bases [base0, base0 + ascii.length) of the packed blob; returns the ambiguous bases seen. */
	private static native long packBases(byte[] ascii, long base0, ByteBuffer twoBit, ByteBuffer nmask);
	private static native int alignBatchPacked(long handle, int n, ByteBuffer seqTwoBit, ByteBuffer seqMask, ByteBuffer seqOff,
			ByteBuffer refTwoBit, ByteBuffer refMask, ByteBuffer refOff, int[] startOut, String[] cigarOut, int[] statusOut);
	private static native int loadReference(long handle, byte[] bases, long[] contigOff);
	private static native int alignBatchWindows(long handle, int n, byte[] seqBlob, long[] seqOff, int[] segContig, int[] segFlags,
			long[] segStart, long[] segEnd, long[] winOff, int[] startOut, String[] cigarOut, int[] statusOut);
	private static native int breakpointHomology(long handle, int n, int[] contig1, int[] dir1, long[] pos1, int[] contig2, int[] dir2,
			long[] pos2, byte[] insBlob, long[] insOff, int maxBreakendLength, int margin, int[] localOut, int[] remoteOut);
}
