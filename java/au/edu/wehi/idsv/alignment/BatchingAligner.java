package au.edu.wehi.idsv.alignment;

import java.util.ArrayList;
import java.util.List;
import java.util.concurrent.CompletableFuture;

/**
 * Batch-collect layer for the worker-pool callers (ParallelTransformIterator workers in
 * AnnotateInexactHomology.java:17-21, SAMRecordUtil.realign): each worker parks on a future, one
 * JNI round trip is made per batch (size- or time-triggered), modelled on BwaStreamingAligner.java:79-127.
 */
public class BatchingAligner implements Aligner {
	private final CudaAligner aligner;
	private final int batchSize;
	private final long maxWaitNanos;
	private final List<byte[]> seqs = new ArrayList<>(), refs = new ArrayList<>();
	private final List<CompletableFuture<Alignment>> futures = new ArrayList<>();
	private long oldest;
	public BatchingAligner(CudaAligner aligner, int batchSize, long maxWaitMillis) {
		this.aligner = aligner; this.batchSize = batchSize; this.maxWaitNanos = maxWaitMillis * 1_000_000L;
	}
	public CompletableFuture<Alignment> submit(byte[] seq, byte[] ref) {
		CompletableFuture<Alignment> f = new CompletableFuture<>();
		List<byte[]> s = null, r = null; List<CompletableFuture<Alignment>> fs = null;
		synchronized (this) {
			if (seqs.isEmpty()) oldest = System.nanoTime();
			seqs.add(seq); refs.add(ref); futures.add(f);
			if (seqs.size() >= batchSize || System.nanoTime() - oldest > maxWaitNanos) {
				s = new ArrayList<>(seqs); r = new ArrayList<>(refs); fs = new ArrayList<>(futures);
				seqs.clear(); refs.clear(); futures.clear();
			}
		}
		if (s != null) dispatch(s, r, fs);
		return f;
	}
	public void flush() {
		List<byte[]> s, r; List<CompletableFuture<Alignment>> fs;
		synchronized (this) {
			if (seqs.isEmpty()) return;
			s = new ArrayList<>(seqs); r = new ArrayList<>(refs); fs = new ArrayList<>(futures);
			seqs.clear(); refs.clear(); futures.clear();
		}
		dispatch(s, r, fs);
	}
	private void dispatch(List<byte[]> s, List<byte[]> r, List<CompletableFuture<Alignment>> fs) {
		try {
			Alignment[] out = aligner.alignBatch(s.toArray(new byte[0][]), r.toArray(new byte[0][]));
			for (int i = 0; i < out.length; i++) fs.get(i).complete(out[i]);
		} catch (RuntimeException e) {
			for (CompletableFuture<Alignment> f : fs) f.completeExceptionally(e);
		}
	}
	@Override
	public Alignment align_smith_waterman(byte[] seq, byte[] ref) {
		CompletableFuture<Alignment> f = submit(seq, ref);
		if (!f.isDone()) flush();
		return f.join();
	}
}
