package au.edu.wehi.idsv.alignment;

import java.util.ArrayList;
import java.util.List;
import java.util.concurrent.CompletableFuture;

/**
 * Batch-collect layer for the worker-pool callers (ParallelTransformIterator workers in
 * AnnotateInexactHomology.java:17-21, SAMRecordUtil.realign), modelled on BwaStreamingAligner.java:79-127: a single
 * driver thread turns the stream of submitted pairs into GPU batches, flushing when `batchSize` pairs are pending or the
 * oldest pending pair has waited `maxWaitMillis`. Callers only submit and park on their future - they never flush, so a
 * worker's call does not cut the batch short - and a pair that fails (an empty sequence, an oversize query) completes
 * only ITS future exceptionally: calls to the reference's aligners are independent, and stay so here.
 */
public class BatchingAligner implements Aligner, AutoCloseable {
	private final CudaAligner aligner;
	private final int batchSize;
	private final long maxWaitNanos;
	private List<byte[]> seqs = new ArrayList<>(), refs = new ArrayList<>();
	private List<CompletableFuture<Alignment>> futures = new ArrayList<>();
	private long oldest;
	private boolean closed;
	private final Thread driver;
	public BatchingAligner(CudaAligner aligner, int batchSize, long maxWaitMillis) {
		this.aligner = aligner; this.batchSize = Math.max(1, batchSize); this.maxWaitNanos = Math.max(0, maxWaitMillis) * 1_000_000L;
		driver = new Thread(this::drive, "cudasw-batcher");
		driver.setDaemon(true);
		driver.start();
	}
	public CompletableFuture<Alignment> submit(byte[] seq, byte[] ref) {
		CompletableFuture<Alignment> f = new CompletableFuture<>();
		synchronized (this) {
			if (closed) throw new IllegalStateException("aligner closed");
			if (seqs.isEmpty()) oldest = System.nanoTime();
			seqs.add(seq); refs.add(ref); futures.add(f);
			// wake the driver for the first pair (it starts its deadline) and for a full batch
			if (seqs.size() == 1 || seqs.size() >= batchSize) notifyAll();
		}
		return f;
	}
	/** Dispatches what is pending now (end of input); does not wait for the results. */
	public synchronized void flush() {
		if (!seqs.isEmpty()) { oldest = System.nanoTime() - maxWaitNanos - 1; notifyAll(); }
	}
	private void drive() {
		for (;;) {
			List<byte[]> s, r; List<CompletableFuture<Alignment>> fs;
			synchronized (this) {
				try {
					for (;;) {
						if (closed && seqs.isEmpty()) return;
						if (seqs.isEmpty()) { wait(); continue; }
						long left = maxWaitNanos - (System.nanoTime() - oldest);
						if (closed || seqs.size() >= batchSize || left <= 0) break;
						wait(left / 1_000_000L, (int) (left % 1_000_000L));
					}
				} catch (InterruptedException e) {
					Thread.currentThread().interrupt();
					closed = true;
				}
				s = seqs; r = refs; fs = futures;
				seqs = new ArrayList<>(); refs = new ArrayList<>(); futures = new ArrayList<>();
			}
			if (!s.isEmpty()) dispatch(s, r, fs);
			if (Thread.currentThread().isInterrupted()) return;
		}
	}
	private void dispatch(List<byte[]> s, List<byte[]> r, List<CompletableFuture<Alignment>> fs) {
		try {
			int[] status = new int[s.size()];
			Alignment[] out = aligner.alignBatch(s.toArray(new byte[0][]), r.toArray(new byte[0][]), status);
			for (int i = 0; i < out.length; i++) {
				if (status[i] == CudaAligner.STATUS_OK) fs.get(i).complete(out[i]);
				else fs.get(i).completeExceptionally(CudaAligner.failure(status[i]));
			}
		} catch (RuntimeException | Error e) {  // the whole native call failed (CUDA error, out of memory)
			for (CompletableFuture<Alignment> f : fs) f.completeExceptionally(e);
		}
	}
	@Override
	public Alignment align_smith_waterman(byte[] seq, byte[] ref) {
		if (seq == null || seq.length == 0) throw new IllegalArgumentException("seq must be non-zero size");  // SswJniAligner.java:45-50
		if (ref == null || ref.length == 0) throw new IllegalArgumentException("ref must be non-zero size");
		try {
			return submit(seq, ref).join();  // parks; the driver thread flushes by size or age
		} catch (java.util.concurrent.CompletionException e) {
			if (e.getCause() instanceof RuntimeException) throw (RuntimeException) e.getCause();
			throw e;
		}
	}
	@Override
	public void close() {
		synchronized (this) { closed = true; notifyAll(); }
		try { driver.join(); } catch (InterruptedException e) { Thread.currentThread().interrupt(); }
		aligner.close();
	}
}
