package au.edu.wehi.idsv.alignment;

import org.junit.Assume;
import org.junit.Test;

import static org.junit.Assert.assertEquals;
import static org.junit.Assert.assertTrue;

/**
 * Runs the reference's aligner contract (SmithWatermanAlignerTest: deletion, zero-based offset, soft clips) against the
 * CUDA backend, the way SswJniAlignerTest (src/test/java/au/edu/wehi/idsv/alignment/SswJniAlignerTest.java:5-10) runs it
 * against sswjni. Needs -Dgridss.cudaaligner=true, a GPU and libcudaswjni.so / libcudasw.so on java.library.path;
 * skipped (not failed) otherwise, like the GKL tests on machines without the library.
 */
public class CudaAlignerTest extends SmithWatermanAlignerTest {
    @Override
    protected Aligner create(int match, int mismatch, int ambiguous, int gapOpen, int gapExtend) {
        Assume.assumeTrue("cudasw not loaded (-Dgridss.cudaaligner=true and a GPU are needed)", AlignerFactory.isCudaswLoaded());
        return new CudaAligner(match, mismatch, ambiguous, gapOpen, gapExtend, 0);
    }
    @Test
    public void factory_should_hand_out_the_batching_front_end() {
        Assume.assumeTrue(AlignerFactory.isCudaswLoaded());
        assertTrue(AlignerFactory.create() instanceof BatchingAligner);
    }
    @Test
    public void batch_should_match_single_calls_and_ssw() {
        Assume.assumeTrue(AlignerFactory.isCudaswLoaded());
        byte[][] seqs = { B("AAACCCCCCCCCCCCCTTAATTAATTAATTTT"), B("AACCCTTTTTT"), B("GGGGGGTTTTTT"), B("aaccctttttt"), B("ACGTNNACGT") };
        byte[][] refs = { B("AAACCCCCCCCCCCCCGTTAATTAATTAATTTT"), B("AAACCCTTTTTT"), B("AAACCCTTTTTT"), B("AAACCCTTTTTT"), B("ACGTNNACGT") };
        try (CudaAligner cuda = new CudaAligner(1, -4, -4, 6, 1, 0)) {
            Alignment[] batch = cuda.alignBatch(seqs, refs);
            for (int i = 0; i < seqs.length; i++) {
                Alignment single = cuda.align_smith_waterman(seqs[i], refs[i]);
                assertEquals(single.getCigar(), batch[i].getCigar());
                assertEquals(single.getStartPosition(), batch[i].getStartPosition());
                if (AlignerFactory.isSswjniLoaded()) {
                    Alignment ssw = new SswJniAligner(1, -4, -4, 6, 1).align_smith_waterman(seqs[i], refs[i]);
                    assertEquals(ssw.getCigar(), batch[i].getCigar());
                    assertEquals(ssw.getStartPosition(), batch[i].getStartPosition());
                }
            }
        }
    }
    @Test(expected = IllegalArgumentException.class)
    public void empty_sequence_should_throw_like_sswjni() {
        Assume.assumeTrue(AlignerFactory.isCudaswLoaded());
        create().align_smith_waterman(new byte[0], B("ACGT"));
    }
    @Test
    public void one_bad_pair_should_not_fail_its_batch() {
        Assume.assumeTrue(AlignerFactory.isCudaswLoaded());
        try (CudaAligner cuda = new CudaAligner(1, -4, -4, 6, 1, 0)) {
            int[] status = new int[2];
            Alignment[] out = cuda.alignBatch(new byte[][] { B("AACCCTTTTTT"), new byte[0] }, new byte[][] { B("AAACCCTTTTTT"), B("ACGT") }, status);
            assertEquals(0, status[0]);
            assertEquals("11M", out[0].getCigar());
            assertEquals(-1, status[1]);
        }
    }
}
