"""Regenerates tests/golden/*.json. Run in the build container (needs /root/reference).

golden_reference_tests.json : (seq, ref) -> (start, cigar) vectors copied from the reference's own
    JUnit tests at the Aligner boundary (file:line in each entry), plus the scores/outputs the
    reference's shipped libsswjni.so produces for them here.
golden_sswjni_fuzz.json     : small adversarial + synthetic sets with the outputs of the reference's
    shipped libsswjni.so (oracle/_ref), so the oracle restatement stays pinned on machines where
    /root/reference is absent.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from tools import synth as S  # noqa: E402

CHR9 = (
    "cctccttaggatgcaaggttggtccagcatgtgcaaatcaataaatgtaatacatcacat"
    "aaacagaactaaaaacaaaaaacagacgattatcccaatagatgcagaaaataataaaag"
    "ccgtgtataacaaacccacagccaacaatctgttgaatgggcaaaagttgggagcgtttt"
    "ttttttgaaaaccagcacaaggcaagaatgccctctctcaccatttctattcaacatagt"
    "attgaaagtcctgaccagggcaatcaggctagagaaagaaatacaaaaagcatccaaata"
    "ggaagagagaaagccaaactatttctgtttgcagataacataattctatatctagaaaac"
    "cctgtagtttcagctacaaagcttctttagttaacaaacaacttcagcgaattttcaaga"
    "tactaaatcaatgtgcaaagatcactaacatttctctacaccaatgataaccacactgac"
    "agcaaaatcagaaaggccatc"
)
POLY = "A" * 100
# (cite, seq, ref, expected start or None, expected cigar)
REFERENCE_TESTS = [
    ("SmithWatermanAlignerTest.java:15-19", "AAACCCCCCCCCCCCCTTAATTAATTAATTTT", "AAACCCCCCCCCCCCCGTTAATTAATTAATTTT", 0, "16M1D16M"),
    ("SmithWatermanAlignerTest.java:23-25", "AACCCTTTTTT", "AAACCCTTTTTT", 1, "11M"),
    ("SmithWatermanAlignerTest.java:29-33", "GGGGGGTTTTTT", "AAACCCTTTTTT", 6, "6S6M"),
    ("AlignmentHelperTest.java:12-16", "GTACC", "gtacC", None, "5M"),
    ("AlignmentHelperTest.java:19-23", "TTTGGCATGTTTTTGCAGTGGCTGGGGGGGGGTGGTTTTT",
     "tgcagtttcttcctagcattgatggtctttacaatttggcatgtttttgcagtggctgggaccagttgttcctttccatgtttagtgcttccttcaggagc", None, "26M14S"),
    ("AlignmentHelperTest.java:26-40",
     "TAGCAAGAGAGAAAGCCAAACTATTTCTGTTTGCAGATAACATAATTCTATATCTAGAAAACCCTGTAGTTTCAGGTACAAAACCCTGTAGTTTCAGGTACAATGTGCAAAGATCACTAACATTTCT",
     CHR9, None, "4S78M45S"),
    ("AlignerFactory.java:74-84 (self-test)", "A", "A", 0, "1M"),
    ("SURVEY.md §8a A5 (observed on libsswjni.so)", "AAAA", "CCCCCC", -1, "1M3S"),
    ("SURVEY.md §8a A5 (N matches N)", "ACGTNNACGT", "ACGTNNACGT", 0, "10M"),
]


def main():
    O.build()
    seqs = [t[1].encode() for t in REFERENCE_TESTS]
    refs = [t[2].encode() for t in REFERENCE_TESTS]
    r = O.ref_align(seqs, refs)
    out = []
    for i, t in enumerate(REFERENCE_TESTS):
        score, start, cigar = r.tuple(i)
        assert cigar == t[4], (t, cigar)
        assert t[3] is None or start == t[3], (t, start)
        out.append({"cite": t[0], "seq": t[1], "ref": t[2], "score": score, "start": start, "cigar": cigar})
    json.dump(out, open(os.path.join(HERE, "golden_reference_tests.json"), "w"), indent=1)

    fuzz = []
    for kind in S.ADVERSARIAL_KINDS:
        seqs, refs = S.adversarial(kind, 40, seed=101, qmax=120, wmax=240)
        for sc in ((1, -4, -4, 6, 1), (2, -6, -1, 5, 3)):
            r = O.ref_align(seqs, refs, sc)
            for i in range(len(seqs)):
                score, start, cigar = r.tuple(i)
                fuzz.append({"kind": kind, "scoring": sc, "seq": seqs[i].decode("latin1"), "ref": refs[i].decode("latin1"),
                             "score": score, "start": start, "cigar": cigar})
    for cfg, n in ((1, 60), (2, 40), (3, 12)):
        sb, so, rb, ro = S.generate(cfg, n)
        r = O.ref_align_blobs(sb, so, rb, ro)
        for i in range(n):
            score, start, cigar = r.tuple(i)
            fuzz.append({"kind": f"C{cfg}", "scoring": (1, -4, -4, 6, 1),
                         "seq": bytes(sb[so[i]:so[i + 1]]).decode("latin1"), "ref": bytes(rb[ro[i]:ro[i + 1]]).decode("latin1"),
                         "score": score, "start": start, "cigar": cigar})
    json.dump(fuzz, open(os.path.join(HERE, "golden_sswjni_fuzz.json"), "w"))
    print(len(out), "reference-test vectors;", len(fuzz), "sswjni fuzz vectors")


if __name__ == "__main__":
    main()
