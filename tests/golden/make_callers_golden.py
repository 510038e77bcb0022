"""Regenerates tests/golden/golden_callers.json. Run in the build container (needs /root/reference).

Known answers of the reference's caller-level tests, transcribed with their inline references:
  BreakpointHomologyTest.java:43-224   (expected local / remote homology lengths)
  SAMRecordUtilTest.java:375-382       (realign: expected start and CIGAR)
The `random` contig of the reference's test genome (src/test/resources/small.fa) is test data the realign case needs;
only its first 600 bases are stored, and the 251-base chr1 excerpt of issue #778. Every expected value is re-derived here by a plain string restatement of
BreakpointHomology.calculate / SAMRecordUtil.realign driven through the reference's shipped libsswjni.so
(oracle/_ref), so a transcription slip fails at generation time.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from oracle.callers_restatement import homology_strings, realign_strings  # noqa: E402

SMALL_FA = "/root/reference/src/test/resources/small.fa"


def read_fasta(path):
    out, name = {}, None
    for line in open(path):
        line = line.strip()
        if line.startswith(">"):
            name = line[1:].split()[0]
            out[name] = []
        elif name:
            out[name].append(line)
    return {k: "".join(v) for k, v in out.items()}


TWO = ["CATTAATCGCAAGAGCGGGTTGTATTCGACGCCAAGTCAGCTGAAGCACCATTACC", "CACCGCCTATTCGAACGGGCGAATCTACCTAGGTCGCTCAGAACCGGCACCCTTAA"]
HOM = ["AGAACCGGCACCCTTAACATTAATCGCAAGAGCGGGTTGTATTCGACGCCAAGTCAGCTGAAGCACCATTACC",
       "AGAACCGGCACCCTTAACACCGCCTACAAGAGCGGGTTGTATTCGACGCCAAGTCAGCTGAAGCACCATTACC"]
POLYA = ["@polyA10000"]  # expanded by the test to b"A" * 10000 (small.fa contig 0)

# (cite, contigs, (ref1, dir1, pos1, ref2, dir2, pos2), inserted, maxBreakendLength, margin, expected (local, remote))
HOMOLOGY = [
    ("BreakpointHomologyTest.java:44-56", ["CCCAATGGGCCC", "TTTAATGGGAAA"], (0, "f", 6, 1, "b", 7), "", 100, 0, (3, 3)),
    ("BreakpointHomologyTest.java:58-69", ["CCCCCCCCCCCC", "TTTTTTTTTTTTTTTTTTTTTTTT"], (0, "f", 6, 1, "b", 7), "", 100, 10, (0, 0)),
    ("BreakpointHomologyTest.java:71-82", ["CCCAATGGGCCCTTTTTTTTTTTTTTTTTTT", "TTTAATGGGAAAGGGGGGGGGGGGGGGGGGG"], (0, "f", 6, 1, "b", 7), "", 1, 5, (1, 1)),
    ("BreakpointHomologyTest.java:84-88", POLYA, (0, "f", 100, 0, "b", 103), "", 100, 0, (2, 2)),
    ("BreakpointHomologyTest.java:90-107", ["CCCAAAATTTGGGAAAAAATTTTTTTTTTTTTTTTTTT", "TTTAAAATTTGGGAAAAAAGGGGGGGGGGGGGGGGGGG"],
     (0, "f", 10, 1, "b", 11), "c", 20, 10, (7, 9)),
    ("BreakpointHomologyTest.java:133-145", TWO, (0, "b", 10, 0, "f", 20), "", 20, 10, (0, 0)),
    ("BreakpointHomologyTest.java:147-159", TWO, (0, "f", 10, 1, "b", 10), "GGGGG", 20, 10, (0, 0)),
    ("BreakpointHomologyTest.java:161-174", HOM, (0, "f", 27, 1, "b", 28), "GGGGG", 20, 5, (0, 20)),
    ("BreakpointHomologyTest.java:176-178 (remoteBreakpoint)", HOM, (1, "b", 28, 0, "f", 27), "GGGGG", 20, 5, (20, 0)),
    ("BreakpointHomologyTest.java:201-214",
     ["CATTAATCGCAAGAGCGGGCAAGAGCGGGCAAGAGCGGGCAAGAGCGGGCAAGAGCGGGTTGTATTCGACGCCAAGTCAGCTGAAGCACCATTACC"],
     (0, "f", 30, 0, "b", 41), "", 20, 10, (10, 10)),
    ("BreakpointHomologyTest.java:216-220", POLYA, (0, "f", 100, 0, "b", 500), "AAAAA", 10, 50, (10, 10)),
    ("BreakpointHomologyTest.java:221-223 (remoteBreakpoint)", POLYA, (0, "b", 500, 0, "f", 100), "AAAAA", 10, 50, (10, 10)),
]


def align_one(seq, ref):
    r = O.ref_align([seq], [ref])
    score, start, cigar = r.tuple(0)
    return start, cigar


def main():
    O.build()
    assert O.have_reference(), "oracle/_ref is needed to regenerate"
    fa = read_fasta(SMALL_FA)
    assert set(fa["polyA"]) == {"A"} and len(fa["polyA"]) == 10000
    cases = []
    for cite, contigs, bp, ins, mbl, margin, expect in HOMOLOGY:
        real = [b"A" * 10000] if contigs is POLYA else [c.encode() for c in contigs]
        got = homology_strings(align_one, real, bp, ins, mbl, margin)
        assert tuple(got) == tuple(expect), (cite, got, expect)
        cases.append({"cite": cite, "contigs": contigs, "breakpoint": list(bp), "inserted": ins, "maxBreakendLength": mbl,
                      "margin": margin, "local": expect[0], "remote": expect[1]})
    # realign. A read is (referenceIndex, alignmentStart, cigar, bases); AssemblyFactory.createAnchoredBreakend
    # (AssemblyFactory.java:186-197) makes a FWD assembly "<anchored>M<rest>S" starting at pos - anchored + 1 and a
    # BWD one "<rest>S<anchored>M" starting at pos; the tests' breakend position is the alignment end (FWD) or
    # start (BWD) of the realigned record.
    R = fa["random"]
    random600 = R[:600]
    chr1 = read_fasta("/root/reference/src/test/resources/chr1_170849600_170849850.fa")
    chr1 = next(iter(chr1.values()))
    asm778 = ("ATCCATCCCTATGACCCAAACATCTCCCACCAGGCCTCATGTTCAATATTAAAGATCACATTTCAACTTGAGATTTGGAGGGGACAAACATACAAATCATATCATT"
              "ATCTCTCTCCCCACTTCTCTCTTTATCAATCCCTCCCTCTTTGTCAATCTTAGCCTTGGCCTTCAGATTTTACCACTTGATTTTTCACATTTTCTGTATTCTTAAT"
              "GATTATTATATTTTCATGTTCTTGCTAATCTATATCATGGTTAGAAATCAAAGCATGCCGAAATTTCTCTCTTACTTTTTTTGCTGTT")
    assert len(asm778) == 300
    small = ["@polyA10000", "ACGT" * 4, "@random600"]
    realign = [
        ("SAMRecordUtilTest.java:375-382", small, (2, 101, "100M", R[100:199] + "N"), 0, True, 101, "99M1S"),
        ("SAMRecordAssemblyEvidenceTest.java:166-180 (#778)", [chr1], (0, 103, "203S97M", asm778), 50, True, 114, "212S88M"),
        ("SAMRecordAssemblyEvidenceTest.java:182-188", small, (0, 1, "10M12S", "AAAAAAAAAATTAAAAAAAAAA"), 50, True, None, "10M2I10M"),
        ("SAMRecordAssemblyEvidenceTest.java:229-237", small, (0, 1, "5M3S", "AAAATTTT"), 50, True, 1, "4M4S"),
        ("SAMRecordAssemblyEvidenceTest.java:261-266 (BWD)", small, (2, 300, "101S100M", "N" + R[179:279] + R[299:399]), 50, True, None, "1S100M20D100M"),
        ("SAMRecordAssemblyEvidenceTest.java:268-272 (FWD)", small, (2, 300, "100M101S", R[299:399] + R[419:519] + "N"), 50, True, None, "100M20D100M1S"),
        ("SAMRecordAssemblyEvidenceTest.java:274-281", small, (2, 1, "90S210M", "N" * 100 + R[0:100] + R[110:210]), 50, True, None, "100S100M10D100M"),
        ("SAMRecordAssemblyEvidenceTest.java:283-290", small, (2, 1, "110S200M", "N" * 100 + R[0:100] + "N" * 10 + R[100:200]), 50, True, None, "100S100M10I100M"),
    ]
    for startpos in range(250, 351):   # SAMRecordAssemblyEvidenceTest.java:239-258: +-50 bp sweep, both directions
        realign.append((f"SAMRecordAssemblyEvidenceTest.java:241-248 (BWD @{startpos})", small,
                        (2, startpos, "50S100M", "N" * 50 + R[299:399]), 50, True, 300, "50S100M"))
        realign.append((f"SAMRecordAssemblyEvidenceTest.java:250-257 (FWD @{startpos})", small,
                        (2, startpos, "100M50S", R[299:399] + "N" * 50), 50, True, 300, "100M50S"))
    out_realign = []
    for cite, contigs, read, window, extend, start, cigar in realign:
        real = [b"A" * 10000 if c == "@polyA10000" else random600.encode() if c == "@random600" else c.encode() for c in contigs]
        got = realign_strings(align_one, real, (read[0], read[1], read[2], read[3].encode()), window, extend)
        assert got[1] == cigar and (start is None or got[0] == start), (cite, got, start, cigar)
        out_realign.append({"cite": cite, "contigs": contigs, "read": list(read), "windowSize": window, "extend": extend,
                            "start": got[0], "cigar": cigar})
    json.dump({"random600": random600, "homology": cases, "realign": out_realign},
              open(os.path.join(HERE, "golden_callers.json"), "w"), indent=1)
    print("wrote golden_callers.json:", len(cases), "homology cases,", len(out_realign), "realign cases")


if __name__ == "__main__":
    main()
