"""Regenerates tests/golden/golden_callers.json. Run in the build container (needs /root/reference).

Known answers of the reference's caller-level tests, transcribed with their inline references:
  BreakpointHomologyTest.java:43-224   (expected local / remote homology lengths)
  SAMRecordUtilTest.java:375-382       (realign: expected start and CIGAR)
The `random` contig of the reference's test genome (src/test/resources/small.fa) is test data the realign case needs;
only its first 400 bases are stored. Every expected value is re-derived here by a plain string restatement of
BreakpointHomology.calculate / SAMRecordUtil.realign driven through the reference's shipped libsswjni.so
(oracle/_ref), so a transcription slip fails at generation time.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle as O  # noqa: E402
from callers_restatement import homology_strings, realign_strings  # noqa: E402

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
    random400 = fa["random"][:400]
    cases = []
    for cite, contigs, bp, ins, mbl, margin, expect in HOMOLOGY:
        real = [b"A" * 10000] if contigs is POLYA else [c.encode() for c in contigs]
        got = homology_strings(align_one, real, bp, ins, mbl, margin)
        assert tuple(got) == tuple(expect), (cite, got, expect)
        cases.append({"cite": cite, "contigs": contigs, "breakpoint": list(bp), "inserted": ins, "maxBreakendLength": mbl,
                      "margin": margin, "local": expect[0], "remote": expect[1]})
    # SAMRecordUtilTest.java:375-382: Read(2, 101, "100M"), bases RANDOM[100:199] + "N", window 0, extend -> 99M1S @ 101
    bases = random400[100:199] + "N"
    contigs = [b"A" * 10, b"ACGT" * 4, random400.encode()]
    got = realign_strings(align_one, contigs, (2, 101, "100M", bases.encode()), 0, True)
    assert got == (101, "99M1S"), got
    realign = [{"cite": "SAMRecordUtilTest.java:375-382", "random400": random400, "read": [2, 101, "100M", bases],
                "windowSize": 0, "extend": True, "start": 101, "cigar": "99M1S"}]
    json.dump({"homology": cases, "realign": realign}, open(os.path.join(HERE, "golden_callers.json"), "w"), indent=1)
    print("wrote golden_callers.json:", len(cases), "homology cases,", len(realign), "realign case")


if __name__ == "__main__":
    main()
