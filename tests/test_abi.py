"""CPU: the C-ABI library loads, exports every symbol include/cudasw.h declares, and refuses to run
without a GPU (no CPU fallback). No compute calls here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "cudasw.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cudasw_[a-z_]+)\s*\(", src)))


def test_header_declares_expected_entry_points():
    names = _declared()
    for must in ("cudasw_create", "cudasw_destroy", "cudasw_align_batch", "cudasw_align_batch_device",
                 "cudasw_submit", "cudasw_flush", "cudasw_poll", "cudasw_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import gridss_b200 as g
    from gridss_b200 import aligner

    lib = g.load_library()
    names = _declared()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cudasw.h but not exported"
    assert sorted(n for n, _, _ in aligner.ABI) == names


def test_result_struct_layout_matches_header():
    import gridss_b200 as g

    assert g.RESULT_DTYPE.itemsize == 32
    assert g.RESULT_DTYPE.names == ("score", "ref_begin", "ref_end", "read_begin", "read_end", "cigar_off", "cigar_len", "status")


def test_cigar_string_host_helper():
    import numpy as np

    import gridss_b200 as g

    lib = g.load_library()
    ops = np.array([(4 << 4) | 4, (78 << 4) | 0, (1 << 4) | 2, (3 << 4) | 1, (45 << 4) | 4], dtype=np.uint32)
    buf = ctypes.create_string_buffer(64)
    n = lib.cudasw_cigar_string(ops.ctypes.data, len(ops), buf, 64)
    assert buf.value[:n] == b"4S78M1D3I45S"
    assert g.cigar_string(ops) == "4S78M1D3I45S"
    assert lib.cudasw_cigar_string(ops.ctypes.data, len(ops), buf, 4) < 0


def test_no_cpu_fallback_without_gpu():
    import gridss_b200 as g

    lib = g.load_library()
    if lib.cudasw_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(g.CudaSwError):
        g.CudaAligner()


def test_scoring_validation_mirrors_reference():
    import gridss_b200 as g

    # ssw.Aligner.align: "Gap open and extension penalties must ..." / "Scoring matrix values must fit into signed 8-bit"
    for bad in ((1, -4, -4, 256, 1), (1, -4, -4, 6, -1), (200, -4, -4, 6, 1)):
        with pytest.raises(ValueError):
            g.CudaAligner(*bad)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "gridss_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in txt.replace("SURVEY", ""), f"{f} mentions the oracle"


def test_jni_glue_type_checks_against_the_header():
    """java/jni/cudaswjni.c (the reference-side binding of INTEGRATION.md) compiles against include/cudasw.h; a
    minimal stand-in for <jni.h> is used because the build image has no JDK."""
    import subprocess

    src = os.path.join(ROOT, "java", "jni", "cudaswjni.c")
    subprocess.check_call(["gcc", "-fsyntax-only", "-Wall", "-Werror", "-Wno-unused-parameter",
                           "-I" + os.path.join(ROOT, "tests", "mock_jni"), "-I" + os.path.join(ROOT, "include"), src])
    # every native method CudaAligner.java declares has its JNI symbol in the glue
    java = open(os.path.join(ROOT, "java", "au", "edu", "wehi", "idsv", "alignment", "CudaAligner.java")).read()
    glue = open(src).read()
    for name in re.findall(r"private static native [\w.]+(?:\[\])? (\w+)\(", java):
        assert f"Java_au_edu_wehi_idsv_alignment_CudaAligner_{name}(" in glue, name


def _np_pack(blob):
    """numpy statement of the packed layout of include/cudasw.h (2 bits per base + N mask)."""
    import numpy as np

    lut = np.full(256, 4, dtype=np.uint32)
    for k, c in enumerate(b"ACGT"):
        lut[c] = k
        lut[c + 32] = k
    code = lut[blob]
    n = len(code)
    pad = (-n) % 32
    c2 = np.concatenate([code & 3, np.zeros(pad, dtype=np.uint32)]).reshape(-1, 16)
    nm = np.concatenate([code >> 2, np.zeros(pad, dtype=np.uint32)]).reshape(-1, 32)
    two = (c2 << (2 * np.arange(16, dtype=np.uint32))).sum(axis=1, dtype=np.uint64).astype(np.uint32)
    msk = (nm.astype(np.uint64) << np.arange(32, dtype=np.uint64)).sum(axis=1, dtype=np.uint64).astype(np.uint32)
    return two, msk, int((code >> 2).sum())


def test_host_packer_matches_the_declared_layout():
    """cudasw_pack_bases is pure host code (what the JNI shim calls while collecting a batch): whole blobs, appends at
    unaligned base offsets, lower case, IUPAC codes and N runs, threads."""
    import numpy as np

    import gridss_b200 as g

    rng = np.random.default_rng(11)
    alphabet = np.frombuffer(b"ACGTacgtNnRYKM-*", dtype=np.uint8)
    for n in (1, 15, 16, 17, 31, 32, 33, 1000, 4099):
        blob = alphabet[rng.integers(0, len(alphabet), size=n)].copy()
        two, msk, amb = g.pack_bases(blob)
        w2, wm, a2 = _np_pack(blob)
        assert np.array_equal(two[: len(w2)], w2) and np.array_equal(msk[: len(wm)], wm) and amb == a2, n
    # pair-by-pair appends into dirty buffers == one shot
    lens = [int(x) for x in rng.integers(1, 90, size=200)]
    blob = alphabet[rng.integers(0, len(alphabet), size=sum(lens))].copy()
    w2, wm = g.packed_words(len(blob))
    two = np.full(w2, 0xFFFFFFFF, dtype=np.uint32)
    msk = np.full(wm, 0xFFFFFFFF, dtype=np.uint32)
    at, amb = 0, 0
    for ln in lens:
        _, _, a = g.pack_bases(blob[at: at + ln], two, msk, base0=at)
        amb += a
        at += ln
    e2, em, ea = _np_pack(blob)
    assert np.array_equal(two[: len(e2)], e2) and np.array_equal(msk[: len(em)], em) and amb == ea
    # threaded == serial
    big = alphabet[rng.integers(0, 8, size=(1 << 21) + 77)].copy()
    a = g.pack_bases(big, threads=1)
    b = g.pack_bases(big, threads=4)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2]


def test_reference_side_patches_apply_and_the_contract_test_exists():
    """N2: the AlignerFactory / Defaults wiring is a real patch against the reference sources, and the
    reference's aligner contract test has a subclass for the new backend. Needs /root/reference (not on the GPU box)."""
    import shutil
    import subprocess
    import tempfile

    ref = "/root/reference"
    jdir = os.path.join(ROOT, "java")
    test_src = open(os.path.join(jdir, "test", "au", "edu", "wehi", "idsv", "alignment", "CudaAlignerTest.java")).read()
    assert "extends SmithWatermanAlignerTest" in test_src and "new CudaAligner(" in test_src
    # every class / member the patched factory refers to exists in the shim sources
    fac = open(os.path.join(jdir, "patches", "AlignerFactory.java.patch")).read()
    dfl = open(os.path.join(jdir, "patches", "Defaults.java.patch")).read()
    cuda = open(os.path.join(jdir, "au", "edu", "wehi", "idsv", "alignment", "CudaAligner.java")).read()
    batching = open(os.path.join(jdir, "au", "edu", "wehi", "idsv", "alignment", "BatchingAligner.java")).read()
    for name in re.findall(r"Defaults\.(CUDASW_\w+|USE_CUDASW)", fac):
        assert f" {name};" in dfl and f"{name} = " in dfl, name
    assert "public CudaAligner(int match, int mismatch, int ambiguous, int gapOpen, int gapExtend, int device)" in cuda
    assert "public BatchingAligner(CudaAligner aligner, int batchSize, long maxWaitMillis)" in batching
    assert "implements Aligner, AutoCloseable" in cuda and "isCudaswLoaded" in fac and "isCudaswLoaded" in test_src
    if not os.path.isdir(ref):
        pytest.skip("reference tree not present")
    with tempfile.TemporaryDirectory() as tmp:
        for rel in ("src/main/java/au/edu/wehi/idsv/alignment/AlignerFactory.java", "src/main/java/au/edu/wehi/idsv/Defaults.java"):
            os.makedirs(os.path.dirname(os.path.join(tmp, rel)), exist_ok=True)
            shutil.copy(os.path.join(ref, rel), os.path.join(tmp, rel))
        for p in ("AlignerFactory.java.patch", "Defaults.java.patch"):
            subprocess.check_call(["patch", "--dry-run", "-p1", "-s", "-d", tmp, "-i", os.path.join(jdir, "patches", p)])
            subprocess.check_call(["patch", "-p1", "-s", "-d", tmp, "-i", os.path.join(jdir, "patches", p)])
        out = open(os.path.join(tmp, "src/main/java/au/edu/wehi/idsv/alignment/AlignerFactory.java")).read()
        # the new branch sits ahead of GKL / sswjni / JAligner and the fall-through chain is intact
        body = out[out.index("public static Aligner create(int match"):]
        assert body.index("cudaswLoaded") < body.index("isw != null") < body.index("sswjniLoaded") < body.index("JAlignerAligner")
        assert out.count("{") == out.count("}")
