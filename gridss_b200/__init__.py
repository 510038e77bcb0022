"""gridss_b200 - B200-native batched Smith-Waterman engine behind GRIDSS's `Aligner` interface.

Host-side mirror of the reference interface for the hot path only
(reference: src/main/java/au/edu/wehi/idsv/alignment/{Aligner,Alignment,AlignerFactory,
SswJniAligner,StreamingAligner}.java). All arithmetic happens in libcudasw.so (hand-written
sm_100a CUDA kernels, C ABI in include/cudasw.h). There is no CPU fallback: importing works
anywhere, constructing an aligner without the CUDA library or without a GPU raises.
"""
from .aligner import (  # noqa: F401
    Aligner,
    AlignerFactory,
    Alignment,
    BREAKPOINT_DTYPE,
    BatchingAligner,
    CudaAligner,
    CudaSwError,
    MultiGpuAligner,
    RESULT_DTYPE,
    WINDOW_DTYPE,
    WINDOW_REVCOMP,
    cigar_string,
    load_library,
    pack_bases,
    packed_words,
)
from .callers import (  # noqa: F401
    BreakendSummary,
    BreakpointSummary,
    SmithWatermanStreamingAligner,
    breakpoint_array,
    breakpoint_homology_batch,
    breakpoint_homology_device,
    realign_batch,
)

__all__ = [
    "Aligner",
    "BreakendSummary",
    "BreakpointSummary",
    "BREAKPOINT_DTYPE",
    "SmithWatermanStreamingAligner",
    "breakpoint_array",
    "breakpoint_homology_batch",
    "breakpoint_homology_device",
    "realign_batch",
    "AlignerFactory",
    "Alignment",
    "BatchingAligner",
    "CudaAligner",
    "CudaSwError",
    "MultiGpuAligner",
    "RESULT_DTYPE",
    "WINDOW_DTYPE",
    "WINDOW_REVCOMP",
    "cigar_string",
    "load_library",
    "pack_bases",
    "packed_words",
]
