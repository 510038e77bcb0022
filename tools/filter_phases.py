"""GPU experiment: phase times of device-resident batches (one chunk: phases do not overlap) with the k-mer filter."""
import os, sys
sys.path.insert(0, ".")
import numpy as np, torch
import gridss_b200 as g
from tools import synth as S

runs = [(2, 400000), (1, 400000), (3, 100000), (4, 400000)]
if len(sys.argv) > 1:
    runs = [tuple(int(x) for x in a.split(":")) for a in sys.argv[1:]]
for cfg, n in runs:
    sb, so, rb, ro = S.generate(cfg, n)
    d = [torch.from_numpy(x).cuda() for x in (sb, so, rb, ro)]
    torch.cuda.synchronize()
    for flt in os.environ.get("EXP_FILTER", "1,0").split(","):
        for est in os.environ.get("EXP_EST", "default").split(","):
            if est == "default":
                os.environ.pop("CUDASW_FILTER_EST", None)
            else:
                os.environ["CUDASW_FILTER_EST"] = est
            os.environ["CUDASW_FILTER"] = flt
            al = g.CudaAligner()
            for rep in range(3):
                al.align_blobs_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), so, ro)
                st = al.stats()
            print(f"C{cfg} n {n} filter {flt} est {est}: total {st['ms_total']:.2f} pack {st['ms_pack']:.2f} fwd {st['ms_forward']:.2f} rev {st['ms_reverse']:.2f} "
                  f"tb {st['ms_traceback']:.2f} d2h {st['ms_d2h']:.2f} padded/cells {st['forward_cells_padded'] / max(st['cells'], 1):.3f}", flush=True)
            al.close()
            if flt == "0":
                break
