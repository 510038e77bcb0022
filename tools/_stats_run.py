import sys; sys.path.insert(0,'.')
import gridss_b200 as g
from tools import synth as S
al=g.CudaAligner()
for cfg,n in ((2,200000),(1,200000),(3,50000)):
    sb,so,rb,ro=S.generate(cfg,n)
    print("cfg",cfg,flush=True)
    al.align_blobs(sb,so,rb,ro)
