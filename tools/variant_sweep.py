"""GPU experiment: forward-kernel throughput of alternative (G, R) variants for a fixed shape."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import sys, json, numpy as np
sys.path.insert(0, %r)
import gridss_b200 as g
from tools import synth as S
cfg, n = int(sys.argv[1]), int(sys.argv[2])
al = g.CudaAligner()
sb, so, rb, ro = S.generate(cfg, n)
for _ in range(3):
    al.align_blobs(sb, so, rb, ro)
st = al.stats()
q = int(so[1]-so[0])
print(json.dumps({"variant": al.variant_for(q), "fwd_ms": round(st["ms_forward"],3), "rev_ms": round(st["ms_reverse"],3), "tb_ms": round(st["ms_traceback"],3),
  "fwd_TCUPS": round(st["cells"]/st["ms_forward"]/1e9,3)}))
''' % ROOT

if __name__ == "__main__":
    cfg, n = sys.argv[1], sys.argv[2]
    for v in sys.argv[3:]:
        env = dict(os.environ)
        if v != "default":
            env["CUDASW_VARIANTS"] = v
        out = subprocess.run([sys.executable, "-c", CODE, cfg, n], env=env, capture_output=True, text=True)
        print(v, out.stdout.strip(), out.stderr.strip()[-300:], flush=True)
