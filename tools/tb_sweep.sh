#!/bin/bash
# sweep of the traceback routing threshold (banded cells above which a pair takes the warp kernel); -1 = band rule only
for t in -1 2048 4096 8192 16384 32768 65536; do
  echo "== CUDASW_TB_WIDE_CELLS=$t"
  CUDASW_TB_WIDE_CELLS=$t timeout 300 python tools/bench_configs.py 2:1000000 3:200000 3:1000000 5:1000 1:400000 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['cfg'],d['n'],'tb',d['ms_traceback'],'total',d['ms_total'])
"
done
