#!/bin/bash
# round 2: routing sweep between the one-thread and the warp-cooperative traceback kernels
for tb in 16 8 4; do for t in -1 16384 65536; do
  echo "== CUDASW_TB_THREAD_BAND=$tb CUDASW_TB_WIDE_CELLS=$t"
  CUDASW_TB_THREAD_BAND=$tb CUDASW_TB_WIDE_CELLS=$t timeout 300 python tools/bench_configs.py 2:1000000 3:1000000 5:1000 1:400000 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['cfg'],d['n'],'tb',d['ms_traceback'],'rev',d['ms_reverse'],'total',d['ms_total'])
"
done; done
