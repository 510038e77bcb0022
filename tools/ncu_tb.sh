#!/bin/bash
# launch list (durations + lane efficiency) of the traceback kernels for the given bench_configs arguments
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,launch__grid_size,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:traceback --csv --log-file gpurun_out/tb_list.csv python tools/bench_configs.py "$@" > gpurun_out/ncu_tb.log 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(l for l in open("gpurun_out/tb_list.csv") if l.startswith('"')))
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
cur={}
for r in rows[1:]:
    cur.setdefault((int(r[ii]),r[ki][:34]),{})[r[mi]]=r[vi]
for k,v in cur.items():
    print(k[0], k[1], "ms=%.3f"%(float(v['gpu__time_duration.sum'].replace(',',''))/1e6), "grid="+v['launch__grid_size'], "lanes/inst="+v['smsp__thread_inst_executed_per_inst_executed.ratio'], "occ%="+v['sm__warps_active.avg.pct_of_peak_sustained_active'])
PY
