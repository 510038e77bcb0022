#!/bin/bash
# One gpurun call: where does C3's forward pass spend its time, with and without quiet chains (serialised per-kernel
# times from an ncu launch list, and whole-batch times with and without spreading the variant launches over streams).
O=gpurun_out; mkdir -p $O
for flags in "" "-DCUSW_NO_CHAIN"; do
  tag=chain; [ -n "$flags" ] && tag=nochain
  CUDASW_NVCC_EXTRA="$flags" python gridss_b200/build.py --force > /dev/null 2>&1
  echo "flags [$flags]"
  for env in "" "CUDASW_NO_SPREAD=1"; do
    echo " env [$env]"
    env $env timeout 400 python tools/bench_configs.py 3:1000000 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('  ',{k:d[k] for k in ('cfg','n','ms_total','ms_forward','ms_reverse','ms_traceback','GCUPS_fwd')})
"
  done
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:sw_score_kernel --csv --log-file $O/c3_launches_$tag.csv python tools/bench_configs.py 3:300000 > /dev/null 2>&1
  python - $O/c3_launches_$tag.csv <<'P' > $O/c3_kernels_$tag.txt
import csv,sys,re,collections
rows=list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h=rows[0]; ki,vi,ui=h.index("Kernel Name"),h.index("Metric Value"),h.index("Metric Unit")
agg=collections.OrderedDict()
for r in rows[1:]:
    k=re.sub(r"\(.*","",r[ki]); v=float(r[vi].replace(",",""))
    v = v/1e6 if r[ui] in ("ns","nsecond") else v/1e3 if r[ui] in ("us","usecond") else v
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=v
for k,a in agg.items(): print(f"{a[1]:.3f}\t{a[0]}\t{k}")
P
done 2>&1 | tee $O/c3_probe.txt
python gridss_b200/build.py --force > /dev/null 2>&1
