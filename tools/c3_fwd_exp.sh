for e in "" "CUDASW_NO_SPREAD=1" "CUDASW_CHAIN=1" "CUDASW_CHAIN=1 CUDASW_NO_SPREAD=1" "CUDASW_VARIANTS=2x16,4x16,8x16,16x16,32x16,32x32" "CUDASW_VARIANTS=2x16,4x16,8x16,16x16,32x16,32x32 CUDASW_CHAIN=1"; do
  echo "== $e"
  env $e python tools/bench_configs.py 3:1000000 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('fwd ms',d['ms_forward'],'TCUPS',d['GCUPS_fwd']/1000,'total',d['ms_total'])
"
done
