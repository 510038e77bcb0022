#!/bin/bash
# multi-GPU evidence (N = $1): bench of record under torchrun, the one-process multi handle against rank-per-GPU rates
N=${1:-2}
TAG=${2:-r02}
O=gpurun_out
mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-worst-case > $O/bench_${N}gpu_$TAG.json 2> $O/bench_${N}gpu_$TAG.err; echo "bench x$N exit $?"
tail -c 300 $O/bench_${N}gpu_$TAG.err
timeout 300 python tools/multi_probe.py 1000000 4 > $O/multi_probe_${N}gpu_1m_$TAG.json 2>&1; echo "probe 1M exit $?"
timeout 300 python tools/multi_probe.py 4000000 4 > $O/multi_probe_${N}gpu_4m_$TAG.json 2>&1; echo "probe 4M exit $?"
timeout 300 python tools/multi_probe.py 10000 1 > $O/multi_probe_${N}gpu_10k_$TAG.json 2>&1; echo "probe 10k exit $?"
cat $O/multi_probe_${N}gpu_*_$TAG.json | cut -c1-900
python - <<PY
import json
b=json.load(open("$O/bench_${N}gpu_$TAG.json"))
print({k:(b[k] if not isinstance(b[k],dict) else {kk:b[k][kk] for kk in ("value","h2d_bytes_per_step","ms_per_step_wall") if kk in b[k]}) for k in ("value","ms_per_step","e2e","e2e_ascii","e2e_windows")})
PY
