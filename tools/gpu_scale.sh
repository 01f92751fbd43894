#!/bin/bash
# Scaling pass on one multi-GPU box: bench.py at N = 1, 2, 4, 8 (as many as the box has), SSSP only.
TAG=${1:-sc1}; O=gpurun_out/$TAG; mkdir -p $O
NG=$(nvidia-smi -L | wc -l); echo "gpus on box: $NG"
timeout 600 python bench.py --no-tick --no-cpu > $O/bench_n1.json 2> $O/bench_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do
  if [ $N -le $NG ]; then
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) \
      bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "n$N rc=$?"
  fi
done
for f in $O/bench_n*.json; do python - "$f" <<'PY'
import json,sys
for line in open(sys.argv[1]):
    line=line.strip()
    if line.startswith('{'):
        d=json.loads(line); print(sys.argv[1], d['n_gpus'], round(d['value']), round(d['ms_per_step'],2), d.get('e2e',{}) and round(d['e2e']['value']))
PY
done
