#!/bin/bash
# N = 8 experiments: NCCL CTA budget and chunk count for the overlapped all-gather.
TAG=${1:-sc2}; O=gpurun_out/$TAG; mkdir -p $O
run() { # name, env..., -- bench args
  name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 8 --steps 5 --warmup 3 $EXTRA > $O/$name.json 2> $O/$name.err
  python - $O/$name.json $name <<'PY'
import json,sys
for line in open(sys.argv[1]):
    if line.startswith('{'):
        d=json.loads(line); print(sys.argv[2], round(d['value']), round(d['ms_per_step'],2))
PY
}
EXTRA="--chunks 4" run base_c4 X=1
EXTRA="--chunks 4" run ctas4_c4 NCCL_MAX_CTAS=4
EXTRA="--chunks 8" run ctas4_c8 NCCL_MAX_CTAS=4
EXTRA="--chunks 8" run ctas8_c8 NCCL_MAX_CTAS=8
EXTRA="--chunks 4 --no-gather" run nogather X=1
