#!/bin/bash
# Short multi-GPU pass: the default bench line and the gather-to-rank-0 variant.  Usage: gpurun --gpus N -- 'bash tools/gpu_scale_r2b.sh N tag'
N=${1:-2}; TAG=${2:-r2t}
O=gpurun_out/$TAG; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nproc > $O/nproc.txt; (lscpu | grep -i "numa\|socket\|model name" > $O/lscpu.txt) 2>/dev/null; nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 600 $RUN --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "all-to-all rc=$?"
timeout 600 $RUN --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --gather-to 0 > $O/bench_n${N}_to0.json 2> $O/bench_n${N}_to0.err; echo "gather-to-0 rc=$?"
tail -c 400 $O/bench_n$N.err
