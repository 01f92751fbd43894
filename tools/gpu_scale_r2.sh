#!/bin/bash
# Multi-GPU pass of round 2 (N = number of GPUs of the lease): the default bench line, the gather-to-rank-0 variant and
# the strong-scaling sweep.  Usage: gpurun --gpus N --timeout 1500 -- 'bash tools/gpu_scale_r2.sh N tag'
N=${1:-2}; TAG=${2:-r2s}
O=gpurun_out/$TAG; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $RUN --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "all-to-all rc=$?"
timeout 600 $RUN --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --gather-to 0 > $O/bench_n${N}_to0.json 2> $O/bench_n${N}_to0.err; echo "gather-to-0 rc=$?"
timeout 600 $RUN --master-port 29513 bench.py --gpus $N --sweep --steps 3 --warmup 3 > $O/sweep_n$N.json 2> $O/sweep_n$N.err; echo "sweep rc=$?"
timeout 600 $RUN --master-port 29514 bench.py --gpus $N --sweep --steps 3 --warmup 3 --gather-to 0 > $O/sweep_n${N}_to0.json 2> $O/sweep_n${N}_to0.err; echo "sweep to0 rc=$?"
tail -c 600 $O/bench_n$N.err; head -c 1500 $O/bench_n$N.json; echo; head -c 600 $O/bench_n${N}_to0.json
