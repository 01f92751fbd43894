#!/bin/bash
# One GPU-box pass of round 2: parity tests, smoke, bench line, sweep, ncu launch lists and --set full captures.
# Usage (from the build container): tools/gpurun_retry.sh 3000 'bash tools/gpu_round2.sh r2d'
TAG=${1:-r2x}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 1200 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/bench.err
timeout 600 python bench.py --sweep --steps 3 > $O/sweep_n1.json 2> $O/sweep.err; echo "sweep rc=$?" | tee -a $O/sweep.err
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 --no-tick > $O/bench_reference.json 2> $O/bench_reference.err
# launch lists (share of the step per kernel; absolute times under ncu are cold-cache and serialised)
timeout 600 python bench.py --steps 1 --warmup 3 --batch 2368 --no-tick --no-cpu > $O/bench_small.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/sssp_launches.csv \
  python bench.py --steps 1 --warmup 3 --batch 2368 --no-tick --no-cpu > $O/ncu_sssp_launches.log 2>&1
TICKS=700 timeout 600 python tools/tick_prof.py > $O/tick_prof_plain.log 2>&1 && \
TICKS=700 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1200 -c 900 --csv \
  --log-file $O/tick_launches.csv python tools/tick_prof.py > $O/ncu_tick_launches.log 2>&1
CASE=c3 TICKS=3400 timeout 600 python tools/tick_prof.py > $O/tick_c3_prof_plain.log 2>&1 && \
CASE=c3 TICKS=3400 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 9000 -c 1200 --csv \
  --log-file $O/tick_c3_launches.csv python tools/tick_prof.py > $O/ncu_tick_c3_launches.log 2>&1
# full captures
BATCH=2368 timeout 300 python tools/sssp_prof.py > $O/sssp_prof_plain.log 2>&1 && \
BATCH=2368 timeout 900 ncu --set full --clock-control none --import-source on -k regex:sssp_warp_kernel -s 1 -c 1 \
  -o $O/sssp_warp_full -f python tools/sssp_prof.py > $O/ncu_sssp_full.log 2>&1
TICKS=2600 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_node_phase -s 2500 -c 1 \
  -o $O/tick_node_full -f python tools/tick_prof.py > $O/ncu_tick_full.log 2>&1
for T0 in 3000 7000; do T0=$T0 timeout 300 python tools/tick_timeline.py > $O/timeline_c4_$T0.json 2> $O/timeline_c4_$T0.err; done
CASE=c2 T0=800 timeout 300 python tools/tick_timeline.py > $O/timeline_c2.json 2> $O/timeline_c2.err
CASE=c3 T0=3000 timeout 300 python tools/tick_timeline.py > $O/timeline_c3.json 2> $O/timeline_c3.err
ls -la $O
tail -3 $O/pytest_gpu.log; tail -2 $O/smoke.log; head -c 3000 $O/bench.json
