#!/bin/bash
# Tick kernel iteration on the GPU box: parity tests of the tick path, then the config-4 run (stepped and graph replay).
TAG=${1:-t1}; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests/test_tick_gpu.py tests/test_queue_class.py -m gpu -x -q > $O/pytest_tick.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_tick.log
TICKS=${TICKS:-3000} GRAPH=1 timeout 600 python tools/tick_prof.py 2>&1 | tee $O/tick_graph.log
if [ -n "$NCU" ]; then
TICKS=700 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1200 -c 900 --csv \
  --log-file $O/tick_launches.csv python tools/tick_prof.py > $O/ncu_tick_launches.log 2>&1
fi
