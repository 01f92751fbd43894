#!/bin/bash
# SSSP kernel iteration on the GPU box: parity tests of the sp path, a tuning sweep, one full ncu capture.
TAG=${1:-s1}; O=gpurun_out/$TAG; mkdir -p $O
if [ -z "$SKIP_TESTS" ]; then
timeout 900 python -m pytest tests/test_sssp_gpu.py -m gpu -x -q > $O/pytest_sssp.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_sssp.log
fi
ALGO=${ALGO:-2} BATCH=${BATCH:-4096} timeout 1200 python tools/sssp_sweep.py > $O/sweep.jsonl 2>&1; echo "sweep rc=$?"
cat $O/sweep.jsonl | cut -c1-260
if [ -n "$NCU_CFG" ]; then
ALGO=${ALGO:-2} CFG="$NCU_CFG" BATCH=2048 timeout 900 ncu --set full --clock-control none --import-source on -k regex:sssp_warp_kernel -s 1 -c 1 \
  -o $O/sssp_warp_full -f python tools/sssp_prof.py > $O/ncu_sssp_full.log 2>&1; tail -5 $O/ncu_sssp_full.log
fi
