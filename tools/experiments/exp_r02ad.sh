#!/bin/bash
# round 2: curved-solid marching kernel (kernel_variant 6): parity, timing at 8192^2 against phase pass + collide
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ad; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_shape.py -x -q -m gpu > $O/pytest_shape.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_shape.log; tail -15 $O/pytest_shape.log
[ $rc -eq 124 ] && exit 1
timeout 400 python tools/time2d.py shape 8192 0 6 6:LBM_TUNE_XC=32 6:LBM_TUNE_XC=128 0 > $O/time_shape.jsonl 2> $O/time_shape.err; echo "time rc=$?"
cat $O/time_shape.jsonl; tail -3 $O/time_shape.err
echo done
