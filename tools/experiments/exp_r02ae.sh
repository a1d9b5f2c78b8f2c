#!/bin/bash
# round 2: binary D2Q9 marching kernel at four blocks per SM (128 registers, two f buffers): parity, timing
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ae; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "binary2d" > $O/pytest_b2.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_b2.log; tail -5 $O/pytest_b2.log
[ $rc -eq 124 ] && exit 1
timeout 300 python tools/time2d.py binary 8192 6 0 0:LBM_TUNE_XC=32 0:LBM_TUNE_XC=128 0 > $O/time_binary.jsonl 2> $O/time_binary.err; echo "time rc=$?"
cat $O/time_binary.jsonl
echo done
