#!/bin/bash
# round 2: Shan-Chen marching kernel (kernel_variant 6): parity, then timing at 8192^2 against the fused tile kernel
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ac; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "shanchen" > $O/pytest_sc.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_sc.log; tail -15 $O/pytest_sc.log
[ $rc -eq 124 ] && exit 1
timeout 300 python tools/time2d.py shanchen 8192 0 6 6:LBM_TUNE_XC=32 6:LBM_TUNE_XC=128 6:LBM_TUNE_XC=16 0 > $O/time_shanchen.jsonl 2> $O/time_shanchen.err; echo "time rc=$?"
cat $O/time_shanchen.jsonl
echo done
