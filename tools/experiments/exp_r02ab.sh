#!/bin/bash
# round 2: binary D2Q9 marching kernel (kernel_variant 6): parity, then timing at 8192^2 against the two-pass default
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ab; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "binary2d" > $O/pytest_b2.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_b2.log; tail -15 $O/pytest_b2.log
[ $rc -eq 124 ] && exit 1
timeout 300 python tools/time2d.py binary 8192 0 6 6:LBM_TUNE_XC=32 6:LBM_TUNE_XC=128 6:LBM_TUNE_XC=256 0 > $O/time_binary.jsonl 2> $O/time_binary.err; echo "time rc=$?"
cat $O/time_binary.jsonl
echo done
