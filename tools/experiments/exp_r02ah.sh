#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ah; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_multi.py -q -m gpu -k "in_place" > $O/pytest_multi_inplace.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_multi_inplace.log; grep -E "passed|failed|FAILED|AssertionError|assert \(" $O/pytest_multi_inplace.log | head -20
echo done
