#!/bin/bash
# round 2: the bench-size property test with the in-place run added
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02an; mkdir -p $O
timeout 400 python -m pytest tests/test_gpu_fullsize.py -q -m gpu -k "translation" > $O/pytest.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest.log; tail -5 $O/pytest.log
echo done
