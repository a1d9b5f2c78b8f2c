#!/bin/bash
# round 2: GPU suite and smoke on the last tree of the round
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ao; mkdir -p $O
timeout 600 python -m pytest tests -q -m gpu > $O/pytest_gpu.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
echo done
