#!/bin/bash
# round 2: after the 2-D variants were pruned: GPU suite, smoke, default bench line
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02al; mkdir -p $O
timeout 1200 python -m pytest tests -q -m gpu > $O/pytest_gpu.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"
echo done
