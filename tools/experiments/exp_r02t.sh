#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02t; mkdir -p $O
timeout 900 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 200 python bench.py --workload full --steps 400 --warmup 40 --no-e2e --no-cpu --no-secondary > $O/bench_full_n1.json 2> $O/bench_full_n1.err
timeout 200 python bench.py --nx 128 --ny 128 --nz-per-gpu 256 --steps 400 --warmup 40 --no-e2e --no-cpu --no-secondary > $O/bench_small.json 2> $O/bench_small.err
timeout 200 python bench.py --steps 50 --warmup 5 --no-e2e --no-cpu --no-secondary > $O/bench_v0.json 2> $O/bench_v0.err
for f in full_n1 small v0; do python - $O/bench_$f.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"]), round(d["ms_per_step"],4), round(d["roofline"]["step_frac"],4), d["gpu_launches"], d["phase_fallbacks"])
PY
done
echo done
