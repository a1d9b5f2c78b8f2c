#!/bin/bash
# round 2: in-place (AA-pattern) D3Q19 step, kernel_variant 2: parity tests, then bench lines at 512^3 and at twice the planes
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02u; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_mrt_d3q19_extension.py tests/test_trt_d3q19_extension.py -x -q -m gpu -k "d3q19 or mrt_equals or trt_equals" > $O/pytest_aa.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_aa.log; tail -5 $O/pytest_aa.log
[ $rc -eq 124 ] && exit 1
timeout 300 python bench.py --variant 2 --no-cpu --no-secondary --no-e2e --steps 30 > $O/bench_v2_512.json 2> $O/bench_v2_512.err; rc=$?; echo "bench rc=$rc"
[ $rc -eq 124 ] && exit 1
timeout 300 python bench.py --variant 2 --nz-per-gpu 1024 --no-cpu --no-secondary --no-e2e --steps 20 > $O/bench_v2_1024.json 2> $O/bench_v2_1024.err; echo "bench rc=$?"
timeout 300 python bench.py --variant 5 --no-cpu --no-secondary --no-e2e --steps 30 > $O/bench_v5_512.json 2> $O/bench_v5_512.err; echo "bench rc=$?"
echo done
