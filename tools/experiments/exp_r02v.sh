#!/bin/bash
# round 2: in-place D3Q19 step in ONE pass (kernel_variant 2 = k_aa_one): parity tests, bench at 512^3 / 1024 planes, lag sweep
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02v; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_mrt_d3q19_extension.py tests/test_trt_d3q19_extension.py -x -q -m gpu -k "d3q19 or mrt_equals or trt_equals" > $O/pytest_aa.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_aa.log; tail -5 $O/pytest_aa.log
[ $rc -ne 0 ] && exit 1
B="--no-cpu --no-secondary --no-e2e"
timeout 200 python bench.py --variant 2 $B --steps 30 > $O/bench_v2_512.json 2> $O/bench_v2_512.err; rc=$?; echo "bench rc=$rc"
[ $rc -eq 124 ] && exit 1
timeout 200 python bench.py --variant 2 --nz-per-gpu 1024 $B --steps 20 > $O/bench_v2_1024.json 2> $O/bench_v2_1024.err; echo "bench rc=$?"
for k in 1.0 2.0 3.0; do
  LBM_TUNE_LAGK=$k timeout 200 python bench.py --variant 2 $B --steps 20 > $O/bench_v2_lag$k.json 2> $O/bench_v2_lag$k.err; echo "lag $k rc=$?"
done
timeout 200 python bench.py --variant 0 $B --steps 30 > $O/bench_v0_512.json 2> $O/bench_v0_512.err; echo "bench rc=$?"
echo done
