#!/bin/bash
# round 2: k_aa_one with the phase sums deferred to the end of the block (asynchronous copies): tests, bench, DRAM traffic
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02w; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "d3q19" > $O/pytest_aa.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_aa.log; tail -5 $O/pytest_aa.log
[ $rc -ne 0 ] && exit 1
B="--no-cpu --no-secondary --no-e2e"
timeout 200 python bench.py --variant 2 $B --steps 30 > $O/bench_v2_512.json 2> $O/bench_v2_512.err; rc=$?; echo "bench rc=$rc"
[ $rc -eq 124 ] && exit 1
for k in 1.0 2.0; do
  LBM_TUNE_LAGK=$k timeout 200 python bench.py --variant 2 $B --steps 20 > $O/bench_v2_lag$k.json 2> $O/bench_v2_lag$k.err; echo "lag $k rc=$?"
done
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_aa_one --launch-skip 3 -c 2 --csv --log-file $O/ncu_aa_one.csv python bench.py --variant 2 $B --steps 3 --warmup 3 > $O/ncu.log 2>&1; echo "ncu rc=$?"
echo done
