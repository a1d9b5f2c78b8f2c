#!/bin/bash
# round 2: evidence run after the in-place variants went in: whole GPU suite, smoke, default bench line, the in-place lines
# (one pass at 512^3 and at twice the planes, two passes), DRAM traffic of k_aa_one / k_aa_core / k_aa_phase
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02aa; mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"
B="--no-cpu --no-secondary"
timeout 300 python bench.py --variant 2 $B > $O/bench_inplace_512.json 2> $O/bench_inplace_512.err; echo "bench rc=$?"
timeout 400 python bench.py --variant 2 --nz-per-gpu 1024 $B --steps 30 > $O/bench_inplace_1024.json 2> $O/bench_inplace_1024.err; echo "bench rc=$?"
timeout 300 python bench.py --variant 6 $B --no-e2e > $O/bench_inplace_twopass_512.json 2> $O/bench_inplace_twopass_512.err; echo "bench rc=$?"
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_aa_ --launch-skip 6 -c 4 --csv --log-file $O/ncu_inplace_v2.csv python bench.py --no-cpu --no-secondary --no-e2e --variant 2 --steps 3 --warmup 3 > $O/ncu2.log 2>&1; echo "ncu rc=$?"
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_aa_ --launch-skip 6 -c 4 --csv --log-file $O/ncu_inplace_v6.csv python bench.py --no-cpu --no-secondary --no-e2e --variant 6 --steps 3 --warmup 3 > $O/ncu6.log 2>&1; echo "ncu rc=$?"
echo done
