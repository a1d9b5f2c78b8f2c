#!/bin/bash
# round 2, last 1-GPU run on the final tree: GPU suite, smoke, default bench line; then (after those have exited 0 without a
# profiler) ncu --set full of the two kernels that are new since the last capture: k_aa_one at 512^3 and k_b2_march at 8192^2
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ak; mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"
timeout 300 python bench.py --variant 2 --no-cpu --no-secondary > $O/bench_inplace_512.json 2> $O/bench_inplace_512.err; echo "bench rc=$?"
timeout 400 ncu --set full --clock-control none -k regex:k_aa_one -s 4 -c 1 --log-file $O/aa_one_512cubed.ncu.txt python bench.py --variant 2 --no-cpu --no-secondary --no-e2e --steps 3 --warmup 3 > $O/ncu1.log 2>&1; echo "ncu rc=$?"
timeout 400 ncu --set full --clock-control none -k regex:k_b2_march -s 8 -c 1 --log-file $O/b2_march_8192.ncu.txt python tools/time2d.py binary 8192 0 > $O/ncu2.log 2>&1; echo "ncu rc=$?"
echo done
