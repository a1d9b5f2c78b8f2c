#!/bin/bash
# round 2: k_aa_one with L2 hints; strip width and lag knobs
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02x; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "d3q19" > $O/pytest_aa.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_aa.log; tail -5 $O/pytest_aa.log
[ $rc -ne 0 ] && exit 1
B="--no-cpu --no-secondary --no-e2e --variant 2 --steps 20"
run() { n=$1; shift; env "$@" timeout 200 python bench.py $B > $O/$n.json 2> $O/$n.err; rc=$?; echo "$n rc=$rc"; [ $rc -eq 124 ] && exit 1; }
run pol1 LBM_TUNE_POL=1
run pol0 LBM_TUNE_POL=0
run pol1_sb128 LBM_TUNE_POL=1 LBM_TUNE_SB=128
run pol1_sb32 LBM_TUNE_POL=1 LBM_TUNE_SB=32
run pol1_lag1.25 LBM_TUNE_POL=1 LBM_TUNE_LAGK=1.25
run pol1_sb128_lag2 LBM_TUNE_POL=1 LBM_TUNE_SB=128 LBM_TUNE_LAGK=2
run pol1_again LBM_TUNE_POL=1
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_aa_one --launch-skip 3 -c 2 --csv --log-file $O/ncu_aa_one.csv python bench.py --no-cpu --no-secondary --no-e2e --variant 2 --steps 3 --warmup 3 > $O/ncu.log 2>&1; echo "ncu rc=$?"
echo done
