#!/bin/bash
# round 2: k_aa_one, which accesses should carry an L2 hint (LBM_TUNE_POL = mask: 1 phase-part g loads stay, 2 collision g loads,
# 4 f loads, 8 stores go first)
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02y; mkdir -p $O
B="--no-cpu --no-secondary --no-e2e --variant 2 --steps 20"
run() { n=$1; shift; env "$@" timeout 200 python bench.py $B > $O/$n.json 2> $O/$n.err; rc=$?; echo "$n rc=$rc"; [ $rc -eq 124 ] && exit 1; }
run h0 LBM_TUNE_POL=0
run h1 LBM_TUNE_POL=1
run h3 LBM_TUNE_POL=3
run h5 LBM_TUNE_POL=5
run h9 LBM_TUNE_POL=9
run h2 LBM_TUNE_POL=2
run h7 LBM_TUNE_POL=7
run h0b LBM_TUNE_POL=0
echo done
