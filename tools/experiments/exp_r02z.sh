#!/bin/bash
# round 2: k_aa_one: phi stores that stay in L2 (16), L2 prefetch with evict_last of the phase part's g lines (32), lag sweep
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02z; mkdir -p $O
B="--no-cpu --no-secondary --no-e2e --variant 2 --steps 20"
run() { n=$1; shift; env "$@" timeout 200 python bench.py $B > $O/$n.json 2> $O/$n.err; rc=$?; echo "$n rc=$rc"; [ $rc -eq 124 ] && exit 1; }
run h0 LBM_TUNE_POL=0
run h16 LBM_TUNE_POL=16
run h32 LBM_TUNE_POL=32
run h48 LBM_TUNE_POL=48
run h17 LBM_TUNE_POL=17
run h0_lag1.25 LBM_TUNE_POL=0 LBM_TUNE_LAGK=1.25
run h0_lag1.75 LBM_TUNE_POL=0 LBM_TUNE_LAGK=1.75
run h0_lag1.0 LBM_TUNE_POL=0 LBM_TUNE_LAGK=1.0
run h0_sb128 LBM_TUNE_POL=0 LBM_TUNE_SB=128
run h0_sb32 LBM_TUNE_POL=0 LBM_TUNE_SB=32
echo done
