#!/bin/bash
# round 2, run H1: whole GPU test suite on the cleaned-up tree + last tuning of the one-pass step (lag, L2 hints)
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02h; mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "d3q19" > $O/pytest_d3q19.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest_d3q19.log
tail -3 $O/pytest_d3q19.log
if [ $rc -ne 0 ]; then echo "parity failed or hung: stopping"; exit 1; fi
B="python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu --no-secondary"
run() { n=$1; v=$2; shift; shift; env "$@" timeout 100 $B --variant $v > $O/bench_$n.json 2> $O/bench_$n.err; r=$?; echo "$n rc=$r"; if [ $r -eq 124 ]; then echo "hung: stopping"; exit 1; fi; }
run v0_p0_k1.5 0 X=1
run v0_p1_k1.5 0 LBM_TUNE_POL=1
run v0_p1_k1.25 0 LBM_TUNE_POL=1 LBM_TUNE_LAGK=1.25
run v0_p1_k1.75 0 LBM_TUNE_POL=1 LBM_TUNE_LAGK=1.75
run v0_p1_k2.0 0 LBM_TUNE_POL=1 LBM_TUNE_LAGK=2.0
run v0_p0_k1.25 0 LBM_TUNE_LAGK=1.25
run v0_p0_k1.75 0 LBM_TUNE_LAGK=1.75
run v5 5 X=1
run v1 1 X=1
timeout 100 $B --variant 0 --precision f32 > $O/bench_v0_f32.json 2> $O/bench_v0_f32.err
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]),round(d["ms_per_step"],3),round(d["roofline"].get("step_frac"),4),{k:round(v["ms"]/v["launches"],3) for k,v in d["roofline"]["kernels"].items()},d.get("phase_fallbacks"))
except Exception as e: print("ERR",e)
PY
done | tee $O/summary.txt
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest_all.log 2>&1; echo "pytest rc=$?" >> $O/pytest_all.log
tail -5 $O/pytest_all.log
echo done
