#!/bin/bash
# round 2, experiment C: one-pass kernel after the index/grid clean-up -- ablation again, carve-out, full ncu capture
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02c; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "d3q19" > $O/pytest_d3q19.log 2>&1; echo "pytest rc=$?" >> $O/pytest_d3q19.log
B="python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu --no-secondary"
run() { n=$1; shift; env "$@" timeout 300 $B --variant 0 > $O/bench_$n.json 2> $O/bench_$n.err; }
timeout 300 $B --variant 5 > $O/bench_v5.json 2> $O/bench_v5.err
run v0_default X=1
run v0_carve30 LBM_TUNE_CARVE=30
run v0_carve50 LBM_TUNE_CARVE=50
run v0_abl5 LBM_TUNE_ABLATE=5
run v0_abl4 LBM_TUNE_ABLATE=4
run v0_abl2 LBM_TUNE_ABLATE=2
run v0_abl7 LBM_TUNE_ABLATE=7
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]),round(d["ms_per_step"],3),round(d["roofline"].get("step_frac"),4),{k:round(v["ms"]/v["launches"],3) for k,v in d["roofline"]["kernels"].items()},d.get("phase_fallbacks"))
except Exception as e: print("ERR",e)
PY
done | tee $O/summary.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"k_collide_lag" -s 3 -c 1 -o $O/lag_full python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu --no-secondary --nz-per-gpu 128 --variant 0 > $O/ncu_full.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"k_collide_interior" -s 3 -c 1 -o $O/interior_full python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu --no-secondary --nz-per-gpu 128 --variant 5 > $O/ncu_full5.log 2>&1
ls -la $O
echo done
