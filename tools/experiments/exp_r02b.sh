#!/bin/bash
# round 2, experiment B: one-pass kernel with the light counter protocol -- parity, ablation, tuning, traffic
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02b; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "d3q19" > $O/pytest_d3q19.log 2>&1; echo "pytest rc=$?" >> $O/pytest_d3q19.log
tail -3 $O/pytest_d3q19.log
B="python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu"
run() { # name, env...
  n=$1; shift
  env "$@" timeout 300 $B --variant 0 > $O/bench_$n.json 2> $O/bench_$n.err
}
timeout 300 $B --variant 5 > $O/bench_v5.json 2> $O/bench_v5.err
run v0_default X=1
run v0_pol1 LBM_TUNE_POL=1
run v0_k0.75 LBM_TUNE_LAGK=0.75
run v0_k1.0 LBM_TUNE_LAGK=1.0
run v0_k2.0 LBM_TUNE_LAGK=2.0
run v0_k3.0 LBM_TUNE_LAGK=3.0
run v0_sb32 LBM_TUNE_SB=32
run v0_sb128 LBM_TUNE_SB=128
run v0_abl4 LBM_TUNE_ABLATE=4
run v0_abl5 LBM_TUNE_ABLATE=5
run v0_abl2 LBM_TUNE_ABLATE=2
run v0_abl1 LBM_TUNE_ABLATE=1
run v0_abl5_onestrip LBM_TUNE_ABLATE=5 LBM_TUNE_SB=100000
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]),round(d["ms_per_step"],3),round(d["roofline"].get("step_frac"),4),{k:round(v["ms"]/v["launches"],3) for k,v in d["roofline"]["kernels"].items()},d.get("phase_fallbacks"))
except Exception as e: print("ERR",e)
PY
done | tee $O/summary.txt
for pol in 0 1; do
LBM_TUNE_POL=$pol timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct --clock-control none -k regex:"k_collide_lag|k_collide_rim|k_phase" -s 10 -c 5 --csv --log-file $O/traffic_v0_p$pol.csv python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu --variant 0 > $O/ncu_v0_p$pol.log 2>&1
done
echo done
