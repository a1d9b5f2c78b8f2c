#!/bin/bash
# round 2, experiment E: one-pass step with g staged through shared memory (k_step_one, 128 registers, 16 warps/SM)
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02e; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "d3q19" > $O/pytest_d3q19.log 2>&1; echo "pytest rc=$?" >> $O/pytest_d3q19.log
tail -3 $O/pytest_d3q19.log
B="python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu --no-secondary"
run() { n=$1; v=$2; shift; shift; env "$@" timeout 300 $B --variant $v > $O/bench_$n.json 2> $O/bench_$n.err; }
run v5 5 X=1
run v0_default 0 X=1
run v2_default 2 X=1
run v2_k1 2 LBM_TUNE_LAGK=1.0
run v2_k2 2 LBM_TUNE_LAGK=2.0
run v2_k3 2 LBM_TUNE_LAGK=3.0
run v2_pol1 2 LBM_TUNE_POL=1
run v2_abl5 2 LBM_TUNE_ABLATE=5
run v2_abl4 2 LBM_TUNE_ABLATE=4
run v2_abl2 2 LBM_TUNE_ABLATE=2
run v2_sb32 2 LBM_TUNE_SB=32
run v2_sb128 2 LBM_TUNE_SB=128
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]),round(d["ms_per_step"],3),round(d["roofline"].get("step_frac"),4),{k:round(v["ms"]/v["launches"],3) for k,v in d["roofline"]["kernels"].items()},d.get("phase_fallbacks"))
except Exception as e: print("ERR",e)
PY
done | tee $O/summary.txt
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:"k_step_one" -s 2 -c 2 --csv --log-file $O/traffic_v2.csv python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu --no-secondary --variant 2 > $O/ncu_v2.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"k_step_one" -s 2 -c 1 -o $O/one_full python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu --no-secondary --nz-per-gpu 128 --variant 2 > $O/ncu_full.log 2>&1
echo done
