#!/bin/bash
# round 2, experiment A: one-pass (lag) kernel -- parity, tuning sweep, DRAM traffic
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/gpu.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "d3q19" > $O/pytest_d3q19.log 2>&1; echo "pytest rc=$?" >> $O/pytest_d3q19.log
tail -5 $O/pytest_d3q19.log
B="python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu"
timeout 300 $B --variant 5 > $O/bench_v5.json 2> $O/bench_v5.err
for pol in 0 1; do for k in 1.0 1.5 2.5; do for sb in 64; do
  LBM_TUNE_POL=$pol LBM_TUNE_LAGK=$k LBM_TUNE_SB=$sb timeout 300 $B --variant 0 > $O/bench_v0_p${pol}_k${k}_s${sb}.json 2> $O/bench_v0_p${pol}_k${k}_s${sb}.err
done; done; done
for sb in 32 128 256; do
  LBM_TUNE_POL=1 LBM_TUNE_LAGK=1.5 LBM_TUNE_SB=$sb timeout 300 $B --variant 0 > $O/bench_v0_p1_k1.5_s${sb}.json 2> $O/bench_v0_p1_k1.5_s${sb}.err
done
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]),round(d["ms_per_step"],3),d["roofline"].get("step_frac"),{k:round(v["ms"]/v["launches"],3) for k,v in d["roofline"]["kernels"].items()},d.get("phase_fallbacks"))
except Exception as e: print("ERR",e)
PY
done | tee $O/summary.txt
# DRAM traffic per launch of the step's kernels (one pass of metrics, no replay needed beyond that)
for pol in 0 1; do
LBM_TUNE_POL=$pol timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:"k_collide_lag|k_collide_rim|k_phase" -s 12 -c 12 --csv --log-file $O/traffic_v0_p$pol.csv python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu --variant 0 > $O/ncu_v0_p$pol.log 2>&1
done
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:"k_collide_interior|k_collide_rim|k_phase" -s 9 -c 9 --csv --log-file $O/traffic_v5.csv python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu --variant 5 > $O/ncu_v5.log 2>&1
echo done
