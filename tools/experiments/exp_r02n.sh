#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/${1:-r02n}; mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "d3q19" > $O/pytest.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest.log
tail -3 $O/pytest.log
if [ $rc -ne 0 ]; then echo "parity failed or hung: stopping"; exit 1; fi
B="python bench.py --steps 50 --warmup 5 --no-e2e --no-cpu --no-secondary"
for i in 1 2; do timeout 100 $B > $O/bench_v0_$i.json 2> $O/bench_v0_$i.err; done
timeout 100 $B --variant 5 > $O/bench_v5.json 2> $O/bench_v5.err
for f in $O/bench_v*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]),round(d["ms_per_step"],3),round(d["roofline"].get("step_frac"),4),{k:round(v["ms"]/v["launches"],3) for k,v in d["roofline"]["kernels"].items()},d.get("phase_fallbacks"),d["clocks"]["sm_mhz"])
except Exception as e: print("ERR",e)
PY
done | tee $O/summary.txt
echo done
