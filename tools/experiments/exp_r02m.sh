#!/bin/bash
# round 2, run M: binary D2Q9 one-pass step (parity, 8192^2 timing), MRT/TRT extensions and guard-band tests on the GPU, tight strips
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02m; mkdir -p $O
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_mrt_d3q19_extension.py tests/test_trt_d3q19_extension.py tests/test_gpu_guard.py -x -q -m gpu -k "binary2d or mrt or trt or guard or buffers" > $O/pytest.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest.log
tail -15 $O/pytest.log
timeout 200 python - > $O/bench_binary2d.txt 2>&1 <<'PY'
import sys, json
sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import numpy as np
from latticeboltzmannmethod_b200 import api
import bench2d
n = 8192
ph = np.tile(api.binary2d_initial_phase().reshape(128, 128), (n // 128, n // 128)).ravel()
for variant in (0, 5, 0, 5):
    cfg = api.binary2d_config(n, n); cfg.kernel_variant = variant
    ctx = api.Context(cfg); ctx.set_field("phase", ph); ctx.init()
    ms = bench2d.timed(ctx, 30)
    print(json.dumps({"variant": variant, "ms": ms, "glups": n * n / ms / 1e6, "frac": n * n * 288 / ms / 1e6 / 6547.2, "fallbacks": ctx.phase_fallbacks}))
    ctx.close()
PY
cat $O/bench_binary2d.txt
B="python bench.py --steps 50 --warmup 5 --no-e2e --no-cpu --no-secondary"
timeout 100 $B > $O/bench_v0.json 2> $O/bench_v0.err
LBM_TUNE_TIGHT=1 timeout 100 $B > $O/bench_v0_tight.json 2> $O/bench_v0_tight.err
LBM_TUNE_LAGK=1.6 timeout 100 $B > $O/bench_v0_k1.6.json 2> $O/bench_v0_k1.6.err
for f in $O/bench_v*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]),round(d["ms_per_step"],3),round(d["roofline"].get("step_frac"),4),{k:round(v["ms"]/v["launches"],3) for k,v in d["roofline"]["kernels"].items()},d.get("phase_fallbacks"),d["clocks"]["sm_mhz"])
except Exception as e: print("ERR",e)
PY
done | tee $O/summary.txt
echo done
