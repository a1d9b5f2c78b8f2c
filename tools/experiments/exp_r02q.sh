#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02q; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "binary2d" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -3 $O/pytest.log
for v in 0 6 7 8 5 0 6; do
timeout 120 python - $v >> $O/bench_binary2d.txt 2>&1 <<'PY'
import sys, json
sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import numpy as np
from latticeboltzmannmethod_b200 import api
import bench2d
n = 8192; variant = int(sys.argv[1])
ph = np.tile(api.binary2d_initial_phase().reshape(128, 128), (n // 128, n // 128)).ravel()
cfg = api.binary2d_config(n, n); cfg.kernel_variant = variant
ctx = api.Context(cfg); ctx.set_field("phase", ph); ctx.init()
ms = bench2d.timed(ctx, 30)
print(json.dumps({"variant": variant, "ms": ms, "glups": n * n / ms / 1e6, "frac": n * n * 288 / ms / 1e6 / 6547.2}))
ctx.close()
PY
done
cat $O/bench_binary2d.txt
echo done
