#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02s; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_hosts.py tests/test_gpu_guard.py -x -q -m gpu -k "mrt" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -3 $O/pytest.log
timeout 200 python - > $O/bench_mrt.txt 2>&1 <<'PY'
import sys, json
sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import numpy as np
from latticeboltzmannmethod_b200 import api
import bench2d
n = 8192
for variant, om, ob in (("gs", 1.2, 0.8), ("hermite", 0.9, 0.1)):
    for kv in (0, 1):
        cfg = api.mrt_config(variant, n, n, om, ob); cfg.kernel_variant = kv
        ctx = api.Context(cfg)
        x = np.arange(n)
        ctx.set_field("rho", np.ones(n * n)); ctx.set_field("ux", np.repeat(1e-3 * np.cos(x / n * 2 * np.pi), n)); ctx.set_field("uy", np.zeros(n * n))
        ctx.init()
        ms = bench2d.timed(ctx, 40)
        import torch
        free, total = torch.cuda.mem_get_info()
        print(json.dumps({"model": variant, "kernel_variant": kv, "ms": ms, "glups": n * n / ms / 1e6, "frac": n * n * 144 / ms / 1e6 / 6547.2, "device_mem_used_gb": (total - free) / 1e9}))
        ctx.close()
PY
cat $O/bench_mrt.txt
echo done
