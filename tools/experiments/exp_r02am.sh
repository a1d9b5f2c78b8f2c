#!/bin/bash
# round 2: after k_aa_one took its plane-step arithmetic from aa_schedule(): D3Q19 parity tests on one GPU, the in-place slab
# ring on two, one in-place bench line
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02am; mkdir -p $O
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py tests/test_gpu_guard.py -q -m gpu -k "d3q19 or in_place" > $O/pytest.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest.log; grep -E "passed|failed|FAILED" $O/pytest.log | head
timeout 200 python bench.py --variant 2 --no-cpu --no-secondary --no-e2e --steps 30 > $O/bench_inplace_512.json 2> $O/bench_inplace_512.err; echo "bench rc=$?"
echo done
