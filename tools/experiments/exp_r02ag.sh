#!/bin/bash
# round 2: the in-place D3Q19 step on several slabs (NCCL, five directions per face): parity on 2 GPUs, then the bench line
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ag; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -m gpu -k "in_place" > $O/pytest_multi_inplace.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_multi_inplace.log; tail -15 $O/pytest_multi_inplace.log
[ $rc -ne 0 ] && exit 1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --variant 2 --no-secondary --steps 30 > $O/bench_inplace_n2.json 2> $O/bench_inplace_n2.err; echo "bench rc=$?"
tail -3 $O/bench_inplace_n2.err
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -m gpu -k "slab_ring_on_gpus and nccl" > $O/pytest_multi_default.log 2>&1; echo "pytest default rc=$?"; tail -3 $O/pytest_multi_default.log
echo done
