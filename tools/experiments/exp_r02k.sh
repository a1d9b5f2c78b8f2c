#!/bin/bash
# round 2, multi-GPU run: tests of the slab ring and of the multi-GPU C++ host, bench lines with the parity check,
# (2 GPUs) memcheck of the two-stream step.  Usage: exp_r02k.sh NGPU
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
NG=${1:-2}
O=gpurun_out/r02k_n$NG; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 150 python -m pytest tests/test_gpu_multi.py tests/test_gpu_hosts.py -x -q -m gpu -k "slab_ring or several_gpus" -v > $O/pytest_multi.log 2>&1; rc=$?; echo "pytest rc=$rc" >> $O/pytest_multi.log
tail -12 $O/pytest_multi.log
if [ $rc -ne 0 ]; then echo "multi-GPU tests failed: stopping"; exit 1; fi
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in ${BENCH_N:-2 4 8}; do
  if [ $n -le $NG ]; then
    timeout 420 $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --steps 50 --warmup 5 > $O/bench_n$n.json 2> $O/bench_n$n.err; echo "bench n=$n rc=$?"
    tail -c 400 $O/bench_n$n.json
  fi
done
echo done
