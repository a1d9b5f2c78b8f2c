#!/bin/bash
# round 2: 8 GPUs, in place: 512 x 512 x 1024 per GPU (twice the planes of the two-array runs), five directions per face
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02aj; mkdir -p $O
N=${1:-8}
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --variant 2 --nz-per-gpu 1024 --no-e2e --no-secondary --steps 20 > $O/bench_inplace_n$N.json 2> $O/bench_inplace_n$N.err; echo "bench rc=$?"
tail -3 $O/bench_inplace_n$N.err
echo done
