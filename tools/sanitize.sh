#!/bin/bash
# compute-sanitizer memcheck + racecheck on small grids for every kernel that uses shared memory or inter-block
# flags, and memcheck of the two-stream multi-GPU step (when 2 GPUs are visible).  Logs under $1 (default gpurun_out/sanitize).
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=${1:-gpurun_out/sanitize}; mkdir -p $O
CS="compute-sanitizer --error-exitcode 9 --print-limit 20"
run() { # name tool cmd...
  n=$1; t=$2; shift; shift
  timeout 900 $CS --tool $t "$@" > $O/${n}_$t.log 2>&1; echo "exit code $?" >> $O/${n}_$t.log
  echo "== $n $t: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|exit code' $O/${n}_$t.log | tr '\n' ' ')"
}
for tool in memcheck racecheck; do
  run shanchen_fused $tool python tools/run2d.py shanchen 256 3 0
  run binary2d_fused $tool python tools/run2d.py binary 256 3 2
  run binary2d_default $tool python tools/run2d.py binary 256 3 0
  run d3q19_one_pass $tool python tools/run3d.py 150 40 24 3 0
  run d3q19_two_pass $tool python tools/run3d.py 150 40 24 3 5
done
NG=$(nvidia-smi -L | wc -l)
if [ "$NG" -ge 2 ]; then
  timeout 1200 $CS --tool memcheck --target-processes all python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/run3d.py 60 30 40 3 0 > $O/d3q19_overlapped_2gpu_memcheck.log 2>&1; echo "exit code $?" >> $O/d3q19_overlapped_2gpu_memcheck.log
  echo "== 2gpu memcheck: $(grep -E 'ERROR SUMMARY|exit code' $O/d3q19_overlapped_2gpu_memcheck.log | sort | uniq -c | tr '\n' ' ')"
fi
