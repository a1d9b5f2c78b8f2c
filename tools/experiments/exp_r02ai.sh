#!/bin/bash
# round 2: 2 GPUs: the whole multi-GPU part of the suite on the final tree (slab ring both transports and precisions, in-place
# variants, the C++ host on several GPUs incl. in place)
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02ai; mkdir -p $O
timeout 500 python -m pytest tests/test_gpu_multi.py tests/test_gpu_hosts.py -q -m gpu -k "slab_ring or several_gpus" > $O/pytest_multi_n2.log 2>&1
rc=$?; echo "pytest rc=$rc" >> $O/pytest_multi_n2.log; grep -E "passed|failed|FAILED|Error" $O/pytest_multi_n2.log | head -20
echo done
