#!/bin/bash
# round 2, evidence run on one GPU: whole GPU test suite, default bench line in full, reference arm, FullChannel at N = 1,
# ncu launch list + DRAM traffic per launch + full capture of the dominant kernel
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02p; mkdir -p $O
timeout 900 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref rc=$?"
timeout 200 python bench.py --workload full --steps 400 --warmup 40 --no-e2e --no-cpu --no-secondary > $O/bench_full_n1.json 2> $O/bench_full_n1.err
timeout 200 python bench.py --precision f32 --no-e2e --no-cpu --no-secondary > $O/bench_f32.json 2> $O/bench_f32.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_default.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-secondary > $O/ncu_launches.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:"k_step_one|k_collide_rim|k_phase" -s 9 -c 9 --csv --log-file $O/traffic_default.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --no-secondary > $O/ncu_traffic.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_step_one|k_collide_rim|k_phase" -s 9 -c 9 --csv --log-file $O/traffic_two_pass.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --no-secondary --variant 5 > $O/ncu_traffic5.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"k_step_one" -s 4 -c 1 -o $O/step_one_512cubed python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-secondary > $O/ncu_full.log 2>&1
ncu -i $O/step_one_512cubed.ncu-rep --page details > $O/step_one_512cubed.ncu.txt 2>/dev/null
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_b2_" -s 6 -c 6 --csv --log-file $O/traffic_binary2d.csv python tools/run2d.py binary 8192 6 0 > $O/ncu_b2.log 2>&1
echo done
