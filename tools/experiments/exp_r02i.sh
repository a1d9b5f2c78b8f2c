#!/bin/bash
# round 2, run H2: the default bench line in full, ncu launch list + DRAM traffic + full capture of the dominant kernel,
# compute-sanitizer logs.  LBM_TUNE_* exported by the caller apply to every command.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/r02i; mkdir -p $O
timeout 60 tools/dispatch_order 64 525 33 > $O/dispatch_order.json 2>&1; cat $O/dispatch_order.json
timeout 60 tools/dispatch_order 64 100 4 >> $O/dispatch_order.json 2>&1
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"
tail -c 600 $O/bench_default.json
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref rc=$?"
# launch list of the very command of the bench (short), then DRAM bytes per launch
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_default.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-secondary > $O/ncu_launches.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:"k_step_one|k_collide_rim|k_phase" -s 9 -c 9 --csv --log-file $O/traffic_default.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --no-secondary > $O/ncu_traffic.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_collide_core|k_collide_rim|k_phase" -s 9 -c 9 --csv --log-file $O/traffic_v5.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --no-secondary --variant 5 > $O/ncu_traffic5.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"k_step_one" -s 4 -c 1 -o $O/step_one_512cubed python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-secondary > $O/ncu_full.log 2>&1
ncu -i $O/step_one_512cubed.ncu-rep --page details > $O/step_one_512cubed.ncu.txt 2>/dev/null
timeout 1500 bash tools/sanitize.sh $O/sanitize > $O/sanitize_summary.txt 2>&1
cat $O/sanitize_summary.txt
echo done
