#!/bin/bash
# one gpurun call that refreshes the conditioning row's evidence: GPU tests, smoke, timings, bench line, ncu launch list and full capture
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; echo "rc=$?" >> gpurun_out/t_all.log; tail -3 gpurun_out/t_all.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python tools/resample_timing.py 1024 2>&1 | tee gpurun_out/resample_timing.log
timeout 600 python bench.py --workload cfg2pcm --steps 20 --warmup 3 > gpurun_out/bench_cfg2pcm.json 2> gpurun_out/bench_cfg2pcm.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_cfg2pcm.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_cfg2pcm.csv python bench.py --workload cfg2pcm --steps 2 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fe_resample -s 2 -c 1 -f -o gpurun_out/rs_final python tools/prof_run.py resample 1024 > gpurun_out/ncu_rs.log 2>&1; tail -1 gpurun_out/ncu_rs.log
