set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_resample.py -x -q -m gpu > gpurun_out/t_resample.log 2>&1; echo "rc=$?" >> gpurun_out/t_resample.log; tail -15 gpurun_out/t_resample.log
timeout 600 python tools/resample_timing.py 1024 > gpurun_out/resample_timing.log 2>&1; cat gpurun_out/resample_timing.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fe_resample -s 2 -c 1 -f -o gpurun_out/rs_hann python tools/prof_run.py resample 256 > gpurun_out/ncu_rs.log 2>&1; tail -3 gpurun_out/ncu_rs.log
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; echo "rc=$?" >> gpurun_out/t_all.log; tail -5 gpurun_out/t_all.log
