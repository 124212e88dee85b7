mkdir -p gpurun_out
timeout 900 python bench.py --workload cfg2pcm --steps 10 --warmup 3 > gpurun_out/bench_cfg2pcm.json 2> gpurun_out/bench_cfg2pcm.err; echo "rc=$?"; cat gpurun_out/bench_cfg2pcm.json | cut -c1-3000; tail -5 gpurun_out/bench_cfg2pcm.err
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "rc=$?"; cut -c1-400 gpurun_out/bench_cfg2.json
