timeout 300 python -m pytest tests -m gpu -q -x -k "host_sweep or ray_marcher" 2>&1 | tail -2
timeout 900 python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"; cut -c1-1800 gpurun_out/bench_c5.json; tail -3 gpurun_out/bench_c5.err
timeout 600 python bench.py --workload c1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "c1 rc=$?"; cut -c1-400 gpurun_out/bench_c1.json
