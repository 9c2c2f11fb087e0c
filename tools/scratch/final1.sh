set -x
python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r2_1gpu_v3.json 2> gpurun_out/bench_r2_1gpu_v3.err; tail -c 300 gpurun_out/bench_r2_1gpu_v3.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_v3.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
python tools/launch_summary.py gpurun_out/r2_launches_v3.csv > gpurun_out/r2_launches_v3_summary.txt 2>&1
