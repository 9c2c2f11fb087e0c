for n in 2 4 8; do python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/bench_r2_${n}gpu.json 2> gpurun_out/bench_r2_${n}gpu.err; tail -c 200 gpurun_out/bench_r2_${n}gpu.err; done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 tools/profile_batch_bench.py --reference 2 > gpurun_out/c5_8gpu.json 2> gpurun_out/c5_8gpu.err; tail -2 gpurun_out/c5_8gpu.err
python -m pytest tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -2
