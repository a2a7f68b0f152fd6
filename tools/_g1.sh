set -x
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02_bench_v6_8gpu.json 2> gpurun_out/r02_bench_v6_8gpu.err; tail -c 300 gpurun_out/r02_bench_v6_8gpu.json; tail -3 gpurun_out/r02_bench_v6_8gpu.err
