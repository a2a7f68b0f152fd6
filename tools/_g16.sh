set -x
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_bench_v2_2gpu.json 2> gpurun_out/r02_bench_v2_2gpu.err; tail -5 gpurun_out/r02_bench_v2_2gpu.err; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_v2_2gpu.json')); print(d['value'], d['e2e']['value'], d['e2e_header_only']['value']); print([ (m['mode'],m['frames_per_sec'],m['planted_found']) for m in d['inspect']['modes']]); print(d['bridge_match']['frames_per_sec'], d['latency'])"
