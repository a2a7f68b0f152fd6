set -x
timeout 900 python -m pytest tests/test_gpu_frame.py tests/test_gpu_adapter_host.py -m gpu -x -q > gpurun_out/r02_gputest9.log 2>&1; tail -5 gpurun_out/r02_gputest9.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r02_bench_v2.json 2> gpurun_out/r02_bench_v2.err; tail -5 gpurun_out/r02_bench_v2.err; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_v2.json')); print(d['value'], d['e2e'], d['e2e_header_only'])"
