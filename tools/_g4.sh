set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest2.log 2>&1; tail -5 gpurun_out/r02_gputest2.log
for w in inspect_ab match bridge plain; do timeout 120 python tools/prof_target.py $w > gpurun_out/r02_t2_$w.log 2>&1; tail -1 gpurun_out/r02_t2_$w.log; done
timeout 300 python tests/bench_configs.py --config 2 > gpurun_out/r02_cfg2_v2.jsonl 2>gpurun_out/cfg2.err; cat gpurun_out/r02_cfg2_v2.jsonl | cut -c1-400; tail -3 gpurun_out/cfg2.err
