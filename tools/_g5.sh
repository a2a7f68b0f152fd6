set -x
for w in match stream; do timeout 120 python tools/prof_target.py $w > gpurun_out/r02_t3_$w.log 2>&1; tail -1 gpurun_out/r02_t3_$w.log; done
timeout 300 python tests/bench_configs.py --config 2 > gpurun_out/r02_cfg2_v3.jsonl 2>gpurun_out/cfg2.err; cat gpurun_out/r02_cfg2_v3.jsonl | cut -c1-500; tail -3 gpurun_out/cfg2.err
python tools/prof_target.py stream > gpurun_out/plain1.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 40 --csv --log-file gpurun_out/r02_launches_stream_v1.csv python tools/prof_target.py stream > gpurun_out/ncu1.log 2>&1
python tools/prof_target.py match > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o gpurun_out/r02_prof_match_v2 python tools/prof_target.py match > gpurun_out/ncu2.log 2>&1
