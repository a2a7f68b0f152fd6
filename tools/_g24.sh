set -x
P=gpurun_out
for w in wire chunk; do timeout 300 python tools/prof_target.py $w > $P/r02_final_$w.log 2>&1; tail -1 $P/r02_final_$w.log; done
timeout 300 python tools/bench_encode.py > $P/r02_final_encode.json 2>$P/enc.err; cat $P/r02_final_encode.json | cut -c1-300
for w in plain ordered inspect_ab match stream bridge; do timeout 120 python tools/prof_target.py $w > $P/r02_final_$w.log 2>&1; tail -1 $P/r02_final_$w.log; done
# full captures
python tools/prof_target.py plain > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o $P/r02_ncu_plain python tools/prof_target.py plain > $P/n.log 2>&1
python tools/prof_target.py ordered > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 8 -c 2 -o $P/r02_ncu_ordered python tools/prof_target.py ordered > $P/n.log 2>&1
python tools/prof_target.py inspect_ab > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o $P/r02_ncu_inspect python tools/prof_target.py inspect_ab > $P/n.log 2>&1
python tools/prof_target.py match > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o $P/r02_ncu_match python tools/prof_target.py match > $P/n.log 2>&1
python tools/prof_target.py stream > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_stream_walk|k_radix_scatter|k_frame_tiles" -s 18 -c 4 -o $P/r02_ncu_stream python tools/prof_target.py stream > $P/n.log 2>&1
python tools/bench_encode.py --steps 3 > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_encode_messages -s 3 -c 1 -o $P/r02_ncu_encode python tools/bench_encode.py --steps 3 > $P/n.log 2>&1
# launch lists
python bench.py --steps 3 --warmup 3 --no-extras > $P/p.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $P/r02_launches_bench.csv python bench.py --steps 3 --warmup 3 --no-extras > $P/n.log 2>&1
python tools/prof_target.py stream > $P/p.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 52 -c 26 --csv --log-file $P/r02_launches_stream.csv python tools/prof_target.py stream > $P/n.log 2>&1
ls -la $P | tail -20
