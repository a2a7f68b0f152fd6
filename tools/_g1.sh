set -x
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_full5.log 2>&1; tail -3 gpurun_out/r02_gputest_full5.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r02_bench_v6.json 2> gpurun_out/r02_bench_v6.err; tail -c 200 gpurun_out/r02_bench_v6.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_v6_reference_arm.json 2>/dev/null
for w in ordered match; do
  timeout 300 python tools/prof_target.py $w > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 6 -c $([ $w = ordered ] && echo 2 || echo 1) -o gpurun_out/r02_ncu_$w -f python tools/prof_target.py $w > gpurun_out/ncu_$w.log 2>&1; tail -1 gpurun_out/ncu_$w.log
done
du -sh gpurun_out
