set -x
mkdir -p gpurun_out/full
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/full/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/full/pytest_gpu.log
tail -3 gpurun_out/full/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/full/smoke.log 2>&1; tail -2 gpurun_out/full/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/full/bench_c2_n1.json 2> gpurun_out/full/bench_c2_n1.err; tail -c 600 gpurun_out/full/bench_c2_n1.json
timeout 600 python bench.py --steps 20 --warmup 3 --track --no-cpu > gpurun_out/full/bench_c2_track.json 2> gpurun_out/full/bench_c2_track.err; tail -c 400 gpurun_out/full/bench_c2_track.json
timeout 600 python bench.py --latency > gpurun_out/full/latency.json 2> gpurun_out/full/latency.err; tail -c 300 gpurun_out/full/latency.json
# ncu evidence of the same step (only after the plain runs above exited): launch list, then one full capture
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 14 -c 16 --csv --log-file gpurun_out/full/launches_c2_b512.csv python tools/profile_step.py --steps 2 > gpurun_out/full/ncu_launches.log 2>&1; tail -1 gpurun_out/full/ncu_launches.log
timeout 900 ncu --set full --import-source on --clock-control none -s 14 -c 16 -f -o gpurun_out/full/full_c2_b512 python tools/profile_step.py --steps 2 > gpurun_out/full/ncu_full.log 2>&1; tail -1 gpurun_out/full/ncu_full.log
