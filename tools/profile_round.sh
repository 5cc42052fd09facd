set -x
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/s7_prof_bench.json 2> gpurun_out/s7_prof_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/s7_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/s7_ncu_bench.log 2>&1
python tools/prof_run.py acq C2 10 > gpurun_out/s7_prof_acq.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"inv_lower|inv_upper|fwd_upper|fwd_lower" -s 4 -c 4 -o gpurun_out/s7_acq_full -f python tools/prof_run.py acq C2 10 > gpurun_out/s7_ncu_acq.log 2>&1
