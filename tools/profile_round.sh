#!/bin/bash
# One GPU-box pass that produces everything profiles/ holds for a round: bench lines (both arms), the launch list of the
# bench command, --set full captures of the acquisition and multicorrelator kernels, the five-config sweep.
# Usage (through gpurun): bash tools/profile_round.sh <tag>      outputs: gpurun_out/<tag>_*
set -x
T=${1:-r01}
O=gpurun_out
python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_reference.json 2> $O/${T}_bench_reference.err || exit 1
python bench.py --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/${T}_bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/${T}_ncu_bench.log 2>&1
python tools/prof_run.py acq C2 10 > $O/${T}_prof_acq.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"inv_lower|inv_upper|fwd_upper|fwd_lower" -s 4 -c 4 -o $O/${T}_acq_full -f python tools/prof_run.py acq C2 10 > $O/${T}_ncu_acq.log 2>&1
python tools/mc_tune.py > $O/${T}_mc_tune.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:multicorr_window -s 3 -c 1 -o $O/${T}_mc_full -f python tools/mc_tune.py > $O/${T}_ncu_mc.log 2>&1
python tools/config_sweep.py 8 10 > $O/${T}_config_sweep.json 2> $O/${T}_config_sweep.err
