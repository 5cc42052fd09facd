#!/bin/bash
# GPU-box passes that produce what profiles/ holds for a round. One ncu run per gpurun call, each after the same command exited 0
# without ncu. Usage (through gpurun): bash tools/profile_round.sh <tag> <pass>      outputs: gpurun_out/<tag>_*
#   pass bench : bench line (both arms) + the ncu launch list of a short bench run
#   pass mc    : multicorrelator kernel time + one --set full capture of it
#   pass sweep : the five-config sweep
set -x
T=${1:-r02}
P=${2:-bench}
O=gpurun_out
case $P in
bench)
    python bench.py > $O/${T}_bench_n1.json 2> $O/${T}_bench_n1.err || exit 1
    python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_reference.json 2> $O/${T}_bench_reference.err || exit 1
    python bench.py --steps 2 --warmup 3 --no-cpu --no-weak > /dev/null 2>&1 || exit 1
    ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${T}_bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-weak > $O/${T}_ncu_bench.log 2>&1
    ;;
mc)
    python tools/mc_tune.py > $O/${T}_mc_tune.log 2>&1 || exit 1
    python tools/mc_tune.py 2048 3 >> $O/${T}_mc_tune.log 2>&1
    python tools/mc_tune.py 2048 5 >> $O/${T}_mc_tune.log 2>&1
    ncu --set full --clock-control none --import-source on -k regex:multicorr_window -s 3 -c 1 -o $O/${T}_mc_full -f python tools/mc_tune.py > $O/${T}_ncu_mc.log 2>&1
    ;;
sweep)
    python tools/config_sweep.py 8 10 > $O/${T}_config_sweep.json 2> $O/${T}_config_sweep.err
    ;;
esac
