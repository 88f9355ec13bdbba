#!/bin/bash
# round-1 final capture: bench lines, launch lists, full-counter captures (run under gpurun from the repo root)
set -x
O=gpurun_out/final
mkdir -p $O
python bench.py --steps 5 --warmup 3 > $O/bench_c2.json 2> $O/bench_c2.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_c2_ref.json 2> $O/bench_c2_ref.err
python bench.py --workload c1 --steps 5 --warmup 3 > $O/bench_c1.json 2> $O/bench_c1.err
timeout 1200 python bench.py --workload c4s --steps 3 --warmup 3 > $O/bench_c4s.json 2> $O/bench_c4s.err
NCU="ncu --clock-control none --nvtx --nvtx-include timed/"
$NCU --metrics gpu__time_duration.sum --csv --log-file $O/launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
$NCU --metrics gpu__time_duration.sum --csv --log-file $O/launches_c4s.csv python bench.py --workload c4s --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
$NCU --set full --import-source on -k regex:"ivfb200" -o $O/full_c2 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_c2.log 2>&1
$NCU --set full --import-source on -k regex:"ivfb200" -o $O/full_c4s -f python bench.py --workload c4s --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_c4s.log 2>&1
free -g > $O/host.txt; nproc >> $O/host.txt
