#!/bin/bash
# final capture of round 1 (after the two-product coarse GEMM became the default)
O=gpurun_out/final2
mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > $O/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1
python bench.py --steps 5 --warmup 3 > $O/bench_c2.json 2> $O/bench_c2.err
python bench.py --workload c1 --steps 5 --warmup 3 > $O/bench_c1.json 2>/dev/null
python bench.py --workload c5s --steps 5 --warmup 3 > $O/bench_c5s.json 2>/dev/null
timeout 600 python bench.py --workload c4s --steps 3 --warmup 3 > $O/bench_c4s.json 2>/dev/null
timeout 300 python bench.py --workload c3 --steps 3 --warmup 3 > $O/bench_c3.json 2>/dev/null
timeout 600 python bench.py --workload c4n --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c4n.json 2>/dev/null
NCU="ncu --clock-control none --nvtx --nvtx-include timed/"
$NCU --metrics gpu__time_duration.sum --csv --log-file $O/launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
$NCU --metrics gpu__time_duration.sum --csv --log-file $O/launches_c4s.csv python bench.py --workload c4s --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
