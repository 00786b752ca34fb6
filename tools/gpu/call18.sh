#!/bin/bash
# evidence for the parity kernel after the footprint changes: ncu captures (C3 4K, C4 960x540), bench lines N = 1 (C3, C4)
mkdir -p gpurun_out/c18; O=gpurun_out/c18
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --workload c4 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c4_n1.json 2> $O/bench_c4_n1.err; echo "bench c4 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f64_kernel -c 1 -o $O/c3_f64 python tools/profile_one.py f64 3840 2160 1 1 room > $O/ncu_c3_f64.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f64_kernel -c 1 -o $O/c4_f64 python tools/profile_one.py f64 960 540 1 1 c4 > $O/ncu_c4_f64.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_n1.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_under_ncu.json 2> $O/bench_under_ncu.err
ls -la $O | head -20
