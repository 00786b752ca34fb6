#!/bin/bash
mkdir -p gpurun_out/c8; O=gpurun_out/c8
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -8 $O/pytest_gpu.txt
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1; tail -12 $O/perf_configs.txt
timeout 900 python tools/sanitize_run.py --checked > $O/checked_run.txt 2>&1; echo "rc=$?" >> $O/checked_run.txt; tail -2 $O/checked_run.txt
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f32_kernel -c 1 -o $O/c4_lane python tools/profile_one.py f32 960 540 1 1 c4 > $O/ncu_c4.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f64_kernel -c 1 -o $O/c3_f64 python tools/profile_one.py f64 3840 2160 1 1 room > $O/ncu_c3_f64.log 2>&1
ls -la $O
