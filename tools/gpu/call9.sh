#!/bin/bash
mkdir -p gpurun_out/c9; O=gpurun_out/c9
timeout 900 python tools/sanitize_run.py --checked > $O/checked_run.txt 2>&1; echo "rc=$?" >> $O/checked_run.txt; tail -6 $O/checked_run.txt
timeout 300 python tools/perf_c4.py 960 540 > $O/perf_c4_small.txt 2>&1; cat $O/perf_c4_small.txt | tail -2
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -8 $O/pytest_gpu.txt
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1; tail -3 $O/perf_configs.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f32_kernel -c 1 -o $O/c4_bvh python tools/profile_one.py f32 960 540 1 1 c4 > $O/ncu_c4.log 2>&1
ls -la $O
