#!/bin/bash
# new dispatch thresholds (parity kernel: hierarchical form from 65 spheres on, 8 blocks per SM; fast mode -> parity kernel beyond 128 shapes)
mkdir -p gpurun_out/c25; O=gpurun_out/c25
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -4 $O/pytest_gpu.txt
timeout 900 python tools/sanitize_run.py --checked > $O/checked_run.txt 2>&1; echo "rc=$?" >> $O/checked_run.txt; tail -3 $O/checked_run.txt
timeout 300 python tools/perf_mid_scenes.py 32 64 100 150 200 400 800 1500 3000 > $O/mid_scenes.txt 2>&1; cat $O/mid_scenes.txt
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1; grep -E "C4|C3 room 3840x2160  |C2" $O/perf_configs.txt
