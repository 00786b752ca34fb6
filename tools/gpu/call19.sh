#!/bin/bash
# zero-numerator normalize: whole GPU suite, kernel times, the checked run, C4 shard balance by tile height
mkdir -p gpurun_out/c19; O=gpurun_out/c19
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -4 $O/pytest_gpu.txt
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1; grep f64 $O/perf_configs.txt
timeout 900 python tools/sanitize_run.py --checked > $O/checked_run.txt 2>&1; echo "rc=$?" >> $O/checked_run.txt; tail -3 $O/checked_run.txt
timeout 600 python tools/perf_c4_shards.py 1 > $O/c4_shards.txt 2>&1; cat $O/c4_shards.txt
