#!/bin/bash
# whole GPU suite (bit-exactness), then kernel times of every config: the check after a kernel change
mkdir -p gpurun_out/suite; O=gpurun_out/suite
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -8 $O/pytest_gpu.txt
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1; cat $O/perf_configs.txt
