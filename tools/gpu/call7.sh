#!/bin/bash
mkdir -p gpurun_out/c7; O=gpurun_out/c7
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone.txt 2>&1; echo "rc=$?" >> $O/pytest_noalloc_alone.txt; tail -4 $O/pytest_noalloc_alone.txt
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -8 $O/pytest_gpu.txt
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1; tail -12 $O/perf_configs.txt
timeout 900 python tools/sanitize_run.py --checked > $O/checked_run.txt 2>&1; echo "rc=$?" >> $O/checked_run.txt; tail -3 $O/checked_run.txt
( compute-sanitizer --tool memcheck python -c "print(1)"; echo "rc=$?" ) > $O/sanitizer_refusal.txt 2>&1
timeout 120 tools/micro/pcie_direct > $O/pcie_direct.txt 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f32_kernel -c 1 -o $O/c4_coop python tools/profile_one.py f32 960 540 1 1 c4 > $O/ncu_c4.log 2>&1
ls -la $O
