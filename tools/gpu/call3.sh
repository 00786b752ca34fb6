#!/bin/bash
mkdir -p gpurun_out/c3; O=gpurun_out/c3
timeout 200 python tools/gpu/dbg_flags.py > $O/dbg_flags.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q --deselect "tests/test_delivery_and_sync.py::test_render_device_synced_waits_for_the_consumer_and_signals" > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -40 $O/pytest_gpu.txt
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1; tail -12 $O/perf_configs.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f32_kernel -c 1 -o $O/c4_refine python tools/profile_one.py f32 960 540 1 1 c4 > $O/ncu_c4.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f64_kernel -c 1 -o $O/c4_f64 python tools/profile_one.py f64 960 540 1 1 c4 > $O/ncu_c4_f64.log 2>&1
ls -la $O
