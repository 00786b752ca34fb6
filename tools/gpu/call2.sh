#!/bin/bash
# round-2 GPU call 2: full GPU suite on the new code (kRefine kernels, direct delivery, flags), bench N=1, C4 timings
mkdir -p gpurun_out/c2; O=gpurun_out/c2
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -30 $O/pytest_gpu.txt
timeout 400 python tools/perf_configs.py > $O/perf_configs.txt 2>&1
tail -14 $O/perf_configs.txt
timeout 300 python tools/e2e_probe.py > $O/e2e_probe.txt 2>&1
timeout 400 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
tail -c 1500 $O/bench_n1.err
timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c4_n1.json 2> $O/bench_c4_n1.err; echo "bench c4 rc=$?"
