#!/bin/bash
# round-2 GPU call 1: sanity, sanitizers on the round-1 code, PCIe delivery probe, baselines on this box
mkdir -p gpurun_out/c1; O=gpurun_out/c1
nvidia-smi --query-gpu=name,clocks.max.sm,pcie.link.gen.current,pcie.link.width.current --format=csv > $O/gpu.txt 2>&1
nproc >> $O/gpu.txt; lscpu | grep -E "Model name|Socket|NUMA" >> $O/gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "smoke rc=$?" >> $O/smoke.txt
timeout 120 tools/micro/pcie_direct > $O/pcie_direct.txt 2>&1
timeout 300 python tools/sanitize_run.py > $O/sanitize_plain.txt 2>&1; echo "rc=$?" >> $O/sanitize_plain.txt
for tool in memcheck racecheck initcheck synccheck; do
  timeout 600 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_run.py > $O/sanitizer_$tool.txt 2>&1; echo "rc=$?" >> $O/sanitizer_$tool.txt
done
timeout 300 python tools/e2e_probe.py > $O/e2e_probe.txt 2>&1
timeout 400 python tools/perf_configs.py > $O/perf_configs.txt 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err
tail -3 $O/*.txt | tail -80
