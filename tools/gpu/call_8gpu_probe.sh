#!/bin/bash
mkdir -p gpurun_out/g8p; O=gpurun_out/g8p
{ nvidia-smi topo -m; lscpu | grep -E "Model name|Socket|NUMA|^CPU\(s\)"; numactl -H 2>/dev/null | head -12; cat /proc/meminfo | grep -E "MemTotal|Huge" ; } > $O/topology.txt 2>&1
for n in 1 2 4 8; do timeout 120 tools/micro/pcie_multi $n; done > $O/pcie_multi.txt 2>&1
cat $O/pcie_multi.txt
