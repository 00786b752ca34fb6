#!/bin/bash
# eight GPUs, one box: host-side DMA ceiling probe, then the bench at N = 8 (frame pages first-touch / interleaved), 4, 2 and C4 at N = 8
mkdir -p gpurun_out/g8; O=gpurun_out/g8
{ nvidia-smi topo -m; lscpu | grep -E "Model name|Socket|NUMA|^CPU\(s\)"; cat /sys/devices/system/node/has_memory; grep -E "MemTotal|Huge" /proc/meminfo; } > $O/topology.txt 2>&1
for n in 1 2 4 8; do timeout 120 tools/micro/pcie_multi $n; done > $O/pcie_multi.txt 2>&1
cat $O/pcie_multi.txt
run() { # name nproc port extra...
  local name=$1 n=$2 port=$3; shift 3
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --warmup 3 "$@" > $O/$name.json 2> $O/$name.err
  echo "$name rc=$?"; tail -c 300 $O/$name.err | grep -v "^$" | tail -3
}
run bench_n8 8 29521 --steps 20
run bench_n8_interleaved 8 29522 --steps 20 --numa-interleave
run bench_n4 4 29523 --steps 20
run bench_n2 2 29524 --steps 20
run bench_c4_n8 8 29525 --steps 5 --workload c4
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/g8/bench*.json")):
    try:
        l = json.loads(open(f).read().strip().splitlines()[-1])
        e = l["e2e"]; sp = l.get("e2e_single_process") or {}
        print(f.split("/")[-1], "ms", round(l["ms_per_step"], 4), "e2e", round(e["ms_per_frame"], 3), e.get("frame_pages"), "other", round(e["other_delivery"]["ms_per_frame"], 3), "single", sp.get("ms_per_frame"))
    except Exception as ex:
        print(f, "unreadable", ex)
PY
