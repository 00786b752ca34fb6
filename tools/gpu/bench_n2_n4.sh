#!/bin/bash
# four GPUs, one box: the bench at N = 2 and N = 4
mkdir -p gpurun_out/g4; O=gpurun_out/g4
for n in 2 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29530 + n)) bench.py --gpus $n --steps 20 --warmup 3 > $O/bench_n$n.json 2> $O/bench_n$n.err; echo "n$n rc=$?"
done
python - <<'PY'
import json
for n in (2, 4):
    l = json.loads(open(f"gpurun_out/g4/bench_n{n}.json").read().strip().splitlines()[-1])
    print("N", n, "ms", round(l["ms_per_step"], 4), "value", round(l["value"]), "kernels", l.get("kernel_ms_per_rank", {}).get("peer_frame_destination"), "e2e", round(l["e2e"]["ms_per_frame"], 3), "f64", round(l["f64"]["ms_per_frame"], 3), l["frame_check"], "launches", l["gpu_launches"])
PY
