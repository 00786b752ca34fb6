#!/bin/bash
# small-scene kernel: an odd count's last shape in the last slot, tested on its own
mkdir -p gpurun_out/c26; O=gpurun_out/c26
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt
tail -4 $O/pytest_gpu.txt
timeout 900 python tools/sanitize_run.py --checked > $O/checked_run.txt 2>&1; echo "rc=$?" >> $O/checked_run.txt; tail -2 $O/checked_run.txt
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1; grep f32 $O/perf_configs.txt
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
l = json.loads(open("gpurun_out/c26/bench_n1.json").read().strip().splitlines()[-1])
print("value", round(l["value"]), "ms", round(l["ms_per_step"], 4), "frac", round(l["roofline"]["frac"], 4), "e2e", round(l["e2e"]["ms_per_frame"], 3), l["frame_check"])
PY
