#!/bin/bash
# mixed delivery: its tests, then the bench line at N = 1 (C3 and C5)
mkdir -p gpurun_out/c11; O=gpurun_out/c11
timeout 900 python -m pytest tests/test_delivery_and_sync.py tests/test_abi.py -m gpu -x -q > $O/pytest_delivery.txt 2>&1; echo "rc=$?" >> $O/pytest_delivery.txt; tail -5 $O/pytest_delivery.txt
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; tail -c 300 $O/bench_n1.err
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c5_n1.json 2> $O/bench_c5_n1.err; echo "bench c5 rc=$?"
python - <<'PY'
import json
for f in ("bench_n1", "bench_c5_n1"):
    try:
        l = json.loads(open(f"gpurun_out/c11/{f}.json").read().strip().splitlines()[-1])
        e = l["e2e"]
        print(f, "kernel", round(l["ms_per_step"], 4), "e2e", round(e["ms_per_frame"], 3), "min", round(e["ms_per_frame_min"], 3), "bands", e["row_bands"], e["row_bands_direct"], "d2h", e["d2h_bytes_per_step"],
              "direct", round(e["other_delivery"]["ms_per_frame"], 3), "staged", round(e["staged_delivery"]["ms_per_frame"], 3), "f64 e2e", round(l["f64"]["e2e"]["ms_per_frame"], 3), l["frame_check"])
    except Exception as ex:
        print(f, "unreadable", ex)
PY
