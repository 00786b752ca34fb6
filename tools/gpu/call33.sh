#!/bin/bash
# small-scene kernel at 7 / 8 blocks per SM: suite, checked run, bench N = 1, ncu of the bench kernel, launch list, other configs
mkdir -p gpurun_out/c33; O=gpurun_out/c33
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt; tail -3 $O/pytest_gpu.txt
timeout 900 python tools/sanitize_run.py --checked > $O/checked_run.txt 2>&1; echo "rc=$?" >> $O/checked_run.txt; tail -2 $O/checked_run.txt
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --workload c4 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c4_n1.json 2> $O/bench_c4_n1.err; echo "bench c4 rc=$?"
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c5_n1.json 2> $O/bench_c5_n1.err; echo "bench c5 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f32_small_kernel -c 1 -o $O/c3_small python tools/profile_one.py f32 3840 2160 1 1 room > $O/ncu_c3_small.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_f64_kernel -c 1 -o $O/c4_f64 python tools/profile_one.py f64 960 540 1 1 c4 > $O/ncu_c4_f64.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_n1.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_under_ncu.json 2> $O/bench_under_ncu.err
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1
python - <<'PY'
import json
for f in ("bench_n1", "bench_c4_n1", "bench_c5_n1"):
    l = json.loads(open(f"gpurun_out/c33/{f}.json").read().strip().splitlines()[-1])
    print(f, "value", round(l["value"]), "ms", round(l["ms_per_step"], 4), "frac", l["roofline"].get("frac"), "e2e", round(l["e2e"]["ms_per_frame"], 3), "f64", round(l["f64"]["ms_per_frame"], 3), "f64 e2e", round(l["f64"]["e2e"]["ms_per_frame"], 3), l["frame_check"].get("f32_meets_99.9pct"), l["clocks"]["reasons"])
PY
