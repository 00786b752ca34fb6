#!/bin/bash
# final single-GPU evidence: suites, smoke, bench lines (C3 with the CPU baseline, C4, C5, reference arm), launch list, other configs
mkdir -p gpurun_out/final; O=gpurun_out/final
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt; tail -3 $O/pytest_gpu.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "rc=$?" >> $O/smoke.txt; tail -2 $O/smoke.txt
timeout 900 python tools/sanitize_run.py --checked > $O/checked_run.txt 2>&1; echo "rc=$?" >> $O/checked_run.txt; tail -2 $O/checked_run.txt
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --workload c4 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c4_n1.json 2> $O/bench_c4_n1.err; echo "bench c4 rc=$?"
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c5_n1.json 2> $O/bench_c5_n1.err; echo "bench c5 rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 0 > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_n1.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_under_ncu.json 2> $O/bench_under_ncu.err
timeout 300 python tools/perf_configs.py > $O/perf_configs.txt 2>&1
python - <<'PY'
import json
for f in ("bench_n1", "bench_c4_n1", "bench_c5_n1", "bench_reference"):
    try:
        l = json.loads(open(f"gpurun_out/final/{f}.json").read().strip().splitlines()[-1])
        if l.get("impl") == "reference":
            print(f, "value", round(l["value"], 2), l["cpu_baseline"]["cores"], "cores"); continue
        print(f, "value", round(l["value"]), "ms", round(l["ms_per_step"], 4), "frac", l["roofline"].get("frac"), "e2e", round(l["e2e"]["ms_per_frame"], 3), "f64", round(l["f64"]["ms_per_frame"], 3), "f64 e2e", round(l["f64"]["e2e"]["ms_per_frame"], 3), l["frame_check"].get("f32_meets_99.9pct"), l["clocks"]["reasons"], "launches", l["gpu_launches"])
    except Exception as e:
        print(f, "unreadable", e)
PY
