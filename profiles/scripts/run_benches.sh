#!/bin/bash
# N=1 bench lines of every BASELINE configuration + the reference arm (run on the GPU box)
set -u
mkdir -p gpurun_out
python -m pytest tests -q -m gpu 2>&1 | tail -4
cp gpurun_out/refdev_report.json gpurun_out/r02_refdev_report.json 2>/dev/null
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r02_bench_ref.json 2> gpurun_out/r02_bench_ref.err; echo ref rc=$?
python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_c2.json 2> gpurun_out/r02_bench_c2.err; echo c2 rc=$?
for spec in "c1 30 --no-cpu-baseline" "c3 10 --no-cpu-baseline" "c4 3 --no-cpu-baseline" "c5-8k 5 --no-cpu-baseline" "c5-batch 2 --no-cpu-baseline --no-cuda-baseline"; do
  set -- $spec; cfg=$1; steps=$2; shift 2
  python bench.py --config $cfg --steps $steps --warmup 3 "$@" > gpurun_out/r02_bench_$cfg.json 2> gpurun_out/r02_bench_$cfg.err; echo $cfg rc=$?
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02_bench_*.json")):
    try:
        d = json.load(open(f))
        cb = d.get("cuda_baseline") or {}
        print(f.split("/")[-1], round(d["value"], 1), d["unit"], round(d["ms_per_step"], 3), "ms e2e", round(d["e2e"]["value"], 1), "frac", d.get("roofline", {}).get("frac"), "ref_ms", cb.get("ms_per_frame"), "speedup", cb.get("speedup_device"))
    except Exception as e:
        print(f, "unreadable", e)
PY
