# quick GPU check: parity tests + a short bench without the CPU baseline
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -8
python bench.py --steps 20 --warmup 3 --no-cpu-baseline ${BENCH_ARGS} > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err || tail -20 gpurun_out/bench_quick.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_quick.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step")}, "recall_batch", d.get("recall_batch"), "e2e", d["e2e"]["value"], "frac", d["roofline"]["frac"], "build", d["build"]["seconds"], d["build"]["phase_seconds"], "gt_ms", d["groundtruth"]["device_ms"], d["config"]["beam"], d["config"]["recall10"], d["clocks"])
PY
