# search parity tests, then the bench at 1M and 10M (no CPU baseline)
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -4
for n in 1000000 10000000; do
  python bench.py --n $n --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/sab_$n.json 2> gpurun_out/sab_$n.err || tail -20 gpurun_out/sab_$n.err
  python - $n <<'PY'
import json, sys
d = json.loads(open("gpurun_out/sab_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "QPS %.3fM ms %.4f e2e %.3fM frac %.3f" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["roofline"]["frac"]), "recall_batch", d.get("recall_batch"), "build", round(d["build"]["seconds"], 2), d["config"]["beam"], d["config"]["recall10"], d["config"]["avg_cmps"])
PY
done
