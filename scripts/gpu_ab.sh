# A/B: bench twice with different environment settings (no tests)
mkdir -p gpurun_out
for cfg in "$@"; do
  env $cfg python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ab.json 2> gpurun_out/ab.err || tail -5 gpurun_out/ab.err
  python - "$cfg" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
print(sys.argv[1], "QPS %.3fM ms %.4f e2e %.3fM" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6), "build", round(d["build"]["seconds"],3))
PY
done
