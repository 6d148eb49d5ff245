mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_groundtruth_gpu.py -x -q -m gpu 2>&1 | tail -2
SEEDS=512 python scripts/gt_shard_diag.py 2>&1 | tail -2
N=10000000 SEEDS=512 REPS=3 python scripts/gt_shard_diag.py 2>&1 | tail -2
GT_RUNS=2 python scripts/gt_scale.py big 2>&1 | tail -4
REPS=1 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/gt_shard_launches.csv python scripts/gt_shard_diag.py > gpurun_out/gt_shard_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/gt_shard_launches.csv')) if len(r)>5 and r[0].isdigit()]
for r in rows:
    if 'pann_gt' in r[4]:
        print(r[4][:40], r[7] if len(r)>7 else '', r[-1], r[-2])
PY
