for r in 1 2; do
for cfg in ${CFGS:-160,512,3 144,256,4 152,256,4 136,256,4 156,256,4}; do
  echo -n "P1=$cfg  "; PANN_PRUNE_P1=$cfg SORT=1 python scripts/build_only.py 10000000 2>&1 | tail -1
done
done
