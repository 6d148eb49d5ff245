# K3 launch-bound A/B: 10M build with the prune kernel compiled without a min-blocks bound (round 1), with 2, and with 3 (in tree); interleaved
for r in 1 2 3 4; do
for lib in ab/prune_minb0.so ab/prune_minb2.so libpann_cuda.so; do
  echo -n "$lib  "; PANN_LIB_PATH=$PWD/parlayann_b200/$lib SORT=1 python scripts/build_only.py 10000000 2>&1 | tail -1
done
done
