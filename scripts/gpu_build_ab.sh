# 10M build A/B of several builds of the library (parlayann_b200/<path>.so), interleaved
for r in 1 2 3; do
for lib in "$@"; do
  echo -n "$lib  "; PANN_LIB_PATH=$PWD/parlayann_b200/$lib SORT=1 python scripts/build_only.py 10000000 2>&1 | tail -1
done
done
