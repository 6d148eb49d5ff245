# A/B of several builds of the library (parlayann_b200/<path>.so), same box, via the batch-size sweep, interleaved
mkdir -p gpurun_out
for r in 1 2; do
for lib in "$@"; do
  echo "=== $lib"
  PANN_LIB_PATH=$PWD/parlayann_b200/$lib TUNES=${TUNES:-auto} python scripts/nq_sweep.py 2>&1 | grep -E "nq +(10000|1000000) "
done
done
