# A/B of several builds of the library (parlayann_b200/ab/*.so), same box, via the batch-size sweep
mkdir -p gpurun_out
for lib in "$@"; do
  echo "=== $lib"
  PANN_LIB_PATH=$PWD/parlayann_b200/ab/$lib TUNES=${TUNES:-auto} python scripts/nq_sweep.py 2>&1 | tail -5
  PANN_LIB_PATH=$PWD/parlayann_b200/ab/$lib TUNES=${TUNES:-auto} N=1000000 BEAM=12 python scripts/nq_sweep.py 2>&1 | tail -5
done
