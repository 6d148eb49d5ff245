# A/B of library builds from several commits (parlayann_b200/ab/lib_<commit>.so), 10M points, beam 24, batch sweep
mkdir -p gpurun_out
for lib in "$@"; do
  echo "=== $lib"
  PANN_LIB_PATH=$PWD/parlayann_b200/ab/$lib TUNES=auto python scripts/nq_sweep.py 2>&1 | tail -5
done 2>&1 | tee gpurun_out/bisect.log
