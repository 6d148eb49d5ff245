mkdir -p gpurun_out
python -m pytest tests/test_build_gpu.py tests/test_dropin_gpu.py -x -q -m gpu 2>&1 | tail -3
echo "=== default"; SORT=1 PANN_BUILD_VERBOSE=1 python scripts/build_only.py ${N:-10000000} 2>&1 | grep -v "^  batch\|prune mode" | tail -3
for lib in "$@"; do echo "=== $lib"; SORT=1 PANN_LIB_PATH=$PWD/parlayann_b200/ab/$lib python scripts/build_only.py ${N:-10000000} 2>&1 | tail -1; done
