mkdir -p gpurun_out
python -m pytest tests/test_build_gpu.py -x -q -m gpu 2>&1 | tail -2
for n in 1000000 10000000; do echo "=== default n=$n"; SORT=1 python scripts/build_only.py $n 2>&1 | tail -1; SORT=1 python scripts/build_only.py $n 2>&1 | tail -1; done
for lib in "$@"; do for n in 1000000 10000000; do echo "=== $lib n=$n"; SORT=1 PANN_LIB_PATH=$PWD/parlayann_b200/ab/$lib python scripts/build_only.py $n 2>&1 | tail -1; done; done
