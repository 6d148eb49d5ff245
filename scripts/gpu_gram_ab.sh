mkdir -p gpurun_out
python -m pytest tests/test_build_gpu.py -x -q -m gpu 2>&1 | tail -4
for g in 0 1 2; do echo "PANN_PRUNE_GRAM=$g"; PANN_PRUNE_GRAM=$g python scripts/build_only.py 10000000 2>&1 | tail -1; done
for w in 3 2; do echo "GRAM=1 warps/CTA=$w"; PANN_PRUNE_GRAM_WARPS=$w python scripts/build_only.py 10000000 2>&1 | tail -1; done
