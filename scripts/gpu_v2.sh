mkdir -p gpurun_out
python -m pytest tests/test_search_gpu.py tests/test_quantized_gpu.py tests/test_c1_gpu.py -x -q -m gpu 2>&1 | tail -5
bash scripts/gpu_bisect.sh lib_r2a.so lib_v2.so lib_r2a.so lib_v2.so
