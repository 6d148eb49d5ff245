# C1 full-size parity test with timings, then one ncu --set full capture of the float K1s launch (staged rows)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_c1_gpu.py -x -q -s -m gpu 2>&1 | tail -6
N=1000000 timeout 300 python scripts/float_search.py > gpurun_out/float_plain.log 2>&1 && tail -3 gpurun_out/float_plain.log &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pann_beam_search_small -s 12 -c 1 -f -o gpurun_out/k1s_float_full \
    python scripts/float_search.py > gpurun_out/ncu_float.log 2>&1
tail -3 gpurun_out/ncu_float.log; ls -la gpurun_out | tail -5
