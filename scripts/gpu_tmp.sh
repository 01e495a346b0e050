timeout 600 python -m pytest tests/test_quantize_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --no-extra --steps 50 > gpurun_out/bench_tma2.log 2>&1; tail -c 1500 gpurun_out/bench_tma2.log
