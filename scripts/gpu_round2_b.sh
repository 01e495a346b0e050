#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_quantize_gpu.py tests/test_linear_gpu.py tests/test_grouped_linear_gpu.py -x -q -m gpu > gpurun_out/r02_pytest_mx.log 2>&1
tail -5 gpurun_out/r02_pytest_mx.log
timeout 600 python bench.py --no-extra --steps 100 --warmup 10 > gpurun_out/r02_bench_short.json 2> gpurun_out/r02_bench_short.err
BENCH_MXFP8_SEPARATE_LAUNCHES=1 timeout 600 python bench.py --no-extra --steps 100 --warmup 10 > gpurun_out/r02_bench_short_sep.json 2>> gpurun_out/r02_bench_short.err
python - <<'PY'
import json
for f in ("gpurun_out/r02_bench_short.json", "gpurun_out/r02_bench_short_sep.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["value"], d["clocks"])
    except Exception as e:
        print(f, "ERR", e)
PY
tail -5 gpurun_out/r02_bench_short.err
bash scripts/gpu_kernel_bench.sh --only mxfp8_quantize
