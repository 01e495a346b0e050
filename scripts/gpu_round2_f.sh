#!/bin/bash
# round-2 checkpoint on one GPU: the no-allocation entry point alone in a fresh process, the whole GPU suite
# (which ends with it), smoke(), the default bench, the reference arm.
mkdir -p gpurun_out
echo "== no-alloc test alone"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -5 | cut -c1-400
echo "== whole GPU suite"
timeout 3000 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_gpu.log | cut -c1-400
echo "== smoke"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
echo "== bench"
timeout 1500 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"; tail -c 1500 gpurun_out/bench_n1.err; head -c 6000 gpurun_out/bench_n1.json
echo "== reference arm"
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?"; cat gpurun_out/bench_ref.json | head -c 1500
