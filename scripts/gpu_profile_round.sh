# Round profile pass: full test-suite, bench, then ncu launch list of the same bench command and
# one --set full capture per hot kernel (each only after the plain command exited 0).
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err || { tail -20 gpurun_out/bench.err; exit 1; }
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err
python bench.py --steps 2 --warmup 3 --no-extra > gpurun_out/b_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-extra > gpurun_out/ncu_bench.log 2>&1
python scripts/ncu_target.py mxfp8_c4 2 > gpurun_out/plain_mx.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:quantize_tma_kernel -s 3 -c 3 -o gpurun_out/prof_mxfp8_c4 -f python scripts/ncu_target.py mxfp8_c4 2 > gpurun_out/ncu_mx.log 2>&1
python scripts/ncu_target.py gemm_mxfp8 2 > gpurun_out/plain_gemm.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm2_kernel -s 1 -c 1 -o gpurun_out/prof_gemm_mxfp8 -f python scripts/ncu_target.py gemm_mxfp8 2 > gpurun_out/ncu_gemm.log 2>&1
python scripts/dev_grouped.py > gpurun_out/grouped.log 2>&1
exit 0
