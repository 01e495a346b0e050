set -x
python scripts/ncu_target.py nvfp4_rht_swz 2 > gpurun_out/plain_nv2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:nvfp4_tc_kernel -s 3 -c 1 -o gpurun_out/prof_nvfp4_tc3 -f python scripts/ncu_target.py nvfp4_rht_swz 2 > gpurun_out/ncu_nv3.log 2>&1
python scripts/ncu_target.py mxfp8_c4 2 > gpurun_out/plain_mx.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:quantize_kernel -s 3 -c 3 -o gpurun_out/prof_mxfp8_c4 -f python scripts/ncu_target.py mxfp8_c4 2 > gpurun_out/ncu_mx.log 2>&1
python bench.py --steps 2 --warmup 3 --no-extra > gpurun_out/b_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-extra > gpurun_out/ncu_bench.log 2>&1
tail -2 gpurun_out/ncu_nv3.log gpurun_out/ncu_mx.log gpurun_out/ncu_bench.log
