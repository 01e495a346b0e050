#!/bin/bash
# round-2 ncu evidence: launch list of the default bench, --set full of the one-launch MXFP8 step,
# the norm -> MXFP8 kernels and the MXFP8 GEMM. Each target first runs without ncu.
mkdir -p gpurun_out
N="ncu --clock-control none"
python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_bench_plain.json 2> gpurun_out/ncu_bench_plain.err || { echo bench failed; exit 1; }
timeout 1500 $N --metrics gpu__time_duration.sum -c 600 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r02_launches_bench.csv
cap() {  # name, kernel regex, skip, target op
  python scripts/ncu_target.py $4 6 > gpurun_out/plain_$1.log 2>&1 || { echo "$1 plain run failed"; tail -3 gpurun_out/plain_$1.log; return; }
  timeout 600 $N --set full --import-source on -k regex:$2 -s $3 -c 1 -o gpurun_out/r02_ncu_$1 -f python scripts/ncu_target.py $4 6 > gpurun_out/ncu_$1.log 2>&1
  echo "$1 rc=$?"; ls -la gpurun_out/r02_ncu_$1.ncu-rep 2>/dev/null
}
cap mxfp8_c4_one_launch quantize_tma_kernel 8 mxfp8_c4_multi
cap norm_rms_mx norm_fwd_kernel 2 norm_rms_mx
cap norm_ln_mx norm_fwd_kernel 2 norm_ln_mx
cap norm_ln_bf16 norm_fwd_kernel 2 norm_ln_bf16
cap gemm_mxfp8 gemm2_kernel 2 gemm_mxfp8
cap gemm_fp8 gemm2_kernel 2 gemm_fp8
