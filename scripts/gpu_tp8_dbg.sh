#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() {  # tag, env...
  tag=$1; shift
  echo "== $tag"
  env "$@" timeout 200 $TR scripts/dev_xch_timing.py > gpurun_out/xch_${N}_$tag.log 2>&1; echo "rc=$?"
  grep "rank 0" gpurun_out/xch_${N}_$tag.log | cut -c1-900
}
run dbg256 NVTE_B200_GEMM_DEBUG=1
run dbg512 NVTE_B200_GEMM_DEBUG=1 NVTE_B200_XCH_ROWS_PER_FLAG=512
run rpf512 NVTE_B200_XCH_ROWS_PER_FLAG=512
run rpf1024 NVTE_B200_XCH_ROWS_PER_FLAG=1024
