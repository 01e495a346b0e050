#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for rpf in 256 1024 ; do
  echo "== rows_per_flag $rpf"
  NVTE_B200_XCH_ROWS_PER_FLAG=$rpf timeout 300 $TR scripts/dev_xch_timing.py > gpurun_out/xch_timing_${N}_$rpf.log 2>&1; echo "rc=$?"
  grep "rank 0" gpurun_out/xch_timing_${N}_$rpf.log | cut -c1-700
done
