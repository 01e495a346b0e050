#!/bin/bash
# 8-GPU (or N-GPU) exchange timing: dynamic tile queue on / off, then the bench's c5 record
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for dyn in 1 0; do
  echo "== exchange phases, dynamic tile queue=$dyn"
  NVTE_B200_XCH_DYNAMIC=$dyn NVTE_B200_GEMM_DEBUG=${DEBUG:-0} timeout 300 $TR scripts/dev_xch_timing.py > gpurun_out/xch_timing_${N}_dyn$dyn.log 2>&1; echo "rc=$?"
  grep "rank 0" gpurun_out/xch_timing_${N}_dyn$dyn.log | cut -c1-700
done
echo "== bench --gpus $N"
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_tp$N.json 2> gpurun_out/bench_tp$N.err; echo "rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_tp$N.json").read().strip().splitlines()[-1])
    c = d["config"].get("c5_fprop")
    print({k: v for k, v in c.items() if k not in ("checks", "default_path")}, c.get("checks_all_ranks_ok"))
    print("value", d["value"], "e2e", d.get("e2e", {}).get("value"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_tp$N.err").read()[-2000:])
PY
