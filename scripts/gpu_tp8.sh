#!/bin/bash
# 8-GPU (or N-GPU) exchange timing: multicast and unicast pushes, then the bench's c5 record
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for mc in 1 0; do
  echo "== exchange phases, multicast=$mc"
  NVTE_B200_XCH_MULTICAST=$mc timeout 300 $TR scripts/dev_xch_timing.py > gpurun_out/xch_timing_${N}_mc$mc.log 2>&1; echo "rc=$?"
  grep "rank 0" gpurun_out/xch_timing_${N}_mc$mc.log | cut -c1-600
done
echo "== bench --gpus $N"
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_tp$N.json 2> gpurun_out/bench_tp$N.err; echo "rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_tp$N.json").read().strip().splitlines()[-1])
    c = d["config"].get("c5_fprop")
    print({k: v for k, v in c.items() if k != "checks"}, c.get("checks_all_ranks_ok"))
    print("value", d["value"], "e2e", d.get("e2e", {}).get("value"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_tp$N.err").read()[-2000:])
PY
