#!/bin/bash
# N-GPU check of the tensor-parallel Linear: parity worker, exchange phase timing, the bench's c5 record.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
echo "== run_tp_linear"
timeout 600 $TR tests/run_tp_linear.py > gpurun_out/tp_linear_$N.log 2>&1; echo "rc=$?"; grep -v Warning gpurun_out/tp_linear_$N.log | tail -12 | cut -c1-300
echo "== exchange phases"
timeout 300 $TR scripts/dev_xch_timing.py > gpurun_out/xch_timing_$N.log 2>&1; echo "rc=$?"; grep "rank 0\|Error\|error" gpurun_out/xch_timing_$N.log | tail -5 | cut -c1-900
if [ "${SKIP_BENCH:-0}" != "1" ]; then
echo "== bench --gpus $N"
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_tp$N.json 2> gpurun_out/bench_tp$N.err; echo "rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_tp$N.json").read().strip().splitlines()[-1])
    print(json.dumps(d["config"].get("c5_fprop"), indent=1)[:3000])
    print("value", d["value"], "e2e", d.get("e2e", {}).get("value"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_tp$N.err").read()[-2000:])
PY
fi
