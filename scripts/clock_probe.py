"""Run one GEMM kind in a ~2 s loop while sampling nvidia-smi clocks/power."""
import os, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transformerengine_b200 import ops
kind = sys.argv[1]
n, m, k = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (8192, 14336, 4096)
a = (torch.randn(m, k, device="cuda") * 0.5).bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
if kind == "mxfp8":
    qa, qb = ops.mxfp8_quantize(a, swizzled_scales=True), ops.mxfp8_quantize(b, swizzled_scales=True)
elif kind == "fp8":
    s = torch.ones(1, device="cuda"); qa, qb = ops.fp8_quantize(a, s), ops.fp8_quantize(b, s)
elif kind == "nvfp4":
    qa = ops.nvfp4_quantize(a, with_rht=False, swizzled_scales=True); qb = ops.nvfp4_quantize(b, with_rht=False, swizzled_scales=True)
out = torch.empty(n, m, dtype=torch.bfloat16, device="cuda")
fn = (lambda: torch.matmul(b, a.t())) if kind == "bf16" else (lambda: ops.general_gemm(qa, qb, "TN", out=out))
for _ in range(5): fn()
torch.cuda.synchronize()
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
iters = int(os.environ.get("ITERS", "3000"))
e0.record()
for _ in range(iters): fn()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
time.sleep(0.05); p.terminate()
rows = [l.strip() for l in p.stdout.read().splitlines() if l.strip()]
print(kind, f"{ms*1e3:.1f} us {2*n*m*k/ms/1e9:.0f} TFLOP/s sustained over {iters} iters; clocks/power samples:", rows[1:-1:2][:10])
