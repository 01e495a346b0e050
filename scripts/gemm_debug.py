"""Where does the MMA thread of the CTA-pair GEMM wait? NVTE_B200_GEMM_DEBUG=1 python scripts/gemm_debug.py <kind>"""
import ctypes, os, sys
os.environ["NVTE_B200_GEMM_DEBUG"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transformerengine_b200 import ops, cabi
kind = sys.argv[1]
n, m, k = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (8192, 14336, 4096)
a = (torch.randn(m, k, device="cuda") * 0.5).bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
if kind == "mxfp8":
    qa, qb = ops.mxfp8_quantize(a, swizzled_scales=True), ops.mxfp8_quantize(b, swizzled_scales=True)
elif kind == "fp8":
    s = torch.ones(1, device="cuda"); qa, qb = ops.fp8_quantize(a, s), ops.fp8_quantize(b, s)
else:
    qa = ops.nvfp4_quantize(a, with_rht=False, swizzled_scales=True); qb = ops.nvfp4_quantize(b, with_rht=False, swizzled_scales=True)
out = torch.empty(n, m, dtype=torch.bfloat16, device="cuda")
for _ in range(3): ops.general_gemm(qa, qb, "TN", out=out)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); ops.general_gemm(qa, qb, "TN", out=out); e1.record()
torch.cuda.synchronize()
L = cabi.lib(); L.nvte_b200_gemm_debug_counters.restype = ctypes.POINTER(ctypes.c_ulonglong)
d = L.nvte_b200_gemm_debug_counters()
import numpy as np
v = np.array([[d[p * 4 + i] for i in range(4)] for p in range(74)], dtype=np.float64)
tiles = ((n + 255) // 256) * ((m + 255) // 256)
kb = (k // (2 if kind == "nvfp4" else 1) + 127) // 128
print(f"{kind} {n}x{m}x{k}: {e0.elapsed_time(e1)*1e3:.0f} us; per pair mean cycles: total {v[:,0].mean():.0f} wait_full {v[:,1].mean():.0f} "
      f"wait_acc_empty {v[:,2].mean():.0f} producer_wait_empty {v[:,3].mean():.0f}; ideal MMA cycles {tiles/74*kb*4*128:.0f} ({tiles/74:.1f} tiles x {kb} kblocks)")
# hand-off trace of pair 0 between its tiles #1 and #2 (leader CTA, raw clock64 of one SM)
t = [d[i] for i in range(512, 560)]
base = t[0]
rel = lambda v: int(v) - int(base) if v else None
print("hand-off trace (cycles after the MMA thread committed tile #1): next tile's MMA may start at", rel(t[1]))
for w in range(8):
    a = [rel(t[8 + 4 * w + j]) for j in range(4)]
    print(f"  epilogue warp {w + 2} (q={(w + 2) & 3}, half={w >> 2}): acc_full seen {a[0]}, shared columns drained+signalled {a[1]}, "
          f"all drained+released {a[2]}, stores done {a[3]}")
