import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transformerengine_b200 import ops
torch.manual_seed(0)
def q4(t): return ops.nvfp4_quantize(t, with_rht=False, swizzled_scales=True)
def qm(t): return ops.mxfp8_quantize(t, swizzled_scales=True)
for kind, q in (("nvfp4", q4), ("mxfp8", qm)):
  for layout in ("TN", "NN", "NT"):
    for K in (512, 384, 640):
        splits, fin = [256, 128, 384, 256], 512
        A, B, refs = [], [], []
        for m in splits:
            if layout == "TN":   a = torch.randn(fin, K); b = torch.randn(m, K)       # out [m, fin]
            elif layout == "NN": a = torch.randn(K, fin); b = torch.randn(m, K)       # out [m, fin]
            else:                a = torch.randn(K, fin); b = torch.randn(K, m)       # out [m, fin]
            a, b = a.bfloat16().cuda(), b.bfloat16().cuda()
            qa, qb = q(a), q(b)
            A.append(qa); B.append(qb)
            da = ops.dequantize(qa, columnwise=layout[0] != "T").float(); db = ops.dequantize(qb, columnwise=layout[1] == "T").float()
            if kind == "nvfp4":
                # columnwise NVFP4 dequantizes to the transposed tensor
                a2 = da if layout[0] == "T" else da            # [fin,K] either way for nvfp4 colwise (stored transposed)
                b2 = db
                refs.append(b2 @ a2.t())
            else:
                a2 = da if layout[0] == "T" else da.t()
                b2 = db.t() if layout[1] == "T" else db
                refs.append(b2 @ a2.t())
        single = [ops.general_gemm(a, b, layout, torch.float32) for a, b in zip(A, B)]
        grouped = ops.multi_tensor_gemm(A, B, layout, torch.float32)
        torch.cuda.synchronize()
        for i, (s, g, r) in enumerate(zip(single, grouped, refs)):
            es = ((s - r).norm() / r.norm()).item() if r.shape == s.shape else float("nan")
            eg = ((g - r).norm() / r.norm()).item() if r.shape == g.shape else float("nan")
            print(f"{kind} {layout} K={K} member {i} (m={splits[i]}): single-vs-ref {es:.2e} grouped-vs-ref {eg:.2e} equal={torch.equal(s, g)}")
