import sys
sys.path.insert(0, '.')
import numpy as np, torch
from transformerengine_b200 import ops
from oracle import oracle as O
from tests.util import bits16, u8, make_input

def ref_mxfp8(q, colwise):
    d = q.columnwise_data if colwise else q.rowwise_data
    s = q.columnwise_scale_inv if colwise else q.rowwise_scale_inv
    return O.mxfp8_dequantize(u8(d), u8(s), colwise=colwise)

def check(n, m, k, layout="TN", iters=0):
    torch.manual_seed(0)
    # TE shapes: TN: A[m,k], B[n,k]; NN: A[k,m], B[n,k]; NT: A[k,m], B[k,n]
    ashape = (m, k) if layout[0] == "T" else (k, m)
    bshape = (k, n) if layout[1] == "T" else (n, k)
    a = (torch.randn(*ashape) * 0.5).bfloat16().cuda()
    b = torch.randn(*bshape).bfloat16().cuda()
    qa_c = ops.mxfp8_quantize(a); qb_c = ops.mxfp8_quantize(b)
    qa = ops.swizzle_scales(qa_c); qb = ops.swizzle_scales(qb_c)
    out = ops.general_gemm(qa, qb, layout, torch.float32)
    torch.cuda.synchronize()
    if n * m * k <= 2**31:
        A = ref_mxfp8(qa_c, colwise=layout[0] != "T"); B = ref_mxfp8(qb_c, colwise=layout[1] == "T")
        A2 = A if layout[0] == "T" else A.T.copy()     # -> [m, k]
        B2 = B.T.copy() if layout[1] == "T" else B     # -> [n, k]
        ref = O.gemm_tn_ref(A2, B2)
        got = out.cpu().numpy()
        err = np.abs(got - ref).max(); rel = err / np.abs(ref).max()
        print(f"{layout} n={n} m={m} k={k}: max abs err {err:.3e} rel {rel:.3e} {'OK' if rel < 1e-5 else 'BAD'}")
    if iters:
        outb = torch.empty(n, m, dtype=torch.bfloat16, device='cuda')
        for _ in range(3): ops.general_gemm(qa, qb, layout, out=outb)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): ops.general_gemm(qa, qb, layout, out=outb)
        e1.record(); torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / iters
        print(f"  {layout} n={n} m={m} k={k}: {t*1e3:.1f} us, {2*n*m*k/t/1e9:.0f} TFLOPS")

which = sys.argv[1] if len(sys.argv) > 1 else "small"
if which == "small":
    check(128, 128, 128)
    check(128, 128, 512)
    check(256, 384, 1024)
    check(200, 136, 96)
    check(128, 256, 256)
    check(512, 1024, 512)
elif which == "layouts":
    for lay in ("NN", "NT"):
        check(128, 128, 128, lay); check(256, 384, 1024, lay); check(512, 1024, 512, lay)
elif which == "perf":
    check(8192, 14336, 4096, "TN", iters=20)
    check(8192, 4096, 14336, "NN", iters=20)
    check(14336, 4096, 8192, "NT", iters=20)
