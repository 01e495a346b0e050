import sys, time
sys.path.insert(0, '.')
import numpy as np, torch
from transformerengine_b200 import cabi as C
from oracle import oracle as O
from tests.util import bits16, u8, make_input

def run(rows, cols, kind="randn", dtype=torch.bfloat16, row=True, col=True, check=True, iters=0, swz=False):
    x = make_input(rows, cols, kind, dtype=dtype).cuda()
    kb, rb = (cols+31)//32, (rows+31)//32
    out = C.TensorHandle(C.MXFP8_1D_SCALING)
    bufs = {}
    if row:
        bufs['qr'] = torch.zeros(rows, cols, dtype=torch.uint8, device='cuda')
        bufs['sr'] = torch.zeros((rows+127)//128*128, (kb+3)//4*4, dtype=torch.uint8, device='cuda')
        out.set(C.kRowwiseData, bufs['qr'], C.kFloat8E4M3); out.set(C.kRowwiseScaleInv, bufs['sr'], C.kFloat8E8M0)
    if col:
        bufs['qc'] = torch.zeros(rows, cols, dtype=torch.uint8, device='cuda')
        bufs['sc'] = torch.zeros((rb+3)//4*4, (cols+127)//128*128, dtype=torch.uint8, device='cuda')
        out.set(C.kColumnwiseData, bufs['qc'], C.kFloat8E4M3); out.set(C.kColumnwiseScaleInv, bufs['sc'], C.kFloat8E8M0)
    out.set_flag(C.kWithGEMMSwizzledScales, swz)
    inp = C.plain(x)
    C.check(C.lib().nvte_quantize_v2_status(inp.ptr, out.ptr, None, C.stream_ptr()))
    torch.cuda.synchronize()
    if check:
        code = O.F32 if dtype == torch.float32 else (O.BF16 if dtype == torch.bfloat16 else O.F16)
        xin = x.cpu().numpy() if dtype == torch.float32 else bits16(x)
        ref = O.mxfp8_quantize(xin, rowwise=row, colwise=col, code=code)
        ok = True
        if row:
            ok &= np.array_equal(u8(bufs['qr']), ref['rowwise'])
            ok &= np.array_equal(u8(bufs['sr'])[:rows,:kb], ref['rowwise_scale'])
        if col:
            ok &= np.array_equal(u8(bufs['qc']), ref['colwise'])
            ok &= np.array_equal(u8(bufs['sc'])[:rb,:cols], ref['colwise_scale'])
        print(f"{rows}x{cols} {kind} {dtype} row={row} col={col}: {'OK' if ok else 'MISMATCH'}")
    if iters:
        flush = torch.empty(256<<20, dtype=torch.uint8, device='cuda')
        ts = []
        for i in range(iters+3):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); C.check(C.lib().nvte_quantize_v2_status(inp.ptr, out.ptr, None, C.stream_ptr())); e1.record()
            torch.cuda.synchronize()
            if i >= 3: ts.append(e0.elapsed_time(e1))
        esz = x.element_size()
        nbytes = rows*cols*esz + (row+col)*(rows*cols + rows*cols//32)
        t = np.median(ts)
        print(f"  {rows}x{cols} {dtype} row={row} col={col}: median {t*1e3:.1f} us, {nbytes/t/1e6:.0f} GB/s (best {nbytes/min(ts)/1e6:.0f})")

for shp in [(128,128),(256,272),(304,304),(993,512),(1,32),(16,48),(64,96)]:
    for kind in ["randn","uniform","outliers"]:
        run(*shp, kind)
run(993,512,"uniform",torch.float32); run(256,272,"randn",torch.float16); run(128,128,"randn",row=False); run(128,256,"randn",col=False)
run(4096,4096,"randn")
run(4096,4096,"randn",iters=20,check=False)
run(8192,4096,"randn",iters=20,check=False)
run(8192,14336,"randn",iters=20,check=False)
run(8192,14336,"randn",iters=20,check=False,col=False)
run(8192,14336,"randn",iters=20,check=False,row=False)
run(8192,4096,"randn",torch.float32,iters=20,check=False)
