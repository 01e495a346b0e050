import os, sys, torch
sys.path.insert(0, "/root/repo")
from transformerengine_b200 import ops
from tests.util import make_input
x = make_input(2048, 4096, "outliers").cuda()
for rht in (False, True):
    both = ops.nvfp4_quantize(x, with_rht=rht, swizzled_scales=True)
    torch.cuda.synchronize()
    torch.save({k: getattr(both, k).cpu() for k in ("rowwise_data", "columnwise_data", "rowwise_scale_inv", "columnwise_scale_inv")},
               f"/tmp/fp4_{os.environ.get('TAG','x')}_{int(rht)}.pt")
    print("ok", rht)
    row = ops.nvfp4_quantize(x, with_rht=rht, swizzled_scales=True, columnwise=False)
    torch.cuda.synchronize()
    print("row-only ok", torch.equal(row.rowwise_data, both.rowwise_data))
