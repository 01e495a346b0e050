import torch, sys
sys.path.insert(0, "/root/repo")
import transformerengine_b200 as te
from transformerengine_b200 import recipe
rec = recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.E4M3)
torch.manual_seed(3)
eager = te.Linear(256, 512, bias=True, params_dtype=torch.bfloat16)
g = te.Linear(256, 512, bias=True, params_dtype=torch.bfloat16)
g.load_state_dict({k: v for k, v in eager.state_dict().items() if "extra_state" not in k}, strict=False)
sample = torch.randn(384, 256, device="cuda", dtype=torch.bfloat16, requires_grad=True)
fn = te.make_graphed_callables(g, (sample,), num_warmup_iters=3, enabled=True, recipe=rec)
g.zero_grad(set_to_none=True)
print("state after capture", g._delayed[True].scale, g._delayed[False].scale, g._delayed[True].amax_history[:2])
for step in range(3):
    xin = torch.randn(384, 256, device="cuda", dtype=torch.bfloat16) * (1.0 + step)
    xe, xg = xin.clone().requires_grad_(True), xin.clone().requires_grad_(True)
    with te.autocast(enabled=True, recipe=rec):
        ye = eager(xe)
    ye.backward(torch.ones_like(ye))
    yg = fn(xg)
    yg.backward(torch.ones_like(yg))
    torch.cuda.synchronize()
    def d(a, b): return ((a.float()-b.float()).abs().max().item(), b.float().abs().max().item())
    print(step, "y", d(ye, yg), "dx", d(xe.grad, xg.grad), "dw", d(eager.weight.grad, g.weight.grad), "db", d(eager.bias.grad, g.bias.grad))
    print("  scales fwd", eager._delayed[True].scale.tolist(), g._delayed[True].scale.tolist(), "bwd", eager._delayed[False].scale.tolist(), g._delayed[False].scale.tolist())
