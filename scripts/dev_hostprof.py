"""cProfile of the eager config-2 step (development aid): where the host time goes."""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import transformerengine_b200 as te  # noqa: E402
from transformerengine_b200 import recipe  # noqa: E402

model = te.Linear(768, 3072, bias=True)
inp = torch.randn(2048, 768, device="cuda")
rec = recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.E4M3)


def step():
    with te.autocast(enabled=True, recipe=rec):
        o = model(inp)
    o.sum().backward()


for _ in range(10):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
