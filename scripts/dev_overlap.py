"""torchrun worker (development aid): column-parallel FP8 fprop, all-gather/GEMM overlap A/B."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import transformerengine_b200 as te  # noqa: E402
from transformerengine_b200 import recipe  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
group = dist.new_group(list(range(world)))
T, fin, fout = 8192, 8192, 28672
lin = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16, parallel_mode="column",
                sequence_parallel=True, tp_group=group)
x = torch.randn(T // world, fin, device="cuda", dtype=torch.bfloat16)
rec = recipe.Float8CurrentScaling()


def step(first):
    with torch.no_grad(), te.autocast(enabled=True, recipe=rec, amax_reduction_group=group):
        return lin(x, is_first_microbatch=first)


def timed(iters=20):
    step(True)
    for _ in range(3):
        step(False)
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    import time
    t0 = time.perf_counter()
    e0.record()
    for _ in range(iters):
        step(False)
    e1.record()
    host = (time.perf_counter() - t0) / iters * 1e3
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters, host


for chunks, margin in ((1, 16), (2, 16), (4, 16), (4, 0), (4, 32), (8, 16)):
    lin.overlap_ag_chunks = chunks
    lin.overlap_ag_gemm_sms = max(2, 148 - margin)
    ms, host = timed()
    if rank == 0:
        print(f"tp={world} pieces={chunks} sm_margin={margin}: {ms * 1e3:7.1f} us/step (host issue {host * 1e3:6.1f} us)", flush=True)
# the same step replayed from a CUDA graph (no host launch cost)
for chunks, margin in ((1, 16), (2, 16), (4, 16), (4, 8)):
    lin.overlap_ag_chunks = chunks
    lin.overlap_ag_gemm_sms = max(2, 148 - margin)
    try:
        step(True)
        for _ in range(3):
            step(False)
        torch.cuda.synchronize()
        dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            y = step(False)
        torch.cuda.synchronize()
        for _ in range(3):
            g.replay()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        if rank == 0:
            print(f"tp={world} pieces={chunks} sm_margin={margin} CUDA graph: {e0.elapsed_time(e1) / 20 * 1e3:7.1f} us/step", flush=True)
        del g
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print("graph capture failed:", repr(e)[:300], flush=True)
        break
# the pieces alone
from transformerengine_b200 import ops  # noqa: E402
q = torch.empty(T // world, fin, dtype=torch.uint8, device="cuda")
for rows in (T // world, T // world // 4):
    src = q[:rows]
    buf = torch.empty(world * rows, fin, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        dist.all_gather_into_tensor(buf, src, group=group)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        dist.all_gather_into_tensor(buf, src, group=group)
    e1.record()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"all_gather of {rows} x {fin} bytes per rank: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us", flush=True)
dist.destroy_process_group()
