"""torchrun worker (dev aid): phase timing of the flag-driven symmetric exchange on the config-5 fprop step.
   torchrun --nproc-per-node N scripts/dev_xch_timing.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from transformerengine_b200 import ops  # noqa: E402
from transformerengine_b200.distributed import SymmetricExchange  # noqa: E402
from transformerengine_b200.quantizers import Float8CurrentScalingQuantizer  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    group = dist.new_group(list(range(world)))
    T, fin, fout = 8192, 8192, 28672 // world
    torch.manual_seed(rank)
    x = torch.randn(T // world, fin, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(fout, fin, device="cuda") * 0.02).bfloat16()
    ws = torch.ones(1, device="cuda")
    wq = ops.fp8_quantize(w, ws)
    q = Float8CurrentScalingQuantizer(torch.float8_e4m3fn, device="cuda", with_amax_reduction=True,
                                      amax_reduction_group=group)
    xch = SymmetricExchange(group, T // world, fin, x.device)
    out = torch.empty(T, fout, device="cuda", dtype=torch.bfloat16)
    iters = 20

    def seq(arrival, marks=None):
        def mark():
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append(e)
        mark()
        ops.compute_amax(x, q.amax, peer_config=xch.cfg)
        mark()
        ops.compute_scale_from_amax(q.amax, q.scale, q.dtype, q.force_pow_2_scales, q.amax_epsilon,
                                    symmetric_exchange=xch.struct, scale_inv=xch._scale_inv)
        mark()
        local = ops.Float8TensorStorage(data=xch.gathered[rank * xch.rows_local:(rank + 1) * xch.rows_local],
                                        scale_inv=xch._scale_inv, fp8_dtype=q.dtype, shape=(xch.rows_local, fin))
        ops.fp8_quantize(x, q.scale, q.dtype, amax=None, out=local, peer_config=xch.cfg if xch.push == "cast" else None)
        mark()
        full = ops.Float8TensorStorage(data=xch.gathered, scale_inv=xch._scale_inv, fp8_dtype=q.dtype,
                                       shape=tuple(xch.gathered.shape))
        ops.general_gemm(wq, full, "TN", out=out, arrival=arrival)
        mark()

    def total(arrival):
        for _ in range(3):
            seq(arrival)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            seq(arrival)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def graphed(arrival):
        for _ in range(3):
            seq(arrival)
        torch.cuda.synchronize()
        dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            seq(arrival)
        for _ in range(3):
            g.replay()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    # plain GEMM over a resident gather buffer (no flags): the floor
    seq(xch.arrival)
    torch.cuda.synchronize()
    full = ops.Float8TensorStorage(data=xch.gathered, scale_inv=xch._scale_inv, fp8_dtype=q.dtype,
                                   shape=tuple(xch.gathered.shape))
    for _ in range(3):
        ops.general_gemm(wq, full, "TN", out=out)
    ref = out.clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.general_gemm(wq, full, "TN", out=out)
    e1.record()
    torch.cuda.synchronize()
    t_gemm = e0.elapsed_time(e1) / iters
    res = {"gemm_plain_ms": round(t_gemm, 4)}
    res["seq_ms"] = round(total(xch.arrival), 4)
    ok1 = torch.equal(out, ref)
    res["seq_overlap_ms"] = None
    ok2 = torch.equal(out, ref)
    res["seq_graph_ms"] = round(graphed(xch.arrival), 4)
    res["seq_overlap_graph_ms"] = None
    ok3 = torch.equal(out, ref)
    # phases (non-overlapped order, events between the launches)
    acc = [0.0] * 4
    for _ in range(iters):
        marks = []
        seq(xch.arrival, marks)
        torch.cuda.synchronize()
        for i in range(4):
            acc[i] += marks[i].elapsed_time(marks[i + 1]) / iters
    res["phases_ms"] = {k: round(v, 4) for k, v in zip(("amax+push", "scale_wait", "cast_exchange", "gemm_arrival"), acc)}
    if os.environ.get("NVTE_B200_GEMM_DEBUG") == "1":
        import ctypes
        from transformerengine_b200 import cabi
        L = cabi.lib()
        L.nvte_b200_gemm_debug_counters.restype = ctypes.POINTER(ctypes.c_ulonglong)
        d = L.nvte_b200_gemm_debug_counters()
        seq(xch.arrival)
        torch.cuda.synchronize()
        flag = [d[600 + i] / 1e3 for i in range(74)]
        push = [d[700 + i] / 1e3 for i in range(74)]
        fence = [d[800 + i] / 1e3 for i in range(74)]
        res["debug_us"] = {"flag_wait_max": round(max(flag), 1), "flag_wait_mean": round(sum(flag) / 74, 1),
                           "pusher_total_max": round(max(push), 1), "pusher_total_mean": round(sum(push) / 74, 1),
                           "pusher_fence_max": round(max(fence), 1), "pusher_fence_mean": round(sum(fence) / 74, 1)}
    res["bits_equal"] = [ok1, ok2, ok3]
    res["multicast"] = xch.multicast
    res["error_word"] = int(xch.error_word.item())
    print(f"rank {rank}: {res}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
