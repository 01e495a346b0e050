"""Tensor-parallel plumbing of the Linear path over `torch.distributed` (NCCL on the GPU box,
gloo in the CPU tests): gather of quantized shards along the token dimension, reduce-scatter /
all-reduce of high-precision partial results.

Mirrors reference pytorch/distributed.py:982-1073 (FP8), :1324-1640 (MXFP8 / NVFP4 gathers),
:1640-1990 (`gather_along_first_dim`, `reduce_scatter_along_first_dim`, `allreduce`).

Why shards compose: MXFP8 / NVFP4 block scales are local to a row block, so a row shard's bytes
equal the corresponding rows of the full tensor's quantization as long as the shard height is a
multiple of the block height (32 / 16) and — for NVFP4 and per-tensor FP8 — every rank used the
same global amax / scale (all-reduced before the cast). Compact scales are gathered and the
GEMM swizzle is applied afterwards, because swizzled column-wise scales interleave 128-token
tiles across ranks (reference module/linear.py:1986-1987 disables pre-swizzling under SP).
"""
import copy
from typing import Optional

import os

import torch
import torch.distributed as dist

from . import ops


def get_distributed_world_size(group=None) -> int:
    if not dist.is_available() or not dist.is_initialized():
        return 1
    return dist.get_world_size(group=group)


def _all_gather_dim0(t: torch.Tensor, group, world: int) -> torch.Tensor:
    out = torch.empty((t.shape[0] * world,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, t.contiguous(), group=group)
    return out


def _all_gather_dim1(t: torch.Tensor, group, world: int) -> torch.Tensor:
    """Gather [a, b] shards into [a, world*b] (rank r owns columns r*b..(r+1)*b)."""
    g = _all_gather_dim0(t.unsqueeze(0), group, world)          # [world, a, b]
    return g.permute(1, 0, 2).reshape(t.shape[0], world * t.shape[1]).contiguous()


def _unpad_rows(sc: torch.Tensor, rows: int) -> torch.Tensor:
    return sc[:rows]


def _pad_to(sc: torch.Tensor, shape) -> torch.Tensor:
    if tuple(sc.shape) == tuple(shape):
        return sc.contiguous()
    out = torch.zeros(shape, dtype=sc.dtype, device=sc.device)
    out[: sc.shape[0], : sc.shape[1]] = sc
    return out


def gather_along_first_dim(inp, group, *, quantizer=None):
    """All-gather along dim 0. `inp` is a torch.Tensor or a quantized storage with COMPACT
    scales; returns the same kind of object holding the gathered tensor.

    Per-tensor FP8: gather bytes, keep scale_inv (identical on all ranks by construction).
    MXFP8: rowwise and columnwise data are both [rows, cols] -> gather along dim 0; rowwise
    scales [rows(pad 128), cols/32] gather their unpadded rows; columnwise scales
    [rows/32(pad 4), cols(pad 128)] gather their unpadded row-blocks.
    NVFP4: rowwise data/scales as MXFP8; columnwise data is transposed [cols, rows/2] and its
    scales are [cols(pad 128), rows/16(pad 4)] -> gather along dim 1.
    """
    world = get_distributed_world_size(group)
    if world == 1:
        return inp
    if isinstance(inp, torch.Tensor):
        return _all_gather_dim0(inp, group, world)

    res = copy.copy(inp)
    res._handle = None
    if isinstance(inp, ops.Float8TensorStorage):
        res.data = _all_gather_dim0(inp.data, group, world)
        res.shape = (inp.shape[0] * world,) + tuple(inp.shape[1:])
        res.transpose = None
        return res

    assert not inp.with_gemm_swizzled_scales, \
        "gather needs compact scales (swizzle after the gather)"
    rows = 1
    for d in inp.shape[:-1]:
        rows *= d
    cols = inp.shape[-1]
    total = rows * world
    if isinstance(inp, ops.MXFP8TensorStorage):
        assert rows % 32 == 0, "MXFP8 shards must be a multiple of 32 rows"
        rs, cs = ops.mxfp8_scale_shapes(total, cols)
        if inp.rowwise_data is not None:
            res.rowwise_data = _all_gather_dim0(inp.rowwise_data.reshape(rows, cols), group, world)
            res.rowwise_scale_inv = _pad_to(
                _all_gather_dim0(_unpad_rows(inp.rowwise_scale_inv, rows), group, world), rs)
        if inp.columnwise_data is not None:
            res.columnwise_data = _all_gather_dim0(inp.columnwise_data.reshape(rows, cols), group, world)
            res.columnwise_scale_inv = _pad_to(
                _all_gather_dim0(_unpad_rows(inp.columnwise_scale_inv, rows // 32), group, world), cs)
        res.shape = (total, cols)
        return res
    if isinstance(inp, ops.NVFP4TensorStorage):
        assert rows % 16 == 0, "NVFP4 shards must be a multiple of 16 rows"
        rs, cs = ops.nvfp4_scale_shapes(total, cols)
        if inp.rowwise_data is not None:
            res.rowwise_data = _all_gather_dim0(inp.rowwise_data, group, world)
            res.rowwise_scale_inv = _pad_to(
                _all_gather_dim0(_unpad_rows(inp.rowwise_scale_inv, rows), group, world), rs)
        if inp.columnwise_data is not None:
            res.columnwise_data = _all_gather_dim1(inp.columnwise_data, group, world)
            res.columnwise_scale_inv = _pad_to(
                _all_gather_dim1(inp.columnwise_scale_inv[:, : rows // 16], group, world), cs)
        res.shape = (total, cols)
        return res
    raise TypeError(f"cannot gather {type(inp)}")


def reduce_scatter_along_first_dim(t: torch.Tensor, group) -> torch.Tensor:
    world = get_distributed_world_size(group)
    if world == 1:
        return t
    assert t.shape[0] % world == 0
    n = t.shape[0] // world
    if dist.get_backend(group) == "gloo":  # gloo has no reduce-scatter (CPU tests only)
        full = t.contiguous().clone()
        dist.all_reduce(full, group=group)
        r = dist.get_rank(group)
        return full[r * n:(r + 1) * n].contiguous()
    out = torch.empty((n,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.reduce_scatter_tensor(out, t.contiguous(), group=group)
    return out


def allreduce(t: torch.Tensor, group) -> torch.Tensor:
    if get_distributed_world_size(group) > 1:
        dist.all_reduce(t, group=group)
    return t


def allreduce_amax(amax: Optional[torch.Tensor], group) -> None:
    """all-reduce(MAX) of amax values, on the current stream, no host wait
    (reference csrc/quantizer.cpp:920-926, 2342-2369)."""
    if amax is not None and get_distributed_world_size(group) > 1:
        dist.all_reduce(amax, op=dist.ReduceOp.MAX, group=group)


class ChunkedFP8Gather:
    """All-gather of a per-tensor FP8 shard along dim 0 in `chunks` pieces on a side stream, so that
    a GEMM over piece c can run while piece c + 1 is still on the wire (the part of the
    reference's Userbuffers all-gather overlap that `te.Linear` fprop uses,
    common/comm_gemm_overlap/comm_gemm_overlap.cpp:600-760 `split_overlap_ag`).

    Piece c holds rows [c * rows_c, (c + 1) * rows_c) of EVERY rank's shard, rank-major:
    buffer c is [world * rows_c, cols]; its GEMM output row r belongs to row
    (r // rows_c) * rows_local + c * rows_c + r % rows_c of the full output
    (`ops.general_gemm(out_row_blocks=(rows_c, rows_local, c * rows_c))`).
    """

    def __init__(self, shard: "ops.Float8TensorStorage", group, chunks: int, cache: dict = None):
        self.world = get_distributed_world_size(group)
        self.group, self.shard = group, shard
        data = shard.data.reshape(-1, shard.data.shape[-1])
        self.rows_local, self.cols = data.shape
        while chunks > 1 and (self.rows_local % chunks != 0 or (self.rows_local // chunks) % 128 != 0):
            chunks -= 1
        self.chunks = chunks
        self.rows_c = self.rows_local // chunks
        self._data = data
        self._stream = _comm_stream(data.device)
        # receive buffers and events live with the module: no allocator traffic on the side stream
        key = (self.world, self.rows_c, self.cols, chunks, data.dtype, data.device)
        cache = cache if cache is not None else {}
        if cache.get("key") != key:
            cache["key"] = key
            cache["bufs"] = [torch.empty(self.world * self.rows_c, self.cols, dtype=data.dtype, device=data.device)
                             for _ in range(chunks)]
            cache["events"] = [torch.cuda.Event() for _ in range(chunks)]
        self._bufs, self._events = cache["bufs"], cache["events"]

    def start(self):
        """Enqueue all pieces on the communication stream, after the work already queued on the
        current stream (the cast that produced the shard; the previous step's GEMMs, which read
        the receive buffers)."""
        cur = torch.cuda.current_stream()
        self._stream.wait_stream(cur)
        self._data.record_stream(self._stream)
        with torch.cuda.stream(self._stream):
            for c in range(self.chunks):
                src = self._data[c * self.rows_c:(c + 1) * self.rows_c]
                torch.distributed.all_gather_into_tensor(self._bufs[c], src, group=self.group)
                self._events[c].record(self._stream)
        return self

    def piece(self, c: int):
        """Storage of piece c; the current stream waits for its arrival."""
        torch.cuda.current_stream().wait_event(self._events[c])
        buf = self._bufs[c]
        return ops.Float8TensorStorage(data=buf, scale_inv=self.shard.scale_inv, fp8_dtype=self.shard.fp8_dtype,
                                       shape=tuple(buf.shape))

    def row_blocks(self, c: int):
        return (self.rows_c, self.rows_local, c * self.rows_c)


_COMM_STREAMS = {}


def _comm_stream(device):
    key = (device.type, device.index)
    if key not in _COMM_STREAMS:
        _COMM_STREAMS[key] = torch.cuda.Stream(device=device)
    return _COMM_STREAMS[key]


class SymmetricFP8Gather:
    """The per-tensor FP8 cast of a sequence-parallel shard IS its all-gather: the cast kernel stores
    every 16-byte output vector into the gather buffer of EVERY rank of the tensor-parallel group
    through peer-mapped (symmetric-memory) addresses, so the bytes cross NVLink / NVSwitch while the
    tensor is being cast, and what remains of the collective is one barrier. Replaces cast ->
    NCCL all-gather (reference pytorch/distributed.py:982-1073; its Userbuffers path gets the same
    effect with copy engines, common/comm_gemm_overlap/userbuffers/).

    The shard is cast in `chunks` row pieces: piece c of every rank lands in buffer c
    ([world * rows_c, cols], rank-major, the layout `ChunkedFP8Gather` uses), piece 0 on the current
    stream and the others on a side stream, each followed by its own barrier -- so the GEMM over piece
    c (`out_row_blocks`) runs while piece c + 1 is still crossing the switch. The cast kernel needs no
    shared memory and few registers, so its CTAs sit next to the persistent GEMM's on the same SMs.

    The current-scaling amax all-reduce goes the same way (`allreduce_amax`): every rank stores its
    amax into its slot of every rank's slot array, one barrier, local max over the slots.

    Two buffer sets alternate between steps: while step k's casts land in set k % 2 a slow peer may
    still be reading set (k - 1) % 2 in its GEMMs; it cannot be further behind, because every barrier
    of step k - 1 sits behind that rank's step k - 2 GEMMs in stream order.
    """

    def __init__(self, group, rows_local: int, cols: int, device, chunks: int = 2):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.rows_local, self.cols = rows_local, cols
        while chunks > 1 and (rows_local % chunks != 0 or (rows_local // chunks) % 128 != 0):
            chunks -= 1
        self.chunks, self.rows_c = chunks, rows_local // chunks
        piece = self.world * self.rows_c * cols            # bytes of one piece buffer
        amax_off = 2 * chunks * piece
        self._buf = symm.empty(amax_off + 2 * 4 * self.world + 64, dtype=torch.uint8, device=device)
        self._hdl = symm.rendezvous(self._buf, self.group)
        self._step = 0
        ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self._full, self._cfg, self._slots, self._slot_cfg = [], [], [], []
        for b in range(2):
            full, cfg = [], []
            for c in range(chunks):
                base = (b * chunks + c) * piece
                full.append(self._buf[base:base + piece].view(self.world * self.rows_c, cols))
                peers = [ptrs[r] + base + self.rank * self.rows_c * cols for r in range(self.world) if r != self.rank]
                cfg.append(ops.C.QuantConfig(peer_rowwise=peers) if peers else None)
            self._full.append(full)
            self._cfg.append(cfg)
            off = amax_off + b * 4 * self.world
            self._slots.append(self._buf[off:off + 4 * self.world].view(torch.float32))
            self._slot_cfg.append(ops.C.QuantConfig(peer_rowwise=[ptrs[r] + off + 4 * self.rank
                                                                  for r in range(self.world)]))
        self._side = torch.cuda.Stream(device=device)
        self._ev_cast0 = torch.cuda.Event()
        self._ev_piece = [torch.cuda.Event() for _ in range(chunks)]

    def allreduce_amax(self, x: torch.Tensor, amax: torch.Tensor) -> None:
        """amax[0] = max over the group of max|x| (call before `cast_and_gather` of the same step)."""
        b = self._step % 2
        ops.compute_amax(x, amax, peer_config=self._slot_cfg[b])
        self._hdl.barrier(channel=self.chunks)
        torch.amax(self._slots[b], dim=0, keepdim=True, out=amax)

    def cast_and_gather(self, x: torch.Tensor, scale: torch.Tensor, fp8_dtype, amax=None):
        """Enqueue the cast of `x` [rows_local, cols] with `scale` into every rank's piece buffers and
        the barriers. Use `piece(c)` / `row_blocks(c)` for the GEMMs and `local_copy()` for backward."""
        b = self._set = self._step % 2
        self._step += 1
        assert tuple(x.shape) == (self.rows_local, self.cols)
        self._scale_inv = torch.empty(1, dtype=torch.float32, device=x.device)
        self._fp8_dtype = fp8_dtype
        cur = torch.cuda.current_stream()
        for c in range(self.chunks):
            rows = slice(c * self.rows_c, (c + 1) * self.rows_c)
            local = ops.Float8TensorStorage(
                data=self._full[b][c][self.rank * self.rows_c:(self.rank + 1) * self.rows_c],
                scale_inv=self._scale_inv, fp8_dtype=fp8_dtype, shape=(self.rows_c, self.cols))
            if c == 0:
                ops.fp8_quantize(x[rows], scale, fp8_dtype, amax=amax, out=local, peer_config=self._cfg[b][c])
                self._ev_cast0.record(cur)
                self._hdl.barrier(channel=0)   # all ranks' piece 0 has landed everywhere behind this
            else:
                if c == 1:
                    self._side.wait_event(self._ev_cast0)   # piece 0 gets the links first
                    x.record_stream(self._side)
                with torch.cuda.stream(self._side):
                    ops.fp8_quantize(x[rows], scale, fp8_dtype, amax=amax, out=local, peer_config=self._cfg[b][c])
                    self._hdl.barrier(channel=c)
                    self._ev_piece[c].record(self._side)
        return self

    def piece(self, c: int):
        """Gathered piece c ([world * rows_c, cols], rank-major); the current stream waits for it."""
        if c > 0:
            torch.cuda.current_stream().wait_event(self._ev_piece[c])
        buf = self._full[self._set][c]
        return ops.Float8TensorStorage(data=buf, scale_inv=self._scale_inv, fp8_dtype=self._fp8_dtype,
                                       shape=tuple(buf.shape))

    def row_blocks(self, c: int):
        return (self.rows_c, self.rows_local, c * self.rows_c)

    def local_copy(self):
        """The rank's own FP8 shard, copied out of the piece buffers (they are recycled two steps later)."""
        r0, r1 = self.rank * self.rows_c, (self.rank + 1) * self.rows_c
        for c in range(1, self.chunks):
            torch.cuda.current_stream().wait_event(self._ev_piece[c])
        data = torch.cat([self._full[self._set][c][r0:r1] for c in range(self.chunks)])
        return ops.Float8TensorStorage(data=data, scale_inv=self._scale_inv, fp8_dtype=self._fp8_dtype,
                                       shape=(self.rows_local, self.cols))


_SHARED = {}


def _shared(cls, make, group, *key):
    """One instance per (class, group, shape...) for all modules of a process: the buffers are as large as the gathered
    activation (or twice the GEMM output), and the modules of a model run one after the other on one stream, which is
    all the protocols of these classes need -- every call is one more step of the same exchange. (Not for callers
    that run two such layers concurrently on different streams: construct private instances there.)"""
    k = (cls.__name__, id(group)) + key
    obj = _SHARED.get(k)
    if obj is None:
        obj = _SHARED[k] = make()
    return obj


class SymmetricGatherMXFP8:
    """All-gather of the sequence-parallel MXFP8 input INSIDE the GEMM that consumes it (the MXFP8 counterpart of
    `SymmetricExchange`; reference `_all_gather_mxfp8`, pytorch/distributed.py:1509-1640: separate NCCL all-gathers of
    data and scales followed by the GEMM). Block scales are local to the shard, so there is no amax exchange:

      1. the shard is quantized straight into the rank's own window of its gather buffers -- rowwise data rows and the
         GEMM-swizzled scale atoms of its 128-row blocks, which are contiguous in the gathered scale buffer;
      2. the GEMM (arrival flags + push, dynamic tile queue) sends both windows through the NVSwitch multicast address
         (`multimem.st`), chunk by chunk, and consumes the peers' chunks as their counters come up.

    What orders the steps without an amax exchange: the GEMM's last CTA pair advances the device-resident step and adds
    one to the `ready` word of every rank ("this GPU has finished reading"); the next launch's push waits until `ready`
    has reached world x step. One buffer set, CUDA-graph replayable. Needs NVSwitch multicast; `create` returns None
    when it is missing (the caller then gathers with NCCL)."""

    UNIT, UNIT2 = 16384, 4096

    @classmethod
    def shared(cls, group, rows_local: int, cols: int, device):
        return _shared(cls, lambda: cls.create(group, rows_local, cols, device), group, rows_local, cols, str(device))

    @classmethod
    def create(cls, group, rows_local: int, cols: int, device):
        if rows_local % 256 != 0 or cols % 512 != 0 or dist.get_world_size(group) > 8:
            return None
        obj = cls(group, rows_local, cols, device)
        return obj if obj.multicast else None

    def __init__(self, group, rows_local: int, cols: int, device, rows_per_flag: int = 256):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.rows_local, self.cols = rows_local, cols
        self.fpr = rows_local // rows_per_flag
        data_rank, sc_rank = rows_local * cols, rows_local * cols // 32
        data, scales = self.world * data_rank, self.world * sc_rank
        off_sc = (data + 1023) // 1024 * 1024
        off_flags = (off_sc + scales + 255) // 256 * 256
        off_step = off_flags + 4 * self.world * self.fpr
        off_step = (off_step + 63) // 64 * 64
        off_ready, off_err = off_step + 64, off_step + 128
        total = off_err + 64
        self._buf = symm.empty(total, dtype=torch.uint8, device=device)
        self._buf[off_flags:].zero_()
        torch.cuda.synchronize(device)
        self._hdl = symm.rendezvous(self._buf, self.group)
        self._hdl.barrier(channel=0)
        torch.cuda.synchronize(device)
        mc = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
        self.multicast = mc != 0 and os.environ.get("NVTE_B200_XCH_MULTICAST", "1") != "0"
        base = int(self._hdl.buffer_ptrs[self.rank])
        T = self.world * rows_local
        self.gathered = self._buf[:data].view(T, cols)
        self.gathered_scales = self._buf[off_sc:off_sc + scales].view(T, cols // 32)
        r0 = self.rank * rows_local
        self._win = self.gathered[r0:r0 + rows_local]
        self._win_sc = self.gathered_scales[r0:r0 + rows_local]
        self.error_word = self._buf[off_err:off_err + 4].view(torch.int32)
        self.step_word = self._buf[off_step:off_step + 4].view(torch.int32)
        self._tile_counter = torch.zeros(2, dtype=torch.int32, device=device)
        a = ops.C.ArrivalFlagsStruct()
        a.flags, a.step, a.error = base + off_flags, base + off_step, base + off_err
        a.world, a.rank, a.rows_per_flag, a.flags_per_rank = self.world, self.rank, rows_per_flag, self.fpr
        a.push_unit_bytes = self.UNIT
        a.push_mc_dst = mc + self.rank * data_rank
        a.push_mc_flags = mc + off_flags
        a.tile_counter = self._tile_counter.data_ptr()
        a.push_src2 = base + off_sc + self.rank * sc_rank
        a.push_mc_dst2 = mc + off_sc + self.rank * sc_rank
        a.push_unit2_bytes = self.UNIT2
        a.push_bytes2_per_flag = rows_per_flag * cols // 32
        a.advance_step = 1
        a.ready, a.ready_mc = base + off_ready, mc + off_ready
        self.arrival = a

    def cast(self, x: torch.Tensor, fp8_dtype, columnwise: bool):
        """Quantize the shard into the gather windows. Returns (storage of the FULL gathered tensor -- complete only
        for the GEMM that honours `self.arrival` --, storage with the shard's column-wise payload for the backward)."""
        assert tuple(x.shape) == (self.rows_local, self.cols)
        cs = ops.mxfp8_scale_shapes(self.rows_local, self.cols)[1]
        local = ops.MXFP8TensorStorage(
            rowwise_data=self._win, rowwise_scale_inv=self._win_sc,
            columnwise_data=torch.empty(self.rows_local, self.cols, dtype=torch.uint8, device=x.device) if columnwise else None,
            columnwise_scale_inv=torch.empty(cs, dtype=torch.uint8, device=x.device) if columnwise else None,
            fp8_dtype=fp8_dtype, shape=(self.rows_local, self.cols), with_gemm_swizzled_scales=True)
        ops.mxfp8_quantize(x, fp8_dtype, rowwise=True, columnwise=columnwise, swizzled_scales=True, out=local)
        full = ops.MXFP8TensorStorage(
            rowwise_data=self.gathered, rowwise_scale_inv=self.gathered_scales, columnwise_data=None,
            columnwise_scale_inv=None, fp8_dtype=fp8_dtype, shape=tuple(self.gathered.shape),
            with_gemm_swizzled_scales=True)
        saved = None
        if columnwise:
            saved = ops.MXFP8TensorStorage(
                rowwise_data=None, rowwise_scale_inv=None, columnwise_data=local.columnwise_data,
                columnwise_scale_inv=local.columnwise_scale_inv, fp8_dtype=fp8_dtype,
                shape=(self.rows_local, self.cols), with_gemm_swizzled_scales=True)
        return full, saved


class SymmetricReduceScatter:
    """Reduce-scatter (along the rows) of a row-parallel GEMM's output with its first half INSIDE the GEMM: the epilogue
    stores every output row straight from the accumulator into its owner rank's buffer of partial results, over
    NVLink (`kNVTEMatmulConfigB200ScatterOutput`); after a barrier each rank sums the `world` partial slabs of its own
    rows in rank order (fp32 accumulation, one rounding: `nvte_b200_reduce_partials`). Replaces GEMM -> NCCL
    reduce-scatter (reference module/linear.py:560-600; its Userbuffers overlap pipelines split GEMMs against ring
    reduce-scatter steps instead, common/comm_gemm_overlap/comm_gemm_overlap.cpp:611-760). The full [T, N] GEMM
    output never exists in HBM; the NVLink traffic runs under the GEMM. Deterministic (fixed summation order).

    Two buffer sets alternate by call: rank r's next GEMM may start writing into rank q's buffers while q still sums
    the previous call's; the barrier of the call in between orders everything older. (Under CUDA-graph capture, where
    a replay repeats the captured set, the call ends with a second barrier instead.)"""

    @classmethod
    def shared(cls, group, rows_total: int, cols: int, dtype: torch.dtype, device):
        return _shared(cls, lambda: cls(group, rows_total, cols, dtype, device), group, rows_total, cols, dtype, str(device))

    def __init__(self, group, rows_total: int, cols: int, dtype: torch.dtype, device):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        assert self.world <= 8 and rows_total % self.world == 0
        self.rows_total, self.cols, self.dtype = rows_total, cols, dtype
        self.rows = rows_total // self.world
        esz = torch.empty(0, dtype=dtype).element_size()
        self.slab = self.rows * cols * esz                     # one rank's partial of my rows
        assert self.slab % 32 == 0, "rows_per_rank * cols * element size must be a multiple of 32 bytes"
        per_set = self.world * self.slab
        self._buf = symm.empty(2 * per_set, dtype=torch.uint8, device=device)
        self._hdl = symm.rendezvous(self._buf, self.group)
        ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self._sets, self._cfgs = [], []
        for k in range(2):
            self._sets.append(self._buf[k * per_set:(k + 1) * per_set].view(dtype).view(self.world, self.rows, cols))
            sc = ops.C.ScatterOutputStruct()
            sc.world, sc.rank, sc.rows_per_rank = self.world, self.rank, self.rows
            for q in range(self.world):
                sc.dst[q] = ptrs[q] + k * per_set + self.rank * self.slab
            self._cfgs.append(sc)
        self._k = 0
        self._hdl.barrier(channel=0)

    def gemm(self, A, B, layout: str) -> torch.Tensor:
        """out[rows of this rank] of sum over ranks of (this GEMM on every rank)."""
        k = self._k
        self._k ^= 1
        shell = self._sets[k].view(self.rows_total, self.cols)   # shape / dtype carrier, not written through
        ops.general_gemm(A, B, layout, out=shell, scatter=self._cfgs[k])
        self._hdl.barrier(channel=1 + k)     # every rank's GEMM of this call has completed
        out = torch.empty(self.rows, self.cols, dtype=self.dtype, device=self._buf.device)
        ops.reduce_partials(self._sets[k], out)
        if torch.cuda.is_current_stream_capturing():
            # The set alternation above is host-side: a REPLAYED graph uses the same set every time, and with an odd
            # number of calls per replay a fast rank's next GEMM could overwrite slabs a slow rank is still summing.
            # Under capture the call therefore closes with a second barrier (every rank has finished its sum).
            self._hdl.barrier(channel=3)
        return out


class SymmetricExchange:
    """The whole exchange of the sequence-parallel column-parallel FP8 Linear -- amax all-reduce(MAX) and all-gather of
    the cast shard -- over peer-mapped (symmetric) memory, with the all-gather INSIDE the GEMM that consumes it. No
    NCCL kernel, no barrier kernel (reference: amax all-reduce `csrc/quantizer.cpp:920-926`, cast, NCCL all-gather
    `pytorch/distributed.py:982-1073`, GEMM; its Userbuffers overlap splits the GEMM instead,
    common/comm_gemm_overlap/comm_gemm_overlap.cpp:600-760).

    Per step, on the compute stream:
      1. `nvte_compute_amax_with_config`: local amax, then (amax, step) into this rank's slot on every rank;
      2. `nvte_compute_scale_from_amax`: waits for the world's slots of this step, max -> amax -> scale (+ scale_inv),
         advances the step counter;
      3. `nvte_quantize_v2`: an ordinary local cast of the shard into the rank's own rows of its gather buffer;
      4. the GEMM (`arrival=`, push): walks the tile rows in arrival order -- own rows, then chunk 0 of every peer,
         chunk 1, ... -- while an otherwise idle warp of every second CTA streams the rank's rows to the peers in
         16 KB bulk copies (chunk 0 first) and bumps the peers' per-chunk counters; the TMA producer loads a remote
         chunk only once its counter has reached this step's value. The NVLink transfer runs under the first tiles.
    `push="cast"` keeps the older form: the cast kernel stores to every rank and raises per-chunk flags (the GEMM then
    starts after that kernel; a system-scope fence behind NVLink stores costs ~9 us per chunk).

    Why one buffer set is enough: step k + 1's scale kernel waits for every rank's amax of step k + 1, and a rank sends
    that only after its step-k GEMM (stream order) -- so nobody overwrites a gather buffer a peer is still reading, and
    nobody's counters run ahead. The amax slots alternate by step parity for the same reason one level down. The step
    counter lives on the device, so a captured CUDA graph replays correctly. A peer that never arrives times out after
    5 s and sets `error`. Current scaling only (the amax exchange is what orders the steps)."""

    PUSH_UNIT = 16384

    @classmethod
    def shared(cls, group, rows_local: int, cols: int, device):
        return _shared(cls, lambda: cls(group, rows_local, cols, device), group, rows_local, cols, str(device))

    def __init__(self, group, rows_local: int, cols: int, device, rows_per_flag: int = 256, push: Optional[str] = None):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        assert self.world <= 8, "symmetric exchange supports up to 8 ranks (one NVSwitch domain)"
        env = os.environ.get("NVTE_B200_XCH_ROWS_PER_FLAG")   # dev knob: coarser arrival flags
        if env and rows_local % int(env) == 0:
            rows_per_flag = int(env)
        self.push = push or os.environ.get("NVTE_B200_XCH_PUSH", "gemm")
        assert self.push in ("gemm", "cast")
        assert rows_local % rows_per_flag == 0 and rows_per_flag % 256 == 0 and cols % 16 == 0
        if self.push == "gemm" and (rows_per_flag * cols) % self.PUSH_UNIT != 0:
            self.push = "cast"
        self.rows_local, self.cols, self.rows_per_flag = rows_local, cols, rows_per_flag
        self.fpr = rows_local // rows_per_flag
        data = self.world * rows_local * cols
        off_flags = (data + 255) // 256 * 256
        off_slots = off_flags + 4 * self.world * self.fpr
        off_slots = (off_slots + 7) // 8 * 8
        off_step = off_slots + 8 * 2 * self.world
        off_done = off_step + 64
        off_err = off_done + 4 * (self.fpr + 1)
        off_err = (off_err + 3) // 4 * 4
        total = off_err + 64
        self._buf = symm.empty(total, dtype=torch.uint8, device=device)
        self._buf[data:].zero_()
        torch.cuda.synchronize(device)
        self._hdl = symm.rendezvous(self._buf, self.group)
        self._hdl.barrier(channel=0)     # every rank's control words are zero before anyone pushes
        torch.cuda.synchronize(device)
        ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self.gathered = self._buf[:data].view(self.world * rows_local, cols)
        self.error_word = self._buf[off_err:off_err + 4].view(torch.int32)
        self.step_word = self._buf[off_step:off_step + 4].view(torch.int32)
        x = ops.C.SymmetricExchangeStruct()
        x.world, x.rank, x.rows_per_flag = self.world, self.rank, rows_per_flag
        x.flags_per_rank = self.fpr if self.push == "cast" else 0   # 0: the scale kernel advances the step
        for r in range(self.world):
            x.gather[r] = ptrs[r] + self.rank * rows_local * cols
            x.flags[r] = ptrs[r] + off_flags
            x.amax_slots[r] = ptrs[r] + off_slots
        x.step, x.done, x.error = ptrs[self.rank] + off_step, ptrs[self.rank] + off_done, ptrs[self.rank] + off_err
        self.struct = x
        self.cfg = ops.C.QuantConfig(symmetric_exchange=x)
        a = ops.C.ArrivalFlagsStruct()
        a.flags, a.step, a.error = ptrs[self.rank] + off_flags, x.step, x.error
        a.world, a.rank, a.rows_per_flag, a.flags_per_rank = self.world, self.rank, rows_per_flag, self.fpr
        if self.push == "gemm":
            a.push_unit_bytes = self.PUSH_UNIT
            for r in range(self.world):
                a.push_dst[r] = ptrs[r] + self.rank * rows_local * cols
                a.push_flags[r] = ptrs[r] + off_flags
            # NVSwitch multicast (NVLS) when the symmetric allocation has it: the rows leave this GPU once
            mc = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
            if mc != 0 and os.environ.get("NVTE_B200_XCH_MULTICAST", "1") != "0":
                a.push_mc_dst = mc + self.rank * rows_local * cols
                a.push_mc_flags = mc + off_flags
        if os.environ.get("NVTE_B200_XCH_DYNAMIC", "1") != "0":
            self._tile_counter = torch.zeros(2, dtype=torch.int32, device=device)
            a.tile_counter = self._tile_counter.data_ptr()
        self.arrival = a
        self.multicast = bool(a.push_mc_dst)
        self._scale_inv = torch.empty(1, dtype=torch.float32, device=device)

    def cast_and_gather(self, x: torch.Tensor, quantizer) -> "ops.Float8TensorStorage":
        """amax -> (exchange) -> scale -> cast into the gather buffer. Returns the storage of the FULL gathered
        tensor; it is complete only for a consumer that honours `self.arrival` (the GEMM, which with push="gemm"
        is also what sends the rows)."""
        assert tuple(x.shape) == (self.rows_local, self.cols)
        q = quantizer
        ops.compute_amax(x, q.amax, peer_config=self.cfg)
        ops.compute_scale_from_amax(q.amax, q.scale, q.dtype, q.force_pow_2_scales, q.amax_epsilon,
                                    symmetric_exchange=self.struct, scale_inv=self._scale_inv)
        local = ops.Float8TensorStorage(
            data=self.gathered[self.rank * self.rows_local:(self.rank + 1) * self.rows_local],
            scale_inv=self._scale_inv, fp8_dtype=q.dtype, shape=(self.rows_local, self.cols))
        ops.fp8_quantize(x, q.scale, q.dtype, amax=None, out=local,
                         peer_config=self.cfg if self.push == "cast" else None)
        self._dtype = q.dtype
        return ops.Float8TensorStorage(data=self.gathered, scale_inv=self._scale_inv, fp8_dtype=q.dtype,
                                       shape=tuple(self.gathered.shape))

    def local_copy(self):
        """The rank's own FP8 shard, copied out of the gather buffer (recycled by the next step)."""
        r0 = self.rank * self.rows_local
        return ops.Float8TensorStorage(data=self.gathered[r0:r0 + self.rows_local].clone(), scale_inv=self._scale_inv.clone(),
                                       fp8_dtype=self._dtype, shape=(self.rows_local, self.cols))

    def wait_all(self):
        """Make the current stream wait until every rank's shard of the last step has fully arrived (for consumers
        other than the flag-aware GEMM): a barrier over the group."""
        self._hdl.barrier(channel=1)
