"""Helpers shared by the tests: torch <-> numpy bit views and seeded inputs."""
import numpy as np
import torch


def bits16(t: torch.Tensor) -> np.ndarray:
    """bf16/fp16 torch tensor (any device) -> uint16 numpy bit pattern."""
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


def u8(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.uint8).numpy()


def make_input(rows, cols, kind="randn", seed=0, dtype=torch.bfloat16):
    """Synthetic inputs S1-S3 of SURVEY.md section 8(d)."""
    g = torch.Generator().manual_seed(seed)
    if kind == "randn":
        x = torch.randn(rows, cols, generator=g)
    elif kind == "uniform":
        x = torch.rand(rows, cols, generator=g) * 3.0 - 2.0
    elif kind == "outliers":
        x = torch.randn(rows, cols, generator=g)
        mask = torch.rand(rows, cols, generator=g) < 1e-3
        x = torch.where(mask, x * 1024.0, x)
        x[: min(32, rows), : min(32, cols)] = 0.0
        if rows > 40:
            x[40, :] = torch.where(torch.rand(cols, generator=g) < 0.5, -3.3895e38, 3.3895e38)
    elif kind == "zeros":
        x = torch.zeros(rows, cols)
    else:
        raise ValueError(kind)
    return x.to(dtype)
