"""Host <-> device plumbing for the numpy-facing drop-in classes.

numpy arrays owned by the caller come in, fresh numpy arrays go out (the
reference's ownership contract). Large inputs are streamed in
row batches through two pinned staging buffers on a copy stream so the PCIe
transfer of batch i+1 overlaps the kernels of batch i. Arrays that already live
in pinned memory (bench.py's e2e leg) are copied directly.
"""
from __future__ import annotations

from typing import Callable

import numpy as np
import torch

from .ops import require_cuda

DEFAULT_BATCH_BYTES = 512 << 20  # per staging buffer


def device() -> torch.device:
    require_cuda()
    return torch.device("cuda", torch.cuda.current_device())


def as_f32_2d(X) -> np.ndarray:
    X = np.ascontiguousarray(X, dtype=np.float32)
    if X.ndim != 2:
        raise ValueError(f"expected a 2-D array, got shape {X.shape}")
    return X


def to_device(a: np.ndarray, dtype: torch.dtype | None = None) -> torch.Tensor:
    """Copy a (small or pinned) numpy array to the current device."""
    dev = device()
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint16:
        a = a.view(np.int16)
    elif a.dtype == np.uint32:
        a = a.view(np.int32)
    t = torch.from_numpy(a)
    if t.numel() and not t.is_pinned() and t.numel() * t.element_size() > (1 << 20):
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t)
        t = h
    out = t.to(dev, non_blocking=True)
    if dtype is not None and out.dtype != dtype:
        out = out.to(dtype)
    return out


def to_host(t: torch.Tensor) -> np.ndarray:
    """Device tensor -> fresh numpy array (through pinned memory)."""
    if t.numel() == 0:
        return t.cpu().numpy()
    h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    h.copy_(t, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return h.numpy()


def batch_rows_for(row_bytes: int, n: int, batch_bytes: int = DEFAULT_BATCH_BYTES) -> int:
    return int(max(1, min(n, batch_bytes // max(1, row_bytes))))


def stream_rows(X: np.ndarray, fn: Callable[[torch.Tensor, int, int], None],
                batch_rows: int | None = None) -> None:
    """Call fn(x_dev, start, end) for consecutive row batches of the host array X.

    x_dev is only valid during the call (the buffer is recycled two batches later).
    """
    dev = device()
    n, D = X.shape
    if n == 0:
        return
    src = torch.from_numpy(X)
    if batch_rows is None:
        batch_rows = batch_rows_for(D * X.itemsize, n)
    pinned_src = src.is_pinned()
    main = torch.cuda.current_stream()
    if n <= batch_rows:
        if pinned_src:
            x = src.to(dev, non_blocking=True)
        else:
            x = to_device(X)
        fn(x, 0, n)
        return
    copy = torch.cuda.Stream(device=dev)
    dbuf = [torch.empty((batch_rows, D), dtype=src.dtype, device=dev) for _ in range(2)]
    hbuf = None if pinned_src else [torch.empty((batch_rows, D), dtype=src.dtype, pin_memory=True)
                                    for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]      # compute on dbuf[b] finished
    copied = [torch.cuda.Event() for _ in range(2)]    # H2D into dbuf[b] finished
    used = [False, False]
    for i, s in enumerate(range(0, n, batch_rows)):
        e = min(n, s + batch_rows)
        b = i & 1
        m = e - s
        if used[b]:
            done[b].synchronize()  # host may overwrite hbuf[b]; device may overwrite dbuf[b]
        if pinned_src:
            h = src[s:e]
        else:
            hbuf[b][:m].copy_(src[s:e])
            h = hbuf[b][:m]
        with torch.cuda.stream(copy):
            dbuf[b][:m].copy_(h, non_blocking=True)
            copied[b].record(copy)
        main.wait_event(copied[b])
        fn(dbuf[b][:m], s, e)
        done[b].record(main)
        used[b] = True
    main.synchronize()
