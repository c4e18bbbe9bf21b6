"""The two device primitives the path is built on, through the C ABI.

    radix_argsort == brush_sort::radix_argsort      (brush-sort/src/lib.rs:32-37)
    prefix_sum    == brush_prefix_sum::prefix_sum   (brush-prefix-sum/src/lib.rs:17)"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .render import _stream_ptr, get_context


def radix_argsort(input_keys: torch.Tensor, input_values: torch.Tensor, n_sort: torch.Tensor, sorting_bits: int):
    """Stable key-value sort of the first n_sort[0] entries (device-side count) on the low
    4*ceil(sorting_bits/4) key bits.  keys: int32 (treated as u32) or int64 (treated as u64); values int32."""
    if input_keys.shape[0] != input_values.shape[0]:
        raise ValueError("keys and values must have the same length")  # brush-sort/src/lib.rs:38
    wide = input_keys.dtype == torch.int64
    if sorting_bits > (64 if wide else 32):
        raise ValueError("sorting_bits too large")  # brush-sort/src/lib.rs:39
    if not input_keys.is_cuda:
        raise ValueError("radix_argsort needs CUDA tensors (no CPU fallback exists)")
    dev = input_keys.device
    ctx = get_context(dev)
    keys, vals = input_keys.contiguous(), input_values.contiguous().to(torch.int32)
    n_sort = n_sort.to(device=dev, dtype=torch.int32).contiguous()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    fn = _lib.lib().brush_radix_argsort_u64 if wide else _lib.lib().brush_radix_argsort_u32
    with torch.cuda.device(dev):
        rc = fn(ctx.handle, keys.data_ptr(), vals.data_ptr(), n_sort.data_ptr(), keys.shape[0], int(sorting_bits),
                ko.data_ptr(), vo.data_ptr(), C.c_void_p(_stream_ptr(dev)))
    ctx.check(rc)
    return ko, vo


def prefix_sum(x: torch.Tensor) -> torch.Tensor:
    """Inclusive scan of an int32 (u32, wrapping) tensor."""
    if not x.is_cuda:
        raise ValueError("prefix_sum needs a CUDA tensor (no CPU fallback exists)")
    dev = x.device
    ctx = get_context(dev)
    x = x.contiguous().to(torch.int32)
    out = torch.empty_like(x)
    with torch.cuda.device(dev):
        rc = _lib.lib().brush_prefix_sum_u32(ctx.handle, x.data_ptr(), x.shape[0], out.data_ptr(),
                                             C.c_void_p(_stream_ptr(dev)))
    ctx.check(rc)
    return out
