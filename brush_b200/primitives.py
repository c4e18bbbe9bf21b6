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


NOALLOC_SCAN_MAX_N = 1 << 24  # BRUSH_NOALLOC_SCAN_MAX_N


def prefix_sum_noalloc(x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """Inclusive scan of `x` into the caller's `out` through brush_prefix_sum_u32_noalloc: no ctx, no allocation (the
    kernel's scratch is a static device array inside the library), at most NOALLOC_SCAN_MAX_N elements.  No
    counterpart in the reference.  Calls on a device share the scratch: keep them on one stream or order them."""
    for t in (x, out):
        if not t.is_cuda or t.dtype != torch.int32 or not t.is_contiguous():
            raise ValueError("prefix_sum_noalloc needs contiguous int32 CUDA tensors (no CPU fallback exists)")
    if out.shape != x.shape or out.device != x.device or x.dim() != 1:
        raise ValueError("x and out must be 1-d, of one shape, on one device")
    err = C.create_string_buffer(256)
    with torch.cuda.device(x.device):
        rc = _lib.lib().brush_prefix_sum_u32_noalloc(x.data_ptr(), x.shape[0], out.data_ptr(),
                                                     C.c_void_p(_stream_ptr(x.device)), err, len(err))
    if rc != 0:
        raise (ValueError if rc == -1 else RuntimeError)(err.value.decode() or f"brush error {rc}")
    return out
