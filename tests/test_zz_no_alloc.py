"""brush_prefix_sum_u32_noalloc: the allocation-free variant of the scan (scratch in a static __device__ array inside
the library; include/brush_b200.h).  No counterpart in the reference: the checker is the existing entry point
brush_prefix_sum_u32 on the same input, bit-exact (u32 work), on calls in a row that share the static scratch.
Named test_zz_* so pytest collects it last."""
import ctypes as C

import numpy as np
import pytest
import torch

gpu = pytest.mark.gpu


def _both(x):
    from brush_b200 import primitives as P

    out = torch.full_like(x, -12345)
    P.prefix_sum_noalloc(x, out)
    ref = P.prefix_sum(x)
    return out, ref


@gpu
def test_noalloc_scan_matches_ctx_scan_on_calls_in_a_row():
    from brush_b200 import primitives as P

    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(7)
    # a large call first, smaller ones after it: tickets and status words of the earlier call are still in the static
    # array and must not leak into the later ones; sizes straddle the 4096-element tile and the 16-byte vector path
    sizes = [1_000_003, 5000, 1, 4096, 4097, 16, 1_000_003, P.NOALLOC_SCAN_MAX_N]
    for rep, n in enumerate(sizes):
        x = torch.randint(0, 1 << 20, (n,), generator=g, dtype=torch.int32).to(dev)
        out, ref = _both(x)
        assert torch.equal(out, ref), f"call {rep} (n={n}) differs from brush_prefix_sum_u32"
        if n <= 5000:
            assert np.array_equal(out.cpu().numpy().view(np.uint32),
                                  np.cumsum(x.cpu().numpy().view(np.uint32), dtype=np.uint64).astype(np.uint32))
    # two calls back to back on a side stream, no host synchronisation in between
    s = torch.cuda.Stream(dev)
    xa = torch.randint(0, 255, (300_000,), generator=g, dtype=torch.int32).to(dev)
    xb = torch.randint(0, 255, (77_777,), generator=g, dtype=torch.int32).to(dev)
    ra, rb = P.prefix_sum(xa), P.prefix_sum(xb)
    oa, ob = torch.empty_like(xa), torch.empty_like(xb)
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        P.prefix_sum_noalloc(xa, oa)
        P.prefix_sum_noalloc(xb, ob)
    s.synchronize()
    assert torch.equal(oa, ra) and torch.equal(ob, rb)
    # unaligned views take the scalar path
    out, ref = _both(xa[1:100_000].clone()[1:].contiguous())
    assert torch.equal(out, ref)


@gpu
def test_noalloc_scan_allocates_nothing_and_rejects_oversize():
    from brush_b200 import _lib, primitives as P

    dev = torch.device("cuda:0")
    x = torch.arange(1, 200_001, dtype=torch.int32, device=dev)
    out = torch.empty_like(x)
    P.prefix_sum_noalloc(x, out)  # first call on the device: symbol lookup, module load
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(3):
        P.prefix_sum_noalloc(x, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, "the no-alloc entry point changed the device's free memory"
    assert np.array_equal(out.cpu().numpy().view(np.uint32),
                          np.cumsum(np.arange(1, 200_001, dtype=np.uint64)).astype(np.uint32))
    # beyond the static scratch: a clear error, nothing launched; no buffer that large is needed to ask
    err = C.create_string_buffer(256)
    rc = _lib.lib().brush_prefix_sum_u32_noalloc(x.data_ptr(), P.NOALLOC_SCAN_MAX_N + 1, out.data_ptr(), None, err, 256)
    assert rc == -1 and b"BRUSH_NOALLOC_SCAN_MAX_N" in err.value
    # n == 0 and a missing error buffer are fine
    assert _lib.lib().brush_prefix_sum_u32_noalloc(None, 0, None, None, None, 0) == 0
    assert _lib.lib().brush_prefix_sum_u32_noalloc(None, 5, None, None, None, 0) == -1
