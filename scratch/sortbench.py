import sys, torch, numpy as np
sys.path.insert(0, '.')
from brush_b200 import radix_argsort, prefix_sum
dev = 'cuda'
def bench(n, max_n, bits, reps=50):
    g = torch.Generator(device=dev); g.manual_seed(1)
    keys = torch.randint(0, 2**31 - 1, (max_n,), device=dev, dtype=torch.int32, generator=g)
    if bits < 31:
        keys = keys & ((1 << bits) - 1)
    vals = torch.arange(max_n, device=dev, dtype=torch.int32)
    nd = torch.tensor([n], device=dev, dtype=torch.int32)
    for _ in range(5): radix_argsort(keys, vals, nd, bits)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): radix_argsort(keys, vals, nd, bits)
    e1.record(); torch.cuda.synchronize()
    print(f"n={n} max_n={max_n} bits={bits}: {e0.elapsed_time(e1)/reps*1e3:.1f} us")
for bits in (8, 16, 24, 32):
    bench(857_000, 1_000_000, bits)
for bits in (8, 13):
    bench(3_100_000, 16_000_000, bits)
    bench(3_100_000, 3_200_000, bits)
import ctypes as C
from brush_b200 import _lib
L = _lib.lib()
if hasattr(L, "brush_debug_sort_prof"):
    for n, mx, bits in ((857_000, 1_000_000, 8), (3_100_000, 16_000_000, 8)):
        g = torch.Generator(device=dev); g.manual_seed(1)
        keys = torch.randint(0, 255, (mx,), device=dev, dtype=torch.int32, generator=g)
        vals = torch.arange(mx, device=dev, dtype=torch.int32)
        nd = torch.tensor([n], device=dev, dtype=torch.int32)
        for _ in range(3): radix_argsort(keys, vals, nd, bits)
        torch.cuda.synchronize()
        buf = (C.c_ulonglong * 16)()
        L.brush_debug_sort_prof(buf)
        t = [buf[i] for i in range(8)]
        print(n, "phase cycles:", [t[i + 1] - t[i] for i in range(7)], "total", t[7] - t[0])
