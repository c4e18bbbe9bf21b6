"""Times brush_image_loss_backward alone (1920x1080, 3-channel target).  usage: python scratch/lossbench.py [reps]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from brush_b200 import _lib, get_context

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
dev = torch.device("cuda", 0)
H, W = 1080, 1920
g = torch.Generator(device=dev)
g.manual_seed(1)
pred = torch.rand((H, W, 4), device=dev, generator=g)
gt = torch.rand((H, W, 3), device=dev, generator=g)
v_out = torch.empty((H, W, 4), device=dev)
loss = torch.zeros(1, device=dev)
ctx = get_context(dev)
L = _lib.lib()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def call():
    ctx.check(L.brush_image_loss_backward(ctx.handle, pred.data_ptr(), gt.data_ptr(), H, W, 3, 0.2, 11,
                                          v_out.data_ptr(), loss.data_ptr(), st))


for _ in range(3):
    call()
torch.cuda.synchronize()
tot = 0.0
for _ in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    call()
    e1.record()
    torch.cuda.synchronize()
    tot += e0.elapsed_time(e1)
print(f"image loss fwd+bwd: {tot / reps * 1000:.1f} us per call (L2 flushed), loss {float(loss):.6f}")
