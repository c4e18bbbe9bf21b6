"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the image loss of the train step.

numpy restatement of crates/brush-train/src/ssim.rs (Ssim::new :34-41, gaussian :7-15, Ssim::ssim :43-112) and of
the loss line crates/brush-train/src/train.rs:249-268.  Only tests/ may import this module.

Parity status: UNPINNED -- the reference holds no golden vector or known-answer test for SSIM or the loss; the
restatement is cross-checked in tests/ against an independent torch.nn.functional.conv2d evaluation of the same
definition (grouped conv, zero padding 6), and the CUDA gradient against torch autograd of that evaluation."""
from __future__ import annotations

import numpy as np


def gaussian_window(window_size: int = 11, sigma: float = 1.5) -> np.ndarray:
    """ssim.rs:7-15: exp(-(x - window/2)^2 / (2 sigma^2)), normalised; f32 like the reference."""
    extent = np.float32(window_size // 2)
    x = np.arange(window_size, dtype=np.float32)
    vals = np.exp(-((x - extent) ** 2) / np.float32(2.0 * sigma * sigma)).astype(np.float32)
    return (vals / vals.sum(dtype=np.float32)).astype(np.float32)


def _conv_same_plus(img: np.ndarray, w1: np.ndarray) -> np.ndarray:
    """Grouped conv2d with the 2-D window w1 (x) w1 and zero padding ceil(window/2) on every side (ssim.rs:47-49):
    [H, W] -> [H + 2*pad - window + 1, W + ...] = [H+2, W+2] for window 11.  float64 accumulation."""
    win = len(w1)
    pad = (win + 1) // 2
    h, w = img.shape
    xp = np.zeros((h + 2 * pad, w + 2 * pad), dtype=np.float64)
    xp[pad:pad + h, pad:pad + w] = img
    oh, ow = h + 2 * pad - win + 1, w + 2 * pad - win + 1
    tmp = np.zeros((h + 2 * pad, ow), dtype=np.float64)
    for k in range(win):
        tmp += float(w1[k]) * xp[:, k:k + ow]
    out = np.zeros((oh, ow), dtype=np.float64)
    for k in range(win):
        out += float(w1[k]) * tmp[k:k + oh, :]
    return out


def ssim(img1: np.ndarray, img2: np.ndarray, window_size: int = 11) -> float:
    """ssim.rs:43-112.  img: [H, W, 3].  Mean over the 3 x (H+2) x (W+2) map."""
    w1 = gaussian_window(window_size)
    c1, c2 = float(np.float32(0.01) ** 2), float(np.float32(0.03) ** 2)
    total, count = 0.0, 0
    for c in range(3):
        x, y = img1[..., c].astype(np.float64), img2[..., c].astype(np.float64)
        mu_x, mu_y = _conv_same_plus(x, w1), _conv_same_plus(y, w1)
        mu_xx, mu_yy, mu_xy = mu_x * mu_x, mu_y * mu_y, mu_x * mu_y
        s_xx = np.maximum(_conv_same_plus(x * x, w1) - mu_xx, 0.0)   # clamp_min(0.0), ssim.rs:84
        s_yy = np.maximum(_conv_same_plus(y * y, w1) - mu_yy, 0.0)   # ssim.rs:92
        s_xy = _conv_same_plus(x * y, w1) - mu_xy
        m = ((mu_xy * 2.0 + c1) * (s_xy * 2.0 + c2)) / ((mu_xx + mu_yy + c1) * (s_xx + s_yy + c2))
        total += float(m.sum())
        count += m.size
    return total / count


def image_loss(pred: np.ndarray, gt: np.ndarray, ssim_weight: float, window_size: int = 11) -> float:
    """train.rs:249-268.  pred [H,W,4]; gt [H,W,3] (compare RGB) or [H,W,4] (compare RGBA)."""
    compare = pred if gt.shape[-1] == 4 else pred[..., :3]
    l1 = float(np.abs(compare.astype(np.float64) - gt.astype(np.float64)).mean())
    if ssim_weight > 0.0:
        return l1 * (1.0 - ssim_weight) - ssim(pred[..., :3], gt[..., :3], window_size) * ssim_weight
    return l1
