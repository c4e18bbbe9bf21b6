"""CPU checks of the image-loss oracle (oracle/ssim_oracle.py) against an independent torch evaluation of the
definition in crates/brush-train/src/ssim.rs (grouped conv2d, zero padding ceil(11/2) = 6)."""
import numpy as np
import torch

from oracle import ssim_oracle as so


def torch_ssim(x, y):  # x, y: [H, W, 3] double tensors; mirrors Ssim::ssim with torch's own conv2d
    w1 = torch.as_tensor(so.gaussian_window(11), dtype=torch.float64)
    w = (w1[:, None] @ w1[None, :])[None, None].repeat(3, 1, 1, 1)
    conv = lambda t: torch.nn.functional.conv2d(t.permute(2, 0, 1)[None], w, padding=6, groups=3)
    mu_x, mu_y = conv(x), conv(y)
    s_xx = (conv(x * x) - mu_x * mu_x).clamp_min(0.0)
    s_yy = (conv(y * y) - mu_y * mu_y).clamp_min(0.0)
    s_xy = conv(x * y) - mu_x * mu_y
    c1, c2 = float(np.float32(0.01) ** 2), float(np.float32(0.03) ** 2)
    return (((mu_x * mu_y * 2 + c1) * (s_xy * 2 + c2)) / ((mu_x * mu_x + mu_y * mu_y + c1) * (s_xx + s_yy + c2))).mean()


def test_window_is_normalised_and_symmetric():
    w = so.gaussian_window(11)
    assert w.dtype == np.float32 and abs(float(w.sum()) - 1.0) < 1e-6 and np.allclose(w, w[::-1])
    assert np.argmax(w) == 5


def test_ssim_matches_torch_conv_definition():
    rng = np.random.default_rng(3)
    for h, w in ((17, 23), (40, 33), (9, 5)):
        a, b = rng.uniform(0, 1, (h, w, 3)).astype(np.float32), rng.uniform(0, 1, (h, w, 3)).astype(np.float32)
        ref = float(torch_ssim(torch.as_tensor(a, dtype=torch.float64), torch.as_tensor(b, dtype=torch.float64)))
        assert abs(so.ssim(a, b) - ref) < 1e-12
    a = rng.uniform(0, 1, (30, 30, 3)).astype(np.float32)
    assert abs(so.ssim(a, a) - 1.0) < 1e-6  # identical images, padding included
    assert -1.0 < so.ssim(a, 1.0 - a) < 0.0  # anti-correlated pair


def test_loss_line():
    rng = np.random.default_rng(4)
    pred = rng.uniform(0, 1, (20, 24, 4)).astype(np.float32)
    for c in (3, 4):
        gt = rng.uniform(0, 1, (20, 24, c)).astype(np.float32)
        l1 = np.abs((pred if c == 4 else pred[..., :3]).astype(np.float64) - gt).mean()
        assert abs(so.image_loss(pred, gt, 0.0) - l1) < 1e-12
        assert abs(so.image_loss(pred, gt, 0.2) - (0.8 * l1 - 0.2 * so.ssim(pred[..., :3], gt[..., :3]))) < 1e-12
