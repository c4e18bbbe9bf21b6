"""Host-side checks of the .ply export/import (brush_b200/splat_io.py) against the format the reference writes
(crates/brush-dataset/src/splat_export.rs) and the rules it reads with (splat_import.rs)."""
import struct

import numpy as np
import pytest
import torch

from brush_b200 import Splats
from brush_b200 import splat_io as sio


def _splats(n, k, seed=0):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    return Splats(r(n, 3), r(n, k, 3), r(n, 4), r(n), r(n, 3))


def test_export_header_and_layout():
    s = _splats(3, 4)
    data = sio.splat_to_ply(s)
    head, body = data.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[:5] == ["ply", "format binary_little_endian 1.0", "comment Exported from Brush",
                         "comment Vertical axis: y", "element vertex 3"]
    props = [l.split()[-1] for l in lines[5:]]
    assert props == list(sio.MIN_PROPS) + [f"f_rest_{i}" for i in range(9)] and all("float" in l for l in lines[5:])
    row = struct.unpack("<23f", body[:92])
    assert len(body) == 3 * 92
    sh = s.sh_coeffs.detach().numpy()
    assert np.allclose(row[:3], s.means[0].detach().numpy()) and np.allclose(row[3:6], s.log_scales[0].detach().numpy())
    assert row[6] == pytest.approx(float(s.raw_opacity[0])) and np.allclose(row[7:11], s.rotation[0].detach().numpy())
    assert np.allclose(row[11:14], sh[0, 0, :])                                   # f_dc = coefficient 0 of r, g, b
    assert np.allclose(row[14:17], sh[0, 1:, 0]) and np.allclose(row[17:20], sh[0, 1:, 1])  # channel-major rest
    assert sio.ply_count(data) == 3


@pytest.mark.parametrize("k", [1, 4, 16])
def test_round_trip(k):
    s = _splats(1234, k, seed=k)
    back = sio.load_splat_from_ply(sio.splat_to_ply(s))
    assert torch.equal(back.means, s.means) and torch.equal(back.sh_coeffs, s.sh_coeffs)
    assert torch.equal(back.raw_opacity, s.raw_opacity) and torch.equal(back.log_scales, s.log_scales)
    norm = s.rotation / s.rotation.norm(dim=1, keepdim=True).clamp_min(1e-6)      # import normalises rotations
    assert torch.allclose(back.rotation, norm, atol=1e-6)


def test_import_rules():
    s = _splats(7, 25, seed=3)                                                      # degree 4 in the file ...
    back = sio.load_splat_from_ply(sio.splat_to_ply(s))
    assert back.sh_coeffs.shape == (7, 16, 3)                                       # ... truncated to degree 3
    flat = torch.cat([s.sh_coeffs[:, :1, :].reshape(7, 3), s.sh_coeffs[:, 1:, :].reshape(7, -1)], 1)[:, :48]
    assert torch.equal(back.sh_coeffs.reshape(7, 48), flat)
    # ascii and big-endian encodings, shuffled property order, an extra non-float property that is ignored
    names = list(sio.MIN_PROPS) + ["f_rest_0", "f_rest_1", "f_rest_2"]
    rng = np.random.default_rng(0)
    vals = rng.normal(size=(2, len(names))).astype(np.float32)
    perm = rng.permutation(len(names))
    hdr = ["ply", "format ascii 1.0", "element vertex 2"] + [f"property float {names[i]}" for i in perm] + \
          ["property uchar red", "end_header"]
    body = "\n".join(" ".join(repr(float(v)) for v in vals[r, perm]) + " 7" for r in range(2)) + "\n"
    a = sio.load_splat_from_ply(("\n".join(hdr) + "\n" + body).encode())
    hdr[1] = "format binary_big_endian 1.0"
    bin_body = b"".join(struct.pack(">17f", *vals[r, perm]) + b"\x07" for r in range(2))
    b = sio.load_splat_from_ply(("\n".join(hdr) + "\n").encode() + bin_body)
    for got in (a, b):
        assert np.allclose(got.means.detach().numpy(), vals[:, :3]) and got.sh_coeffs.shape == (2, 2, 3)
        assert np.allclose(got.sh_coeffs.detach().numpy()[:, 1, :], vals[:, 14:17])  # one rest coeff per channel
    with pytest.raises(ValueError, match="Missing properties"):
        sio.load_splat_from_ply(b"ply\nformat ascii 1.0\nelement vertex 0\nproperty float x\nend_header\n")


def test_stream_updates_every_50000():
    s = _splats(120_001, 1)
    sizes = [p.num_splats() for p in sio.load_splat_from_ply_stream(sio.splat_to_ply(s))]
    assert sizes == [50_000, 100_000, 120_001]
