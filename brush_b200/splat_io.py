"""Splat files either side of the path: the Inria-layout .ply Brush exports and imports.

    splat_to_ply         == crates/brush-dataset/src/splat_export.rs:11-106
    load_splat_from_ply  == crates/brush-dataset/src/splat_import.rs:179-330 (header checks, SH interleave,
                            degree-3 truncation, rotation normalisation, 50 000-splat progress updates)
    ply_count            == splat_import.rs:155-165

Host-side numpy; tensors land on the device the caller names.  SH layout: the path keeps coefficients
[n, coeff, channel]; the file stores f_dc_{0,1,2} = coefficient 0 per channel, then f_rest_* channel-major
([channel][coeff-1], "Inria format", splat_export.rs:17-23, splat_import.rs:167-177)."""
from __future__ import annotations

import io
from typing import Iterator, List, Tuple

import numpy as np
import torch

from .render import Splats

MIN_PROPS = ("x", "y", "z", "scale_0", "scale_1", "scale_2", "opacity", "rot_0", "rot_1", "rot_2", "rot_3",
             "f_dc_0", "f_dc_1", "f_dc_2")  # splat_export.rs:71-74, splat_import.rs:194-197
UPDATE_EVERY = 50_000                        # splat_import.rs:187
MAX_SH_FLOATS = 16 * 3                       # sh_coeffs_for_degree(3) * 3, splat_import.rs:236-241

_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
              "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
              "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


def splat_to_ply(splats: Splats) -> bytes:
    """splat_export.rs:66-106: binary little-endian, one `vertex` element, every property `float`."""
    f = lambda t: t.detach().to("cpu", torch.float32).numpy()
    means, scales, rot, opac = f(splats.means), f(splats.log_scales), f(splats.rotation), f(splats.raw_opacity)
    sh = f(splats.sh_coeffs)                          # [n, K, 3]
    n, k = sh.shape[0], sh.shape[1]
    sh_inria = np.transpose(sh, (0, 2, 1))            # [n, channel, coeff]   (splat_export.rs:20)
    sh_dc = sh_inria[:, :, 0]                         # [n, 3]
    sh_rest = sh_inria[:, :, 1:].reshape(n, 3 * (k - 1))  # red rest | green rest | blue rest
    names = list(MIN_PROPS) + [f"f_rest_{i}" for i in range(3 * (k - 1))]
    header = ["ply", "format binary_little_endian 1.0", "comment Exported from Brush", "comment Vertical axis: y",
              f"element vertex {n}"] + [f"property float {p}" for p in names] + ["end_header"]
    table = np.concatenate([means, scales, opac[:, None], rot, sh_dc, sh_rest], axis=1).astype("<f4")
    return ("\n".join(header) + "\n").encode("ascii") + table.tobytes()


def _read_header(buf: io.BytesIO):
    """Returns (encoding, [(element name, count, [(property name, numpy dtype | None for lists)])])."""
    if buf.readline().strip() != b"ply":
        raise ValueError("Invalid ply file")
    encoding, elements = None, []
    while True:
        line = buf.readline()
        if not line:
            raise ValueError("Invalid ply file: no end_header")
        tok = line.decode("ascii", "replace").split()
        if not tok or tok[0] in ("comment", "obj_info"):
            continue
        if tok[0] == "format":
            encoding = tok[1]
        elif tok[0] == "element":
            elements.append((tok[1], int(tok[2]), []))
        elif tok[0] == "property":
            if tok[1] == "list":
                elements[-1][2].append((tok[4], None))
            else:
                elements[-1][2].append((tok[2], _PLY_TYPES[tok[1]]))
        elif tok[0] == "end_header":
            break
    if encoding not in ("ascii", "binary_little_endian", "binary_big_endian"):
        raise ValueError("Invalid ply file: unknown format")
    return encoding, elements


def ply_count(ply_data: bytes) -> int:
    """splat_import.rs:155-165."""
    _, elements = _read_header(io.BytesIO(ply_data))
    for name, count, _ in elements:
        if name == "vertex":
            return count
    raise ValueError("Invalid ply file")


def _vertex_table(ply_data: bytes) -> Tuple[np.ndarray, List[str]]:
    """The vertex element as float32 [n, n_props] plus the property names (only scalar properties are supported,
    like the reference's GaussianData, which ignores every non-float property)."""
    buf = io.BytesIO(ply_data)
    encoding, elements = _read_header(buf)
    for name, count, props in elements:
        if any(t is None for _, t in props):
            raise ValueError("list properties are not supported in a splat ply")
        if name != "vertex":
            if encoding == "ascii":
                for _ in range(count):
                    buf.readline()
            else:
                buf.seek(count * sum(np.dtype(t).itemsize for _, t in props), io.SEEK_CUR)
            continue
        names = [p for p, _ in props]
        if not all(p in names for p in MIN_PROPS):
            raise ValueError("Invalid splat ply. Missing properties!")  # splat_import.rs:199-204
        if encoding == "ascii":
            rows = [buf.readline().split() for _ in range(count)]
            table = np.array(rows, dtype=np.float64).reshape(count, len(props)).astype(np.float32)
            is_float = [True] * len(props)
        else:
            order = "<" if encoding == "binary_little_endian" else ">"
            dt = np.dtype([(f"p{i}", order + t) for i, (_, t) in enumerate(props)])
            raw = np.frombuffer(buf.read(count * dt.itemsize), dtype=dt, count=count)
            table = np.stack([raw[f"p{i}"].astype(np.float32) for i in range(len(props))], axis=1) \
                if count else np.zeros((0, len(props)), np.float32)
            is_float = [t == "f4" for _, t in props]
        # set_property only accepts Property::Float (splat_import.rs:43): other types leave the field at 0
        for i, ok in enumerate(is_float):
            if not ok:
                table[:, i] = 0.0
        return table, names
    raise ValueError("Invalid ply file")


def _columns_to_arrays(table: np.ndarray, names: List[str]):
    col = {p: i for i, p in enumerate(names)}
    pick = lambda keys: table[:, [col[k] for k in keys]]
    means, scales = pick(("x", "y", "z")), pick(("scale_0", "scale_1", "scale_2"))
    opac, rot = table[:, col["opacity"]], pick(("rot_0", "rot_1", "rot_2", "rot_3"))
    sh_dc = pick(("f_dc_0", "f_dc_1", "f_dc_2"))
    rest_ids = [int(p[len("f_rest_"):]) for p in names if p.startswith("f_rest_") and p[len("f_rest_"):].isdigit()]
    n_rest = (1 + max(rest_ids)) if rest_ids else 0  # missing indices stay 0 (Vec::resize, splat_import.rs:61-63)
    rest = np.zeros((table.shape[0], n_rest), np.float32)
    for p in names:
        if p.startswith("f_rest_") and p[len("f_rest_"):].isdigit():
            rest[:, int(p[len("f_rest_"):])] = table[:, col[p]]
    per_channel = n_rest // 3                          # interleave_coeffs, splat_import.rs:167-177
    if per_channel:
        inter = rest[:, :3 * per_channel].reshape(-1, 3, per_channel).transpose(0, 2, 1).reshape(-1, 3 * per_channel)
    else:
        inter = np.zeros((table.shape[0], 0), np.float32)
    sh = np.concatenate([sh_dc, inter], axis=1)[:, :MAX_SH_FLOATS]  # degree-3 truncation, :236-241
    n_coeffs = sh.shape[1] // 3
    return means, sh[:, :3 * n_coeffs].reshape(-1, n_coeffs, 3), rot, opac, scales


def _make_splats(arrs, device) -> Splats:
    means, sh, rot, opac, scales = (torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32, device=device)
                                    for a in arrs)
    s = Splats(means, sh, rot, opac, scales)
    s.norm_rotations()  # splat_import.rs:140,151
    return s


def load_splat_from_ply_stream(ply_data: bytes, device="cpu") -> Iterator[Splats]:
    """splat_import.rs:179-330: yields the splats read so far after every 50 000 and once more at the end."""
    table, names = _vertex_table(ply_data)
    arrs = _columns_to_arrays(table, names)
    n = table.shape[0]
    for end in list(range(UPDATE_EVERY, n, UPDATE_EVERY)) + [n]:
        yield _make_splats([a[:end] for a in arrs], device)


def load_splat_from_ply(ply_data: bytes, device="cpu") -> Splats:
    last = None
    for last in load_splat_from_ply_stream(ply_data, device):
        pass
    return last
