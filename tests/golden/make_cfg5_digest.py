#!/usr/bin/env python
"""Writes tests/golden/cfg5_digest.json: the ORACLE's integer outputs at BASELINE.json configs[4] (6 M Gaussians, SH
degree 3, 3840x2160, ~77 M intersections -- the sort / scan stress case), reduced to SHA-256 digests so that the GPU
test can demand bit-exact lists at 1e8 keys without shipping them (the reference's own large sort test demands exact
key AND value equality, crates/brush-sort/src/lib.rs:219-265).

    python tests/golden/make_cfg5_digest.py          # minutes of CPU on all cores; run offline, commit the json

Digested, in the canonical form of SURVEY.md 8a note (vi):
    num_visible, num_intersections
    global_from_compact_gid[:V]                                  (depth order of the visible splats)
    cum_tiles_hit[:V]                                            (inclusive scan of the tile counts)
    tile_bins                                                    (per-tile [first, one-past-last) intersection)
    global_from_compact_gid[compact_gid_from_isect[:I]]          (the tile-sorted, depth-ordered intersection list)
The scene is brush_b200.scene.make_config_scene("cfg5_6m_4k") (seeded numpy generator) seen from the
reference bench camera (benches/render_bench.rs:163-174)."""
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

WORKLOAD = "cfg5_6m_4k"
CAPACITY = 160 * 1024 * 1024


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint32).tobytes()).hexdigest()


def digest(num_visible, num_intersections, gfc, cum, tile_bins, cg_from_isect):
    V, I = int(num_visible), int(num_intersections)
    gfc = np.asarray(gfc).view(np.uint32) if np.asarray(gfc).dtype != np.uint32 else np.asarray(gfc)
    cg = np.asarray(cg_from_isect).view(np.uint32) if np.asarray(cg_from_isect).dtype != np.uint32 else np.asarray(cg_from_isect)
    return {
        "num_visible": V,
        "num_intersections": I,
        "sha256_global_from_compact_gid": sha(gfc[:V]),
        "sha256_cum_tiles_hit": sha(np.asarray(cum).reshape(-1)[:V]),
        "sha256_tile_bins": sha(np.asarray(tile_bins).reshape(-1)),
        "sha256_global_gid_from_isect": sha(gfc[cg[:I]]),
    }


def main():
    from brush_b200 import scene
    from oracle import pyoracle as O

    O.build()
    O.set_threads(len(os.sched_getaffinity(0)))  # the PARITY build; thread count does not change integer outputs
    sc = scene.make_config_scene(WORKLOAD)
    W, H = sc.width, sc.height
    vm = np.eye(4, dtype=np.float32)
    vm[2, 3] = 8.0
    u = O.make_uniforms(vm, (0.5 * W, 0.5 * W), (W, H), (0.5 * W, 0.5 * H))
    t0 = time.time()
    _, aux = O.render_forward(u, sc.means, sc.log_scales, sc.quats, sc.sh_coeffs, sc.raw_opacity, capacity=CAPACITY)
    assert aux.status == 0, "capacity too small"
    d = digest(aux.num_visible, aux.num_intersections, aux.global_from_compact_gid, aux.cum_tiles_hit, aux.tile_bins,
               aux.compact_gid_from_isect)
    d.update({"workload": WORKLOAD, "n": int(sc.means.shape[0]), "width": W, "height": H, "capacity": CAPACITY,
              "camera": "bench camera (0,0,-8), identity, fov_x 90 deg",
              "generator": "tests/golden/make_cfg5_digest.py (oracle parity build)",
              "oracle_seconds": round(time.time() - t0, 1)})
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cfg5_digest.json")
    with open(out, "w") as f:
        json.dump(d, f, indent=1)
        f.write("\n")
    print(json.dumps(d, indent=1))


if __name__ == "__main__":
    main()
