"""Prints SHA-256 of the forward's image and final_index on the bench scene (A/B of compositor variants: same bits?)."""
import hashlib, sys
sys.path.insert(0, '.')
import torch
import bench
from brush_b200 import Splats, scene
dev = torch.device("cuda", 0)
for wl in ("cfg4_1m_1080p", "cfg2_100k_1024"):
    sc = scene.make_config_scene(wl, None)
    t = lambda a: torch.as_tensor(a, device=dev)
    sp = Splats(t(sc.means), t(sc.sh_coeffs), t(sc.quats), t(sc.raw_opacity), t(sc.log_scales))
    cams = bench.view_cameras(sc.width, sc.height)
    for v in (0, 5):
        with torch.no_grad():
            img, aux = sp.render(cams[v], (sc.width, sc.height), False, isect_capacity=16777216)
        torch.cuda.synchronize()
        h = lambda x: hashlib.sha256(x.contiguous().cpu().numpy().tobytes()).hexdigest()[:16]
        print(wl, v, "img", h(img), "final_index", h(aux.final_index), flush=True)
