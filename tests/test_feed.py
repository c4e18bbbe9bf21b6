"""Host-side checks of the view feed (brush_b200/feed.py): image conversion, scene bounds, the NeRF-synthetic
camera convention (crates/brush-dataset/src/formats/nerf_synthetic.rs:55-68) and the prefetching loader."""
import math

import numpy as np
import torch

from brush_b200 import Camera
from brush_b200.feed import Scene, SceneLoader, SceneView, camera_from_nerf_transform, image_to_tensor


def test_image_to_tensor():
    img = np.arange(2 * 3 * 4, dtype=np.uint8).reshape(2, 3, 4)
    t = image_to_tensor(img)
    assert t.dtype == torch.float32 and t.shape == (2, 3, 4) and float(t[1, 2, 3]) == np.float32(23) / np.float32(255)
    assert image_to_tensor(img[..., :3]).shape == (2, 3, 3)


def test_nerf_camera_convention():
    # identity camera-to-world in the NeRF (OpenGL) frame: camera at the origin looking down -z with +y up.
    cam = camera_from_nerf_transform(np.eye(4), 0.8, 800, 800)
    # flip the camera's y and z axes: its forward axis is world (0,0,-1), its down axis world (0,-1,0); then rotate
    # the world +90 degrees about x, (x, y, z) -> (x, -z, y): forward -> (0,1,0), down -> (0,0,-1)
    r = cam._rotmat()
    assert np.allclose(r @ [0, 0, 1], [0, 1, 0], atol=1e-12) and np.allclose(r @ [0, 1, 0], [0, 0, -1], atol=1e-12)
    assert np.allclose(r @ [1, 0, 0], [1, 0, 0], atol=1e-12)
    assert abs(cam.fov_y - 0.8) < 1e-12 and cam.center_uv == (0.5, 0.5)
    # translation rotates with the world: (1, 2, 3) -> (1, -3, 2)
    m = np.eye(4)
    m[:3, 3] = [1, 2, 3]
    cam = camera_from_nerf_transform(m, 0.8, 800, 400)
    assert np.allclose(cam.position, [1, -3, 2], atol=1e-12)
    assert abs(cam.fov_y - 2 * math.atan(400 / (2 * (0.5 * 800 / math.tan(0.4))))) < 1e-12
    # a general rotation survives the round trip through the quaternion
    a = 0.7
    rot = np.array([[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1.0]])
    m = np.eye(4)
    m[:3, :3] = rot
    cam = camera_from_nerf_transform(m, 0.8, 64, 64)
    rx = np.array([[1, 0, 0], [0, 0, -1], [0, 1, 0.0]])
    assert np.allclose(cam._rotmat(), rx @ (rot * np.array([1, -1, -1.0])), atol=1e-9)
    assert np.allclose(cam.world_to_local().astype(np.float64) @ cam.local_to_world(), np.eye(4), atol=1e-6)


def _scene(n=6, h=8, w=12, c=3):
    rng = np.random.default_rng(1)
    views = []
    for i in range(n):
        cam = Camera((float(i), 0.0, -2.0 * i), (0.0, 0.0, 0.0, 1.0), 1.0, 1.0)
        views.append(SceneView(f"v{i}", cam, rng.integers(0, 256, (h, w, c), dtype=np.uint8)))
    return Scene(views)


def test_scene_bounds_and_extent():
    sc = _scene()
    center, extent = sc.bounds(0.0, 0.0)
    assert np.allclose(center, [2.5, 0, -5.0]) and np.allclose(extent, [2.5, 0, 5.0]) and sc.scene_extent() == 5.0
    _, ext = sc.bounds(0.0, 4.0)  # identity rotation looks down +z
    assert np.allclose(ext, [2.5, 0, 7.0])


def test_scene_loader_prefetch_seed_and_rank_slices():
    sc = _scene()
    a, b = SceneLoader(sc, 2, 7, pin=False), SceneLoader(sc, 2, 7, pin=False)
    r0, r1 = SceneLoader(sc, 1, 7, rank=0, world=2, pin=False), SceneLoader(sc, 1, 7, rank=1, world=2, pin=False)
    try:
        for _ in range(8):  # more than the 5 prefetched
            ba, bb, b0, b1 = a.next_batch(), b.next_batch(), r0.next_batch(), r1.next_batch()
            assert ba.gt_images.shape == (2, 8, 12, 3) and ba.gt_images.dtype == torch.float32
            assert torch.equal(ba.gt_images, bb.gt_images) and ba.scene_extent == 5.0 and len(ba.gt_views) == 2
            # two ranks with batch 1 see the two views one rank with batch 2 sees
            assert torch.equal(b0.gt_images[0], ba.gt_images[0]) and torch.equal(b1.gt_images[0], ba.gt_images[1])
    finally:
        for l in (a, b, r0, r1):
            l.close()


def test_scene_loader_quantized_batches_keep_bytes():
    """quantized=True hands the 8-bit pixels through unchanged (the trainer divides by 255 after the upload)."""
    import numpy as np
    import torch

    from brush_b200 import Camera
    from brush_b200.feed import Scene, SceneLoader, SceneView, image_to_tensor

    rng = np.random.default_rng(0)
    views = [SceneView(f"v{i}", Camera((0.0, 0.0, -float(i + 1)), (0.0, 0.0, 0.0, 1.0), 0.8, 0.6),
                       rng.integers(0, 256, (6, 8, 3), dtype=np.uint8)) for i in range(3)]
    loader = SceneLoader(Scene(views), 1, seed=1, pin=False, quantized=True)
    try:
        b = loader.next_batch()
    finally:
        loader.close()
    assert b.gt_images.dtype == torch.uint8 and tuple(b.gt_images.shape) == (1, 6, 8, 3)
    match = [v for v in views if np.array_equal(v.image, b.gt_images[0].numpy())]
    assert len(match) == 1
    assert torch.equal(image_to_tensor(match[0].image), b.gt_images[0].to(torch.float32) / 255.0)
