"""Provenance of tests/golden/ (run in the build container, where /root/reference exists).

tiny_case.safetensors, basic_case.safetensors: byte-for-byte copies of the reference's own golden
vectors crates/brush-render/test_cases/{tiny_case,basic_case}.safetensors (gsplat outputs consumed by
`test_reference`, crates/brush-render/src/render.rs:695-833).  mix_case.safetensors is absent from the
reference checkout (.MISSING_LARGE_BLOBS) and cannot be regenerated without gsplat.
crab_rgb8.npy: crates/brush-render/test_cases/crab.png decoded to RGB8 [82,123,3] (the loss target).
sort_vectors / scan vectors are generated in-test from the formulas in brush-sort/src/lib.rs:164-217 and
brush-prefix-sum/src/lib.rs:110-175 (deterministic, no file needed).
"""
import shutil
import numpy as np
from PIL import Image

SRC = "/root/reference/crates/brush-render/test_cases/"
DST = __file__.rsplit("/", 1)[0] + "/"

if __name__ == "__main__":
    for n in ("tiny_case", "basic_case"):
        shutil.copyfile(SRC + n + ".safetensors", DST + n + ".safetensors")
    np.save(DST + "crab_rgb8.npy", np.asarray(Image.open(SRC + "crab.png").convert("RGB"), dtype=np.uint8))
