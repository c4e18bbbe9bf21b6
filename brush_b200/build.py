"""Builds brush_b200/lib/libbrush_b200.so (the C-ABI library, include/brush_b200.h) with nvcc for sm_100a.

Each translation unit is compiled separately because project.cu needs -fmad=false (its float math decides
integer outputs that are bit-exact against the oracle) while the rasterizer wants FMA contraction.
nvcc cross-compiles without a GPU, so this runs in the build container; the .so travels to the GPU box."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(HERE, "build")
SO = os.path.join(LIBDIR, "libbrush_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]
# translation unit -> extra flags
UNITS = {
    "project.cu": ["-fmad=false"],
    "scan.cu": [],
    "sort.cu": [],
    "raster.cu": [],
    "backward.cu": [],
    "optim.cu": [],
    "dp_adam.cu": [],
    "ssim.cu": [],
    "refine.cu": [],
    "peaks.cu": [],
    "api.cu": [],
}


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _sources():
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "brush_b200.h"))
    deps.append(os.path.abspath(__file__))
    return deps


def needs_build() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(d) > t for d in _sources())


def build(force: bool = False, verbose: bool = False, tag: str = "") -> str:
    """tag: build a VARIANT (extra flags from BRUSH_NVCC_EXTRA) to lib/libbrush_b200_<tag>.so next to the product
    library, for A/B and profiling runs (select it with BRUSH_B200_LIB); the product build has no tag."""
    global SO, OBJDIR
    if tag:
        so, objdir = os.path.join(LIBDIR, f"libbrush_b200_{tag}.so"), os.path.join(HERE, f"build_{tag}")
        saved = (SO, OBJDIR)
        SO, OBJDIR = so, objdir
        try:
            return build(force=True, verbose=verbose)
        finally:
            SO, OBJDIR = saved
    if not force and not needs_build():
        return SO
    nvcc = _nvcc()
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    units = {u: f for u, f in UNITS.items() if os.path.exists(os.path.join(CSRC, u))}

    def compile_one(item):
        unit, extra = item
        obj = os.path.join(OBJDIR, unit.replace(".cu", ".o"))
        cmd = [nvcc, *ARCH, *COMMON, *extra, *os.environ.get("BRUSH_NVCC_EXTRA", "").split(), "-c",
               os.path.join(CSRC, unit), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {unit}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(compile_one, units.items()))
    # -cudart shared: one CUDA runtime per process (the host framework's), and no runtime code inside the product .so
    cmd = [nvcc, *ARCH, "-shared", "-cudart", "shared", "-o", SO, *objs, "-Xcompiler", "-fPIC", "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return SO


if __name__ == "__main__":
    _tag = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--tag=")), "")
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, tag=_tag))
