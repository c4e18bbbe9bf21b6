"""ctypes binding of libbrush_b200.so -- the same C ABI a Rust `brush-b200-sys` crate would bind
(include/brush_b200.h).  There is NO fallback: if the library is missing or a call fails, this raises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# BRUSH_B200_LIB: another build of the SAME library (e.g. a profiling build made with BRUSH_NVCC_EXTRA); never a fallback
SO_PATH = os.environ.get("BRUSH_B200_LIB") or os.path.join(_HERE, "lib", "libbrush_b200.so")

BRUSH_OK, BRUSH_E_INVALID, BRUSH_E_CAPACITY, BRUSH_E_CUDA = 0, -1, -2, -3
STAGE_NAMES = ["ProjectSplats", "DepthSort", "ProjectVisible", "PrefixSum", "MapGaussiansToIntersect", "TileSort",
               "GetTileBinEdges", "Rasterize", "RasterizeBackwards", "GatherProjectBackwards"]


class BrushUniforms(C.Structure):
    """helpers.wgsl:7-30"""

    _fields_ = [("viewmat", C.c_float * 16), ("focal", C.c_float * 2), ("img_size", C.c_uint32 * 2),
                ("tile_bounds", C.c_uint32 * 2), ("pixel_center", C.c_float * 2), ("sh_degree", C.c_uint32),
                ("num_visible", C.c_uint32), ("total_splats", C.c_uint32), ("padding", C.c_uint32)]


class BrushAux(C.Structure):
    """lib.rs:20-33 as raw device pointers"""

    _fields_ = [("projected_splats", C.c_void_p), ("uniforms_buffer", C.c_void_p),
                ("num_intersections", C.c_void_p), ("num_visible", C.c_void_p), ("final_index", C.c_void_p),
                ("cum_tiles_hit", C.c_void_p), ("tile_bins", C.c_void_p), ("compact_gid_from_isect", C.c_void_p),
                ("global_from_compact_gid", C.c_void_p), ("tile_id_from_isect", C.c_void_p),
                ("isect_capacity", C.c_uint64)]


class BrushAdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p),
                ("count", C.c_uint64), ("lr", C.c_float), ("sh_coeffs_per_channel", C.c_uint32),
                ("sh_rest_lr_scale", C.c_float)]


class BrushTrainState(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "means", "log_scales", "rotation", "sh_coeffs", "raw_opacity", "m_means", "v_means", "m_log_scales",
        "v_log_scales", "m_rotation", "v_rotation", "m_sh_coeffs", "v_sh_coeffs", "m_raw_opacity",
        "v_raw_opacity")] + [(n, C.c_float) for n in (
            "lr_mean", "lr_scale", "lr_rotation", "lr_coeffs_dc", "lr_opac", "sh_rest_lr_scale")]


class BrushSplatParams(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("means", "rotation", "sh_coeffs", "raw_opacity", "log_scales")]


class BrushRefineConfig(C.Structure):
    _fields_ = [("densify_grad_thresh", C.c_float), ("densify_size_thresh", C.c_float),
                ("cull_alpha_thresh", C.c_float), ("cull_scale_thresh", C.c_float), ("reset_alpha", C.c_int32),
                ("reset_alpha_value", C.c_float)]


class BrushRefineCounts(C.Structure):
    _fields_ = [("num_split", C.c_uint32), ("num_cloned", C.c_uint32), ("alpha_pruned", C.c_uint32),
                ("scale_pruned", C.c_uint32), ("num_out", C.c_uint32), ("reserved", C.c_uint32 * 3)]


DP_MAX_WORLD = 8


class BrushViewExchange(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("rows", C.c_void_p * DP_MAX_WORLD),
                ("flags", C.c_void_p * DP_MAX_WORLD), ("epoch", C.c_uint32)]


class BrushDpSegment(C.Structure):
    _fields_ = [("src_offset", C.c_uint64), ("dst_offset", C.c_uint64), ("count", C.c_uint64), ("adam", C.c_int32),
                ("lr", C.c_float), ("sh_coeffs_per_channel", C.c_uint32), ("sh_rest_lr_scale", C.c_float)]


class BrushDpStep(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("src", C.c_void_p * DP_MAX_WORLD),
                ("dst", C.c_void_p * DP_MAX_WORLD), ("src_multicast", C.c_void_p), ("dst_multicast", C.c_void_p),
                ("flags", C.c_void_p * DP_MAX_WORLD), ("m", C.c_void_p), ("v", C.c_void_p),
                ("loss_offset", C.c_uint64), ("loss_out", C.c_void_p), ("epoch", C.c_uint32), ("lane", C.c_int32),
                ("ctas_per_sm", C.c_int32)]


assert C.sizeof(BrushUniforms) == 112


class BrushError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"brush_b200 error {code}: {msg}")
        self.code = code


class CapacityError(BrushError):
    pass


_lib = None

_SIGS = {
    "brush_ctx_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "brush_ctx_destroy": (None, [C.c_void_p]),
    "brush_last_error": (C.c_char_p, [C.c_void_p]),
    "brush_version": (C.c_char_p, []),
    "brush_ctx_set_sort_mode": (C.c_int, [C.c_void_p, C.c_int]),
    "brush_ctx_set_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "brush_ctx_set_sh_ready_event": (C.c_int, [C.c_void_p, C.c_void_p]),
    "brush_ctx_set_split_colors": (C.c_int, [C.c_void_p, C.c_int]),
    "brush_stage_times_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.c_int, C.c_void_p]),
    "brush_ctx_check": (C.c_int, [C.c_void_p, C.c_void_p]),
    "brush_measure_peaks": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.c_void_p]),
    "brush_ctx_set_pair_counting": (C.c_int, [C.c_void_p, C.c_int]),
    "brush_ctx_read_pair_counts": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p]),
    "brush_render_forward": (C.c_int, [C.c_void_p, C.POINTER(BrushUniforms), C.c_uint32, C.c_uint32] +
                             [C.c_void_p] * 5 + [C.c_int, C.c_void_p, C.POINTER(BrushAux), C.c_void_p]),
    "brush_render_backward": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32] + [C.c_void_p] * 6 +
                              [C.POINTER(BrushAux)] + [C.c_void_p] * 6 + [C.c_void_p]),
    "brush_render_backward_adam": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(BrushTrainState),
                                             C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(BrushAux), C.c_void_p,
                                             C.c_void_p, C.c_void_p,
                                             C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_float, C.c_void_p]),
    "brush_render_backward_rows": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                             C.POINTER(BrushAux), C.c_void_p, C.POINTER(BrushViewExchange),
                                             C.c_void_p]),
    "brush_render_backward_hybrid": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32] + [C.c_void_p] * 6 +
                                     [C.POINTER(BrushAux)] + [C.c_void_p] * 6 + [C.POINTER(BrushViewExchange), C.c_void_p]),
    "brush_sh_views_adam_step": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32] + [C.c_void_p] * 4 +
                                 [C.c_float, C.c_float, C.POINTER(BrushViewExchange), C.c_void_p, C.c_float, C.c_float,
                                  C.c_float, C.c_uint32, C.c_float, C.c_void_p]),
    "brush_views_adam_step": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(BrushTrainState), C.c_void_p,
                                        C.POINTER(BrushViewExchange), C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_float, C.c_void_p]),
    "brush_radix_argsort_u32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32,
                                          C.c_void_p, C.c_void_p, C.c_void_p]),
    "brush_radix_argsort_u64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32,
                                          C.c_void_p, C.c_void_p, C.c_void_p]),
    "brush_prefix_sum_u32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]),
    "brush_prefix_sum_u32_noalloc": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_char_p, C.c_uint32]),
    "brush_l1_loss_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                         C.c_void_p, C.c_void_p, C.c_void_p]),
    "brush_image_loss_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                            C.c_float, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "brush_adam_step": (C.c_int, [C.c_void_p, C.POINTER(BrushAdamGroup), C.c_int, C.c_float, C.c_float, C.c_float,
                                  C.c_uint32, C.c_float, C.c_void_p]),
    "brush_launch_count": (C.c_uint64, []),
    "brush_refine_plan": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(BrushSplatParams),
                                    C.POINTER(BrushSplatParams), C.POINTER(BrushRefineConfig),
                                    C.POINTER(BrushRefineCounts), C.c_void_p]),
    "brush_refine_apply": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(BrushSplatParams),
                                     C.POINTER(BrushSplatParams), C.c_void_p, C.POINTER(BrushRefineConfig),
                                     C.POINTER(BrushSplatParams), C.c_void_p]),
    "brush_dp_adam_step": (C.c_int, [C.c_void_p, C.POINTER(BrushDpStep), C.POINTER(BrushDpSegment), C.c_int,
                                     C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_float, C.c_void_p]),
    "brush_dp_owned_range": (None, [C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "brush_image_u8_to_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "brush_quat_normalize": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]),
    "brush_quat_normalize_vjp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]),
    "brush_xy_grad_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(BrushAux), C.c_uint32, C.c_uint32,
                                      C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "brush_render_forward_host": (C.c_int, [C.c_void_p, C.POINTER(BrushUniforms), C.c_uint32, C.c_uint32] +
                                  [C.c_void_p] * 5 + [C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
}


def lib():
    """Loads the CUDA library.  Raises (never falls back) if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(f"{SO_PATH} not found: run `python -m brush_b200.build` (needs nvcc). "
                               "brush_b200 has no CPU or PyTorch fallback.")
        l = C.CDLL(SO_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(l, name)  # AttributeError if the .so does not export the ABI
            fn.restype, fn.argtypes = res, args
        _lib = l
    return _lib


def exported_symbols():
    return sorted(_SIGS)


class Context:
    """One brush_ctx per (device); owns only scratch."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        rc = lib().brush_ctx_create(int(device), C.byref(self._h))
        if rc != 0:
            raise BrushError(rc, "brush_ctx_create failed (is a CUDA device present?)")
        self.device = device
        if os.environ.get("BRUSH_SPLIT_COLORS") == "1":  # A/B: every forward takes the deferred-colour route
            self.set_split_colors(True)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().brush_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def check(self, rc: int):
        if rc != 0:
            msg = lib().brush_last_error(self._h).decode()
            raise (CapacityError if rc == BRUSH_E_CAPACITY else BrushError)(rc, msg)

    def set_sort_mode(self, fused64: bool):
        self.check(lib().brush_ctx_set_sort_mode(self._h, int(fused64)))

    def set_sh_ready_event(self, event_ptr):
        """The next forward reads sh_coeffs only behind this CUDA event (a raw cudaEvent_t; None cancels)."""
        self.check(lib().brush_ctx_set_sh_ready_event(self._h, C.c_void_p(event_ptr) if event_ptr else None))

    def set_split_colors(self, enabled: bool):
        self.check(lib().brush_ctx_set_split_colors(self._h, int(enabled)))

    def set_timing(self, enabled: bool):
        self.check(lib().brush_ctx_set_timing(self._h, int(enabled)))

    def stage_times_ms(self, stream=0):
        arr = (C.c_float * len(STAGE_NAMES))()
        self.check(lib().brush_stage_times_ms(self._h, arr, len(STAGE_NAMES), C.c_void_p(stream)))
        return dict(zip(STAGE_NAMES, list(arr)))

    def measure_peaks(self, stream=0):
        """Measured by micro-kernels on this device (synchronises): FP32 TFLOP/s of scalar FFMA with constant-bank
        operands, of scalar FFMA with three register operands and of packed FFMA2, and 1e12 MUFU.EX2 per second."""
        out = (C.c_float * 4)()
        self.check(lib().brush_measure_peaks(self._h, out, C.c_void_p(stream)))
        return {"fp32_ffma_const_tflops": float(out[0]), "fp32_ffma_3reg_tflops": float(out[1]),
                "fp32_ffma2_packed_tflops": float(out[2]), "mufu_ex2_tops": float(out[3])}

    def set_pair_counting(self, enabled: bool):
        self.check(lib().brush_ctx_set_pair_counting(self._h, int(enabled)))

    def read_pair_counts(self, stream=0):
        """{forward, backward} (pixel, splat) pairs the rasterizers evaluated since the last read; synchronises."""
        arr = (C.c_uint64 * 2)()
        self.check(lib().brush_ctx_read_pair_counts(self._h, arr, C.c_void_p(stream)))
        return int(arr[0]), int(arr[1])

    def sync_check(self, stream=0):
        """Synchronises and raises CapacityError if any forward since the last check overflowed."""
        self.check(lib().brush_ctx_check(self._h, C.c_void_p(stream)))
