"""brush_b200 -- B200-native (sm_100a CUDA) implementation of Brush's differentiable splat pipeline,
behind the reference's own operator surface.  See DESIGN.md / INTEGRATION.md."""
from .render import (Camera, RenderAux, Splats, fov_to_focal, focal_to_fov, get_context, make_uniforms,  # noqa: F401
                     render_backward, render_forward, render_splats, sh_coeffs_for_degree, sh_degree_from_coeffs)
from .primitives import prefix_sum, radix_argsort  # noqa: F401

__all__ = ["Camera", "RenderAux", "Splats", "render_splats", "render_forward", "render_backward", "radix_argsort",
           "prefix_sum", "fov_to_focal", "focal_to_fov", "get_context", "make_uniforms"]
