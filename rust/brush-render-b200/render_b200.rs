//! `brush_render::Backend` on the B200-native kernels (libbrush_b200.so, sm_100a) -- compiled as a module of
//! brush-render behind the `b200` feature (see README.md next to this file for why and how).
//!
//! Mirrors, one to one:
//!   crates/brush-render/src/render.rs:55-323    render_forward            -> [`render_forward_b200`]
//!   crates/brush-render/src/render.rs:325-350   impl Backend for PrimaryBackend
//!   crates/brush-render/src/render.rs:366-463   impl Backend for Autodiff<PrimaryBackend, C>
//!   crates/brush-render/src/render.rs:465-626   impl Backward<PrimaryBackend, 6> for RenderBackwards
//! Source only: never compiled in the build image (no Rust toolchain).
use std::os::raw::c_void;
use std::sync::{Mutex, OnceLock};

use brush_b200_sys::{self as sys, BrushAux, BrushUniforms, Ctx, GradPtrs, SplatPtrs};
use burn::backend::autodiff::{
    checkpoint::{base::Checkpointer, strategy::CheckpointStrategy},
    grads::Gradients,
    ops::{Backward, Ops, OpsKind},
    NodeID,
};
use burn::backend::Autodiff;
use burn::tensor::ops::{FloatTensor, FloatTensorOps, IntTensorOps};
use burn::tensor::{Tensor, TensorPrimitive};
use burn_cuda::CudaDevice;
use burn_jit::{tensor::JitTensor, JitBackend};
use cubecl::cuda::CudaRuntime;

use crate::camera::Camera;
use crate::dim_check::DimCheck;
use crate::render::{sh_coeffs_for_degree, sh_degree_from_coeffs};
use crate::shaders;
use crate::{Backend, RenderAux};

/// Burn's CUDA JIT backend with Brush's element types (the reference's is `JitBackend<WgpuRuntime, f32, i32>`,
/// lib.rs:91).
pub type CudaBackend = JitBackend<CudaRuntime, f32, i32>;

/// One library context per CUDA device (the analogue of the per-device cubecl client, render.rs:91).
pub fn ctx_for(device: &CudaDevice) -> &'static Mutex<Ctx> {
    static CTXS: OnceLock<Mutex<Vec<(usize, &'static Mutex<Ctx>)>>> = OnceLock::new();
    let mut all = CTXS.get_or_init(|| Mutex::new(Vec::new())).lock().unwrap();
    if let Some((_, c)) = all.iter().find(|(i, _)| *i == device.index) {
        return c;
    }
    let ctx = Ctx::new(device.index as i32).expect("brush_ctx_create");
    let leaked: &'static Mutex<Ctx> = Box::leak(Box::new(Mutex::new(ctx)));
    all.push((device.index, leaked));
    leaked
}

/// Raw CUDA device pointer of a Burn CUDA tensor: the cubecl handle's storage resource (cubecl-cuda
/// `CudaResource { ptr: CUdeviceptr, .. }`) already offset to the tensor's first element.  Tensors handed to the
/// operator are contiguous (`into_contiguous` below), as the reference's kernels assume (render.rs binds
/// `handle.binding()` directly).
pub fn dptr<E>(t: &JitTensor<CudaRuntime, E>) -> *mut c_void {
    let res = t.client.get_resource(t.handle.clone().binding());
    res.resource().ptr as *mut c_void
}

/// The CUDA stream cubecl enqueues this device's kernels on.  cubecl 0.3 keeps it private; until it grows an
/// accessor the binding orders itself against cubecl explicitly: flush + wait for cubecl's queue, run on the legacy
/// default stream, wait for it.  With the accessor (`feature = "cubecl-stream"`) every call is enqueue-and-return,
/// like the reference's `client.execute_unchecked`.
#[cfg(feature = "cubecl-stream")]
pub fn stream_of(t: &JitTensor<CudaRuntime, f32>) -> *mut c_void {
    t.client.cuda_stream() as *mut c_void
}
#[cfg(not(feature = "cubecl-stream"))]
pub fn stream_of(t: &JitTensor<CudaRuntime, f32>) -> *mut c_void {
    t.client.sync(); // cubecl's pending work is complete before the library touches the buffers
    std::ptr::null_mut()
}
fn after_call(ctx: &Ctx, stream: *mut c_void) {
    if stream.is_null() {
        ctx.sync_check(stream).expect("brush_b200"); // default-stream fallback: finish before cubecl continues
    }
}

fn int_empty(shape: &[usize], device: &CudaDevice) -> JitTensor<CudaRuntime, i32> {
    // create_tensor (crates/brush-kernel/src/lib.rs:125-147): uninitialised; the library zero-fills what the
    // reference's int_zeros calls zero (render.rs:124,165,241) in its own single init launch.
    CudaBackend::int_empty(shape.to_vec().into(), device)
}

/// The C view of a forward's auxiliary tensors (`RenderAux`, lib.rs:20-33) for the backward entry points; also used by
/// the trainer glue (`brush-train-b200/train_b200.rs`).
pub fn aux_of(aux: &RenderAux<CudaBackend>) -> BrushAux {
    BrushAux {
        projected_splats: dptr(&aux.projected_splats) as *mut f32,
        uniforms_buffer: dptr(&aux.uniforms_buffer) as *mut i32,
        num_intersections: dptr(&aux.num_intersections) as *mut i32,
        num_visible: dptr(&aux.num_visible) as *mut i32,
        final_index: dptr(&aux.final_index) as *mut i32,
        cum_tiles_hit: dptr(&aux.cum_tiles_hit) as *mut i32,
        tile_bins: dptr(&aux.tile_bins) as *mut i32,
        compact_gid_from_isect: dptr(&aux.compact_gid_from_isect) as *mut i32,
        global_from_compact_gid: dptr(&aux.global_from_compact_gid) as *mut i32,
        tile_id_from_isect: std::ptr::null_mut(),
        isect_capacity: aux.compact_gid_from_isect.shape.dims[0] as u64,
    }
}

/// == render_forward, crates/brush-render/src/render.rs:55-323.
pub(crate) fn render_forward_b200(
    camera: &Camera,
    img_size: glam::UVec2,
    means: JitTensor<CudaRuntime, f32>,
    log_scales: JitTensor<CudaRuntime, f32>,
    quats: JitTensor<CudaRuntime, f32>,
    sh_coeffs: JitTensor<CudaRuntime, f32>,
    raw_opacities: JitTensor<CudaRuntime, f32>,
    raster_u32: bool,
) -> (JitTensor<CudaRuntime, f32>, RenderAux<CudaBackend>) {
    let _span = tracing::trace_span!("render_forward", sync_burn = true).entered();
    let device = means.device.clone();

    // render.rs:74-79
    DimCheck::new()
        .check_dims(&means, &["D".into(), 3.into()])
        .check_dims(&log_scales, &["D".into(), 3.into()])
        .check_dims(&quats, &["D".into(), 4.into()])
        .check_dims(&sh_coeffs, &["D".into(), "C".into(), 3.into()])
        .check_dims(&raw_opacities, &["D".into()]);

    let means = burn_jit::kernel::into_contiguous(means);
    let log_scales = burn_jit::kernel::into_contiguous(log_scales);
    let quats = burn_jit::kernel::into_contiguous(quats);
    let sh_coeffs = burn_jit::kernel::into_contiguous(sh_coeffs);
    let raw_opacities = burn_jit::kernel::into_contiguous(raw_opacities);

    let tile_bounds = glam::uvec2(
        img_size.x.div_ceil(shaders::helpers::TILE_WIDTH),
        img_size.y.div_ceil(shaders::helpers::TILE_WIDTH),
    );
    let num_points = means.shape.dims[0];
    let coeffs_per_channel = sh_coeffs.shape.dims[1] as u32;
    let _ = sh_degree_from_coeffs(coeffs_per_channel); // panics "Invalid nr. of sh bases" like render.rs:51

    // render.rs:102-116.  tile_bounds, sh_degree, num_visible and total_splats are (re)computed by the library.
    let uniforms = BrushUniforms {
        viewmat: camera.world_to_local().to_cols_array(),
        focal: camera.focal(img_size).into(),
        img_size: img_size.into(),
        tile_bounds: tile_bounds.into(),
        pixel_center: camera.center(img_size).into(),
        sh_degree: 0,
        num_visible: 0,
        total_splats: num_points as u32,
        padding: 0,
    };

    // render.rs:204-206: the reference's hard-coded intersection buffer length
    let max_intersects = sys::reference_max_intersects(num_points as u32, tile_bounds.into()) as usize;
    let (h, w) = (img_size.y as usize, img_size.x as usize);

    let projected_splats = CudaBackend::float_empty([num_points, 9].into(), &device);
    let uniforms_buffer = int_empty(&[28], &device);
    let num_intersections = int_empty(&[1], &device);
    let num_visible = int_empty(&[1], &device);
    let final_index = int_empty(&[h, w], &device);
    let cum_tiles_hit = int_empty(&[num_points], &device);
    let tile_bins = int_empty(&[tile_bounds.y as usize, tile_bounds.x as usize, 2], &device);
    let compact_gid_from_isect = int_empty(&[max_intersects.max(1)], &device);
    let global_from_compact_gid = int_empty(&[num_points], &device);
    let out_img: JitTensor<CudaRuntime, f32> = if raster_u32 {
        CudaBackend::float_empty([h, w, 1].into(), &device) // packed RGBA8 in a 4-byte element (render.rs:269-280)
    } else {
        CudaBackend::float_empty([h, w, 4].into(), &device)
    };

    let aux_ffi = BrushAux {
        projected_splats: dptr(&projected_splats) as *mut f32,
        uniforms_buffer: dptr(&uniforms_buffer) as *mut i32,
        num_intersections: dptr(&num_intersections) as *mut i32,
        num_visible: dptr(&num_visible) as *mut i32,
        final_index: dptr(&final_index) as *mut i32,
        cum_tiles_hit: dptr(&cum_tiles_hit) as *mut i32,
        tile_bins: dptr(&tile_bins) as *mut i32,
        compact_gid_from_isect: dptr(&compact_gid_from_isect) as *mut i32,
        global_from_compact_gid: dptr(&global_from_compact_gid) as *mut i32,
        tile_id_from_isect: std::ptr::null_mut(),
        isect_capacity: max_intersects as u64,
    };
    let ptrs = SplatPtrs {
        means: dptr(&means) as *const f32,
        log_scales: dptr(&log_scales) as *const f32,
        quats: dptr(&quats) as *const f32,
        sh_coeffs: dptr(&sh_coeffs) as *const f32,
        raw_opacity: dptr(&raw_opacities) as *const f32,
    };
    let stream = stream_of(&means);
    let ctx = ctx_for(&device).lock().unwrap();
    unsafe {
        ctx.render_forward(
            &uniforms,
            num_points as u32,
            coeffs_per_channel,
            ptrs,
            raster_u32,
            dptr(&out_img),
            &aux_ffi,
            stream,
        )
    }
    .unwrap_or_else(|e| panic!("{e}"));
    after_call(&ctx, stream);

    (
        out_img,
        RenderAux {
            projected_splats,
            uniforms_buffer,
            num_intersections,
            num_visible,
            final_index,
            cum_tiles_hit,
            tile_bins,
            compact_gid_from_isect,
            global_from_compact_gid,
        },
    )
}

// == render.rs:325-350
impl Backend for CudaBackend {
    fn render_splats(
        camera: &Camera,
        img_size: glam::UVec2,
        means: Tensor<Self, 2>,
        _xy_dummy: Tensor<Self, 2>,
        log_scales: Tensor<Self, 2>,
        quats: Tensor<Self, 2>,
        sh_coeffs: Tensor<Self, 3>,
        raw_opacity: Tensor<Self, 1>,
        render_u32_buffer: bool,
    ) -> (Tensor<Self, 3>, RenderAux<Self>) {
        let (out_img, aux) = render_forward_b200(
            camera,
            img_size,
            means.into_primitive().tensor(),
            log_scales.into_primitive().tensor(),
            quats.into_primitive().tensor(),
            sh_coeffs.into_primitive().tensor(),
            raw_opacity.into_primitive().tensor(),
            render_u32_buffer,
        );
        (Tensor::from_primitive(TensorPrimitive::Float(out_img)), aux)
    }
}

// == render.rs:352-363
#[derive(Debug, Clone)]
struct GaussianBackwardStateB200 {
    means: NodeID,
    log_scales: NodeID,
    quats: NodeID,
    raw_opac: NodeID,
    sh_degree: u32,
    out_img: FloatTensor<CudaBackend>,
    aux: RenderAux<CudaBackend>,
}

#[derive(Debug)]
struct RenderBackwardsB200;

// == render.rs:366-463
impl<C: CheckpointStrategy> Backend for Autodiff<CudaBackend, C> {
    fn render_splats(
        camera: &Camera,
        img_size: glam::UVec2,
        means: Tensor<Self, 2>,
        xy_dummy: Tensor<Self, 2>,
        log_scales: Tensor<Self, 2>,
        quats: Tensor<Self, 2>,
        sh_coeffs: Tensor<Self, 3>,
        raw_opacity: Tensor<Self, 1>,
        render_u32_buffer: bool,
    ) -> (Tensor<Self, 3>, RenderAux<Self>) {
        use burn::backend::autodiff::ops::BasicAutodiffOps; // wrapping of the aux's float tensor, as render.rs:400-404
        let means = means.into_primitive().tensor();
        let xy_dummy = xy_dummy.into_primitive().tensor();
        let log_scales = log_scales.into_primitive().tensor();
        let quats = quats.into_primitive().tensor();
        let sh_coeffs = sh_coeffs.into_primitive().tensor();
        let raw_opacity = raw_opacity.into_primitive().tensor();

        let (out_img, aux) = render_forward_b200(
            camera,
            img_size,
            means.clone().into_primitive(),
            log_scales.clone().into_primitive(),
            quats.clone().into_primitive(),
            sh_coeffs.clone().into_primitive(),
            raw_opacity.clone().into_primitive(),
            render_u32_buffer,
        );

        let diff_proj = <burn::tensor::Float as BasicAutodiffOps<Self>>::from_inner(TensorPrimitive::Float(
            aux.projected_splats.clone(),
        ))
        .tensor();
        let auxc = aux.clone();
        let wrapped_aux = RenderAux::<Self> {
            projected_splats: diff_proj,
            uniforms_buffer: aux.uniforms_buffer,
            num_intersections: aux.num_intersections,
            num_visible: aux.num_visible,
            final_index: aux.final_index,
            cum_tiles_hit: aux.cum_tiles_hit,
            tile_bins: aux.tile_bins,
            compact_gid_from_isect: aux.compact_gid_from_isect,
            global_from_compact_gid: aux.global_from_compact_gid,
        };

        // six tracked parents, in the reference's order (render.rs:419-428)
        let prep_nodes = RenderBackwardsB200
            .prepare::<C>([
                means.clone().node,
                xy_dummy.clone().node,
                log_scales.clone().node,
                quats.clone().node,
                sh_coeffs.clone().node,
                raw_opacity.clone().node,
            ])
            .compute_bound()
            .stateful();

        let sh_degree = sh_degree_from_coeffs(sh_coeffs.primitive.shape.dims[1] as u32);

        match prep_nodes {
            OpsKind::Tracked(mut prep) => {
                let state = GaussianBackwardStateB200 {
                    means: prep.checkpoint(&means),
                    log_scales: prep.checkpoint(&log_scales),
                    quats: prep.checkpoint(&quats),
                    raw_opac: prep.checkpoint(&raw_opacity),
                    sh_degree,
                    aux: auxc,
                    out_img: out_img.clone(),
                };
                let finish = prep.finish(state, out_img);
                (Tensor::from_primitive(TensorPrimitive::Float(finish)), wrapped_aux)
            }
            OpsKind::UnTracked(prep) => (
                Tensor::from_primitive(TensorPrimitive::Float(prep.finish(out_img))),
                wrapped_aux,
            ),
        }
    }
}

// == render.rs:465-626.  The reference zero-fills nine accumulators and six outputs and dispatches
// RasterizeBackwards, GatherGrads and ProjectBackwards; here ONE call writes the six dense gradients completely
// (zeros for splats no pixel saw), so they are allocated uninitialised.
impl Backward<CudaBackend, 6> for RenderBackwardsB200 {
    type State = GaussianBackwardStateB200;

    fn backward(self, ops: Ops<Self::State, 6>, grads: &mut Gradients, checkpointer: &mut Checkpointer) {
        let _span = tracing::trace_span!("render_gaussians backwards").entered();

        let state = ops.state;
        let aux = state.aux;
        let v_output = burn_jit::kernel::into_contiguous(grads.consume::<CudaBackend>(&ops.node));
        let device = v_output.device.clone();

        let means = checkpointer.retrieve_node_output::<FloatTensor<CudaBackend>>(state.means);
        let quats = checkpointer.retrieve_node_output::<FloatTensor<CudaBackend>>(state.quats);
        let log_scales = checkpointer.retrieve_node_output::<FloatTensor<CudaBackend>>(state.log_scales);
        let raw_opac = checkpointer.retrieve_node_output::<FloatTensor<CudaBackend>>(state.raw_opac);

        let num_points = means.shape.dims[0];
        let k = sh_coeffs_for_degree(state.sh_degree) as usize;

        let v_means = CudaBackend::float_empty([num_points, 3].into(), &device);
        let v_xys_global = CudaBackend::float_empty([num_points, 2].into(), &device);
        let v_scales = CudaBackend::float_empty([num_points, 3].into(), &device);
        let v_quats = CudaBackend::float_empty([num_points, 4].into(), &device);
        let v_coeffs = CudaBackend::float_empty([num_points, k, 3].into(), &device);
        let v_opacities = CudaBackend::float_empty([num_points].into(), &device);

        let aux_ffi = aux_of(&aux);
        let ptrs = SplatPtrs {
            means: dptr(&means) as *const f32,
            log_scales: dptr(&log_scales) as *const f32,
            quats: dptr(&quats) as *const f32,
            sh_coeffs: std::ptr::null(), // not an input of the backward (render.rs checkpoints four tensors)
            raw_opacity: dptr(&raw_opac) as *const f32,
        };
        let out = GradPtrs {
            v_means: dptr(&v_means) as *mut f32,
            v_xy: dptr(&v_xys_global) as *mut f32,
            v_log_scales: dptr(&v_scales) as *mut f32,
            v_quats: dptr(&v_quats) as *mut f32,
            v_sh_coeffs: dptr(&v_coeffs) as *mut f32,
            v_raw_opacity: dptr(&v_opacities) as *mut f32,
        };
        let stream = stream_of(&v_output);
        {
            let ctx = ctx_for(&device).lock().unwrap();
            unsafe {
                ctx.render_backward(
                    num_points as u32,
                    k as u32,
                    ptrs,
                    dptr(&state.out_img) as *const f32,
                    dptr(&v_output) as *const f32,
                    &aux_ffi,
                    out,
                    stream,
                )
            }
            .unwrap_or_else(|e| panic!("{e}"));
            after_call(&ctx, stream);
        }

        // render.rs:596-624
        let [mean_parent, xys_parent, log_scales_parent, quats_parent, coeffs_parent, raw_opacity_parent] = ops.parents;
        if let Some(node) = mean_parent {
            grads.register::<CudaBackend>(node.id, v_means);
        }
        if let Some(node) = xys_parent {
            grads.register::<CudaBackend>(node.id, v_xys_global);
        }
        if let Some(node) = log_scales_parent {
            grads.register::<CudaBackend>(node.id, v_scales);
        }
        if let Some(node) = quats_parent {
            grads.register::<CudaBackend>(node.id, v_quats);
        }
        if let Some(node) = coeffs_parent {
            grads.register::<CudaBackend>(node.id, v_coeffs);
        }
        if let Some(node) = raw_opacity_parent {
            grads.register::<CudaBackend>(node.id, v_opacities);
        }
    }
}
