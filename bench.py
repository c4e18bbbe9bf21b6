#!/usr/bin/env python
"""bench.py -- fwd+bwd training iterations/s of the splat path (BASELINE.json metric).

    python bench.py [--gpus N --steps K --warmup W]            our arm: B200 kernels through the C ABI
    python bench.py --impl reference [--steps K --warmup W]    the reference's CPU path (oracle port), host cores

One "step" = one complete training iteration of brush_b200.train.SplatTrainer.step on one view per GPU:
rotation normalise -> render forward -> image loss (0.8 L1 - 0.2 SSIM, the reference default) + d loss/d image
-> render backward -> refine statistics
-> per-splat backward + normalisation VJP + Adam in one kernel.  N > 1 (--exchange views, default): each rank
publishes its view's 48-byte per-splat rows in NVLink peer memory and every rank redoes the per-splat backward of
every view and updates its own replica; --exchange peer: dense gradient sum + Adam + parameter broadcast as one
peer-memory kernel; --exchange nccl: all-reduce of the flat dense gradient, then replicated Adam.
Workload = BASELINE.json configs[3]: 1M random Gaussians, SH degree 3, 1920x1080 (synthetic, seeded;
generator in brush_b200/scene.py).  `value` times steps whose ground-truth view is already in HBM; `e2e`
times the same steps with the view's image coming from pinned host memory (H2D inside the timed region)
and the loss read back to the host every step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "fwd+bwd train it/s at 1M Gaussians 1080p"
UNIT = "it/s"
WORKLOAD = "cfg4_1m_1080p"
N_VIEWS = 16
ISECT_CAPACITY = 16 * 1024 * 1024


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def view_cameras(width, height):
    """16 training views: the reference bench camera (benches/render_bench.rs:163-174) plus small yaw steps on
    the radius-8 ring around the origin, so every view sees about the same number of splats."""
    from brush_b200 import Camera
    from brush_b200.scene import ring_camera_params

    cams = []
    for k in range(N_VIEWS):
        j = (k + 1) // 2 * (1 if k % 2 else -1)  # 0, +1, -1, +2, -2 ...
        ang_steps = 2.0 * np.pi / 0.02
        pos, rot, fx, fy = ring_camera_params(width, height, j, ang_steps)
        cams.append(Camera(pos, rot, fx, fy))
    return cams


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe).  NVML in a thread
    (a sample every 20 ms: the default timed region lasts about 0.2 s, less than `nvidia-smi -lms` needs to print its
    first line); falls back to the nvidia-smi loop when the NVML binding is missing."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
               ("sw_power_cap", 0x4))

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index
        self.nvml, self.handle, self.samples, self.run = None, None, [], False
        try:
            import pynvml

            pynvml.nvmlInit()
            try:
                import torch

                uuid = str(torch.cuda.get_device_properties(gpu_index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n, h = self.nvml, self.handle
        while self.run:
            try:
                sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
                mx = n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)
                get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
                self.samples.append((float(sm), float(mx), int(get(h))))
            except Exception:
                pass
            time.sleep(0.02)

    def start(self):
        if self.nvml is not None:
            self.run = True
            self.thread = threading.Thread(target=self._sample_nvml, daemon=True)
            self.thread.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        sm, mx, reasons = [], None, set()
        if self.nvml is not None:
            self.run = False
            self.thread.join(timeout=1.0)
            for s_, m_, bits in self.samples:
                sm.append(s_)
                mx = m_
                reasons.update(name for name, bit in self.REASONS if bits & bit)
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            time.sleep(0.15)
            self.proc.terminate()
            for r in self.rows:
                c = [x.strip() for x in r.split(",")]
                if len(c) < 7:
                    continue
                try:
                    sm.append(float(c[0]))
                    mx = float(c[1])
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        busy = [x for x in sm if mx and x > 0.5 * mx] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def algorithmic_bytes(stage, N, V, I, P, T, K):
    """algorithmic bytes per launch of each stage."""
    return {
        "ProjectSplats": 40 * N + 8 * V,
        "DepthSort": 68 * V,
        "ProjectVisible": (48 + 12 * K) * V + 40 * V,
        "PrefixSum": 8 * V,
        "MapGaussiansToIntersect": 40 * V + 8 * I,
        "TileSort": 36 * I,
        "GetTileBinEdges": 4 * I + 8 * T,
        "Rasterize": 40 * I + 20 * P,
        "RasterizeBackwards": 40 * I + 36 * P + 36 * V,
        "GatherProjectBackwards": (84 + 52 + 12 * K) * V + (52 + 12 * K) * N,
        # the kernel that runs in the single-GPU step is the FUSED tail (gather_grads + project_backwards +
        # normalisation VJP + Adam + refine statistics): it reads the 48-byte rows and reads + writes parameters and
        # both moments once, (44 + 12 K) B each, plus 9 B of statistics per splat; no dense gradient exists
        "GatherProjectBackwards+Adam (fused)": 48 * N + 6 * (44 + 12 * K) * N + 9 * N,
    }[stage]


def run_ours(args):
    import torch

    from brush_b200 import Splats, _lib, get_context, scene
    from brush_b200.parallel import view_for_rank
    from brush_b200.train import SceneBatch, SplatTrainer, TrainConfig

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    pg = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
        pg = dist.group.WORLD
    if args.gpus != world and rank == 0:
        print(f"[bench] note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)

    sc = scene.make_config_scene(args.workload, args.n_override)
    W, H, N, K = sc.width, sc.height, sc.means.shape[0], sc.sh_coeffs.shape[1]
    P, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    t = lambda a: torch.as_tensor(a, device=dev)
    splats = Splats(t(sc.means), t(sc.sh_coeffs), t(sc.quats), t(sc.raw_opacity), t(sc.log_scales))
    cams = view_cameras(W, H)
    ctx = get_context(dev)
    if args.sort_mode == "fused64":
        ctx.set_sort_mode(True)
    # Targets = the scene's own render of each view; training starts from perturbed colours and opacities.  The
    # loss gradient is dense and non-zero, but its only consistent direction is back towards the scene, so the
    # workload (visible splats, intersections, pairs) stays what BASELINE.json names for the whole run.  (Noise
    # targets drive every opacity down: the per-step cost falls by a third within 300 steps, and faster with
    # more views per step, which would flatter the multi-GPU numbers.)
    gen = torch.Generator(device=dev)
    gen.manual_seed(4242)
    gt_dev = torch.empty((N_VIEWS, H, W, 3), dtype=torch.float32, device=dev)
    with torch.no_grad():
        for v, cam in enumerate(cams):
            img, _ = splats.render(cam, (W, H), False, isect_capacity=args.capacity)
            gt_dev[v] = img[..., :3].clamp_(0.0, 1.0)
        # the data set holds 8-bit images (as the reference's does before image_to_tensor); both timed loops train on
        # the same quantised targets, the end-to-end loop uploads the bytes and converts on the device
        gt_u8 = (gt_dev * 255.0).round_().to(torch.uint8)
        gt_dev.copy_(gt_u8.to(torch.float32) / 255.0)
        splats.sh_coeffs.add_(torch.randn(splats.sh_coeffs.shape, device=dev, generator=gen) * 0.1)
        splats.raw_opacity.add_(torch.randn(splats.raw_opacity.shape, device=dev, generator=gen) * 0.3)
    gt_host = gt_u8.cpu().pin_memory()
    # measured on 8xB200 (profiles/): publishing 48-byte rows and replicating the tail wins up to 4 ranks (its
    # inbound NVLink traffic grows with world - 1), the sharded sum + Adam + multicast broadcast wins at 8
    exchange = args.exchange
    if exchange == "auto":
        exchange = "views" if world <= 4 else "peer"
    if world == 1:
        exchange = "nccl"  # single device: the fused tail, no exchange
    # reference defaults, except that refinement is off: the benchmark times a fixed number of splats
    # (max_refine_step = 0), and warmup_steps = 0 so the refine statistics (train.rs:284-316) are accumulated in every
    # timed step, as in the reference's steady state after its 500 warm-up iterations
    train_cfg = TrainConfig(max_refine_step=0, warmup_steps=0)
    trainer = SplatTrainer(N, train_cfg, dev, sh_coeffs_per_channel=K, process_group=pg,
                           isect_capacity=args.capacity, optimize=not args.no_optim,
                           exchange=exchange)
    start_params = [p.detach().clone() for p in (splats.means, splats.sh_coeffs, splats.rotation, splats.raw_opacity,
                                                 splats.log_scales)] if world > 1 else None
    splats = trainer.adopt(splats)  # peer exchange: parameters live in the peer-mapped replica
    if world > 1 and trainer.exchange == "views":
        exchange_desc = ("no gradient exchange: every rank publishes its view's nine per-splat sums (48 B rows) in NVLink "
                         "peer memory, then redoes the per-splat backward of all views + Adam on its own replica in ONE "
                         "kernel (brush_views_adam_step)")
    elif world > 1 and trainer.exchange == "hybrid":
        exchange_desc = ("hybrid exchange: the SH group travels as 16-byte colour rows (v_rgb + visibility) per splat and view, "
                         "rebuilt and Adam-stepped on every rank (brush_sh_views_adam_step, replicated SH state); the 11 "
                         "small gradient floats per splat are summed + stepped + broadcast by brush_dp_adam_step over "
                         f"NVLink peer memory ({'NVSwitch multimem.ld_reduce/st' if trainer.peer.grad_mc else 'peer loads/stores'})")
    elif world > 1 and trainer.exchange == "peer":
        exchange_desc = ("gradient sum + Adam + parameter broadcast fused in brush_dp_adam_step over NVLink peer memory "
                         f"({'NVSwitch multimem.ld_reduce/st' if trainer.peer.grad_mc else 'peer loads/stores'}); "
                         + ("two launches per step: the small groups on the step's stream, the SH group on a side "
                            "stream under the next step's projection / sorts / binning (exact parameters: the next "
                            "forward evaluates colours last and waits for the SH exchange there)"
                            if trainer.overlap_exchange else "one launch per step, not overlapped"))
    elif world > 1:
        exchange_desc = "NCCL all-reduce of the flat dense gradient, replicated Adam"
    else:
        exchange_desc = "single GPU"
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()

    def make_batch(step_idx, host_inputs):
        v = view_for_rank(step_idx, world, rank, N_VIEWS)
        src = gt_host if host_inputs else gt_dev
        return SceneBatch(src[v:v + 1], [cams[v]], 1.0)

    def one_step(step_idx, host_inputs):
        _, stats = trainer.step(make_batch(step_idx, host_inputs), splats)
        return stats

    loss_slots = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_events = [torch.cuda.Event(), torch.cuda.Event()]
    step_done = [torch.cuda.Event(), torch.cuda.Event()]
    readback = torch.cuda.Stream(dev)

    def e2e_steps(n_steps, first_idx):
        """The user-facing loop: every step's view comes from pinned host memory (copy started one step ahead on
        a side stream, as a prefetching feed does) and every step's loss is read back to the host; the host waits
        for the loss of step t only after it has launched step t+1, so the device never idles on the host."""
        dbg = os.environ.get("BRUSH_E2E_DEBUG", "")
        host_in = "nohost" not in dbg
        nxt = trainer.prefetch(make_batch(first_idx, host_in))
        for s in range(n_steps):
            _, stats = trainer.step(nxt, splats)
            if "noloss" in dbg:
                if s + 1 < n_steps:
                    nxt = trainer.prefetch(make_batch(first_idx + s + 1, host_in))
                continue
            # the device->host read of the loss runs on its own stream behind the step (a copy-engine operation
            # on the compute stream costs a bubble between two kernels: measured 0.02 ms per step); the trainer
            # keeps a step's loss scalar valid while the next step runs
            step_done[s & 1].record()
            with torch.cuda.stream(readback):
                readback.wait_event(step_done[s & 1])
                loss_slots[s & 1].copy_(stats.loss, non_blocking=True)
                loss_events[s & 1].record(readback)
            if s + 1 < n_steps:
                nxt = trainer.prefetch(make_batch(first_idx + s + 1, host_in))
            if s >= 1:
                loss_events[(s - 1) & 1].synchronize()
                loss_host.copy_(loss_slots[(s - 1) & 1])
        loss_events[(n_steps - 1) & 1].synchronize()
        loss_host.copy_(loss_slots[(n_steps - 1) & 1])

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(n_steps, host_inputs, first_idx):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = _lib.lib().brush_launch_count()
        e0.record()
        if host_inputs:
            e2e_steps(n_steps, first_idx)
        else:
            for s in range(n_steps):
                one_step(first_idx + s, False)
        trainer.finish()  # an overlapped SH exchange of the last step belongs to the timed region
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = _lib.lib().brush_launch_count() - l0
        if world > 1:
            tt = torch.tensor([ms], device=dev)
            torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
            ms = float(tt.item())
        return ms, launches

    # ---- warm-up, then the device-resident timed region ----
    for s in range(args.warmup):
        one_step(s, False)
    ctx.sync_check()  # raises if the intersection capacity was too small
    if args.profile_step:  # ncu --profile-from-start off: exactly one steady-state step is captured
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        one_step(args.warmup, False)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total, launches = timed(args.steps, False, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    # ---- the same steps end to end: view image from pinned host memory, loss read back each step ----
    e2e_steps(2, 0)
    ms_e2e, _ = timed(args.steps, True, args.warmup + args.steps)
    ms_e2e_step = ms_e2e / args.steps
    ctx.sync_check()

    # ---- per-stage device times (CUDA events around each stage, on the launching stream) ----
    # Every rank runs these steps (the step holds a collective when world > 1); only rank 0's times are reported.
    stage_ms, counts = {}, None
    ctx.set_timing(True)
    acc = {}
    reps = max(3, min(args.steps, 10))
    for s in range(reps):
        stats = one_step(1000 + s, False)
        tm = ctx.stage_times_ms(torch.cuda.current_stream().cuda_stream)
        for k_, v_ in tm.items():
            acc[k_] = acc.get(k_, 0.0) + v_ / reps
    ctx.set_timing(False)
    stage_ms = acc
    # one more step with the rasterizers counting the (pixel, splat) pairs they evaluate after culling (diagnostic
    # build of the forward kernel; outside every timed region)
    ctx.set_pair_counting(True)
    ctx.read_pair_counts(torch.cuda.current_stream().cuda_stream)
    stats = one_step(2000, False)
    pairs_eval_fwd, pairs_eval_bwd = ctx.read_pair_counts(torch.cuda.current_stream().cuda_stream)
    ctx.set_pair_counting(False)
    if rank == 0:
        counts = (stats.auxes[0].read_num_visible(), stats.auxes[0].read_num_intersections())
    check = None
    xt = None
    if world > 1:
        barrier()
        xt = exchange_timing(trainer, one_step, world, rank, dev, N)
    if world > 1:
        barrier()
        check = multi_gpu_self_check(args, trainer, splats, start_params, cams, gt_dev, N, K, dev, pg, world, rank)
        barrier()

    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return
    V, I = counts
    peak, peak_src = load_peaks()
    peaks_dev = ctx.measure_peaks(torch.cuda.current_stream().cuda_stream)
    # the FP32 roofline denominator: the best the device does with scalar FFMA (what these kernels issue)
    fp32_peak_measured = max(peaks_dev["fp32_ffma_const_tflops"], peaks_dev["fp32_ffma_3reg_tflops"])
    mufu_peak_measured = peaks_dev["mufu_ex2_tops"]
    fused_tail = world == 1 and not args.no_optim
    stages = []
    for name, ms in stage_ms.items():
        label = name
        if name == "GatherProjectBackwards" and fused_tail:
            label = "GatherProjectBackwards+Adam (fused)"
        b = algorithmic_bytes(label, N, V, I, P, T, K)
        stages.append({"stage": label, "ms": round(ms, 4), "alg_bytes": b,
                       "gbs": round(b / (ms * 1e-3) / 1e9, 1) if ms > 0 else None,
                       "frac_hbm": round(b / (ms * 1e-3) / 1e9 / peak, 4) if ms > 0 else None})
    # ---- (pixel, splat) pairs of the two compositors as the REFERENCE loops evaluate them (R_fwd, R_bwd),
    # from the last step's aux tensors; outside every timed region ----
    aux = stats.auxes[0]
    fi = aux.final_index.to(torch.int64)
    tb = aux.tile_bins.to(torch.int64)
    ty = (torch.arange(H, device=dev) // 16)[:, None].expand(H, W)
    tx = (torch.arange(W, device=dev) // 16)[None, :].expand(H, W)
    start, end = tb[ty, tx, 0], tb[ty, tx, 1]
    nonempty = end > start
    r_bwd = int(torch.where(nonempty & (fi >= start), fi - start + 1, torch.zeros_like(fi)).sum().item())
    opaque = stats.pred_images[0][..., 3] > 0.9998  # the pixel stopped early (T <= 1e-4): walked to final_index + 1
    r_fwd = int(torch.where(nonempty, torch.where(opaque, torch.minimum(fi - start + 2, end - start), end - start),
                            torch.zeros_like(fi)).sum().item())
    fp32_nominal = 148 * 128 * 2 * ((clocks or {}).get("sm_max_mhz") or 1965.0) * 1e6 / 1e12  # TFLOP/s, FMA = 2
    fp32_peak = fp32_peak_measured
    flops = {"Rasterize": 26 * r_fwd, "RasterizeBackwards": 64 * r_bwd}  # SURVEY 8d per-pair figures
    pairs_ref = {"Rasterize": r_fwd, "RasterizeBackwards": r_bwd}
    pairs_eval = {"Rasterize": pairs_eval_fwd, "RasterizeBackwards": pairs_eval_bwd}
    # executed FFMA / FMUL / FADD thread-instructions per launch from the committed ncu capture of this workload
    # (profiles/executed.json, written by profiles/ncu_executed.py); FFMA counts two flops
    executed = {}
    try:
        with open(os.path.join(ROOT, "profiles", "executed.json")) as f:
            executed = json.load(f)
    except (OSError, ValueError):
        pass
    for s_ in stages:
        nm = s_["stage"]
        if nm in flops and s_["ms"] > 0:
            sec = s_["ms"] * 1e-3
            s_["alg_flops"] = flops[nm]
            s_["tflops"] = round(flops[nm] / sec / 1e12, 2)
            s_["frac_fp32"] = round(s_["tflops"] / fp32_peak, 4)            # algorithmic: reference pairs x 26 / 64
            s_["pairs_reference"] = pairs_ref[nm]
            s_["pairs_evaluated"] = pairs_eval[nm]                           # after the kernel's culling
            s_["culled_pair_ratio"] = round(pairs_eval[nm] / max(pairs_ref[nm], 1), 4)
            ex = executed.get(nm) if executed.get("workload") == args.workload else None
            if ex:
                ex_flop = 2 * ex["ffma"] + ex["fmul"] + ex["fadd"]
                s_["executed_flops"] = ex_flop
                s_["frac_executed"] = round(ex_flop / sec / 1e12 / fp32_peak, 4)  # what the FP32 pipe really did
                if ex.get("mufu"):
                    s_["frac_mufu"] = round(ex["mufu"] / sec / 1e12 / mufu_peak_measured, 4)
    dom = max(stages, key=lambda s_: s_["ms"])
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(dom["stage"])
    except OSError:
        pass
    limits = None
    try:
        with open(os.path.join(ROOT, "profiles", "limits.json")) as f:
            limits = json.load(f).get(dom["stage"])
    except (OSError, ValueError):
        pass
    if "tflops" in dom:
        roofline = {"kernel": dom["stage"], "bound": "fp32", "achieved": dom["tflops"], "peak": round(fp32_peak, 2),
                    "unit": "TFLOP/s", "frac": dom["frac_fp32"], "frac_executed": dom.get("frac_executed"),
                    "traffic": traffic,
                    "peak_source": "measured: FFMA micro-kernel on this device in this run (brush_measure_peaks); "
                                   f"nominal 148 SMs x 128 lanes x 2 x max clock = {fp32_nominal:.1f}",
                    "mufu_peak_tops": round(mufu_peak_measured, 3), "frac_mufu": dom.get("frac_mufu"),
                    "pairs_reference": dom["pairs_reference"], "pairs_evaluated": dom["pairs_evaluated"],
                    "culled_pair_ratio": dom["culled_pair_ratio"],
                    # pipe utilisations of the kernel from the committed ncu --set full capture (profiles/limits.json):
                    # which unit binds first (for rasterize_backwards: the shared-memory data pipe, not the FP32 pipe)
                    "ncu_limits": limits,
                    # the same kernel against the HBM roofline, in the base contract's shape
                    "hbm_view": {"bound": "hbm", "achieved": dom["gbs"], "peak": peak, "unit": "GB/s",
                                 "frac": dom["frac_hbm"], "traffic": traffic, "peak_source": peak_src},
                    "note": "frac = ALGORITHMIC flops (64 bwd / 26 fwd per pair the reference loop evaluates) / kernel "
                            "time / measured FP32 peak: pairs the kernel culls before evaluating still count, so it can "
                            "exceed the hardware fraction; frac_executed = executed FFMA*2 + FMUL + FADD thread "
                            "instructions (ncu, profiles/executed.json) / kernel time / the same peak"}
    else:
        roofline = {"kernel": dom["stage"], "bound": "hbm", "achieved": dom["gbs"], "peak": peak, "unit": "GB/s",
                    "frac": dom["frac_hbm"], "traffic": traffic, "peak_source": peak_src}
    out = {
        "metric": METRIC, "value": round(world * 1000.0 / ms_step, 3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {N} random Gaussians, SH degree {sc.sh_degree}, {W}x{H}, one view per "
                               f"GPU per step ({N_VIEWS} views; targets = the scene's own renders, start = colours and opacities perturbed), loss 0.8 L1 - 0.2 SSIM "
                               f"(reference default), {'fused Adam' if not args.no_optim else 'no optimizer'}, refine "
                               f"statistics accumulated every step (warmup_steps = 0), {exchange_desc}",
                   "num_visible": V, "num_intersections": I, "views": N_VIEWS, "isect_capacity": args.capacity,
                   "sort_mode": args.sort_mode,
                   "l2": "inputs larger than L2 (236 MB of parameters + 472 MB of Adam moments per step vs 126 MB L2)"},
        "e2e": {"value": round(world * 1000.0 / ms_e2e_step, 3), "unit": UNIT, "ms_per_step": round(ms_e2e_step, 4),
                "h2d_bytes_per_step": int(gt_host[0].numel() * gt_host.element_size() + 112), "d2h_bytes_per_step": 4},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "peaks_measured": {"fp32_tflops": round(fp32_peak_measured, 2), "mufu_tops": round(mufu_peak_measured, 3),
                           "micro_kernels": {k: round(v, 3) for k, v in peaks_dev.items()},
                           "fp32_nominal_tflops": round(fp32_nominal, 1), "hbm_gbs": peak, "hbm_source": peak_src},
        "stages": stages,
    }
    if check is not None:
        out.update(check)
    if xt is not None:
        out["exchange_timing"] = xt
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline_sample(args, sc)
    print(json.dumps(out))
    if world > 1:
        torch.distributed.destroy_process_group()


def exchange_timing(trainer, one_step, world, rank, dev, N, reps=12):
    """world > 1, outside every timed region: where a multi-GPU step's extra time goes.  Every step starts from a
    barrier on all ranks; CUDA events on the step's stream bracket the exchange launch (one launch: the overlapped
    two-launch form is switched off for these steps).  A rank that arrives early waits inside the exchange kernel for
    the slowest rank's gradients, so per step the MINIMUM over ranks of the kernel's duration is the exchange itself and
    the spread of the time-to-exchange over ranks is the cost of the views being unequal work (each rank renders a
    different camera; a synchronous step lasts as long as its slowest view)."""
    import torch
    import torch.distributed as dist

    name = "_views_step" if trainer.exchange == "views" else "_peer_step"
    if trainer.exchange == "nccl" or not hasattr(trainer, name):
        return None
    orig = getattr(trainer, name)
    saved_overlap = trainer.overlap_exchange
    trainer.finish()
    trainer.overlap_exchange = False
    ev = []

    def wrapped(*a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = orig(*a, **k)
        e1.record()
        ev.append((e0, e1))
        return r

    setattr(trainer, name, wrapped)
    starts = []
    try:
        for s in range(reps + 2):
            torch.cuda.synchronize(dev)
            dist.barrier()
            torch.cuda.synchronize(dev)
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            starts.append(e)
            one_step(3000 + s, False)
        torch.cuda.synchronize(dev)
    finally:
        setattr(trainer, name, orig)
        trainer.overlap_exchange = saved_overlap
    mine = torch.tensor([[starts[i].elapsed_time(ev[i][0]), ev[i][0].elapsed_time(ev[i][1])] for i in range(2, reps + 2)],
                        device=dev)
    every = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    t = torch.stack(every).cpu()  # [rank, step, (to_exchange, exchange)]
    to_x, x = t[..., 0], t[..., 1]
    pure = float(x.min(dim=0).values.mean())
    gl, pl = trainer.layout, trainer.param_layout
    G, Pm = 4 * gl.total, 4 * pl.total
    mode = trainer.exchange
    if mode == "views":
        b_in = b_out = 48 * N * (world - 1)
    else:
        mc = bool(trainer.peer is not None and trainer.peer.grad_mc)
        if mode == "hybrid":  # SH leaves the flat exchange; 16-byte colour rows of every other view come in instead
            sh = 4 * pl.padded_size("sh")
            G, Pm = G - sh, Pm - sh
        b_out = G * (world - 1) // world + (Pm // world if mc else Pm * (world - 1) // world)
        b_in = (G // world if mc else G * (world - 1) // world) + Pm * (world - 1) // world
        if mode == "hybrid":
            b_in += 16 * N * (world - 1)
            b_out += 16 * N * (world - 1)
    return {
        "mode": mode, "steps": reps, "launches_timed": name,
        "to_exchange_ms_per_rank": [round(float(v), 4) for v in to_x.mean(dim=1)],
        "to_exchange_ms_mean": round(float(to_x.mean()), 4),
        "to_exchange_ms_slowest_rank_per_step": round(float(to_x.max(dim=0).values.mean()), 4),
        "imbalance_ms": round(float((to_x.max(dim=0).values - to_x.mean(dim=0)).mean()), 4),
        "exchange_kernel_ms_per_rank": [round(float(v), 4) for v in x.mean(dim=1)],
        "exchange_ms": round(pure, 4),
        "nvlink_bytes_in_per_rank": int(b_in), "nvlink_bytes_out_per_rank": int(b_out),
        "gbs_in": round(b_in / (pure * 1e-3) / 1e9, 1), "gbs_out": round(b_out / (pure * 1e-3) / 1e9, 1),
        "note": "exchange_ms = mean over steps of the minimum over ranks of the exchange kernel's duration (the last "
                "rank to arrive does not wait); imbalance_ms = mean over steps of (slowest rank's time to the exchange - "
                "mean over ranks): the part of a synchronous step that no exchange kernel can remove",
    }


def _checksum64(t):
    """Two order-sensitive 64-bit sums of a tensor's bit pattern (device side; wraps mod 2^64)."""
    import torch

    bits = t.detach().contiguous().view(torch.int32).reshape(-1).to(torch.int64) & 0xFFFFFFFF
    idx = torch.arange(bits.numel(), device=bits.device, dtype=torch.int64) % 65521 + 1
    return torch.stack([bits.sum(), (bits * idx).sum()])


def multi_gpu_self_check(args, trainer, splats, start_params, cams, gt_dev, N, K, dev, pg, world, rank):
    """world > 1, outside every timed region.  (a) The replicas the timed run left behind must be bit-identical on
    every rank (64-bit checksums of every parameter group, all-gathered).  (b) The exchange mode under test and the
    plain NCCL all-reduce + replicated Adam are each run for the same few steps from the same start (the parameters
    the bench started from) on fresh trainers; the maximum relative parameter difference, the fraction of parameters
    within 2e-5 relative + 2e-6 absolute (the tolerance of tests/test_gpu_peer.py: Adam normalises the step, so a
    gradient within atomics-order noise of zero may flip its sign) and the loss difference are reported.  The
    reference is single-device (train.rs:216-219 asserts one view), so the NCCL path -- the mean over the batch of
    train.rs:229-268 done with a library collective -- is the oracle of the exchange kernels."""
    import torch
    import torch.distributed as dist

    from brush_b200 import Splats
    from brush_b200.parallel import view_for_rank
    from brush_b200.train import SceneBatch, SplatTrainer, TrainConfig

    names = ("means", "sh_coeffs", "rotation", "raw_opacity", "log_scales")
    mine = torch.stack([_checksum64(p) for p in (splats.means, splats.sh_coeffs, splats.rotation, splats.raw_opacity,
                                                 splats.log_scales)])
    every = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(every, mine, group=pg)
    identical = all(bool(torch.equal(every[0], e)) for e in every)
    mode = trainer.exchange
    steps = max(2, min(args.steps, 6))
    results = {}
    for m in (mode, "nccl"):
        sp = Splats(*[p.clone() for p in start_params])
        tr = SplatTrainer(N, TrainConfig(max_refine_step=0, warmup_steps=0), dev, sh_coeffs_per_channel=K,
                          process_group=pg, isect_capacity=args.capacity, exchange=m)
        sp = tr.adopt(sp)
        loss = None
        for s in range(steps):
            v = view_for_rank(s, world, rank, N_VIEWS)
            sp, st = tr.step(SceneBatch(gt_dev[v:v + 1], [cams[v]], 1.0), sp)
            loss = st.loss
        torch.cuda.synchronize(dev)
        results[m] = ([p.detach().clone() for p in sp.params()], float(loss),
                      tr.grad_2d_accum.clone(), tr.xy_grad_counts.clone())
        del tr, sp
    a, b = results[mode], results["nccl"]
    max_rel, close_frac, n_all = 0.0, 0.0, 0
    per_group = {}
    for name, pa, pb in zip(names, a[0], b[0]):
        d = (pa - pb).abs()
        rel = d / pb.abs().clamp_min(1e-3)
        close = d <= 2e-5 * pb.abs() + 2e-6
        per_group[name] = {"max_rel": float(rel.max()), "frac_close": float(close.float().mean())}
        max_rel = max(max_rel, float(rel.max()))
        close_frac += float(close.sum())
        n_all += pa.numel()
    mine2 = torch.stack([_checksum64(p) for p in a[0]])
    every2 = [torch.zeros_like(mine2) for _ in range(world)]
    dist.all_gather(every2, mine2, group=pg)
    identical2 = all(bool(torch.equal(every2[0], e)) for e in every2)
    out = {
        "replicas_identical": bool(identical and identical2),
        "exchange_check": {
            "mode": mode, "against": "nccl all-reduce + replicated Adam", "steps": steps,
            "replicas_identical_after_timed_run": bool(identical),
            "replicas_identical_after_check_run": bool(identical2),
            "max_rel_param_diff": max_rel, "frac_params_close": close_frac / max(n_all, 1),
            "loss": a[1], "loss_nccl": b[1], "loss_abs_diff": abs(a[1] - b[1]),
            "visible_counts_equal": bool(torch.equal(a[3], b[3])),
            "grad_norm_stat_frac_close": float(torch.isclose(a[2], b[2], rtol=1e-4, atol=1e-9).float().mean()),
            "per_group": per_group,
        },
    }
    return out


def _oracle_iteration(O, u, sc, v_out, capacity):
    """One fwd+bwd of the reference's CPU path: render.rs:55-323 then render.rs:468-625 (oracle restatement)."""
    img, aux = O.render_forward(u, sc.means, sc.log_scales, sc.quats, sc.sh_coeffs, sc.raw_opacity,
                                capacity=capacity)
    O.render_backward(aux, sc.means, sc.log_scales, sc.quats, sc.raw_opacity, sc.sh_coeffs.shape[1], img, v_out)
    return aux


def _cpu_threads():
    """Every host core this process may run on -- NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1 to
    its workers, which would silently turn the reference arm into a one-thread run at N > 1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def _oracle_setup(args, sc):
    from oracle import pyoracle as O  # the ONLY place bench.py touches oracle/: the CPU baseline legs

    O.build()
    O.use_timed()  # -O3 -march=native build made on THIS machine (oracle/Makefile `timed`)
    threads = _cpu_threads()
    O.set_threads(threads)  # num_threads(...) clause on every parallel region: overrides OMP_NUM_THREADS
    W, H = sc.width, sc.height
    vm = np.eye(4, dtype=np.float32)
    vm[2, 3] = 8.0  # bench camera (0,0,-8), identity (benches/render_bench.rs:163-174)
    u = O.make_uniforms(vm, (0.5 * W, 0.5 * W), (W, H), (0.5 * W, 0.5 * H))
    v_out = np.full((H, W, 4), 1.0 / (4 * H * W), np.float32)  # loss = mean(out) (render_bench.rs:180)
    return O, u, v_out, threads


def _timed_samples(fn, n_samples, inner):
    """divan's shape (benches/render_bench.rs:29-30,199): n_samples samples of `inner` back-to-back iterations;
    returns the per-iteration seconds of every sample."""
    out = []
    for _ in range(n_samples):
        t0 = time.perf_counter()
        for _ in range(inner):
            fn()
        out.append((time.perf_counter() - t0) / inner)
    return out


def cpu_baseline_sample(args, sc):
    O, u, v_out, threads = _oracle_setup(args, sc)
    t0 = time.perf_counter()
    _oracle_iteration(O, u, sc, v_out, args.capacity)  # warm (page-in, allocator) + probe
    probe = time.perf_counter() - t0
    inner = 4 if probe < 0.5 else 1               # small configs: the reference's 4 inner iterations per sample
    n_samples = 5 if probe * 5 * inner <= 40.0 else max(1, int(40.0 / (probe * inner)))
    per_it = _timed_samples(lambda: _oracle_iteration(O, u, sc, v_out, args.capacity), n_samples, inner)
    dt = statistics.median(per_it)
    return {"value": round(1.0 / dt, 4), "unit": UNIT, "cores": threads, "kind": "port",
            "stage_ms": {k: round(v * 1e3, 2) for k, v in O.stage_seconds().items()},
            "sample": f"median of {n_samples} sample(s) x {inner} full fwd+bwd iteration(s) of the whole workload "
                      f"({sc.means.shape[0]} Gaussians, {sc.width}x{sc.height}, bench camera, v_out = 1/(4P)) by the "
                      f"timed build of the C++ oracle port (-O3 -march=native, thread-private f32 accumulators) with "
                      f"OpenMP on {threads} threads; {dt:.3f} s per iteration (min {min(per_it):.3f}, max {max(per_it):.3f})"}


def run_reference(args):
    """The reference arm: the reference's own implementation of the path is WGSL through wgpu, which cannot run on
    this box (no Rust toolchain, no graphics stack), so this times the oracle port of it on all host cores.  Always
    the WHOLE workload: if K + W full iterations do not fit the time budget, fewer iterations are timed (and the
    line says so); the workload is never made smaller."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from brush_b200 import scene

    sc = scene.make_config_scene(args.workload, args.n_override)
    O, u, v_out, threads = _oracle_setup(args, sc)
    t0 = time.perf_counter()
    aux = _oracle_iteration(O, u, sc, v_out, args.capacity)  # untimed: page-in + the probe that sizes the run
    probe = time.perf_counter() - t0
    budget = float(os.environ.get("BRUSH_REF_BUDGET_S", "240"))
    affordable = max(1, int(budget / max(probe, 1e-6)))
    warm = max(0, min(args.warmup - 1, affordable - 1))  # the probe was the first warm-up iteration
    steps = max(1, min(args.steps, affordable - warm))
    for _ in range(warm):
        _oracle_iteration(O, u, sc, v_out, args.capacity)
    per_it = []
    for _ in range(steps):
        t0 = time.perf_counter()
        aux = _oracle_iteration(O, u, sc, v_out, args.capacity)
        per_it.append(time.perf_counter() - t0)
    dt = sum(per_it) / len(per_it)
    val = round(1.0 / dt, 4)
    N, K = sc.means.shape[0], sc.sh_coeffs.shape[1]
    reduced = "" if steps == args.steps else (f"; only {steps} of the requested {args.steps} steps were timed to stay "
                                              f"within {budget:.0f} s (the workload itself is never reduced)")
    out = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "steps_timed": steps, "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 2),
        "ms_per_step_median": round(statistics.median(per_it) * 1e3, 2), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {N} random Gaussians, SH degree {sc.sh_degree}, {sc.width}x"
                               f"{sc.height}, fwd+bwd of one view per step on the host CPU (no optimizer; the "
                               "reference bench shape, benches/render_bench.rs:176-184)",
                   "num_visible": aux.num_visible, "num_intersections": aux.num_intersections},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": (f"every step = one full fwd+bwd of the whole workload by the timed build of the C++ "
                                    f"oracle port (-O3 -march=native, OpenMP, {threads} threads = every core this "
                                    f"process may use); the reference's wgpu path is unavailable on this box" + reduced)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "stage_ms": {k: round(v * 1e3, 2) for k, v in O.stage_seconds().items()},
    }
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--n-override", type=int, default=None, help="debug: fewer Gaussians (not a bench number)")
    ap.add_argument("--capacity", type=int, default=ISECT_CAPACITY)
    ap.add_argument("--no-optim", action="store_true", help="skip Adam (pure fwd+bwd, the reference bench shape)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sort-mode", choices=["two-stage", "fused64"], default="two-stage",
                    help="tile sort: depth sort then stable 32-bit tile-id sort (default), or ONE sort on the fused "
                         "64-bit (tile | depth) key as render.rs:227-237 names it; identical output (A/B)")
    ap.add_argument("--profile-step", action="store_true",
                    help="run the warm-up, then ONE step between cudaProfilerStart/Stop and exit (for ncu "
                         "--profile-from-start off); prints no bench line")
    ap.add_argument("--exchange", choices=["auto", "views", "peer", "hybrid", "nccl"], default="auto",
                    help="multi-GPU step: per-view rows in peer memory + replicated fused tail (default), dense "
                         "gradient sum fused with Adam over peer memory, or NCCL all-reduce")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
