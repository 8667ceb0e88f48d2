#!/usr/bin/env python
"""bench.py -- 2DGS surfel rasterizer fwd+bwd view-iterations/s on the Mip-NeRF-360-bicycle-shaped
synthetic scene (BASELINE.json config C4: 5 M surfels, 1237x822, SH degree 3).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

A "step" is one forward + backward pass of the hot path over one camera per GPU (view-parallel, weak
scaling: every rank rasterises its own view of the replicated surfel set, then the per-surfel
gradients are sum-all-reduced over NCCL).  Prints ONE JSON line on rank 0.

  value     device-resident throughput (inputs in HBM when the clock starts), CUDA events, max over ranks
  e2e       same metric through the public GaussianRasterizer API with HOST buffers: per step the
            camera matrices and the ground-truth image are copied from pinned host memory, the loss
            is read back (train.py:104-152 shape)
  roofline  dominant kernel, timed live with cudaEvent pairs on the launching stream (gsm_stage_timing_*)
  cpu_baseline  the CPU oracle (oracle/surfel_oracle.py) on a bounded sample, rank 0 at N=1 only

--impl reference times the UNMODIFIED reference on the same workload: its CUDA build (oracle/_ref,
the reference has no CPU implementation of this path) when present, else the CPU oracle port.
Nothing here reads /root/reference at run time.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "2dgs_fwd_bwd_view_iterations_per_sec"
UNIT = "iter/s"
N_VIEWS = 8           # cameras on a circle (C5: 8 views/step); rank r, step s uses view (s*N + r) % 8
WORKLOAD = "C4"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--views-per-step", type=int, default=None,
                    help="views in the step's global batch (default: one per GPU = weak scaling; BASELINE.json configs[4] is "
                         "--workload C5 --views-per-step 8: a FIXED batch of 8 views shared by the N GPUs = strong scaling)")
    ap.add_argument("--P", type=int, default=None, help="override the surfel count (debug only; flagged in config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-knn", action="store_true")
    ap.add_argument("--no-render", action="store_true", help="skip the render-FPS leg (forward only, workload size and 4K)")
    ap.add_argument("--no-drop-in", action="store_true", help="skip the leg that runs the reference's unchanged render()")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi during the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        # 200 ms between polls, as in the profiling recipe: every NVML poll of power and event reasons holds up work
        # submission on the sampled GPU for a few ms.  At 20 ms the device-resident loop measured 0.07-0.15 ms per step
        # slower than the same loop without a sampler (outliers of several ms in a 20-step region); at 100 ms a 20-step
        # region still carried 2-4 ms of stall in total
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:
            pass
        try:
            rows = [r.strip().split(",") for r in open(self.path) if r.strip()]
            os.unlink(self.path)
        except Exception:
            return out
        sm, reasons, mx, pw = [], set(), None, []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                pw.append(float(r[2]))
                for n, v in zip(names, r[3:7]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm),
                       power_w_max=max(pw) if pw else None)
        return out


# ------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------
def make_workload(args, device):
    from gaussmart_b200 import synthetic
    P_cfg, W, H, deg, _ = synthetic.CONFIGS[args.workload]
    P = P_cfg if args.P is None else args.P
    idx = list(synthetic.CONFIGS).index(args.workload) + 1
    scene = synthetic.make_scene(P, W, deg, synthetic.BASE_SEED + idx)   # same seed on every rank: replicas
    cams = synthetic.make_cameras(W, H, N_VIEWS)
    params = dict(means3D=scene.means3D.to(device), shs=scene.shs.to(device), opacities=scene.opacities.to(device),
                  scales=scene.scales.to(device), rotations=scene.rotations.to(device))
    cams_d = [synthetic.Camera(*[x.to(device) if isinstance(x, torch.Tensor) else x for x in c]) for c in cams]
    g = torch.Generator(device="cpu").manual_seed(4711)
    gt_host = [torch.rand(3, H, W, generator=g).pin_memory() for _ in range(2)]
    cot = synthetic.make_cotangents(W, H, 777)
    cot_d = (cot[0].to(device), cot[1].to(device))
    return dict(P=P, W=W, H=H, deg=deg, scene=scene, cams=cams, cams_d=cams_d, params=params, gt_host=gt_host,
                cot=cot, cot_d=cot_d)


LAMBDA_DSSIM = 0.2  # arguments/__init__.py:86 (OptimizationParams.lambda_dssim)
DEPTH_RATIO, LAMBDA_NORMAL, LAMBDA_DIST = 0.0, 0.05, 100.0  # arguments/__init__.py:70,88; lambda_dist: 2DGS's setting for
#                                                             bounded scenes (the reference's default 0.0 leaves that branch dead)


def loss_fn(color, allmap, gt, cam, photometric, regularizer):
    """train.py:113-143: (1 - lambda) L1 + lambda (1 - SSIM) on the colour (train.py:113-114) + normal-consistency and
    distortion regularisers on the allmap (gaussian_renderer/__init__.py:117-147 post-processing, utils/point_utils.py
    depth_to_normal, train.py:131-140).  `photometric` / `regularizer` are each arm's own implementation: the
    reference's torch ops for the reference arm, the fused gsm_photometric_loss / gsm_surface_regularizer for ours."""
    return photometric(color, gt, LAMBDA_DSSIM) + regularizer(allmap, cam, DEPTH_RATIO, LAMBDA_NORMAL, LAMBDA_DIST)


class OursArm:
    """gaussmart_b200: bucket path for `value` (K6 writes the flat gradient bucket, NCCL all-reduce at N>1),
    public autograd API for `e2e`."""
    name = "ours"

    def __init__(self, wl, device, world, views_per_step=None):
        import gaussmart_b200 as gb
        from gaussmart_b200 import view_parallel as vp
        from gaussmart_b200.loss import photometric_loss
        self.gb, self.vp, self.wl, self.device, self.world = gb, vp, wl, device, world
        self.V = views_per_step or world
        local = (self.V + world - 1) // world
        self.photometric = photometric_loss
        from gaussmart_b200.regularizer import surface_regularizer
        self.regularizer = surface_regularizer
        self.rank = int(os.environ.get("RANK", "0"))
        M = wl["params"]["shs"].shape[1]
        self.bucket = vp.GradBucket(wl["P"], M, device, with_stats=True) if world == 1 else None
        # N > 1: factored exchange (all-reduce 40 B/surfel + all-gather dL/dRGB 12 B/surfel/view + local SH expand)
        self.ex, self.peer = None, False
        if world > 1:
            # the fused exchange over NVLink peer memory (K6 pushes dL/dRGB into every rank's slot, one two-shot kernel
            # reduces the rest); NCCL collectives (FactoredExchange) where symmetric memory is not available
            if vp.peer_exchange_available():
                try:
                    self.ex, self.peer = vp.PeerExchange(wl["P"], M, device, None, local, with_stats=True), True
                except Exception as ex_:  # pragma: no cover
                    print(f"[bench] peer exchange unavailable ({ex_!r}); falling back to NCCL collectives", file=sys.stderr)
            if self.ex is None:
                self.ex = vp.FactoredExchange(wl["P"], M, device, world, local, with_stats=True)
        self.bg = torch.zeros(3, device=device)
        self.leaves = {k: v.clone().requires_grad_(True) for k, v in wl["params"].items()}
        self.means2D = torch.zeros_like(self.leaves["means3D"], requires_grad=True)
        self.leaves_data = {k: v.detach() for k, v in self.leaves.items()}

    def value_step(self, step):
        """One step = the fwd+bwd of the step's V views (V = one per GPU by default): views (step*V + i) mod 8."""
        wl = self.wl
        from gaussmart_b200.rasterizer import rasterize_into_bucket
        views = [(step * self.V + i) % N_VIEWS for i in range(self.V)]
        if self.world == 1:
            for j, v in enumerate(views):   # the views of the step accumulate in the bucket
                color, radii, allmap, backward = rasterize_into_bucket(wl["params"], wl["cams_d"][v], self.bucket, j > 0,
                                                                       bg=self.bg)
                backward(wl["cot_d"][0], wl["cot_d"][1])
            return
        # rank r renders views {r, r+N, ...} of the batch; every rank knows all cameras of the step
        cams = [wl["cams_d"][v] for v in views]
        self._step(wl["params"], cams, lambda v, c, a: wl["cot_d"])

    def _step(self, params, cams, cot):
        from gaussmart_b200.rasterizer import rasterize_into_bucket
        if self.peer:
            self.vp.view_parallel_step_peer(
                params, cams, self.ex, cot,
                raster_fn=lambda p_, c_, b_, a_, rgb_peers=None: rasterize_into_bucket(p_, c_, b_, a_, bg=self.bg, rgb_peers=rgb_peers,
                                                                                       fused_stats=True))
            return
        self.vp.view_parallel_step_factored(
            params, cams, self.ex, cot,
            raster_fn=lambda p_, c_, b_, a_, rgb_out=None: rasterize_into_bucket(p_, c_, b_, a_, bg=self.bg, rgb_out=rgb_out,
                                                                                fused_stats=True),
            fused_stats=True)

    def api_step(self, cam_d, gt, cam_h=None, gt_ready=None):
        """Public API: GaussianRasterizationSettings / GaussianRasterizer + autograd (gr.py:37-97 shape)."""
        from gaussmart_b200 import synthetic
        gb = self.gb
        S = gb.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, self.wl["deg"], self.bg))
        for v in self.leaves.values():
            v.grad = None
        self.means2D.grad = None
        L = self.leaves
        color, radii, allmap = gb.GaussianRasterizer(S)(
            means3D=L["means3D"], means2D=self.means2D, opacities=L["opacities"], shs=L["shs"], scales=L["scales"],
            rotations=L["rotations"])
        if gt_ready is not None:
            gt_ready()   # the ground-truth upload runs on a copy stream under the render
        loss = loss_fn(color, allmap, gt, cam_d if cam_h is None else cam_h, self.photometric, self.regularizer)
        loss.backward()
        return loss

    def raw_step(self, raw, m2d, cam_d, gt, cam_h, gt_ready):
        """render() level (gaussian_renderer/__init__.py:60-106) from the STORED parameters: the raw-parameter entry, the
        getters of gaussian_model.py:104-123 fused into the preprocess kernels."""
        from gaussmart_b200 import synthetic
        from gaussmart_b200.rasterizer import rasterize_gaussians_raw
        S = self.gb.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, self.wl["deg"], self.bg))
        for v in raw.values():
            v.grad = None
        m2d.grad = None
        color, radii, allmap = rasterize_gaussians_raw(raw["xyz"], m2d, raw["f_dc"], raw["f_rest"], raw["opacity"],
                                                       raw["scaling"], raw["rotation"], S)
        gt_ready()
        loss = loss_fn(color, allmap, gt, cam_h, self.photometric, self.regularizer)
        loss.backward()
        self.last_radii = radii
        return loss

    def render(self, cam_d):
        """Inference render (render.py / gr.py:37-97 under torch.no_grad): forward only through the public API."""
        from gaussmart_b200 import synthetic
        S = self.gb.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, self.wl["deg"], self.bg))
        L = self.leaves_data
        with torch.no_grad():
            return self.gb.GaussianRasterizer(S)(
                means3D=L["means3D"], means2D=self.means2D, opacities=L["opacities"], shs=L["shs"], scales=L["scales"],
                rotations=L["rotations"])

    def api_step_view_parallel(self, cams_step, gt, cam_h=None, gt_ready=None):
        """N > 1: the public multi-GPU call, view_parallel_step_factored, with the loss of train.py's shape
        differentiated w.r.t. the rendered maps inside cotangent_fn (same work as api_step per rank)."""
        from gaussmart_b200.rasterizer import rasterize_into_bucket
        box = {}

        def cot(v, color, allmap):
            c = color.detach().requires_grad_(True)
            a = allmap.detach().requires_grad_(True)
            if gt_ready is not None:
                gt_ready()
            loss = loss_fn(c, a, gt, cams_step[self.rank] if cam_h is None else cam_h, self.photometric, self.regularizer)
            loss.backward()
            box["loss"] = loss.detach()
            return c.grad, a.grad

        self._step(self.leaves_data, cams_step, cot)
        return box["loss"]


class ReferenceArm:
    """The UNMODIFIED reference CUDA build (oracle/_ref) through its own public API and stock code path."""
    name = "reference"

    def __init__(self, wl, device, world, views_per_step=None):
        import ref_loader
        self.V = views_per_step or 1
        self.ref = ref_loader.load_dsr()
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import loss_oracle  # the reference's own torch loss (utils/loss_utils.py restated; pinned by tests/golden/loss_ref_*)
        import regularizer_oracle  # render()'s allmap post-processing (restated: it sits inside render()) + train.py:131-140
        import ref_callers
        self.modules = "restated (oracle/loss_oracle.py, oracle/regularizer_oracle.py)"
        self.photometric = loss_oracle.photometric_loss
        self.regularizer = regularizer_oracle.surface_regularizer
        if ref_callers.available():
            # the reference's OWN modules (install copy oracle/_ref/gaussmart/utils/{loss_utils,point_utils}.py)
            C = ref_callers.load("reference")
            l1_loss, ssim, d2n = C.loss_utils.l1_loss, C.loss_utils.ssim, C.point_utils.depth_to_normal
            self.photometric = lambda x, gt, lam: (1.0 - lam) * l1_loss(x, gt) + lam * (1.0 - ssim(x, gt))   # train.py:113-114

            def regularizer(allmap, view, depth_ratio, lambda_normal, lambda_dist):
                n, d, _ = regularizer_oracle.regularizer_terms(allmap, view, depth_ratio, lambda_normal, lambda_dist, normal_fn=d2n)
                return n + d
            self.regularizer = regularizer
            self.modules = ("the reference's own utils/loss_utils.py (l1_loss, ssim) and utils/point_utils.py (depth_to_normal) "
                            "from the install copy oracle/_ref/gaussmart; render()'s 12 glue lines around them restated")
        self.wl, self.device, self.world = wl, device, 1
        self.bg = torch.zeros(3, device=device)
        self.leaves = {k: v.clone().requires_grad_(True) for k, v in wl["params"].items()}
        self.means2D = torch.zeros_like(self.leaves["means3D"], requires_grad=True)

    def _render(self, cam_d):
        from gaussmart_b200 import synthetic
        ref = self.ref
        S = ref.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, self.wl["deg"], self.bg))
        for v in self.leaves.values():
            v.grad = None
        self.means2D.grad = None
        L = self.leaves
        return ref.GaussianRasterizer(S)(
            means3D=L["means3D"], means2D=self.means2D, opacities=L["opacities"], shs=L["shs"], scales=L["scales"],
            rotations=L["rotations"])

    def render(self, cam_d):
        with torch.no_grad():
            return self._render(cam_d)

    def raw_step(self, raw, m2d, cam_d, gt, cam_h, gt_ready):
        """render() level from the STORED parameters, the reference's way: the getters of gaussian_model.py:104-123 as
        torch ops (exp, F.normalize, sigmoid, torch.cat) in front of its rasterizer, autograd behind."""
        from gaussmart_b200 import synthetic
        ref = self.ref
        S = ref.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, self.wl["deg"], self.bg))
        for v in raw.values():
            v.grad = None
        m2d.grad = None
        color, radii, allmap = ref.GaussianRasterizer(S)(
            means3D=raw["xyz"], means2D=m2d, shs=torch.cat((raw["f_dc"], raw["f_rest"]), dim=1),
            opacities=torch.sigmoid(raw["opacity"]), scales=torch.exp(raw["scaling"]),
            rotations=torch.nn.functional.normalize(raw["rotation"]))
        gt_ready()
        loss = loss_fn(color, allmap, gt, cam_d, self.photometric, self.regularizer)
        loss.backward()
        self.last_radii = radii
        return loss

    def value_step(self, step):
        for i in range(self.V):   # the reference is single-GPU: the step's V views one after the other
            color, radii, allmap = self._render(self.wl["cams_d"][(step * self.V + i) % N_VIEWS])
            torch.autograd.backward([color, allmap], [self.wl["cot_d"][0], self.wl["cot_d"][1]])

    def api_step(self, cam_d, gt, cam_h=None, gt_ready=None):
        color, radii, allmap = self._render(cam_d)
        if gt_ready is not None:
            gt_ready()
        loss = loss_fn(color, allmap, gt, cam_d, self.photometric, self.regularizer)
        loss.backward()
        return loss


# ------------------------------------------------------------------------------------------------
# timing helpers
# ------------------------------------------------------------------------------------------------
def barrier(world):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed_loop(fn, steps, warmup, world, rank):
    """W untimed + K timed steps, barrier + synchronize on both sides, CUDA events on the launching
    stream, MAX over ranks.  Returns total ms of the K steps."""
    for s in range(warmup):
        fn(s)
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(steps):
        fn(warmup + s)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    barrier(world)
    return ms.item()


def peaks():
    p = dict(hbm_gbs=6650.0, source="fallback (B200_PROFILING.md)", sm_max_mhz=1965.0)
    try:
        j = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        p.update(hbm_gbs=float(j["hbm_gbs"]), source="measured (MEASURED_PEAKS.json)",
                 sm_max_mhz=float(j.get("sm_max_mhz", 1965.0)))
    except Exception:
        pass
    p["fp32_tflops"] = 148 * 128 * 2 * p["sm_max_mhz"] * 1e6 / 1e12   # SIMT FP32 FMA peak at max clock
    p["fp32_source"] = "derived: 148 SMs x 128 FP32 lanes x 2 x sm_max_mhz (no tensor-core work on this path)"
    return p


def measure_fp32_peak(device):
    """MEASURED SIMT FP32 FMA rate (gsm_fp32_peak_probe: 64 resident warps per SM, 8 independent register-resident FMA
    chains per thread), best of 5 launches of ~2 ms, CUDA events: the denominator of the compositing kernels' roofline
    (MEASURED_PEAKS.json has HBM and tensor-core peaks only)."""
    import ctypes
    from gaussmart_b200 import _lib
    lib = _lib.load()
    sms = torch.cuda.get_device_properties(device).multi_processor_count
    out = torch.empty(sms * 4 * 512, device=device)
    flops = ctypes.c_double(0.0)
    stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    best = 0.0
    for rep in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.gsm_fp32_peak_probe(20000, ctypes.c_void_p(out.data_ptr()), out.numel(), ctypes.byref(flops), stream),
                   "fp32_peak_probe")
        e1.record()
        torch.cuda.synchronize()
        if rep:
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


def work_counts(wl, view, device):
    """Realised P_vis, R, E (list entries evaluated), C (contributing pairs) of one view (un-timed)."""
    from gaussmart_b200.rasterizer import _forward_impl
    cam = wl["cams_d"][view]
    e = torch.empty(0, device=device)
    p = wl["params"]
    ctr = torch.zeros(4, dtype=torch.int64, device=device)
    out = _forward_impl(torch.zeros(3, device=device), p["means3D"], e, p["opacities"], p["scales"], p["rotations"],
                        1.0, e, cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy,
                        cam.image_height, cam.image_width, p["shs"], wl["deg"], cam.camera_center, False, False,
                        work_counters=ctr)
    from gaussmart_b200 import _lib
    lib = _lib.load()
    W, H = wl["W"], wl["H"]
    no = lib.gsm_image_ncontrib_offset(W, H)
    ncontrib = out[6][no:no + 4 * W * H].view(torch.int32).long().sum().item()
    c = ctr.tolist()
    return dict(P_vis=int((out[3] > 0).sum().item()), R=int(out[0]), E_fwd=c[0], C=c[1], R_staged=c[2], E_fwd_culled=c[3],
                E_bwd=int(ncontrib))


def stage_model(wl, wc, deg, world=1):
    """ALGORITHMIC bytes / flops per launch of every stage (restated in DESIGN.md)."""
    P, W, H = wl["P"], wl["W"], wl["H"]
    Npx, Nt = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    K = (deg + 1) ** 2
    Pv, R, C = wc["P_vis"], wc["R"], wc["C"]
    m = {}
    m["preprocess_fwd"] = ("hbm", 48 * P + (79 + 8 + 12 * K) * Pv)
    m["tile_scan"] = ("hbm", 20 * Nt)
    m["scatter"] = ("hbm", 4 * P + 20 * Pv + 8 * R)          # radii + (centre, depth) in, 8-B pairs out
    m["tile_sort"] = ("hbm", 8 * R + 13 * R + 8 * R)          # pairs in; list + keys + mask out; box gathered
    m["render_fwd"] = ("fp32", 75 * C + 35 * (wc["E_fwd"] - C))
    m["grad_zero"] = ("hbm", 80 * P)
    m["render_bwd"] = ("fp32", 210 * C + 35 * (wc["E_bwd"] - C))
    m["preprocess_bwd"] = ("hbm", (147 + 12 * K) * Pv + (44 + 12 * K) * Pv + 4 * P)
    if world > 1:  # factored exchange: K6 writes dL/dRGB instead of the SH rows; the expand reads N of them, writes the rows
        m["preprocess_bwd"] = ("hbm", (147 + 12 * K) * Pv + 44 * Pv + 16 * P)
        m["sh_grad_expand"] = ("hbm", (12 + 12 * world) * P + 12 * 16 * P)
    return m


# ------------------------------------------------------------------------------------------------
# CPU baseline: the oracle port on a bounded sample (rank 0, N=1)
# ------------------------------------------------------------------------------------------------
def load_oracle():
    import importlib.util
    spec = importlib.util.spec_from_file_location("surfel_oracle", os.path.join(ROOT, "oracle", "surfel_oracle.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def cpu_baseline(full_entries_per_iter, deg):
    """Times the CPU oracle fwd+bwd on a C1-sized crop of the workload (10 k surfels, 128x128, same surfel
    distribution, same SH degree) and scales to the metric's unit by list entries evaluated:
    value = (entries/s of the oracle) / (entries per iteration of the full workload).  Extrapolated: flagged."""
    from gaussmart_b200 import synthetic
    so = load_oracle()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    scene = synthetic.make_scene(10_000, 128, deg, synthetic.BASE_SEED + 1)
    cam = synthetic.make_camera(128, 128, 0.0)
    cot = synthetic.make_cotangents(128, 128, 777)
    t0 = time.time()
    entries, reps = 0, 0
    while True:  # >= 10 s of CPU work (bounded: each repetition is a fresh crop of the same distribution)
        scene = synthetic.make_scene(10_000, 128, deg, synthetic.BASE_SEED + 1 + reps)
        f = so.rasterize_forward(scene.means3D, scene.opacities, cam.world_view_transform, cam.full_proj_transform,
                                 cam.camera_center, 128, 128, cam.tanfovx, cam.tanfovy, torch.zeros(3),
                                 sh_degree=deg, shs=scene.shs, scales=scene.scales, rotations=scene.rotations)
        so.rasterize_backward(f, cot[0], cot[1])
        entries += int(f["n_contrib"].reshape(2, -1)[0].long().sum().item())
        reps += 1
        if time.time() - t0 >= 10.0 or reps >= 20:
            break
    dt = time.time() - t0
    per_s = entries / dt
    return dict(value=per_s / max(full_entries_per_iter, 1), unit=UNIT, cores=cores, kind="port",
                sample=f"oracle fwd+bwd of {reps} x (10k surfels @128x128 SH{deg}) in {dt:.1f} s ({entries} list entries); "
                       f"extrapolated per list entry to the full workload ({full_entries_per_iter} entries/iter)",
                sample_seconds=dt)


def knn_leg(wl, impl, reps=3):
    """distCUDA2 (the scale initialiser of the path, run once per training: C4 "incl. distCUDA2 init") on the
    workload's 5 M points: one warm-up + `reps` timed calls, CUDA events; reported beside the iteration metric."""
    pts = wl["params"]["means3D"].contiguous()
    if impl == "reference":
        import ref_loader
        fn = ref_loader.load_knn().distCUDA2
    else:
        import gaussmart_b200 as gb
        fn = gb.distCUDA2
    fn(pts)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn(pts)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    P = pts.shape[0]
    return {"ms": ms, "points": P, "points_per_s": P / (ms * 1e-3), "calls": reps,
            "compulsory_gbs": 16.0 * P / (ms * 1e-3) / 1e9,   # 12 B in + 4 B out per point
            "checksum": float(out.double().sum().item())}


def render_leg(arm, wl, world, rank, device, resolutions, reps=10):
    """Render FPS (BASELINE.json metric, second half; configs[4] "render-only 4K sweep"): forward only through the
    public API under no_grad on the workload's scene, rank r renders view (s*N + r) mod 8, no communication;
    aggregate FPS = N * frames per rank / max-over-ranks device time."""
    from gaussmart_b200 import synthetic
    out = []
    for (W, H) in resolutions:
        cams = [synthetic.Camera(*[x.to(device) if isinstance(x, torch.Tensor) else x for x in c])
                for c in synthetic.make_cameras(W, H, N_VIEWS)]
        box = {}

        def frame(s):
            box["c"] = arm.render(cams[(s * world + rank) % N_VIEWS])[0]
        ms = timed_loop(frame, reps, 3, world, rank)
        out.append({"width": W, "height": H, "fps": world * reps / (ms * 1e-3), "ms_per_frame": ms / reps, "frames": reps * world,
                    "checksum": float(box["c"].double().sum())})
        if arm.name == "ours":  # one more pass with the stage timers on (not part of the timed frames)
            from gaussmart_b200 import _lib
            _lib.stage_timing_enable(True)
            _lib.stage_timing_read(reset=True)
            for s_ in range(4):
                frame(s_)
            torch.cuda.synchronize()
            out[-1]["stages_ms"] = {k: v[0] / v[1] for k, v in _lib.stage_timing_read(reset=True).items() if v[1]}
            _lib.stage_timing_enable(False)
        del cams, box
        torch.cuda.empty_cache()
    return out


def adam_leg(wl, impl, reps=10):
    """`gaussians.optimizer.step()` of the iteration (train.py:212) on the workload's six parameter
    groups (gaussian_model.py:283-290; arguments/__init__.py:77-84 learning rates): the reference's own choice,
    torch.optim.Adam(l, lr=0.0, eps=1e-15), for the reference arm; FusedAdam (one gsm_adam_step launch) for ours.
    compulsory bytes: 16 read + 12 written per parameter element."""
    P, M = wl["P"], wl["params"]["shs"].shape[1]
    shapes = [("xyz", (P, 3), 0.00016), ("f_dc", (P, 1, 3), 0.0025), ("f_rest", (P, M - 1, 3), 0.0025 / 20.0),
              ("opacity", (P, 1), 0.05), ("scaling", (P, 2), 0.005), ("rotation", (P, 4), 0.001)]
    g = torch.Generator(device="cuda").manual_seed(5)
    groups = []
    for name, shape, lr in shapes:
        p = torch.nn.Parameter(torch.randn(shape, device="cuda", generator=g))
        p.grad = torch.randn(shape, device="cuda", generator=g) * 1e-3
        groups.append({"params": [p], "lr": lr, "name": name})
    if impl == "reference":
        opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    else:
        from gaussmart_b200.optim import FusedAdam
        opt = FusedAdam(groups, lr=0.0, eps=1e-15)
    for _ in range(3):
        opt.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        opt.step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    n = sum(gr["params"][0].numel() for gr in groups)
    out = {"ms": ms, "elements": n, "calls": reps, "compulsory_gbs": 28.0 * n / (ms * 1e-3) / 1e9,
           "frac_of_hbm_peak": 28.0 * n / (ms * 1e-3) / 1e9 / peaks()["hbm_gbs"],
           "checksum": float(sum(gr["params"][0].detach().double().sum() for gr in groups))}
    del opt, groups
    torch.cuda.empty_cache()
    return out


def densification_leg(wl, impl, reps=20):
    """Densification bookkeeping of the iteration (train.py:199-203 + gaussian_model.py:551-553; SURVEY.md 8f row 3,
    second half) at the workload's surfel count with ~70 % of the surfels visible: the reference's own indexed torch
    statements for the reference arm, one gsm_densification_stats launch for ours.
    compulsory bytes: 4 (radii) + visible x (12 gradient + 3 x 8 read-modify-write) per surfel."""
    P = wl["P"]
    g = torch.Generator(device="cuda").manual_seed(7)
    radii = torch.randint(1, 200, (P,), device="cuda", generator=g, dtype=torch.int32)
    radii[torch.rand(P, device="cuda", generator=g) < 0.3] = 0
    grad = torch.randn(P, 3, device="cuda", generator=g) * 1e-4 * (radii > 0).unsqueeze(1)
    mr, acc, den = torch.zeros(P, device="cuda"), torch.zeros(P, 1, device="cuda"), torch.zeros(P, 1, device="cuda")
    if impl == "reference":
        def step():
            vf = radii > 0  # gaussian_renderer/__init__.py: visibility_filter
            mr[vf] = torch.max(mr[vf], radii[vf])
            acc[vf] += torch.norm(grad[vf], dim=-1, keepdim=True)
            den[vf] += 1
    else:
        from gaussmart_b200.optim import add_densification_stats

        def step():
            add_densification_stats(mr, acc, den, radii, grad)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    vis = int((radii > 0).sum())
    nbytes = 4.0 * P + 36.0 * vis
    return {"ms": ms, "surfels": P, "visible": vis, "calls": reps, "compulsory_gbs": nbytes / (ms * 1e-3) / 1e9,
            "frac_of_hbm_peak": nbytes / (ms * 1e-3) / 1e9 / peaks()["hbm_gbs"],
            "checksum": [float(mr.double().sum()), float(acc.double().sum()), float(den.double().sum())]}


def densify_leg(wl, impl, reps=3):
    """densify_and_prune of the training loop (train.py:205-208; scene/gaussian_model.py:535-549) on a model of the
    workload's size with an Adam history and densification statistics around the thresholds: the reference's own
    GaussianModel.densify_and_prune (install copy oracle/_ref/gaussmart) for the reference arm, gaussmart_b200.densify's
    kernels on the same GaussianModel class for ours.  One call per fresh model (the call changes the surfel count)."""
    import ref_callers
    if not ref_callers.available():
        return {"unavailable": "oracle/_ref/gaussmart (install copy of the reference's callers) not present"}
    from argparse import ArgumentParser
    C = ref_callers.load("reference" if impl == "reference" else "ours")
    P, extent = wl["P"], 4.4
    names = ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation")

    def model(seed):
        g = torch.Generator(device="cuda").manual_seed(seed)
        r = lambda *s_: torch.randn(*s_, device="cuda", generator=g)
        gm = C.GaussianModel(3)
        gm._xyz, gm._features_dc = torch.nn.Parameter(r(P, 3) * 2.0), torch.nn.Parameter(r(P, 1, 3) * 0.5)
        gm._features_rest, gm._opacity = torch.nn.Parameter(r(P, 15, 3) * 0.05), torch.nn.Parameter(r(P, 1) * 2.5 - 1.0)
        gm._scaling = torch.nn.Parameter(r(P, 2) * 0.9 + float(torch.log(torch.tensor(0.01 * extent))))
        gm._rotation = torch.nn.Parameter(r(P, 4))
        gm._segments = torch.zeros(P, dtype=torch.long, device="cuda")
        gm.max_radii2D = torch.zeros(P, device="cuda")
        gm.spatial_lr_scale = 1.0
        gm.training_setup(C.arguments.OptimizationParams(ArgumentParser()))
        for n in names:
            p_ = getattr(gm, n)
            p_.grad = torch.full_like(p_, 1e-3)
        gm.optimizer.step()
        gm.optimizer.zero_grad(set_to_none=True)
        seen = torch.rand(P, device="cuda", generator=g) > 0.2
        gm.denom = (seen.float() * torch.randint(1, 40, (P,), device="cuda", generator=g).float())[:, None].contiguous()
        gm.xyz_gradient_accum = (gm.denom * torch.rand(P, 1, device="cuda", generator=g) * 4e-4).contiguous()
        return gm

    if impl == "reference":
        run = lambda gm: gm.densify_and_prune(0.0002, 0.05, extent, 20)
    else:
        from gaussmart_b200.densify import densify_and_prune
        run = lambda gm: densify_and_prune(gm, 0.0002, 0.05, extent, 20)
    total, after, chk = 0.0, None, None
    for rep in range(reps + 1):
        gm = model(11)
        torch.manual_seed(99)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(gm)
        e1.record()
        torch.cuda.synchronize()
        if rep:   # (the first call warms the allocator)
            total += e0.elapsed_time(e1)
        after = gm._xyz.shape[0]
        chk = [float(gm._scaling.detach().double().sum()), float(gm._features_rest.detach().double().abs().sum())]
        del gm
        torch.cuda.empty_cache()
    moved = 4.0 * after * (58 * 3 + 2) * 2   # every surviving row read and written once: 58 params + 116 moments + segments
    ms = total / reps
    return {"ms": ms, "surfels_before": P, "surfels_after": after, "calls": reps, "compulsory_gbs": moved / (ms * 1e-3) / 1e9,
            "checksum": chk}


def regularizer_leg(arm, wl, reps=20):
    """The allmap regularisers of the iteration (render() post-processing + depth_to_normal + train.py:131-140), value +
    gradient w.r.t. allmap, on a rendered allmap of the workload: each arm's own implementation, CUDA events
    (SURVEY.md 8f row 4).  compulsory bytes: read 7 maps, write 7 gradient maps = 56 B per pixel."""
    cam_d, cam_h = wl["cams_d"][0], wl["cams"][0]
    allmap = arm.render(cam_d)[2].detach().contiguous()
    cam = cam_h if arm.name == "ours" else cam_d

    def once():
        x = allmap.detach().requires_grad_(True)
        loss = arm.regularizer(x, cam, DEPTH_RATIO, LAMBDA_NORMAL, LAMBDA_DIST)
        loss.backward()
        return loss, x.grad
    for _ in range(3):
        once()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        loss, grad = once()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    n = wl["H"] * wl["W"]
    return {"ms": ms, "pixels": n, "calls": reps, "compulsory_gbs": 56.0 * n / (ms * 1e-3) / 1e9,
            "loss": float(loss.detach()), "grad_checksum": float(torch.nan_to_num(grad, 0.0, 0.0, 0.0).double().abs().sum())}


def loss_leg(wl, photometric, reps=20):
    """The photometric loss of the iteration (train.py:113-114), value + gradient w.r.t. the image, at the
    workload's resolution: each arm's own implementation (fused gsm_photometric_loss / the reference's torch ops),
    CUDA events; reported beside the iteration metric.  compulsory bytes: read image + gt,
    write the gradient = 12 B per pixel-channel."""
    g = torch.Generator(device="cpu").manual_seed(99)
    gt = torch.rand(3, wl["H"], wl["W"], generator=g).cuda()
    img = (gt + 0.05 * torch.randn(3, wl["H"], wl["W"], generator=g).cuda()).clamp(0, 1)

    def once():
        x = img.detach().requires_grad_(True)
        loss = photometric(x, gt, LAMBDA_DSSIM)
        loss.backward()
        return loss, x.grad
    for _ in range(3):
        once()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        loss, grad = once()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    n = img.numel()
    return {"ms": ms, "pixels": wl["H"] * wl["W"], "calls": reps, "compulsory_gbs": 12.0 * n / (ms * 1e-3) / 1e9,
            "loss": float(loss.detach()), "grad_checksum": float(grad.double().abs().sum())}


def drop_in_leg(args, wl, device, view_of):
    """What a user of the reference gets by switching the two packages and NOTHING else: the reference's UNCHANGED
    gaussian_renderer.render() (install copy oracle/_ref/gaussmart, tests/ref_callers.py) on a GaussianModel holding the
    workload's surfels, the reference's own L1 + SSIM (utils/loss_utils.py) and train.py:131-140's regularisers as torch
    ops, loss.backward(), loss read-back -- with `diff_surfel_rasterization` resolving to this repo (--impl ours) or to
    the reference build (--impl reference).  Per step the ground truth goes host -> device and the loss comes back."""
    import ref_callers
    if not ref_callers.available():
        return {"unavailable": "oracle/_ref/gaussmart (install copy of the reference's callers) not present"}
    from argparse import ArgumentParser
    C = ref_callers.load("ours" if args.impl == "ours" else "reference")
    prm = wl["params"]
    gm = C.GaussianModel(wl["deg"])
    o_ = prm["opacities"].clamp(1e-6, 1 - 1e-6)
    mk = lambda t: torch.nn.Parameter(t.clone().contiguous().requires_grad_(True))
    gm._xyz, gm._features_dc, gm._features_rest = mk(prm["means3D"]), mk(prm["shs"][:, :1]), mk(prm["shs"][:, 1:])
    gm._opacity, gm._scaling, gm._rotation = mk(torch.log(o_ / (1 - o_))), mk(torch.log(prm["scales"])), mk(prm["rotations"])
    gm.active_sh_degree = wl["deg"]
    parser = ArgumentParser()
    opt, pipe = C.arguments.OptimizationParams(parser), C.arguments.PipelineParams(parser)
    W, H = wl["W"], wl["H"]
    cams = [C.MiniCam(W, H, c.FoVy, c.FoVx, 0.01, 100.0, c.world_view_transform, c.full_proj_transform) for c in wl["cams_d"]]
    background = torch.tensor([0, 0, 0], dtype=torch.float32, device=device)
    gt_dev = torch.empty(3, H, W, device=device)
    loss_host = torch.empty((), dtype=torch.float32).pin_memory()
    l1_loss, ssim = C.loss_utils.l1_loss, C.loss_utils.ssim
    params = [gm._xyz, gm._features_dc, gm._features_rest, gm._opacity, gm._scaling, gm._rotation]

    def step(s):
        gt_dev.copy_(wl["gt_host"][s % 2], non_blocking=True)
        for p_ in params:
            p_.grad = None
        render_pkg = C.render(cams[view_of(s)], gm, pipe, background)             # train.py:104
        image = render_pkg["render"]
        Ll1 = l1_loss(image, gt_dev)                                              # train.py:113-114
        loss = (1.0 - opt.lambda_dssim) * Ll1 + opt.lambda_dssim * (1.0 - ssim(image, gt_dev))
        rend_dist, rend_normal, surf_normal = render_pkg["rend_dist"], render_pkg["rend_normal"], render_pkg["surf_normal"]
        normal_error = (1 - (rend_normal * surf_normal).sum(dim=0))[None]         # train.py:135-140
        total_loss = loss + LAMBDA_DIST * (rend_dist).mean() + LAMBDA_NORMAL * (normal_error).mean()
        total_loss.backward()                                                     # train.py:144
        loss_host.copy_(total_loss.detach(), non_blocking=False)                  # train.py:149 (.item())
    ms = timed_loop(step, args.steps, max(3, args.warmup), 1, 0)
    out = {"value": 1000.0 * args.steps / ms, "unit": UNIT, "ms_per_step": ms / args.steps, "loss": float(loss_host),
           "h2d_bytes_per_step": int(3 * H * W * 4), "d2h_bytes_per_step": 4,
           "what": "the reference's unchanged render() + its torch L1/SSIM/regulariser loss + backward; only the two "
                   "packages diff_surfel_rasterization / simple_knn differ between the arms"}
    del gm, params
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference" and rank != 0:
        return 0  # the reference is single-GPU: rank 0 alone runs it
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback of the product path)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if args.impl == "reference":
        world = 1
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    wl = make_workload(args, device)
    ref_kind = None
    if args.impl == "reference":
        import ref_loader
        if not ref_loader.available():
            # the reference CUDA build did not travel: time the CPU oracle port instead (bounded sample)
            wc = work_counts(wl, 0, device)
            cb = cpu_baseline(int(wc["E_bwd"]), wl["deg"])
            print(json.dumps({
                "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": 1,
                "steps": 1, "warmup": 0, "ms_per_step": 1000.0 / cb["value"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": args.workload, "surfels": wl["P"], "width": wl["W"], "height": wl["H"]},
                "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
            return 0
        arm = ReferenceArm(wl, device, world, args.views_per_step)
        ref_kind = "reference CUDA build (oracle/_ref, unmodified sources, sm_100a) on the GPU"
    else:
        arm = OursArm(wl, device, world, args.views_per_step)
        from gaussmart_b200 import _lib
        _lib.stage_timing_enable(True)
        _lib.stage_timing_read(reset=True)

    view_of = lambda s: (s * world + rank) % N_VIEWS

    # ---- value: device-resident
    sampler = ClockSampler(local_rank)
    # nvidia-smi takes ~0.2 s to start sampling and the timed region is only ~0.1 s at the default 20 steps: the
    # sampler runs from before the warm-up (the same load, kept on for ~0.6 s) to the end of the timed region
    sampler.start()
    V = args.views_per_step or world           # views per step over all GPUs
    for s in range(max(args.warmup, 3)):
        arm.value_step(s)
    barrier(world)
    # keep the same load on for ~0.6 s more (untimed): with 200 ms between polls at least two samples fall under load
    # before the timed region starts, and the median clock is a clock under load.  Same step count on every rank.
    t_w = time.time()
    for s in range(3):
        arm.value_step(s)
    torch.cuda.synchronize()
    step_s = torch.tensor([(time.time() - t_w) / 3.0], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(step_s, op=dist.ReduceOp.MAX)
    n_load = int(min(400, max(0, math.ceil(0.6 / max(step_s.item(), 1e-4)))))
    for s in range(n_load):
        arm.value_step(s)
        if s % 8 == 7:
            torch.cuda.synchronize()
    barrier(world)
    if args.impl == "ours":
        _lib.stage_timing_read(reset=True)   # the stage timers cover the timed region only
    if getattr(arm, "peer", False):
        arm.ex.probe = []                       # per-step event pairs around the parts of the peer exchange
    ms_total = timed_loop(lambda s: arm.value_step(max(args.warmup, 3) + s), args.steps, 0, world, rank)
    exchange_probe = None
    if getattr(arm, "peer", False):
        acc = {}
        for step_marks in arm.ex.probe:
            for name, a_, b_ in step_marks:
                acc[name] = acc.get(name, 0.0) + a_.elapsed_time(b_)
        exchange_probe = {k: v / max(1, len(arm.ex.probe)) for k, v in acc.items()}
        arm.ex.probe = None
    clocks = sampler.stop()
    stages = _lib.stage_timing_read(reset=True) if args.impl == "ours" else {}
    if args.impl == "ours":
        _lib.stage_timing_enable(False)
    ms_per_step = ms_total / args.steps
    value = V * 1000.0 / ms_per_step           # view-iterations per second over all GPUs

    # ---- e2e: public API, host buffers
    e2e = None
    if not args.no_e2e and V == world:   # (one view per GPU; the fixed-batch strong-scaling runs report `value` only)
        W, H = wl["W"], wl["H"]
        gt_dev = torch.empty(3, H, W, device=device)
        mats_host = [torch.cat([c.world_view_transform.reshape(-1), c.full_proj_transform.reshape(-1),
                                c.camera_center.reshape(-1)]).pin_memory() for c in wl["cams"]]
        mats_dev = torch.empty(35, device=device)
        loss_host = torch.empty((), dtype=torch.float32).pin_memory()
        from gaussmart_b200 import synthetic

        copy_stream = torch.cuda.Stream(device=device)
        gt_ready = lambda: torch.cuda.current_stream().wait_stream(copy_stream)

        def e2e_step(s):
            v = view_of(s)
            # this step's ground truth goes up on a copy stream while the render runs (the loss waits for it); the previous
            # step ended with a blocking read of its loss, so gt_dev is free.  The camera is needed at once: main stream.
            copy_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(copy_stream):
                gt_dev.copy_(wl["gt_host"][s % 2], non_blocking=True)
            mats_dev.copy_(mats_host[v], non_blocking=True)
            c = wl["cams"][v]
            cam_d = synthetic.Camera(c.image_width, c.image_height, c.FoVx, c.FoVy, c.tanfovx, c.tanfovy,
                                     mats_dev[0:16].view(4, 4), mats_dev[16:32].view(4, 4), mats_dev[32:35])
            if world > 1 and args.impl == "ours":
                base = (v - rank) % N_VIEWS
                cams_step = [cam_d if r == rank else wl["cams_d"][(base + r) % N_VIEWS] for r in range(world)]
                loss = arm.api_step_view_parallel(cams_step, gt_dev, cam_h=c, gt_ready=gt_ready)
            else:
                loss = arm.api_step(cam_d, gt_dev, cam_h=c, gt_ready=gt_ready)
            loss_host.copy_(loss.detach(), non_blocking=False)   # the per-iteration .item() of train.py:149

        ms_e2e = timed_loop(e2e_step, args.steps, max(3, args.warmup), world, rank)
        e2e = {"value": world * 1000.0 * args.steps / ms_e2e, "unit": UNIT,
               "h2d_bytes_per_step": int(3 * H * W * 4 + 35 * 4), "d2h_bytes_per_step": 4,
               "ms_per_step": ms_e2e / args.steps,
               "api": "GaussianRasterizationSettings/GaussianRasterizer + train.py's loss (L1+SSIM, normal, distortion; each "
                      "arm's own implementation: fused kernels here, the reference's torch ops there) + loss.backward()"}

    value_api = None
    if world == 1 and V == 1 and args.impl == "ours":
        # `value` above runs the bucket path (no autograd; the path the view-parallel step uses); the reference arm's
        # `value` goes through ITS autograd front end.  The same thing for this arm, for a like-for-like reading:
        def autograd_step(s):
            cam_d = wl["cams_d"][view_of(s)]
            S = arm.gb.GaussianRasterizationSettings(**synthetic_settings(cam_d, wl["deg"], arm.bg))
            for v_ in arm.leaves.values():
                v_.grad = None
            arm.means2D.grad = None
            L = arm.leaves
            color, radii, allmap = arm.gb.GaussianRasterizer(S)(
                means3D=L["means3D"], means2D=arm.means2D, opacities=L["opacities"], shs=L["shs"], scales=L["scales"],
                rotations=L["rotations"])
            torch.autograd.backward([color, allmap], [wl["cot_d"][0], wl["cot_d"][1]])
        from gaussmart_b200.synthetic import settings_kwargs as synthetic_settings
        ms_api = timed_loop(autograd_step, args.steps, max(3, args.warmup), world, rank)
        value_api = {"value": 1000.0 * args.steps / ms_api, "unit": UNIT, "ms_per_step": ms_api / args.steps,
                     "path": "GaussianRasterizer + torch.autograd.backward with resident cotangents (what the reference "
                             "arm's `value` times); `value` itself is rasterize_into_bucket (no autograd)"}

    drop_in = None
    if world == 1 and V == 1 and not args.no_drop_in and not args.no_e2e:
        try:
            drop_in = drop_in_leg(args, wl, device, view_of)
        except Exception as ex:  # pragma: no cover
            drop_in = {"error": repr(ex)[:300]}

    render = None
    if not args.no_render:
        torch.cuda.empty_cache()
        render = render_leg(arm, wl, world, rank, device, [(wl["W"], wl["H"]), (3840, 2160)])

    # ---- the full training iteration (BASELINE.json configs[3]): the e2e step + the optimizer step, each arm with its
    #      own optimizer (torch.optim.Adam as gaussian_model.py:292 builds it / FusedAdam).  Run last: it moves the surfels.
    train_iter = None
    if e2e is not None and world == 1:
        prm = wl["params"]
        o_ = prm["opacities"].clamp(1e-6, 1 - 1e-6)
        raw = {"xyz": prm["means3D"].clone(), "f_dc": prm["shs"][:, :1].contiguous(), "f_rest": prm["shs"][:, 1:].contiguous(),
               "opacity": torch.log(o_ / (1 - o_)), "scaling": torch.log(prm["scales"]), "rotation": prm["rotations"].clone()}
        raw = {k: v.requires_grad_(True) for k, v in raw.items()}   # the STORED parameters of gaussian_model.py:50-56
        m2d = torch.zeros_like(raw["xyz"], requires_grad=True)
        # tiny learning rates: the optimizer's work does not depend on lr, and the workload stays the workload
        groups = [{"params": [raw[k]], "lr": 1e-8, "name": k} for k in ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")]
        if args.impl == "reference":
            opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
        else:
            from gaussmart_b200.optim import FusedAdam
            opt = FusedAdam(groups, lr=0.0, eps=1e-15)

        P_ = raw["xyz"].shape[0]  # max_radii2D, xyz_gradient_accum, denom (gaussian_model.py:58-60)
        dens = [torch.zeros(P_, device=device), torch.zeros(P_, 1, device=device), torch.zeros(P_, 1, device=device)]
        if args.impl != "reference":
            from gaussmart_b200.optim import add_densification_stats

        def train_step(s):
            v = view_of(s)
            copy_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(copy_stream):
                gt_dev.copy_(wl["gt_host"][s % 2], non_blocking=True)
            mats_dev.copy_(mats_host[v], non_blocking=True)
            c = wl["cams"][v]
            cam_d = synthetic.Camera(c.image_width, c.image_height, c.FoVx, c.FoVy, c.tanfovx, c.tanfovy,
                                     mats_dev[0:16].view(4, 4), mats_dev[16:32].view(4, 4), mats_dev[32:35])
            loss = arm.raw_step(raw, m2d, cam_d, gt_dev, c, gt_ready)
            loss_host.copy_(loss.detach(), non_blocking=False)
            radii = arm.last_radii
            if args.impl == "reference":  # train.py:199-203 + gaussian_model.py:551-553, the reference's statements
                with torch.no_grad():
                    vf = radii > 0
                    dens[0][vf] = torch.max(dens[0][vf], radii[vf])
                    dens[1][vf] += torch.norm(m2d.grad[vf], dim=-1, keepdim=True)
                    dens[2][vf] += 1
            else:
                add_densification_stats(dens[0], dens[1], dens[2], radii, m2d.grad)
            opt.step()
        for s_ in range(3):
            train_step(s_)
        if args.impl == "ours":
            _lib.stage_timing_enable(True)
            _lib.stage_timing_read(reset=True)   # after the warm-up: the stage sums cover exactly the timed steps
        ms_train = timed_loop(lambda s: train_step(3 + s), args.steps, 0, world, rank)
        train_stages = None
        if args.impl == "ours":
            train_stages = {k: v[0] / args.steps for k, v in _lib.stage_timing_read(reset=True).items() if v[1]}
            _lib.stage_timing_enable(False)
        train_iter = {"value": 1000.0 * args.steps / ms_train, "unit": UNIT, "ms_per_step": ms_train / args.steps,
                      "includes": "H2D of camera + ground truth, render() from the STORED parameters (getters of "
                                  "gaussian_model.py:104-123: torch ops for the reference, fused raw entry here), "
                                  "L1+SSIM+regulariser loss, backward, loss read-back, densification bookkeeping "
                                  "(train.py:199-203), Adam step over all 58 floats/surfel",
                      "densification_checksum": [float(t.double().sum()) for t in dens]}
        if train_stages:
            train_iter["stages_ms_per_step"] = train_stages
        del opt

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    pk = peaks()
    if args.impl == "ours":
        try:
            pk["fp32_tflops_derived"] = pk["fp32_tflops"]
            pk["fp32_tflops"] = measure_fp32_peak(device)
            pk["fp32_source"] = ("measured: gsm_fp32_peak_probe (register-resident FMA chains, 64 warps/SM), best of 5; "
                                 "derived 148 x 128 x 2 x sm_max_mhz = %.1f TFLOP/s" % pk["fp32_tflops_derived"])
        except Exception as ex:  # pragma: no cover
            pk["fp32_source"] += " (probe failed: %s)" % repr(ex)[:80]
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak" if args.views_per_step is None else "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {wl['P']} surfels, {wl['W']}x{wl['H']}, SH degree {wl['deg']}, "
                               + (f"fwd+bwd of one view per GPU per step, {N_VIEWS} cameras cycled" if args.views_per_step is None
                                  else f"fwd+bwd of a fixed batch of {V} views per step shared by the GPUs, {N_VIEWS} cameras cycled"),
                   "surfels": wl["P"], "width": wl["W"], "height": wl["H"], "sh_degree": wl["deg"],
                   "views_per_step": V, "parallelism": f"view-parallel x{world}" if world > 1 else "single GPU",
                   "l2": "inputs larger than L2 (surfel parameters %.2f GB >> 126 MB)" %
                         (wl["P"] * 58 * 4 / 1e9),
                   "exchange": ("none (single GPU)" if world == 1 else
                                ("fused over NVLink peer memory: K6 pushes dL/dRGB (12 B/surfel/view) into every rank's slot, one "
                                 "kernel reduces 48 B/surfel + max radii (%s), local SH expand alongside" %
                                 ("NVLS multimem.ld_reduce / multimem.st" if getattr(arm.ex, "_flat_mc", 0) else "two-shot peer loads / stores")
                                 if getattr(arm, "peer", False) else
                                 "factored over NCCL: all-reduce 40 B/surfel + all-gather dL/dRGB 12 B/surfel/view + local SH expand")),
                   "debug_P_override": args.P is not None},
        "clocks": clocks, "e2e": e2e,
    }
    if exchange_probe is not None:
        line["exchange_ms_rank0"] = exchange_probe   # cudaEvent pairs on rank 0: barrier waits, two-shot all-reduce, SH expand
    if value_api is not None:
        line["value_public_api"] = value_api
    if drop_in is not None:
        line["drop_in"] = drop_in
    if render is not None:
        line["render"] = render
    if train_iter is not None:
        line["train_iteration"] = train_iter
    if args.impl == "reference":
        line["impl"] = "reference"
        line["gpu_launches"] = None
        line["reference_modules"] = arm.modules
        line["cpu_baseline"] = {"value": value, "unit": UNIT, "cores": 0, "kind": "reference",
                                "sample": f"full workload, every step; {ref_kind}. The reference has no CPU "
                                          "implementation of this path, so its own implementation is timed where "
                                          "it runs: on the same B200"}
        if e2e is None:
            line["e2e"] = {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
        if not args.no_knn:
            try:
                line["knn"] = knn_leg(wl, "reference", reps=1)
            except Exception as ex:  # pragma: no cover
                line["knn"] = {"error": repr(ex)[:200]}
            if hasattr(arm, "photometric"):
                line["loss"] = loss_leg(wl, arm.photometric)
            line["adam"] = adam_leg(wl, "reference")
            try:
                line["densification_stats"] = densification_leg(wl, "reference")
            except Exception as ex:  # pragma: no cover
                line["densification_stats"] = {"error": repr(ex)[:200]}
            if hasattr(arm, "regularizer"):
                line["regularizer"] = regularizer_leg(arm, wl)
            try:
                line["densify_and_prune"] = densify_leg(wl, "reference")
            except Exception as ex:  # pragma: no cover
                line["densify_and_prune"] = {"error": repr(ex)[:200]}
        print(json.dumps(line))
        return 0

    # ---- roofline of every stage, dominant kernel first
    wcs = [work_counts(wl, v, device) for v in sorted({(s * V + i * world + rank) % N_VIEWS for s in range(args.steps)
                                                      for i in range(max(1, V // world))})][:N_VIEWS]
    wc = {k: sum(w[k] for w in wcs) / len(wcs) for k in wcs[0]}
    model = stage_model(wl, wc, wl["deg"], world)
    rows = []
    for name, (bound, work) in model.items():
        ms, calls = stages.get(name, (0.0, 0))
        if calls == 0:
            continue
        avg_ms = ms / calls
        if bound == "hbm":
            ach, peak, unit = work / (avg_ms * 1e-3) / 1e9, pk["hbm_gbs"], "GB/s"
        else:
            ach, peak, unit = work / (avg_ms * 1e-3) / 1e12, pk["fp32_tflops"], "TFLOP/s"
        rows.append({"kernel": name, "bound": bound, "ms": avg_ms, "launches": calls, "algorithmic": work,
                     "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak,
                     "share_of_step": ms / ms_total})
    rows.sort(key=lambda r: -r["ms"])
    top = dict(rows[0]) if rows else {}
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get(top.get("kernel"))
        traffic_src = "ncu --set full dram__bytes_read+write per launch, %s of commit %s" % (tj.get("_source", "profiles/"), tj.get("_commit", "?"))
    except Exception:
        pass
    line["roofline"] = {"kernel": top.get("kernel"), "bound": top.get("bound"), "achieved": top.get("achieved"),
                        "peak": top.get("peak"), "unit": top.get("unit"), "frac": top.get("frac"),
                        "traffic": traffic, "traffic_source": traffic_src, "ms": top.get("ms"),
                        "peak_source": pk["source"] if top.get("bound") == "hbm" else pk["fp32_source"]}
    line["stages"] = rows
    line["work"] = wc
    kernels_per_stage = {"preprocess_fwd": 1, "tile_scan": 1, "scatter": 1, "tile_sort": 3,
                         "render_fwd": 1, "grad_zero": 0, "render_bwd": 1, "preprocess_bwd": 1, "sh_grad_expand": 1}
    line["gpu_launches"] = int(sum(kernels_per_stage.get(n, 0) * c for n, (_, c) in stages.items()))
    if world == 1 and not args.no_knn:
        line["knn"] = knn_leg(wl, "ours")
        line["loss"] = loss_leg(wl, arm.photometric)
        line["adam"] = adam_leg(wl, "ours")
        line["densification_stats"] = densification_leg(wl, "ours")
        line["regularizer"] = regularizer_leg(arm, wl)
        try:
            line["densify_and_prune"] = densify_leg(wl, "ours")
        except Exception as ex:  # pragma: no cover
            line["densify_and_prune"] = {"error": repr(ex)[:200]}
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(int(wc["E_bwd"]), wl["deg"])
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    # stdout carries the ONE JSON line and nothing else: libraries that write to file descriptor 1 (NCCL prints its
    # version banner there when NCCL_DEBUG is set on the box) are sent to stderr for the duration of the run.
    sys.stdout.flush()
    _json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = _json_out
    try:
        _rc = main()
    finally:
        _json_out.flush()
    sys.exit(_rc)
