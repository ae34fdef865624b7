#!/usr/bin/env python
"""Benchmark of the render hot path: rays/s for a full coarse+fine render.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--precision fp32|fp16]

Workload (BASELINE.json configs[1]): SRN car 128x128, 1 source view, 64 coarse + 32 fine
(16 importance + 16 depth) samples, whole frame per step (16 384 rays, 2.62 M MLP rows),
synthetic scene and conditioned random weights (pixelnerf_b200.synthetic).  A step is one
frame.  With N > 1 (torchrun) one call carries N frames (N x 16 384 rays) and goes through the
product's own multi-GPU wrapper, `renderer.bind_parallel(net, gpus)` -> `parallel.RaySharded`: contiguous ray
ranges per rank, encoded scene broadcast from rank 0, one all-gather of rgb+depth per call (weak scaling: one
frame's worth of rays per GPU).

`--impl reference` times the CPU oracle port of the reference's torch path (oracle/, the
reference is Python and cannot travel to the GPU box) on all host cores, same metric.

Besides the contract's keys the headline line carries: `sharded` (BASELINE configs 3 and 4 -- the 40-frame NS=2
video and the DTU 300x400 NS=3 frame -- through the same ray-sharded wrapper, strong scaling), and at N=1 `parity`
(max-abs errors of this build against the oracle on 256 rays of the frame), `roofline_kernels` (achieved GB/s of the
HBM-bound kernels) and `gpu_baseline` (the reference's op sequence as torch-CUDA eager on the same GPU).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

SCENE = "srn_car"
N_COARSE, N_FINE, N_DEPTH = 64, 32, 16
WORKLOADS = {"srn_car": "SRN car 128x128 NS=1", "srn_chair": "SRN chair 128x128 NS=2 (BASELINE config 3, one frame)",
             "dtu": "DTU 300x400 NS=3 (BASELINE config 4)", "sn64": "NMR 64x64 NS=1 (BASELINE config 1)"}
MLP_FLOP_PER_POINT = {1: 2 * (1 * 2380800 + 1050624), 2: 2 * (2 * 2380800 + 1050624), 3: 2 * (3 * 2380800 + 1050624)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons while the timed region runs (NVML, ~5 ms period; the
    nvidia-smi CLI takes longer per query than a whole frame)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False
        self.max_mhz = None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(visible.split(",")[self.index]) if visible and visible.split(",")[self.index].isdigit() else self.index
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            while not self.stop_flag:
                self.samples.append((nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
                                     nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
                time.sleep(0.005)
        except Exception as e:  # noqa: BLE001
            self.error = repr(e)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unsampled: " + getattr(self, "error", "")]}
        mhz = sorted(s[0] for s in self.samples)
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                "hw_power_brake_slowdown": 0x80}
        seen = 0
        for s in self.samples:
            seen |= s[1]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_min_mhz": mhz[0], "sm_max_mhz": self.max_mhz,
                "reasons": [n for n, b in bits.items() if seen & b], "samples": len(self.samples)}


def cpu_oracle_rays_per_s(sc, wc, wf, n_rays, repeats=1):
    """The reference's torch-CPU path (oracle port) on a bounded ray sample."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import oracle_render
    torch.set_num_threads(os.cpu_count())
    idx = torch.linspace(0, sc["rays"].shape[0] - 1, n_rays).long()
    rays = sc["rays"][idx]
    draws = syn.draws(7, n_rays, N_COARSE, N_FINE - N_DEPTH, N_DEPTH)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        oracle_render(sc, wc, wf, rays, draws, N_COARSE, N_FINE, N_DEPTH)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n_rays / best, best


def run_reference(args):
    """--impl reference: CPU oracle port, all host threads, one bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from pixelnerf_b200 import synthetic as syn
    sc = syn.scene(SCENE)
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    sample = 2048
    for _ in range(args.warmup):
        cpu_oracle_rays_per_s(sc, wc, wf, sample)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_oracle_rays_per_s(sc, wc, wf, sample)
    dt = time.perf_counter() - t0
    val = sample * args.steps / dt
    cores = os.cpu_count()
    line = {"impl": "reference", "metric": "rays/sec coarse+fine", "value": val, "unit": "rays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"SRN car 128x128 NS=1, {N_COARSE}+{N_FINE} samples; {sample}-ray sample per step "
                                   "(CPU cost is linear in rays)"},
            "cpu_baseline": {"value": val, "unit": "rays/s", "cores": cores, "kind": "port",
                             "sample": f"{sample} rays/step x {args.steps} steps, torch {torch.__version__} CPU, "
                                       f"{cores} threads"},
            "e2e": {"value": val, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def run_torch_cuda(args):
    """--impl torch-cuda: the reference's torch op sequence (the oracle port) run as torch-CUDA eager on the same
    B200 -- the "existing Blackwell path" (cuBLAS/ATen library kernels) the new kernels replace.  Whole
    frame per step; fp32 (TF32 off), TF32 allowed, and bf16 autocast.  A baseline measurement, not a product path."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import scene_camera
    import oracle
    from oracle import model as omodel
    dev = torch.device("cuda", 0)
    sc = syn.scene(args.scene)
    wc = {k: v.to(dev) for k, v in syn.mlp_weights(1).items()}
    wf = {k: v.to(dev) for k, v in syn.mlp_weights(2).items()}
    cam = {k: v.to(dev) for k, v in scene_camera(sc).items()}
    lat = sc["latent"].to(dev)
    rays = sc["rays"].to(dev)
    n = rays.shape[0]
    draws = {k: v.to(dev) for k, v in syn.draws(7, n, N_COARSE, N_FINE - N_DEPTH, N_DEPTH).items()}
    chunk = 4096   # rays per call: 0.4 M points x NS views, like the reference's eval_batch_size chunks

    def net_fn(xyz, dirs, coarse):
        return omodel.net_forward(wc if coarse else wf, lat, cam, sc["ns"], xyz, dirs)

    def frame():
        with torch.no_grad():
            for s0 in range(0, n, chunk):
                d = {k: v[s0:s0 + chunk] for k, v in draws.items()}
                oracle.render_rays(net_fn, rays[s0:s0 + chunk], 1, N_COARSE, N_FINE, N_DEPTH, d, depth_std=0.01,
                                   white_bkgd=sc["white"])

    if args.train:   # config 5 on the same comparator: 4 objects x --train-rays rays, fwd + autograd bwd + Adam
        sb = 4
        params = [v.requires_grad_(True) for v in list(wc.values()) + list(wf.values())]
        opt = torch.optim.Adam(params, lr=1e-4)
        lat4 = syn.latent_maps(21, sb, sc["hl"], sc["wl"]).to(dev)
        cam4 = dict(cam, poses=cam["poses"].expand(sb, -1, -1).contiguous())
        res = {}
        for n_obj in args.train_rays:
            g = torch.Generator().manual_seed(n_obj)
            idx = torch.randint(0, n, (sb, n_obj), generator=g)
            rr = sc["rays"][idx].reshape(-1, 8).to(dev)
            tgt = torch.rand(sb * n_obj, 3, generator=g).to(dev)
            dd = {k: v.to(dev) for k, v in syn.draws(7, sb * n_obj, N_COARSE, N_FINE - N_DEPTH, N_DEPTH).items()}

            def net4(xyz, dirs, coarse):
                return omodel.net_forward(wc if coarse else wf, lat4, cam4, sc["ns"], xyz, dirs)

            def tstep():
                opt.zero_grad(set_to_none=True)
                o = oracle.render_rays(net4, rr, sb, N_COARSE, N_FINE, N_DEPTH, dd, depth_std=0.01, white_bkgd=sc["white"])
                loss = ((o["rgb_coarse"] - tgt) ** 2).mean() + ((o["rgb_fine"] - tgt) ** 2).mean()
                loss.backward()
                opt.step()

            for mode in ("fp32", "tf32", "bf16_autocast"):
                torch.backends.cuda.matmul.allow_tf32 = mode == "tf32"
                ctx = torch.autocast("cuda", dtype=torch.bfloat16, enabled=mode == "bf16_autocast")
                with ctx:
                    for _ in range(3):
                        tstep()
                    torch.cuda.synchronize()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    steps = max(1, min(args.steps, 10))
                    a.record()
                    for _ in range(steps):
                        tstep()
                    b.record()
                    b.synchronize()
                ms = a.elapsed_time(b) / steps
                res[f"{sb * n_obj}_rays_{mode}"] = {"ms_per_step": ms, "rays/s": sb * n_obj / (ms * 1e-3)}
        print(json.dumps({"impl": "torch-cuda-eager", "metric": "training step rays/sec (config 5)", "unit": "rays/s",
                          "config": {"workload": "4 objects, fwd + autograd bwd + Adam on both MLPs; oracle port of the "
                                                 "reference's torch op sequence on cuda:0, torch " + torch.__version__},
                          "modes": res}))
        return
    res = time_torch_cuda_modes(frame, n, max(args.warmup, 2), max(1, min(args.steps, 5)))
    print(json.dumps({"impl": "torch-cuda-eager", "metric": "rays/sec coarse+fine", "unit": "rays/s",
                      "config": {"workload": f"{WORKLOADS[args.scene]} full frame, {n} rays, {chunk}-ray chunks",
                                 "code": "oracle port of the reference's torch op sequence on cuda:0, torch "
                                         + torch.__version__},
                      "modes": res}))


def time_torch_cuda_modes(frame, n, warmup, steps):
    res = {}
    tf32_was = torch.backends.cuda.matmul.allow_tf32
    for mode in ("fp32", "tf32", "bf16_autocast"):
        torch.backends.cuda.matmul.allow_tf32 = mode == "tf32"
        torch.backends.cudnn.allow_tf32 = mode == "tf32"
        ctx = torch.autocast("cuda", dtype=torch.bfloat16) if mode == "bf16_autocast" else torch.autocast("cuda", enabled=False)
        with ctx:
            for _ in range(warmup):
                frame()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(steps):
                frame()
            b.record()
            b.synchronize()
        ms = a.elapsed_time(b) / steps
        res[mode] = {"ms_per_frame": ms, "rays/s": n / (ms * 1e-3)}
    torch.backends.cuda.matmul.allow_tf32 = tf32_was
    return res


def torch_cuda_frame(sc, scene, steps=2, warmup=1):
    """`gpu_baseline` of the headline line: the reference's torch op sequence (oracle port) as torch-CUDA eager on
    this GPU (cuBLAS / ATen library kernels), whole frame, fp32 / TF32 / bf16 autocast -- the existing "Blackwell path"
    the kernels replace.  A comparator, not a product path."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import scene_camera
    import oracle
    from oracle import model as omodel
    dev = torch.device("cuda", torch.cuda.current_device())
    wc = {k: v.to(dev) for k, v in syn.mlp_weights(1).items()}
    wf = {k: v.to(dev) for k, v in syn.mlp_weights(2).items()}
    cam = {k: v.to(dev) for k, v in scene_camera(sc).items()}
    lat = sc["latent"].to(dev)
    rays = sc["rays"].to(dev)
    n = rays.shape[0]
    draws = {k: v.to(dev) for k, v in syn.draws(7, n, N_COARSE, N_FINE - N_DEPTH, N_DEPTH).items()}
    chunk = 4096

    def net_fn(xyz, dirs, coarse):
        return omodel.net_forward(wc if coarse else wf, lat, cam, sc["ns"], xyz, dirs)

    def frame():
        with torch.no_grad():
            for s0 in range(0, n, chunk):
                d = {k: v[s0:s0 + chunk] for k, v in draws.items()}
                oracle.render_rays(net_fn, rays[s0:s0 + chunk], 1, N_COARSE, N_FINE, N_DEPTH, d, depth_std=0.01,
                                   white_bkgd=sc["white"])

    return {"what": "reference torch op sequence (oracle port) as torch-CUDA eager on this GPU, whole frame, "
                    f"{chunk}-ray chunks, torch {torch.__version__}; {warmup} warm-up + {steps} timed frames per mode",
            "unit": "rays/s", "modes": time_torch_cuda_modes(frame, n, warmup, steps)}


def run_kernels(args):
    torch.cuda.set_device(0)
    print(json.dumps(kernel_table()))


def kernel_table(r=1 << 20):
    """Every HBM-bound kernel of the path alone at a size that exceeds L2 (2^20 rays), as achieved GB/s of its
    ALGORITHMIC bytes against the measured copy bandwidth."""
    from pixelnerf_b200 import _lib, synthetic as syn
    from pixelnerf_b200.testing import build_net
    dev = torch.device("cuda", torch.cuda.current_device())
    pk = peaks()
    g = torch.Generator(device="cpu").manual_seed(0)
    rays = torch.zeros(r, 8, device=dev)
    rays[:, 3:6] = torch.nn.functional.normalize(torch.randn(r, 3, device=dev), dim=-1)
    rays[:, 2] = 1.3
    rays[:, 6], rays[:, 7] = 0.8, 1.8
    res = {}

    def timeit(name, fn, nbytes, iters=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        b.synchronize()
        ms = a.elapsed_time(b) / iters
        gbs = nbytes / (ms * 1e-3) / 1e9
        res[name] = {"ms": ms, "algorithmic_bytes": nbytes, "GB/s": gbs, "frac_of_measured_hbm": gbs / pk["hbm"]}

    kc, kf, kd = N_COARSE, N_FINE - N_DEPTH, N_DEPTH
    k2 = kc + N_FINE
    timeit("sample_coarse", lambda: _lib.sample_coarse(rays, kc, False, seed=1), r * (8 + 4 * kc))
    z_c = _lib.sample_coarse(rays, kc, False, seed=1)
    w_c = torch.rand(r, kc, device=dev)
    d_c = torch.full((r,), 1.3, device=dev)
    timeit("sample_fine", lambda: _lib.sample_fine(rays, z_c, w_c, d_c, kf, kd, 0.01, seed=2),
           r * (4 * kc + 4 * kc + 4 + 8 + 4 * k2))
    z_f = _lib.sample_fine(rays, z_c, w_c, d_c, kf, kd, 0.01, seed=2)
    for k, z, tag in ((kc, z_c, "coarse"), (k2, z_f, "fine")):
        rs = torch.rand(r, k, 4, device=dev)
        timeit(f"composite_fwd_{tag}", lambda: _lib.composite_fwd(rs, z, rays, True), r * (k * 20 + 4 + k * 4 + 16))
        g_rgb = torch.rand(r, 3, device=dev)
        timeit(f"composite_bwd_{tag}", lambda: _lib.composite_bwd(rs, z, rays, True, g_rgb, want_gz=False),
               r * (k * 20 + 4 + 12 + k * 16))
        del rs
    # point features at the bench frame's size (rows = 16384 rays x 96 samples), fp16 out
    sc = syn.scene(SCENE)
    net, _ = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision="fp16")
    fr = sc["rays"].to(dev)
    zf = z_f[: fr.shape[0]].contiguous()
    lat = net.latent_nhwc(torch.float16)
    rows = fr.shape[0] * k2
    timeit("point_features_fp16", lambda: _lib.point_features(net._cam, lat, rows, rays=fr, z=zf, ldx=64),
           rows * (512 + 64) * 2 + fr.shape[0] * (32 + 4 * k2) + lat.numel() * 2)
    # what the tensor-core path runs instead: positional code + taps record (128 B/row), no latent traffic
    timeit("point_taps", lambda: _lib.point_taps(net._cam, lat.shape[1], lat.shape[2], rows, rays=fr, z=zf),
           rows * 128 + fr.shape[0] * (32 + 4 * k2))
    del rays, z_c, z_f, w_c, d_c
    torch.cuda.empty_cache()
    return {"kernels": res, "hbm_peak_GB/s": pk["hbm"], "peak_source": pk["source"], "rays": r,
            "bound": "hbm", "timing": "CUDA events around 10 back-to-back launches after 3 warm-up launches; working "
                                      "sets of 0.3-2 GB exceed the 126 MB L2"}


def frame_rays(sc, target_az, per_ray_bounds=False):
    """One target frame of scene `sc` at azimuth `target_az`: (H*W, 8) host tensor."""
    from pixelnerf_b200 import synthetic as syn
    rays = syn.gen_rays(syn.pose_spherical(target_az, -10.0, sc["radius"]), sc["w"], sc["h"], sc["focal"], sc["near"],
                        sc["far"], sc.get("c")).reshape(-1, 8).contiguous()
    if per_ray_bounds:   # near / far columns differ per ray
        g = torch.Generator().manual_seed(4)
        rays[:, 6] = 0.1 + torch.rand(rays.shape[0], generator=g) * 0.4
        rays[:, 7] = 5.0 - torch.rand(rays.shape[0], generator=g) * 0.5
    return rays


class Job:
    """One process-group-wide measurement context: device, ranks, timing helpers."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback; see --impl reference)"
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def timed(self, fn, steps):
        """K steps, CUDA events on the launching stream, L2 flushed between steps (untimed); MAX over ranks."""
        total = 0.0
        for _ in range(steps):
            self.flush.zero_()
            if self.world > 1:
                self.dist.barrier()      # every step starts together: a step ends with the slowest rank's gather
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            total += a.elapsed_time(b)
        t = torch.tensor([total], device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def bound(self, sc, precision):
        """Model + renderer of scene `sc` on this rank, the encoded scene coming from rank 0 only (one NCCL broadcast
        per encode), bound over every GPU of the job through the product's own bind_parallel."""
        from pixelnerf_b200 import parallel, synthetic as syn
        from pixelnerf_b200.testing import build_net
        wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
        net, ren = build_net(sc, wc, wf, self.dev, mlp_precision=precision)
        if self.world > 1:
            if self.rank != 0:   # only rank 0 "encoded": wipe the others so the broadcast is what they render from
                net.encoder.latent = torch.zeros(1, 1, 1, 1, device=self.dev)
                net.poses = torch.zeros(1, 3, 4, device=self.dev)
            parallel.broadcast_encoded(net, src=0)
        gpus = list(range(self.world)) if self.world > 1 else [self.local]
        par = ren.bind_parallel(net, gpus, simple_output=True).eval()
        if self.world > 1:
            assert type(par).__name__ == "RaySharded", type(par).__name__
        return net, ren, par, (wc, wf)


def sharded_extras(job, args):
    """BASELINE configs 3 and 4 through the same ray-sharded wrapper, STRONG scaling (the work is fixed, every call's
    rays are split over the job's GPUs): one DTU 300x400 NS=3 frame with per-ray bounds (120 000 rays) and the
    40-frame SRN-chair NS=2 video (655 360 rays in one call: contiguous ranges = whole frames per GPU)."""
    from pixelnerf_b200 import _lib, synthetic as syn
    out = {}
    reps = 3
    for key, scene, n_frames in (("dtu_frame", "dtu", 1), ("srn_chair_video40", "srn_chair", 40)):
        sc = syn.scene(scene)
        net, ren, par, _ = job.bound(sc, args.precision)
        az = [0.0] if n_frames == 1 else [float(a) for a in torch.linspace(-180, 180, n_frames + 1)[:-1]]
        rays = torch.cat([frame_rays(sc, a, per_ray_bounds=(scene == "dtu")) for a in az])[None].to(job.dev)
        n = rays.shape[1]
        host = torch.empty(n, 3, dtype=torch.uint8).pin_memory()

        def call():
            with torch.no_grad():
                rgb, depth = par(rays)
            return rgb

        def call_e2e():      # + the frame output path: 8-bit conversion on the device, then the copy
            host.copy_(_lib.frame_to_u8(call()[0]), non_blocking=True)
            torch.cuda.current_stream().synchronize()

        for _ in range(2):
            call()
        job.barrier()
        ms = job.timed(call, reps) / reps
        ms_e2e = job.timed(call_e2e, reps) / reps
        ns = sc["ns"]
        flop = n * (N_COARSE + N_COARSE + N_FINE) * MLP_FLOP_PER_POINT[ns]
        out[key] = {"workload": WORKLOADS[scene] + (f" x {n_frames} frames in one call" if n_frames > 1 else ""),
                    "rays": n, "scaling": "strong", "n_gpus": job.world, "rays_per_gpu": n // job.world,
                    "ms_per_call": ms, "rays/s": n / (ms * 1e-3), "e2e_rays/s": n / (ms_e2e * 1e-3),
                    "mlp_algorithmic_TFLOP/s_per_gpu": flop / (ms * 1e-3) / 1e12 / job.world, "reps": reps}
        del net, ren, par, rays
        torch.cuda.empty_cache()
    return out


def live_parity(sc, weights, precision, n_rays=256):
    """Max-abs error of this build on `n_rays` rays of the bench frame against the CPU oracle, stage-wise (coarse
    pass on the same draws; fine pass teacher-forced with the oracle's coarse weights / depth): rgb, weights and
    depth in WORLD UNITS.  The checker, not the product path."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import build_net, oracle_render
    wc, wf = weights
    idx = torch.linspace(0, sc["rays"].shape[0] - 1, n_rays).long()
    rays = sc["rays"][idx]
    draws = syn.draws(11, n_rays, N_COARSE, N_FINE - N_DEPTH, N_DEPTH)
    ref = oracle_render(sc, wc, wf, rays, draws, N_COARSE, N_FINE, N_DEPTH)
    dev = torch.device("cuda", torch.cuda.current_device())
    net, ren = build_net(sc, wc, wf, dev, mlp_precision=precision)
    ren.injected_draws = {k: v.to(dev) for k, v in draws.items()}
    ren.injected_z = dict(w_coarse=ref["w_coarse"].to(dev), depth_coarse=ref["depth_coarse"].to(dev))
    with torch.no_grad():
        out = ren(net, rays[None].to(dev), want_weights=True)
    z_c, z_f = ren.last_z
    res = {"rays": n_rays, "against": "CPU oracle (pinned to the live reference by tests/golden)", "stagewise": True,
           "sample_positions_bit_exact": bool(torch.equal(z_c[0].cpu(), ref["z_coarse"]) and torch.equal(z_f[0].cpu(), ref["z_fine"])),
           "tolerance": 2e-3 if precision in ("fp16", "fp16x2") else 1e-4, "depth_unit": "world units (absolute)"}
    for part, tag in ((out.coarse, "coarse"), (out.fine, "fine")):
        res[f"rgb_{tag}"] = float((part.rgb[0].cpu() - ref["rgb_" + tag]).abs().max())
        res[f"weights_{tag}"] = float((part.weights[0].cpu() - ref["w_" + tag]).abs().max())
        res[f"depth_{tag}"] = float((part.depth[0].cpu() - ref["depth_" + tag]).abs().max())
    return res


def run_gpu(args):
    from pixelnerf_b200 import _lib, synthetic as syn
    job = Job()
    world, rank, dev = job.world, job.rank, job.dev
    sc = syn.scene(args.scene)
    net, ren, par, (wc, wf) = job.bound(sc, args.precision)
    # Headline workload: `world` target frames of the scene in ONE call (one frame's worth of rays per GPU: weak
    # scaling), split over the GPUs by the product's own wrapper (RaySharded: contiguous ray ranges in, one
    # all-gather of the rendered values out).  At N = 1 this is exactly BASELINE configs[1]: one 128x128 frame.
    n_frame = sc["rays"].shape[0]
    n_rays = world * n_frame
    rays_host = torch.cat([frame_rays(sc, 10.0 * i, per_ray_bounds=(args.scene == "dtu")) for i in range(world)])
    rays_host = rays_host.reshape(1, n_rays, 8).contiguous().pin_memory()
    rays_dev = rays_host.to(dev)
    out_host = torch.empty(n_rays, 4).pin_memory()

    def step(rays):
        with torch.no_grad():
            rgb, depth = par(rays)
        return rgb, depth

    for _ in range(max(args.warmup, 3)):
        step(rays_dev)
    job.barrier()
    sampler = ClockSampler(job.local)
    sampler.start()
    _lib.PROFILE = {}
    launches0 = _lib.launch_count()
    ms = job.timed(lambda: step(rays_dev), args.steps)
    job.barrier()
    launches = _lib.launch_count() - launches0
    prof = {k: sum(a.elapsed_time(b) for a, b in v) for k, v in _lib.PROFILE.items()}
    n_mlp = len(_lib.PROFILE.get("mlp", []))
    _lib.PROFILE = None

    def e2e_step():
        # public API with HOST buffers: the rays start in pinned host memory (under torchrun each rank copies only
        # its own range: RaySharded slices before the H2D copy), rgb + depth end in pinned host memory on every rank
        rgb, depth = step(rays_host if world > 1 else rays_host.to(dev, non_blocking=True))
        out_host[:, :3].copy_(rgb[0], non_blocking=True)
        out_host[:, 3].copy_(depth[0], non_blocking=True)
        torch.cuda.current_stream().synchronize()

    e2e_step()
    job.barrier()
    ms_e2e = job.timed(e2e_step, args.steps)
    job.barrier()
    sampler.stop_flag = True
    sampler.join(timeout=2)

    extras = sharded_extras(job, args) if (args.extras and args.scene == SCENE) else None

    if rank == 0:
        pk = peaks()
        ns = sc["ns"]
        rays_total = n_rays * args.steps
        flop_rank = n_frame * (N_COARSE + N_COARSE + N_FINE) * MLP_FLOP_PER_POINT[ns]     # this rank's share of a step
        mlp_ms = prof.get("mlp", 0.0) / max(1, args.steps)
        achieved = flop_rank / (mlp_ms * 1e-3) / 1e12 if mlp_ms > 0 else 0.0
        # share of the algorithmic FLOPs the fp16 chain executes as MMAs (lin_z goes through tables): MMA counts
        # per 128 rows at NS=1 (padding included), MAC counts for NS > 1
        exe = 1.0
        if args.precision == "fp16x3":      # three MMAs per product of the ten 512-wide GEMMs (+ the discarded lin_in)
            exe = 3.0 * (ns * 6 * 262144 + 4 * 262144) / (ns * 2380800 + 1050624)
        if args.precision == "fp16x2":      # two MMAs per product of the ten 512-wide GEMMs
            exe = 2.0 * (ns * 6 * 262144 + 4 * 262144) / (ns * 2380800 + 1050624)
        if args.precision == "fp16":
            exe = 686.0 / 878.0 if ns == 1 else (ns * (6 * 262144 + 21504) + 1050624) / (ns * 2380800 + 1050624)
        launches_per_step = n_mlp // max(1, args.steps)
        points_rank = n_frame * (N_COARSE + N_COARSE + N_FINE)
        line = {
            "metric": "rays/sec coarse+fine", "value": rays_total / (ms * 1e-3), "unit": "rays/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp32": "f32", "fp16": "f16", "fp16x2": "f16x2", "fp16x3": "f16x3"}[args.precision], "data": "synthetic",
            "config": {"workload": f"{WORKLOADS[args.scene]}: {world} frame(s) per step = {n_frame} rays per GPU, "
                                   f"{N_COARSE} coarse + {N_FINE} fine ({N_DEPTH} depth) samples, ResnetFC 5x512 x2",
                       "mlp_precision": args.precision,
                       "parallelism": ("one call of %d rays through renderer.bind_parallel(net, gpus) -> RaySharded: "
                                       "contiguous ray ranges per rank, encoded scene broadcast from rank 0, one "
                                       "all_gather_into_tensor of rgb+depth per call" % n_rays) if world > 1 else
                                      "single GPU: renderer.bind_parallel(net, [0])",
                       "cache": "256 MiB L2 flush between steps",
                       "prepared_once": "weights packed, latent transposed, per-texel lin_z tables (0.3 ms) at encode time"},
            "e2e": {"value": rays_total / (ms_e2e * 1e-3), "unit": "rays/s",
                    "h2d_bytes_per_step": n_rays * 32, "d2h_bytes_per_step": world * n_rays * 16,
                    "note": "rays from pinned host memory (each rank copies its own range), rgb+depth of the whole "
                            "call copied to pinned host memory on every rank"},
            "gpu_launches": launches,
            "clocks": sampler.summary(),
            # `achieved` counts the ALGORITHMIC FLOPs of the reference MLP, lin_z included; the fp16
            # chain applies lin_z through per-texel tables and executes 686 of those 878 MMAs per
            # 128 rows, so `achieved` can exceed the hardware peak: `executed` is what the tensor pipe really did
            "roofline": {"bound": "tensor", "kernel": "ResnetFC chain (%s), rank 0" % args.precision, "achieved": achieved,
                         "executed": achieved * exe,
                         "executed_frac": achieved * exe / pk["tf_sustained"],
                         "peak": pk["tf_sustained"], "unit": "TFLOP/s", "frac": achieved / pk["tf_sustained"],
                         "traffic": chain_traffic(ns, points_rank, launches_per_step) if args.precision == "fp16" else None,
                         "traffic_source": "profiles/r02_chain_traffic.json (from profiles/r02m_chain_ncu_summary.json): ncu dram read+write bytes of the chain "
                                           "launches of one pass at THIS view count, per point, x points per step / "
                                           "launches per step (a live run cannot read dram counters)",
                         "peak_source": pk["source"] + ", sustained bf16",
                         "launches_per_step": launches_per_step, "ms_per_step": mlp_ms},
            "stage_ms_per_step": {k: v / args.steps for k, v in prof.items()},
        }
        if extras is not None:
            line["sharded"] = extras
        if world == 1 and not args.no_cpu_baseline:
            sample = 8192
            v, secs = cpu_oracle_rays_per_s(sc, wc, wf, sample)
            line["cpu_baseline"] = {"value": v, "unit": "rays/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"{sample} evenly spaced rays of the same frame, {secs:.1f} s, torch "
                                              f"{torch.__version__} CPU"}
        if world == 1 and args.extras:
            line["parity"] = live_parity(sc, (wc, wf), args.precision)
            line["roofline_kernels"] = kernel_table()
            line["gpu_baseline"] = torch_cuda_frame(sc, args.scene, steps=2, warmup=1)
        print(json.dumps(line))
    if world > 1:
        job.dist.destroy_process_group()


def run_in_process(args):
    """--in-process: ONE process driving `--gpus N` devices through renderer.bind_parallel(net, [0..N-1]) ->
    parallel.DeviceSharded -- what the reference's scripts do with several --gpu_id values (eval/gen_video.py:110: a
    single process, DataParallel in the reference).  Same workloads as the torchrun form: N SRN-car frames per call
    (weak), one DTU frame and the 40-frame NS=2 video (strong).  Wall-clock around a synchronise of every device (the
    devices are driven by host threads; CUDA events of one stream do not span them).  Not the headline line."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import build_net
    n_dev = args.gpus
    assert torch.cuda.device_count() >= n_dev, f"--in-process --gpus {n_dev} needs {n_dev} visible devices"
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)

    def sync_all():
        for i in range(n_dev):
            torch.cuda.synchronize(i)

    out = {}
    for key, scene, n_frames in (("srn_car_frames", "srn_car", n_dev), ("dtu_frame", "dtu", 1), ("srn_chair_video40", "srn_chair", 40)):
        sc = syn.scene(scene)
        net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision=args.precision)
        par = ren.bind_parallel(net, list(range(n_dev)), simple_output=True).eval()
        az = [10.0 * i for i in range(n_frames)] if scene != "srn_chair" else [float(a) for a in torch.linspace(-180, 180, n_frames + 1)[:-1]]
        rays = torch.cat([frame_rays(sc, a, per_ray_bounds=(scene == "dtu")) for a in az])[None].to(dev)
        n = rays.shape[1]
        with torch.no_grad():
            for _ in range(3):
                par(rays)
            sync_all()
            reps = max(3, min(args.steps, 10))
            t0 = time.perf_counter()
            for _ in range(reps):
                rgb, depth = par(rays)
            sync_all()
            dt = (time.perf_counter() - t0) / reps
        assert rgb.device == dev and torch.isfinite(rgb).all()
        out[key] = {"workload": WORKLOADS[scene] + f" x {n_frames} frame(s) in one call", "rays": n, "n_gpus": n_dev,
                    "ms_per_call": dt * 1e3, "rays/s": n / dt, "reps": reps}
        del net, ren, par, rays
        torch.cuda.empty_cache()
    print(json.dumps({"metric": "rays/sec coarse+fine, one process driving N GPUs (bind_parallel(net, gpu_ids) -> DeviceSharded)",
                      "unit": "rays/s", "n_gpus": n_dev, "timing": "wall clock around a synchronise of every device",
                      "mlp_precision": args.precision, "workloads": out}))


def run_train(args):
    """--train: BASELINE config 5, one training step = render 4 objects x `--train-rays` rays (want_weights, train
    mode) -> MSE on coarse+fine rgb vs U(0,1) targets -> backward through composite / ResnetFC / latent gather
    (hand-written kernels behind torch.autograd.Function) -> Adam step on both MLPs.  Not the headline line."""
    from pixelnerf_b200 import parallel, synthetic as syn
    from pixelnerf_b200.testing import build_net, scene_camera, set_encoded
    # under torchrun (--gpus N): the same batch ray-sharded over N GPUs (RaySharded: every rank renders and
    # back-propagates its own ray range), weight gradients summed by parallel.allreduce_gradients (one flat NCCL
    # all-reduce per dtype), the same Adam step on every rank -- strong scaling of one training step
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        assert not args.graph, "--graph is the single-GPU form"
    sb = 4
    sc = syn.scene(SCENE)
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision=args.precision)
    cam = scene_camera(sc)
    lat = syn.latent_maps(21, sb, sc["hl"], sc["wl"])
    set_encoded(net, lat, cam["poses"].expand(sb, -1, -1).contiguous(), cam["focal"], cam["c"], cam["image_shape"], 1)
    net.train()
    ren.train()
    for q in net.parameters():
        q.requires_grad_(True)
    opt = torch.optim.Adam([q for n, q in net.named_parameters() if n.startswith("mlp_")], lr=1e-4,
                           capturable=bool(args.graph))
    if args.graph:
        ren.use_device_seed(dev)      # the draws' call seed in device memory, advanced by a captured add
    res = []
    for n_obj_rays in args.train_rays:
        g = torch.Generator().manual_seed(n_obj_rays)
        idx = torch.randint(0, sc["rays"].shape[0], (sb, n_obj_rays), generator=g)
        rays = sc["rays"][idx].to(dev)                       # (SB, B, 8)
        target = torch.rand(sb, n_obj_rays, 3, generator=g).to(dev)
        par = ren.bind_parallel(net, list(range(world))) if world > 1 else None

        def step():
            opt.zero_grad(set_to_none=True)
            if par is not None:
                out = par(rays, want_weights=True)
                loss = ((out["coarse"]["rgb"] - target) ** 2).mean() + ((out["fine"]["rgb"] - target) ** 2).mean()
                loss.backward()
                parallel.allreduce_gradients([net.mlp_coarse, net.mlp_fine])
            else:
                out = ren(net, rays, want_weights=True)
                loss = ((out.coarse.rgb - target) ** 2).mean() + ((out.fine.rgb - target) ** 2).mean()
                loss.backward()
            opt.step()
            return loss

        for _ in range(2):
            step()
        torch.cuda.synchronize()
        steps = max(1, min(args.steps, 5))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        a.record()
        for _ in range(steps):
            loss = step()
        b.record()
        b.synchronize()
        ms = a.elapsed_time(b) / steps
        if world > 1:      # the slowest rank's time
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        n = sb * n_obj_rays
        flop = 3.0 * n * (N_COARSE + N_COARSE + N_FINE) * MLP_FLOP_PER_POINT[1]
        # one more step with CUDA events around every kernel family (not part of the timed region)
        from pixelnerf_b200 import _lib
        _lib.PROFILE = {}
        step()
        torch.cuda.synchronize()
        fam = {k: round(sum(e0.elapsed_time(e1) for e0, e1 in v), 3) for k, v in _lib.PROFILE.items()}
        _lib.PROFILE = None
        res.append({"rays_per_step": n, "ms_per_step": ms, "rays/s": n / (ms * 1e-3), "mlp_TFLOP/s_fwd+bwd": flop / (ms * 1e-3) / 1e12,
                    "loss": float(loss), "kernel_ms": fam})
        if args.graph:
            import gc
            # the same step captured once in a CUDA graph and replayed (tests/test_gpu_graph.py: same loss curve).
            # The eager steps' autograd graph must be gone first: it keeps the parameters' gradient accumulators alive,
            # and those remember the (legacy default) stream they were made on, which may not be touched while capturing
            loss = None
            gc.collect()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    step()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                g_loss = step()
            for _ in range(3):
                graph.replay()
            torch.cuda.synchronize()
            reps = max(5, steps)
            a.record()
            for _ in range(reps):
                graph.replay()
            b.record()
            b.synchronize()
            g_ms = a.elapsed_time(b) / reps
            res[-1]["cuda_graph"] = {"ms_per_step": g_ms, "rays/s": n / (g_ms * 1e-3), "loss": float(g_loss),
                                     "speedup_vs_eager": ms / g_ms}
            del graph
            net.invalidate()
    backward = ("tcgen05 dgrad/wgrad GEMMs, bf16 gradients, forward chain re-run per chunk keeping its operands"
                if args.precision != "fp32" else "fp32 CUDA-core kernels, recompute per chunk")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
        if rank != 0:
            return
    print(json.dumps({"metric": "training step rays/sec (config 5)", "unit": "rays/s", "objects": sb, "n_gpus": world,
                      "parallelism": "single GPU" if world == 1 else
                      f"one batch ray-sharded over {world} GPUs (RaySharded), gradients all-reduced, strong scaling",
                      "forward": "%s chain" % args.precision, "backward": backward,
                      "optimizer": "Adam on both MLPs", "sweep": res}))


def chain_traffic(ns, points_per_step, launches_per_step):
    """DRAM bytes per chain launch: ncu-measured dram read + write bytes per POINT of the chain launches at this view
    count (profiles/r02_chain_traffic.json, one `ncu --set full` capture per NS; a live run cannot read dram
    counters) x the points one step of this run processes / its chain launches; None without the summary."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_chain_traffic.json")
    try:
        with open(path) as f:
            per_point = json.load(f)["dram_bytes_per_point"][str(ns)]
        return per_point * points_per_step / max(1, launches_per_step)
    except (OSError, KeyError, ValueError):
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "torch-cuda"])
    ap.add_argument("--precision", default=os.environ.get("PIXELNERF_B200_MLP", "fp16"), choices=["fp32", "fp16", "fp16x2", "fp16x3"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", dest="extras", action="store_false",
                    help="skip the extra keys of the headline line: `sharded` (DTU frame + 40-frame video, strong "
                         "scaling), and at N=1 `parity`, `roofline_kernels`, `gpu_baseline`")
    ap.add_argument("--scene", default=SCENE, choices=sorted(WORKLOADS),
                    help="srn_car = the headline workload (BASELINE configs[1]); the others are extra measurements")
    ap.add_argument("--kernels", action="store_true", help="per-kernel HBM roofline of the memory-bound kernels")
    ap.add_argument("--in-process", action="store_true",
                    help="one process driving --gpus N devices through bind_parallel(net, gpu_ids) (DeviceSharded); not the headline")
    ap.add_argument("--train", action="store_true", help="training-step timing (BASELINE config 5), not the headline")
    ap.add_argument("--train-rays", type=int, nargs="+", default=[128, 4096], help="rays per object for --train")
    ap.add_argument("--graph", action="store_true",
                    help="--train: also capture the step in a CUDA graph (device-resident call seed) and time replays")
    args = ap.parse_args()
    if args.in_process:
        run_in_process(args)
    elif args.kernels:
        run_kernels(args)
    elif args.train and args.impl == "torch-cuda":
        run_torch_cuda(args)
    elif args.train:
        run_train(args)
    elif args.impl == "reference":
        run_reference(args)
    elif args.impl == "torch-cuda":
        run_torch_cuda(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
