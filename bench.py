#!/usr/bin/env python
"""bench.py - Mrays/s of the ray-tracing hot path on N B200s, with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one full frame of the workload traced through the wavefront kernels
(primary + shadow + reflection rays), from camera/light parameters to the quantised 8-bit
image on GPU 0.  `value` is device-timed (CUDA events) with the scene resident in HBM; `e2e` is
the same frame through the public C ABI call rtp_render() with HOST output buffers (camera and
lights uploaded, image downloaded inside the timed region).  With N > 1 the image is split into
interleaved 32x32 screen tiles (strong scaling: total work fixed), the scene is replicated, and
tiles are gathered to GPU 0 with one NCCL gather per frame.

Workloads (BASELINE.json configs; --workload):
  c3 (default) synthetic terrain, 1 002 528 triangles, 1920x1080, 1 light, max_depth 1
               (primary + shadow rays) - the config the >= 1 Grays/s-per-B200 target is quoted on
  c2           shipped default scene (dodgeColorTest, 16 311 triangles), 1920x1080, 4 lights,
               max_depth 4 (reflection depth 3)
  c1           plane.obj, 256x256, 1 light, max_depth 2
  c4           dodgeColorTest instanced 25x25 = 10 194 375 triangles, 3840x2160, 4 lights, max_depth 6
  c5           incoherent stress: 64 ambient-occlusion rays per pixel of the C3 frame (any-hit queries
               through rtp_trace_rays_device), rays sharded over the GPUs
"""
import argparse
import importlib
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
METRIC = "Mrays/sec (primary+secondary)"
UNIT = "Mrays/s"


def b_alg(n_tris):
    """algorithmic bytes per ray for a BVH2 with 64-B nodes and <= 4-triangle
    leaves of 48 B/triangle: one root-to-leaf descent, one leaf, one 16-B hit record."""
    return 64 * math.ceil(math.log2(max(n_tris / 4.0, 2.0))) + 4 * 48 + 16


def load_golden_mesh(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", f"mesh_{name}.npz"))
    return {k: z[k] for k in ("verts", "normals", "faces", "face_mat", "materials", "shape")}


def build_workload(name, synth):
    if name == "c3":
        m = synth.terrain(708)
        return dict(mesh=m, w=1920, h=1080, max_depth=1, lights=synth.LIGHT_DEFAULT,
                    view=synth.look_at((0.0, 0.30, 0.20), (0.0, -0.05, -0.08)),
                    desc="C3: synthetic fBm terrain 708x708 cells = 1002528 triangles, 1920x1080, 1 point light, "
                         "max_depth 1 (primary+shadow), camera eye (0,.30,.20) -> (0,-.05,-.08), ~93% coverage")
    if name == "c2":
        m = load_golden_mesh("dodgeColorTest")
        c = np.load(os.path.join(ROOT, "tests", "golden", "case_dodge_c2.npz"))
        return dict(mesh=m, w=1920, h=1080, max_depth=4, lights=synth.LIGHTS_C2, view=c["view"],
                    desc="C2: dodgeColorTest.obj (16311 triangles), 1920x1080, 4 point lights, max_depth 4")
    if name == "c1":
        m = load_golden_mesh("plane")
        return dict(mesh=m, w=256, h=256, max_depth=2, lights=synth.LIGHT_DEFAULT, view=None,
                    desc="C1: plane.obj (2 triangles), 256x256, 1 point light, max_depth 2, default camera")
    if name == "c4":
        m = synth.instanced(load_golden_mesh("dodgeColorTest"), 25, 25)
        return dict(mesh=m, w=3840, h=2160, max_depth=6, lights=synth.LIGHTS_C2,
                    view=synth.look_at((0.55, 0.45, 0.7), (0.0, -0.1, 0.0)),
                    desc="C4: dodgeColorTest.obj instanced 25x25 = 10194375 triangles, 3840x2160, 4 point lights, "
                         "max_depth 6 (reflection depth 5), camera eye (.55,.45,.7) -> (0,-.1,0)")
    if name == "c5":
        m = synth.terrain(708)
        return dict(mesh=m, w=1920, h=1080, max_depth=1, lights=synth.LIGHT_DEFAULT, ao=64, eye=(0.0, 0.30, 0.20),
                    view=synth.look_at((0.0, 0.30, 0.20), (0.0, -0.05, -0.08)),
                    desc="C5: incoherent stress on the C3 terrain (1002528 triangles): from every primary hit point of the "
                         "1920x1080 frame (moved 0.003 along the interpolated normal) 64 directions uniform on the "
                         "hemisphere around that normal, Philox4x32-10 keyed by (seed 1234, pixel index): the same ray "
                         "set for every GPU count; any-hit queries (= Flyscene::shadow)")
    raise SystemExit(f"unknown workload {name}")


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def roofline_dict(workload, kernel, kernel_ms, rays, n_tris, sm_count, clocks, peaks, extra=None, alg_extra=0):
    """The line's `roofline` for the dominant kernel.

    Primary bound = ISSUE RATE: the traversal kernels are bound by warp-instruction issue under divergence, not by HBM
    (measured DRAM traffic is ~5 % of the algorithmic bytes: neighbouring rays share nodes through L1/L2).
      achieved = warp instructions of the launch (ncu smsp__inst_executed.sum, profiles/traffic.json, scaled to this
                 launch's ray count) / its CUDA-event duration measured live here;
      peak     = SMs x 4 schedulers x SM clock (one warp instruction per scheduler per cycle), clock = median under load.
    `lane_efficiency` (ncu thread_inst / inst / 32) says how many of those slots carry a live lane; `dram_frac` is the
    measured DRAM traffic over the same duration against MEASURED_PEAKS.json hbm_gbs.
    `contract` keeps SURVEY.md 8(d)'s HBM figure (algorithmic bytes per ray x rays / time / hbm peak): it exceeds 1 for
    coherent rays because the caches serve most of those bytes - it is reported, it does not bound the kernel."""
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"
    alg = b_alg(n_tris) + alg_extra
    sec = kernel_ms / 1e3 if kernel_ms > 0 else float("inf")
    contract = dict(bound="hbm", achieved=rays * alg / sec / 1e9, peak=hbm, unit="GB/s", frac=rays * alg / sec / 1e9 / hbm,
                    alg_bytes_per_ray=alg, peak_source=peak_src,
                    note="SURVEY.md 8(d) definition; not a bound for this kernel (see roofline.dram_frac)")
    sm_clock = float((clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0) * 1e6
    issue_peak = sm_count * 4 * sm_clock
    ent = None
    try:
        ent = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(workload, {}).get(kernel)
    except Exception:
        pass
    if not isinstance(ent, dict):
        ent = None
    r = dict(bound="issue", unit="Gwarp-inst/s", peak=issue_peak / 1e9, kernel=kernel, kernel_ms=kernel_ms,
             rays_per_launch_set=rays, grays_per_s_kernel=rays / sec / 1e9,
             peak_source=f"{sm_count} SMs x 4 schedulers x {sm_clock / 1e6:.0f} MHz (median SM clock under load)")
    if ent and ent.get("warp_inst") and ent.get("rays"):
        scale = rays / float(ent["rays"])
        winst = ent["warp_inst"] * scale
        traffic = ent["dram_bytes"] * scale
        r.update(achieved=winst / sec / 1e9, frac=winst / sec / issue_peak, traffic=traffic,
                 warp_inst_per_launch=winst, lane_efficiency=ent["lanes"] / 32.0,
                 useful_lane_frac=winst / sec / issue_peak * ent["lanes"] / 32.0,
                 dram_gbs=traffic / sec / 1e9, dram_frac=traffic / sec / 1e9 / hbm, dram_peak_source=peak_src,
                 traffic_bytes_per_ray=traffic / rays if rays else None,
                 counters_from=ent.get("from", "profiles/traffic.json") + (f"; scaled x{scale:.4f} to this launch's rays" if abs(scale - 1) > 1e-3 else ""))
    else:
        r.update(achieved=None, frac=None, traffic=None, note="no ncu capture of this kernel for this workload in profiles/traffic.json")
    r["contract"] = contract
    if extra:
        r.update(extra)
    return r


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md).  Sampled in-process
    through NVML every few ms (the timed region of the default run is well under a second, too short
    for `nvidia-smi -lms`); falls back to one nvidia-smi query when NVML is unavailable."""
    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, gpu_index):
        self.rows, self.gpu, self.stop_flag, self.th, self.h, self.nv = [], gpu_index, False, None, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.gpu]) if vis and vis.split(",")[self.gpu].strip().isdigit() else self.gpu
            self.h, self.nv = pynvml.nvmlDeviceGetHandleByIndex(idx), pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None
            return
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()

    def _run(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append((time.perf_counter(), float(mhz), int(reasons)))
            except Exception:
                pass
            time.sleep(0.001)

    def stop(self, t_begin=None, t_end=None):
        self.stop_flag = True
        if self.th:
            self.th.join(timeout=2)
        if self.h is None:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=clocks.sm,clocks.max.sm",
                                      "--format=csv,noheader,nounits"], stdout=subprocess.PIPE, text=True, timeout=20).stdout
                a, b = [float(x) for x in out.strip().split(",")]
                return dict(sm_mhz=a, sm_max_mhz=b, reasons=[], samples=1, how="nvidia-smi after the timed region (NVML unavailable)")
            except Exception:
                return dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        rows = [r for r in self.rows if t_begin is None or t_begin <= r[0] <= t_end] or self.rows
        sm = [r[1] for r in rows]
        mask = 0
        for r in rows:
            mask |= r[2]
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=self.max_mhz,
                    reasons=sorted(k for k, bit in self.BAD.items() if mask & bit), samples=len(sm),
                    how="NVML, 1 ms period, samples inside the timed region")


def cpu_baseline_port(wl, n_rays_target=None, seconds=15.0):
    """The oracle (plain-C port of the reference path, all host threads) on a bounded pixel
    subsample of the same frame.  Reported baseline only."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po
    m = wl["mesh"]
    t0 = time.perf_counter()
    mesh = po.Mesh(m["verts"], m["normals"], m["faces"], m["face_mat"], m["materials"], None, m["shape"]).build_boxes()
    t_boxes = time.perf_counter() - t0
    cores = os.cpu_count() or 1
    w, h = wl["w"], wl["h"]
    rng = np.random.default_rng(1234)

    def sample(n):
        idx = rng.choice(w * h, size=min(n, w * h), replace=False)
        return np.stack([idx % w, idx // w], 1).astype(np.int32)
    px = sample(64 * cores)  # calibration
    t0 = time.perf_counter()
    o = po.render(mesh, w, h, view=wl["view"], max_depth=wl["max_depth"], point_lights=wl["lights"], pixels=px, threads=cores)
    dt = max(time.perf_counter() - t0, 1e-6)
    n = int(min(w * h, max(len(px), len(px) * seconds / dt)))
    px = sample(n)
    t0 = time.perf_counter()
    o = po.render(mesh, w, h, view=wl["view"], max_depth=wl["max_depth"], point_lights=wl["lights"], pixels=px, threads=cores)
    dt = time.perf_counter() - t0
    rays = sum(o["counts"][k] for k in ("primary", "shadow", "reflection"))
    return dict(value=rays / dt / 1e6, unit=UNIT, cores=cores, kind="port",
                sample=f"{len(px)} random pixels of the {w}x{h} frame ({rays} rays) in {dt:.2f} s, oracle/oracle.c "
                       f"with {cores} pthreads; BoundingBox build {t_boxes:.1f} s untimed",
                note="the same all-boxes x all-faces algorithm as the reference arm (--impl reference, oracle/_ref/ref_harness) and ~8x "
                     "faster than it on the same cores: plain arrays instead of a Tucano::Face deep copy per face visit, no matrix "
                     "inverse per intersectBox call")


def run_reference(args, wl, synth):
    """--impl reference: the reference's own CPU implementation (oracle/_ref/ref_harness = the
    unmodified flyscene.cpp + BoundingBox.cpp) on all host cores, bounded pixel sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po
    cores = os.cpu_count() or 1
    w, h = wl["w"], wl["h"]
    rng = np.random.default_rng(1234)
    reps = args.steps + args.warmup
    if po.ref_available():
        import tempfile
        tmp = tempfile.mkdtemp(prefix="rtp_bench_ref_")
        obj = synth.write_obj(wl["mesh"], os.path.join(tmp, "workload.obj"))
        # calibrate on a tiny sample, then size the per-step sample for ~60 s total
        def run(n, repeat):
            idx = rng.choice(w * h, size=min(n, w * h), replace=False)
            px = np.stack([idx % w, idx // w], 1).astype(np.int32)
            return po.ref_run(obj, w, h, view=wl["view"], max_depth=wl["max_depth"], point_lights=wl["lights"],
                              pixels=px, procs=cores, repeat=repeat, chain=True)
        r = run(8 * cores, 1)
        per_px = max(r["pass_seconds"][-1], 1e-6) / (8 * cores)
        n = int(min(w * h, max(8 * cores, 60.0 / reps / per_px)))
        r = run(n, reps)
        rays = sum(r["counts"][k] for k in ("primary", "shadow", "reflection"))
        times = r["pass_seconds"][args.warmup:]
        kind, what = "reference", "oracle/_ref/ref_harness (reference flyscene.cpp + BoundingBox.cpp, -O2)"
    else:
        m = wl["mesh"]
        mesh = po.Mesh(m["verts"], m["normals"], m["faces"], m["face_mat"], m["materials"], None, m["shape"]).build_boxes()
        idx = rng.choice(w * h, size=min(64 * cores, w * h), replace=False)
        px = np.stack([idx % w, idx // w], 1).astype(np.int32)
        t0 = time.perf_counter()
        po.render(mesh, w, h, view=wl["view"], max_depth=wl["max_depth"], point_lights=wl["lights"], pixels=px, threads=cores)
        per_px = max(time.perf_counter() - t0, 1e-6) / len(px)
        n = int(min(w * h, max(len(px), 60.0 / reps / per_px)))
        idx = rng.choice(w * h, size=n, replace=False)
        px = np.stack([idx % w, idx // w], 1).astype(np.int32)
        times = []
        for it in range(reps):
            t0 = time.perf_counter()
            o = po.render(mesh, w, h, view=wl["view"], max_depth=wl["max_depth"], point_lights=wl["lights"], pixels=px, threads=cores)
            if it >= args.warmup:
                times.append(time.perf_counter() - t0)
        rays = sum(o["counts"][k] for k in ("primary", "shadow", "reflection"))
        kind, what = "port", "oracle/oracle.c (reference binary not present on this box)"
    total = sum(times)
    value = rays * len(times) / total / 1e6
    sample = f"{n} random pixels of the {w}x{h} frame per step ({rays} rays), {what}, {cores} processes/threads"
    line = dict(impl="reference", metric=METRIC, value=value, unit=UNIT, n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=1e3 * total / len(times), higher_is_better=True, scaling="strong",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=wl["desc"], impl="reference CPU path on host cores"),
                cpu_baseline=dict(value=value, unit=UNIT, cores=cores, kind=kind, sample=sample),
                e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


def run_ours(args, wl, rtp):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the hot path is CUDA only (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    m, w, h, md = wl["mesh"], wl["w"], wl["h"], wl["max_depth"]
    t0 = time.perf_counter()
    scene = rtp.Scene(m["verts"], m["normals"], m["faces"], m["face_mat"], m["materials"], m["shape"], device=local)
    t_build = time.perf_counter() - t0
    scene.set_lights(wl["lights"], None)
    scene.set_camera(w, h, wl["view"])
    info = scene.info()
    mg = importlib.import_module("ray-tracing-project_b200.multigpu")
    # the frame lands in GPU 0's image: --gather peer (default) = every rank's resolve kernel stores its
    # tiles there over NVLink (CUDA IPC), --gather nccl = tile buffers + one NCCL gather + un-tile kernel
    peer = args.gather == "peer"
    tg = mg.TileGather(w, h, rank, world, dev, rtp)
    img = mg.PeerImage(w, h, rank, world, local, rtp)
    rgb8 = torch.zeros(w * h * 3, dtype=torch.uint8, device=dev) if rank == 0 else None
    rgb8_host = torch.zeros(w * h * 3, dtype=torch.uint8).pin_memory() if rank == 0 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream().cuda_stream

    def step(profile=False):
        flags = rtp.RENDER_PROFILE if profile else 0
        if peer:
            mg.render_frame_peer(scene, img, md, stream, flags=flags, barrier=args.barrier)
        else:
            mg.render_frame(scene, tg, md, stream, d_rgb8=rgb8.data_ptr() if rank == 0 else None, flags=flags)

    # both routes must deliver the same bytes
    mg.render_frame(scene, tg, md, stream, d_rgb8=rgb8.data_ptr() if rank == 0 else None)
    mg.render_frame_peer(scene, img, md, stream, barrier=args.barrier)
    torch.cuda.synchronize()
    same = bool(torch.equal(img.rgb8_tensor(img.frame - 1).reshape(-1), rgb8)) if rank == 0 else True
    if not same:
        raise SystemExit("bench.py: the peer-store image differs from the NCCL-gathered image")

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
    sync_all()
    st = scene.frame_stats()
    counts = torch.tensor([st["rays_primary"], st["rays_shadow"], st["rays_reflection"], st["hits"]],
                          dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(counts)
    rays_frame = int(counts[:3].sum().item())

    t_begin = time.perf_counter()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms_steps, k_closest, k_any, k_shade, k_cam, k_paths = [], [], [], [], [], []
    # rays of the camera level alone (k_camera's launch): one untimed frame at max_depth 1 on this rank's tiles
    scene.render_tiles_device(w, h, tg.tiles.data_ptr(), max_depth=1, rank=rank, world=world, stream=stream)
    torch.cuda.synchronize()
    st1 = scene.frame_stats()
    rays_camera = float(st1["rays_primary"] + st1["rays_shadow"])
    for _ in range(args.steps):
        flush.zero_()          # evict the scene from L2 between timed iterations (untimed)
        sync_all()
        if peer:
            mg.align_start(scene, img, stream)  # N > 1: all GPUs leave the start line together (device-side, untimed)
        ev0.record()
        step()
        ev1.record()
        sync_all()
        t = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_steps.append(float(t.item()))
    clocks = sampler.stop(t_begin, time.perf_counter()) if rank == 0 else None
    fs = scene.frame_stats()
    groups = scene.tile_groups()
    launches_rank0 = (fs["kernel_launches"] + (0 if peer else 1)) * args.steps  # frame kernels (+ un-tile on rank 0)
    # per-kernel durations: the same steps once more with RTP_RENDER_PROFILE, which brackets every launch with
    # CUDA events on the launch stream and runs the frame as ONE tile group, so that an event pair times one
    # kernel alone on the GPU (in the timed region above kernels of different groups overlap)
    ms_serial = []
    for _ in range(args.steps):
        flush.zero_()
        sync_all()
        ev0.record()
        step(profile=True)
        ev1.record()
        sync_all()
        ms_serial.append(ev0.elapsed_time(ev1))
        fs = scene.frame_stats()
        k_closest.append(fs["ms_closest"])
        k_any.append(fs["ms_anyhit"])
        k_shade.append(fs["ms_shade"])
        k_cam.append(fs["ms_camera"])
        k_paths.append(fs["ms_paths"])

    # end to end through the public C ABI with host buffers (inputs uploaded, image downloaded)
    e2e_ms = []
    cam_bytes = 4 * (12 + 4 + 2) + 24 * len(wl["lights"])
    for it in range(args.steps + 1):
        sync_all()
        t0 = time.perf_counter()
        scene.set_lights(wl["lights"], None)
        scene.set_camera(w, h, wl["view"])
        if world == 1:
            scene.render(w, h, max_depth=md, want=("rgb8",), out={"rgb8": rgb8_host.numpy().reshape(h, w, 3)})
        else:
            step()
            if rank == 0:
                src = img.rgb8_tensor(img.frame - 1).reshape(-1) if peer else rgb8
                rgb8_host.copy_(src, non_blocking=True)
        sync_all()
        if it > 0:
            t = torch.tensor([1e3 * (time.perf_counter() - t0)], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms.append(float(t.item()))

    # the same end-to-end frames, pipelined: rtp_render_async with two frames in flight (double-buffered image
    # planes, band copies on their own streams), every frame's image landing in page-locked host memory
    pipe = None
    if world == 1:
        host2 = [rgb8_host, torch.zeros(w * h * 3, dtype=torch.uint8).pin_memory()]
        n_pipe = args.steps
        for it in range(2):  # first pass warms the slots up
            sync_all()
            t0 = time.perf_counter()
            for k in range(n_pipe):
                if k >= 2:
                    scene.render_wait()
                scene.set_lights(wl["lights"], None)
                scene.set_camera(w, h, wl["view"])
                scene.render_async(w, h, host2[k & 1].data_ptr(), None, max_depth=md)
            for _ in range(min(2, n_pipe)):
                scene.render_wait()
            sync_all()
            t_pipe = time.perf_counter() - t0
        pipe = dict(value=rays_frame * n_pipe / t_pipe / 1e6, unit=UNIT, ms_per_step=1e3 * t_pipe / n_pipe,
                    what="rtp_render_async / rtp_render_wait, 2 frames in flight, lights + camera set per frame, "
                         "host uint8 image per frame",
                    image_equals_sync=bool(torch.equal(host2[0], host2[1])))

    if world > 1 and scene.peer_barrier_timeouts() != 0:
        raise SystemExit("bench.py: a frame barrier timed out (a rank did not arrive)")
    if rank == 0:
        total_ms = sum(ms_steps)
        value = rays_frame * args.steps / (total_ms / 1e3) / 1e6
        e2e_val = rays_frame * len(e2e_ms) / (sum(e2e_ms) / 1e3) / 1e6
        # roofline of the dominant kernel (rank 0's launches)
        peaks = load_peaks()
        n_tris = info["n_triangles"]
        mc, ms_, mo = float(np.mean(k_closest)), float(np.mean(k_any)), float(np.mean(k_shade))
        mcam = float(np.mean(k_cam))
        # rays of each kind of launch: with the camera level in k_camera, the per-level closest / any-hit kernels only
        # see the rays below it
        fusedcam = mcam > 0
        own = (float(st["rays_primary"] + st["rays_reflection"]) - (float(st["rays_primary"]) if fusedcam else 0.0),
               float(st["rays_shadow"]) - ((rays_camera - float(st["rays_primary"])) if fusedcam else 0.0))
        mpaths = float(np.mean(k_paths))
        if mpaths > 0:  # levels >= 1 ran in k_paths: the per-level kernels saw the camera level only
            own = (float(st["rays_primary"]) if not fusedcam else 0.0, (rays_camera - float(st["rays_primary"])) if not fusedcam else 0.0)
        rays_paths = float(st["rays_primary"] + st["rays_reflection"] + st["rays_shadow"]) - rays_camera if mpaths > 0 else 0.0
        cands = [("k_camera", mcam, rays_camera if fusedcam else 0.0), ("k_trace<closest>", mc, own[0]), ("k_trace<anyhit>", ms_, own[1]),
                 ("k_paths", mpaths, rays_paths)]
        dom, dom_ms, dom_rays = max(cands, key=lambda c: c[1])
        roof = roofline_dict(args.workload, dom, dom_ms, dom_rays, n_tris, torch.cuda.get_device_properties(local).multi_processor_count,
                             clocks, peaks,
                             extra=dict(kernels_ms=dict(camera=mcam, closest=mc, anyhit=ms_, paths=mpaths, streaming=mo),
                                        rays_by_kind=dict(camera=rays_camera if fusedcam else 0.0, closest=own[0], anyhit=own[1], paths=rays_paths),
                                        kernels_ms_from=f"{args.steps} more steps with RTP_RENDER_PROFILE: per-kernel CUDA events on the launch "
                                                        f"stream, the frame as one tile group (each kernel alone on the GPU, k_coop counted with its "
                                                        f"k_trace, k_coop_camera with k_camera), {float(np.mean(ms_serial)):.4f} ms per step"))
        line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                    ms_per_step=total_ms / args.steps, higher_is_better=True, scaling="strong", vs_baseline=None,
                    dtype="f32", data="synthetic",
                    config=dict(workload=wl["desc"], rays_per_frame=rays_frame, triangles=n_tris, bvh_nodes=info["n_bvh_nodes"],
                                parallelism=f"tile{world}" if world > 1 else "1gpu",
                                gather=("resolve kernel stores tiles into GPU 0's image (CUDA IPC peer memory) + " +
                                        ("own one-thread barrier kernel on a counter in GPU 0's memory" if args.barrier == "peer"
                                         else "4-byte NCCL all-reduce")
                                        if peer else "NCCL gather of tile buffers + un-tile kernel") if world > 1 else
                                       "resolve kernel writes the image (1 GPU)" if peer else "un-tile kernel (1 GPU)",
                                peer_image_equals_nccl_gather=same,
                                tile_groups=f"{groups} bands of this rank's tiles, each on its own CUDA stream",
                                l2="256 MiB memset between timed steps (untimed) flushes the 126 MB L2",
                                timed_region="per step: host barrier + device sync, then on every GPU [start-line kernel (N > 1, untimed): "
                                             "waits until all GPUs are there] - start event - frame (+ frame barrier on N > 1) - end event; "
                                             "max over ranks of the event time",
                                scene_build_s=round(t_build, 2)),
                    clocks=clocks,
                    e2e=dict(value=e2e_val, unit=UNIT, h2d_bytes_per_step=cam_bytes, d2h_bytes_per_step=w * h * 3,
                             ms_per_step=float(np.mean(e2e_ms)),
                             what="rtp_set_lights + rtp_set_camera + rtp_render -> host uint8 image" if world == 1 else
                                  "set lights/camera + tile render into GPU 0's image + D2H of the uint8 image on rank 0",
                             pipelined=pipe),
                    gpu_launches=launches_rank0,
                    roofline=roof)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_port(wl)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
    img.close()
    scene.close()
    if world > 1:
        dist.destroy_process_group()


def run_multi_device(args, wl, rtp):
    """`python bench.py --gpus N` WITHOUT torchrun: one process, N GPUs, through the drop-in's own multi-GPU entry point
    (rtp_scene_create_multi + rtp_render = Flyscene::DEVICES / `raytrace --gpus N`): one host thread and one stream per
    GPU inside the library, every GPU's resolve kernel storing its tiles into GPU 0's image, fork / join by CUDA events.
    `value` = rays of the frame / device time of the whole frame on GPU 0's stream (its end event follows the join with
    every other GPU); `e2e` = the same call by wall clock, image in host memory."""
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the hot path is CUDA only (no CPU fallback)")
    n = args.gpus
    ndev = rtp.device_count()
    devices = [r % ndev for r in range(n)]
    m, w, h, md = wl["mesh"], wl["w"], wl["h"], wl["max_depth"]
    t0 = time.perf_counter()
    scene = rtp.Scene(m["verts"], m["normals"], m["faces"], m["face_mat"], m["materials"], m["shape"], devices=devices)
    t_build = time.perf_counter() - t0
    one = rtp.Scene(m["verts"], m["normals"], m["faces"], m["face_mat"], m["materials"], m["shape"], device=0)
    for sc in (scene, one):
        sc.set_lights(wl["lights"], None)
        sc.set_camera(w, h, wl["view"])
    info = scene.info()
    host = torch.zeros(w * h * 3, dtype=torch.uint8).pin_memory()
    host1 = torch.zeros(w * h * 3, dtype=torch.uint8).pin_memory()
    st1 = one.render(w, h, max_depth=1, want=("rgb8",), out={"rgb8": host1.numpy().reshape(h, w, 3)})["stats"]
    rays_camera = float(st1["rays_primary"] + st1["rays_shadow"])  # rays of the camera level (k_camera's launches), whole frame
    one.render(w, h, max_depth=md, want=("rgb8",), out={"rgb8": host1.numpy().reshape(h, w, 3)})
    one.close()
    flushes = [torch.empty(256 << 20, dtype=torch.uint8, device=torch.device("cuda", d)) for d in sorted(set(devices))]

    def sync_all():
        for d in sorted(set(devices)):
            torch.cuda.synchronize(d)

    def frame(flags=0):
        return scene.render(w, h, max_depth=md, want=("rgb8",), out={"rgb8": host.numpy().reshape(h, w, 3)}, flags=flags)["stats"]

    sampler = ClockSampler(devices[0])
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        st = frame()
    same = bool(torch.equal(host, host1))
    if not same:
        raise SystemExit("bench.py: the multi-device image differs from the one-device image")
    rays_frame = st["rays_primary"] + st["rays_shadow"] + st["rays_reflection"]
    t_begin = time.perf_counter()
    ms_steps, e2e_ms = [], []
    for _ in range(args.steps):
        for f in flushes:
            f.zero_()
        sync_all()
        t0 = time.perf_counter()
        scene.set_lights(wl["lights"], None)
        scene.set_camera(w, h, wl["view"])
        st = frame()
        e2e_ms.append(1e3 * (time.perf_counter() - t0))
        ms_steps.append(st["ms_trace"])
    clocks = sampler.stop(t_begin, time.perf_counter())
    k_closest, k_any, k_shade, k_cam = [], [], [], []
    for _ in range(min(args.steps, 20)):
        for f in flushes:
            f.zero_()
        sync_all()
        ps = frame(flags=rtp.RENDER_PROFILE)
        k_closest.append(ps["ms_closest"]); k_any.append(ps["ms_anyhit"]); k_shade.append(ps["ms_shade"]); k_cam.append(ps["ms_camera"])
    total_ms = sum(ms_steps)
    peaks = load_peaks()
    mc, ms_, mo, mcam = float(np.mean(k_closest)), float(np.mean(k_any)), float(np.mean(k_shade)), float(np.mean(k_cam))
    fusedcam = mcam > 0
    own = (float(st["rays_primary"] + st["rays_reflection"]) - (float(st["rays_primary"]) if fusedcam else 0.0),
           float(st["rays_shadow"]) - ((rays_camera - float(st["rays_primary"])) if fusedcam else 0.0))
    cands = [("k_camera", mcam, rays_camera if fusedcam else 0.0), ("k_trace<closest>", mc, own[0]), ("k_trace<anyhit>", ms_, own[1])]
    dom, dom_ms, dom_rays = max(cands, key=lambda c: c[1])
    dom_rays /= n
    roof = roofline_dict(args.workload, dom, dom_ms, dom_rays, info["n_triangles"],
                         torch.cuda.get_device_properties(devices[0]).multi_processor_count, clocks, peaks,
                         extra=dict(kernels_ms=dict(camera=mcam, closest=mc, anyhit=ms_, streaming=mo),
                                    kernels_ms_from="more steps with RTP_RENDER_PROFILE: per-kernel CUDA events, max over the devices; "
                                                    "rays_per_launch_set = the frame's rays of that kind / devices"))
    cam_bytes = 4 * (12 + 4 + 2) + 24 * len(wl["lights"])
    line = dict(metric=METRIC, value=rays_frame * args.steps / (total_ms / 1e3) / 1e6, unit=UNIT, n_gpus=n, steps=args.steps,
                warmup=max(args.warmup, 3), ms_per_step=total_ms / args.steps, higher_is_better=True, scaling="strong",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=wl["desc"], rays_per_frame=rays_frame, triangles=info["n_triangles"], bvh_nodes=info["n_bvh_nodes"],
                            parallelism=f"tile{n}, ONE process: rtp_scene_create_multi on devices {devices}",
                            gather="every device's resolve kernel stores its tiles into device 0's image (peer access), join by CUDA events",
                            multi_device_image_equals_one_device=same,
                            l2="256 MiB memset per device between timed steps (untimed)", scene_build_s=round(t_build, 2),
                            timed_region="CUDA events on device 0's stream around the frame: fork event - all devices' kernels - join; "
                                         "rtp_stats.ms_trace of rtp_render"),
                clocks=clocks,
                e2e=dict(value=rays_frame * len(e2e_ms) / (sum(e2e_ms) / 1e3) / 1e6, unit=UNIT, h2d_bytes_per_step=cam_bytes,
                         d2h_bytes_per_step=w * h * 3, ms_per_step=float(np.mean(e2e_ms)),
                         what="rtp_set_lights + rtp_set_camera + rtp_render (multi-device scene) -> host uint8 image, wall clock"),
                gpu_launches=int(st["kernel_launches"]) * args.steps, roofline=roof)
    print(json.dumps(line))
    scene.close()


def run_c5(args, wl, rtp):
    """C5: 64 ambient-occlusion rays per primary hit, traced as any-hit queries through
    rtp_trace_rays_device; the hit points are split into one contiguous share per GPU (no data-path
    collective), the per-pixel occlusion sums are gathered to rank 0."""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the hot path is CUDA only (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    m, w, h, n_ao = wl["mesh"], wl["w"], wl["h"], wl["ao"]
    scene = rtp.Scene(m["verts"], m["normals"], m["faces"], m["face_mat"], m["materials"], m["shape"], device=local)
    scene.set_lights(wl["lights"], None)
    scene.set_camera(w, h, wl["view"])
    info = scene.info()
    stream = torch.cuda.current_stream().cuda_stream
    # workload generation (untimed): primary hit points and interpolated normals of the frame, then the rays of this
    # rank's contiguous share of the hit pixels from the counter-based generator (synth.ao_rays_pixels: keyed by the
    # pixel index, so the union over the ranks is the same ray set whatever the GPU count)
    synth = importlib.import_module("ray-tracing-project_b200.synth")
    out = scene.render(w, h, max_depth=1, want=("hit",))
    hit = out["hit"].ravel()
    keep = np.nonzero(hit >= 0)[0]
    share = np.array_split(keep, world)[rank]
    eye, dirs = synth.camera_rays(w, h, wl["view"])
    P_np = synth.primary_hit_points(m, hit[share], eye, dirs[share])
    N_np = synth.shading_normals(m, hit[share], P_np)
    del dirs
    P_t, N_t, pix_t = torch.from_numpy(P_np).to(dev), torch.from_numpy(N_np).to(dev), torch.from_numpy(share.astype(np.int64)).to(dev)
    rays = torch.empty((len(share) * n_ao, 6), dtype=torch.float32, device=dev)
    for c0 in range(0, len(share), 1 << 17):  # chunks bound the generator's int64 temporaries
        c1 = min(len(share), c0 + (1 << 17))
        rays[c0 * n_ao:c1 * n_ao] = synth.ao_rays_pixels(P_t[c0:c1], N_t[c0:c1], pix_t[c0:c1], n_per_point=n_ao, xp=torch)
    ray_crc = int(rays.view(torch.int32).to(torch.int64).sum().item()) & 0xffffffffffff  # order-free checksum of the share
    n_rays = rays.shape[0]
    occ = torch.zeros(n_rays, dtype=torch.uint8, device=dev)
    ao = torch.zeros(len(share), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    cnt = torch.tensor([len(share)], dtype=torch.int64, device=dev)
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    if world > 1:
        dist.all_gather(sizes, cnt)
    maxn = int(max(int(x.item()) for x in sizes)) if world > 1 else len(share)
    pad = torch.zeros(maxn, dtype=torch.float32, device=dev)
    gathered = [torch.zeros(maxn, dtype=torch.float32, device=dev) for _ in range(world)] if (world > 1 and rank == 0) else None

    def step():
        scene.trace_rays_device(rays.data_ptr(), n_rays, rtp.RAYS_ANYHIT, d_occ=occ.data_ptr(), stream=stream)
        torch.sum(occ.view(-1, n_ao), dim=1, dtype=torch.float32, out=ao)
        if world > 1:
            pad[:len(share)] = ao
            dist.gather(pad, gathered, dst=0)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
    sync_all()
    total = torch.tensor([n_rays, ray_crc], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(total)
    rays_step = int(total[0].item())
    ray_crc = int(total[1].item()) & 0xffffffffffff
    t_begin = time.perf_counter()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms_steps = []
    for _ in range(args.steps):
        flush.zero_()
        sync_all()
        ev0.record()
        step()
        ev1.record()
        sync_all()
        t = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_steps.append(float(t.item()))
    clocks = sampler.stop(t_begin, time.perf_counter()) if rank == 0 else None
    k_list = []  # the trace kernel alone (this rank's launch), by the library's CUDA events on the launch stream
    for _ in range(3):
        flush.zero_()
        sync_all()
        k_list.append(scene.trace_rays_device(rays.data_ptr(), n_rays, rtp.RAYS_ANYHIT, d_occ=occ.data_ptr(), stream=stream,
                                              want_stats=True)["ms_trace"])
    k_ms = float(np.mean(k_list))
    # end to end: rays from pinned host memory, occlusion flags back to the host (a bounded slice: the full list is 3 GB)
    n_e2e = min(n_rays, 8 << 20)
    h_rays = rays[:n_e2e].cpu().pin_memory()
    h_occ = torch.zeros(n_e2e, dtype=torch.uint8).pin_memory()
    d_rays2 = torch.empty_like(rays[:n_e2e])
    e2e_ms = []
    for it in range(4):
        sync_all()
        t0 = time.perf_counter()
        d_rays2.copy_(h_rays, non_blocking=True)
        scene.trace_rays_device(d_rays2.data_ptr(), n_e2e, rtp.RAYS_ANYHIT, d_occ=occ.data_ptr(), stream=stream)
        h_occ.copy_(occ[:n_e2e], non_blocking=True)
        sync_all()
        if it > 0:
            e2e_ms.append(1e3 * (time.perf_counter() - t0))
    if rank == 0:
        total_ms = sum(ms_steps)
        value = rays_step * args.steps / (total_ms / 1e3) / 1e6
        peaks = load_peaks()
        n_tris = info["n_triangles"]
        roof = roofline_dict("c5", "k_trace<RaysAny>", k_ms, float(n_rays), n_tris, torch.cuda.get_device_properties(local).multi_processor_count,
                             clocks, peaks, alg_extra=24,
                             extra=dict(kernel_ms_from="3 more steps timed by the library's CUDA events around the trace launch alone "
                                                       "(rtp_trace_rays_device with statistics), L2 flushed before each"))
        line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                    ms_per_step=total_ms / args.steps, higher_is_better=True, scaling="strong", vs_baseline=None,
                    dtype="f32", data="synthetic",
                    config=dict(workload=wl["desc"], rays_per_step=rays_step, triangles=n_tris, bvh_nodes=info["n_bvh_nodes"],
                                ray_set_checksum=f"{ray_crc:012x} (sum of the rays' float bits over all ranks, mod 2^48: equal for every GPU count)",
                                parallelism=f"rays{world}" if world > 1 else "1gpu",
                                l2="256 MiB memset between timed steps (untimed); the 3.2 GB ray list alone exceeds the L2"),
                    clocks=clocks,
                    e2e=dict(value=n_e2e * world * len(e2e_ms) / (sum(e2e_ms) / 1e3) / 1e6, unit=UNIT,
                             h2d_bytes_per_step=n_e2e * 24, d2h_bytes_per_step=n_e2e, ms_per_step=float(np.mean(e2e_ms)),
                             what=f"{n_e2e} rays per GPU from pinned host memory -> rtp_trace_rays_device(any-hit) -> occlusion bytes to the host"),
                    gpu_launches=args.steps,
                    roofline=roof)
        print(json.dumps(line))
    scene.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"])
    ap.add_argument("--barrier", default="peer", choices=["peer", "nccl"],
                    help="frame barrier of the peer-store route: the library's own kernel over peer memory, or a 4-byte NCCL all-reduce")
    args = ap.parse_args()
    rtp = importlib.import_module("ray-tracing-project_b200")
    synth = importlib.import_module("ray-tracing-project_b200.synth")
    wl = build_workload(args.workload, synth)
    if args.impl == "reference":
        run_reference(args, wl, synth)
    elif args.workload == "c5":
        run_c5(args, wl, rtp)
    elif args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        run_multi_device(args, wl, rtp)  # one process, N devices: the C++ multi-GPU path of the drop-in
    else:
        run_ours(args, wl, rtp)


if __name__ == "__main__":
    main()
