#!/usr/bin/env python
"""bench.py — throughput of the render path on B200 (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): Cornell box (Mitsuba scene rebuilt procedurally by the host layer, 36
triangles / 8 instances / 2 triangle lights), 1920x1080, NEE + MIS, max depth 8.  One step = `--spp-per-step`
(default 128) samples for every pixel on every GPU (8 steps = the 1024 spp of the config on one GPU).  With N GPUs
the samples are sharded by index (GPU g renders timestamps base + g, base + g + N, ...): per-GPU work is fixed
(weak scaling) and every step ends with one NCCL reduce of the float4 sum buffers to rank 0.

Printed line: metric = Mpaths/s (pixel-samples per second, whole job), plus Mrays/s, `e2e` (same step driven through
the reference-shaped per-frame API with host buffers: per-frame scene arrays up, HDR framebuffer down), `roofline` of
the dominant kernel (closest-hit traversal, k_extend), `cpu_baseline` (the CPU oracle on the host cores).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "Mpaths/s"
UNIT = "Mpaths/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp-per-step", type=int, default=128)
    ap.add_argument("--max-depth", type=int, default=8)
    ap.add_argument("--scene", default="cornell", help="cornell (configs[1]) | gallery (configs[2]) | field (configs[3])")
    ap.add_argument("--scene-params", default="", help="comma-separated integers for the built-in scene (field: seed,meshes,subdiv,grid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-instrumented", action="store_true", help="skip the untimed fetch-count pass (profiler runs)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-extra", action="store_true", help="skip the short runs of configs[2], [3] and the 4096x2160 frame (N = 1 only)")
    return ap.parse_args()


def workload_name(a):
    """The workload both arms run (BASELINE.json configs[1] by default); how much of it one step covers is in config.spp_per_step."""
    return f"{a.scene}{' box' if a.scene == 'cornell' else ' scene'} {a.width}x{a.height}, NEE+MIS, max_depth {a.max_depth}"


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (the recipe's clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ reference arm
def host_threads():
    """Every host core this process may run on; torchrun exports OMP_NUM_THREADS=1, which the CPU arm must not inherit."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(a, rank, world):
    """The reference has no CPU implementation of this path and its Vulkan path is unavailable on this box.  What of it
    does compile for the host is its own shader text (oracle/_ref/libgsref.so: pt_common.glsl, raygen.rgen, rayhit.rchit,
    miss / shadowmiss compiled as C++ against the reference's glm, OpenMP over pixels; the ray/triangle search, which has
    no source in the reference, is the oracle's intersector): that is the reference arm (kind "reference") when the
    library is there; otherwise the oracle port of the shaders (kind "port").  Each step is a bounded sample of the
    workload: the full frame at 1 spp."""
    if rank != 0:
        return
    from oracle import oracle as orc
    from oracle import ref as oref
    orc.build()
    archive = os.path.join(ROOT, "tests", "golden", f"scene_{a.scene}.npz")
    use_ref = oref.available() and a.scene != "gallery"      # the gallery holds a rough dielectric: no reference source
    if os.path.exists(archive) and not a.scene_params:
        # the scene's flat records from a committed archive (tests/golden/make_scene_npz.py): nothing of the product is
        # imported or loaded on this arm
        o = orc.OracleScene.from_npz(archive)
        r = oref.RefScene.from_npz(archive) if use_ref else None
        scene_source = os.path.relpath(archive, ROOT)
    else:
        import gpuspectral_b200 as g      # no archive for this scene: the host layer builds the arrays (no product compute)
        hs = g.Scene.builtin(a.scene, *scene_params(a))
        o = orc.OracleScene.from_host_scene(hs)
        r = oref.RefScene.from_host_scene(hs) if use_ref else None
        scene_source = "host layer (gpuspectral_b200.Scene.builtin)"
    o.set_params(max_depth=a.max_depth)
    if use_ref:
        oref.set_max_depth(a.max_depth)
    spp = 1
    ncpu = host_threads()
    img = np.zeros((a.height, a.width, 4), np.float32)

    def one(k):
        """One step: timestamp k for every pixel -> (paths, extension rays, shadow rays, threads used)."""
        if use_ref:
            _, e, s_ = r.render(o, a.width, a.height, k, spp, image=img, threads=ncpu)     # the reference's running mean
            return a.width * a.height * spp, e, s_, ncpu
        _, c, th = o.render(a.width, a.height, k, spp, 1, image=img, sum_mode=True, threads=ncpu)
        return c["paths"], c["extension_rays"], c["shadow_rays"], th

    paths = ext = sh = threads = 0
    for k in range(a.warmup):
        one(k)
    t0 = time.perf_counter()
    for k in range(a.steps):
        p_, e_, s_, threads = one(a.warmup + k)
        paths += p_; ext += e_; sh += s_
    dt = time.perf_counter() - t0
    val = paths / dt / 1e6
    # the oracle port on the same sample, for the record (it is the faster of the two CPU codes)
    port = None
    if use_ref:
        t1 = time.perf_counter()
        _, c, _ = o.render(a.width, a.height, a.warmup + a.steps, spp, 1, sum_mode=True, threads=ncpu)
        port = c["paths"] / (time.perf_counter() - t1) / 1e6
    kind = "reference" if use_ref else "port"
    what = ("the reference's shader text compiled as C++ (oracle/_ref), ray/triangle search by the oracle's intersector" if use_ref
            else "CPU oracle (port of the shaders, pinned bit for bit to the reference's shader text compiled as C++)")
    sample = f"{a.width}x{a.height} x {spp} spp per step ({a.steps} steps), same scene/params"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dt / max(a.steps, 1) * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "mrays_per_s": (ext + sh) / dt / 1e6,
            "config": {"workload": workload_name(a), "width": a.width, "height": a.height, "spp_per_step": spp, "max_depth": a.max_depth,
                       "scene": a.scene, "scene_source": scene_source, "parallelism": f"{threads} host threads (OpenMP over pixels)",
                       "note": what + "; the reference's Vulkan path is unavailable on this box; one step = the full frame at 1 spp (bounded sample)"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample, "oracle_port_value": port},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ ours
class CudaBuffer:
    """__cuda_array_interface__ view of the context's accumulation buffer (no copy)."""

    def __init__(self, ptr, nfloats):
        self.__cuda_array_interface__ = {"shape": (nfloats,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}


def scene_params(a):
    return [int(x) for x in a.scene_params.split(",") if x.strip()]


def scene_h2d_bytes(scene):
    import gpuspectral_b200 as g
    n = scene.instances[1] * 88 + scene.lights[1] * 64 + (3 + 16 + 1) * 4 + 16
    for t in range(g.BSDF_TYPE_COUNT):
        n += scene.bsdfs(t)[1] * g.BSDF_STRIDES[t]
    return n


def per_frame_upload(ctx, scene, mesh_map):
    """What PathTracer::createRenderPass / prepareScene do every frame in the reference (PathTracer.cpp:9-93):
    instance / BSDF / light arrays + camera from HOST memory, top-level structure rebuilt.  Meshes (BLAS) are cached,
    as Renderer::getOrCreateBLAS does."""
    import gpuspectral_b200 as g
    for t in range(g.BSDF_TYPE_COUNT):
        raw, n = scene.bsdfs(t)
        ctx.set_bsdfs(t, raw, n)
    lights, nl = scene.lights
    ctx.set_lights(lights, nl)
    inst, ni = scene.instances
    for i in range(ni):
        inst[i].mesh = mesh_map[inst[i].mesh]
    ctx.set_instances(inst, ni)
    ctx.set_camera(*scene.camera)


def load_peaks():
    """Roofline denominators: HBM copy bandwidth from MEASURED_PEAKS.json (driver-written), L2 read bandwidth and FP32 FMA
    rate from profiles/peaks_b200.json (lib/gsb_peaks run on a B200 of this pool; SURVEY 8d: not in MEASURED_PEAKS.json)."""
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        hbm, hbm_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy)"
    else:
        hbm, hbm_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    l2 = fp32 = None
    ppath = os.path.join(ROOT, "profiles", "peaks_b200.json")
    if os.path.exists(ppath):
        pk = json.load(open(ppath))
        l2, fp32 = pk.get("l2_read_gbs"), pk.get("fp32_fma_tflops")
    return {"hbm": hbm, "hbm_src": hbm_src, "l2": l2, "fp32": fp32}


def fetch_counts(ctx, W, H):
    """Untimed instrumented passes: nodes / triangles / instance records fetched per extension ray and per shadow ray,
    on the rays this workload really traces (the instrumented variant of the same kernels)."""
    out = []
    for mode in (1, 2):
        ctx.set_instrumented(mode)
        ctx.reset_counters()
        ctx.render(W, H, 1 << 20, 2, 1)
        c = ctx.counters()
        out.append({k: c[k] / max(c["rays_counted"], 1) for k in ("nodes_visited", "tris_tested", "inst_entered")})
    ctx.set_instrumented(0)
    ctx.clear()
    return out


def ray_bytes(fetch, stream_bytes):
    """80 B per 8-wide node, 48 B per triangle, 64 B per instance record + the wavefront streams of the ray."""
    return 80 * fetch["nodes_visited"] + 48 * fetch["tris_tested"] + 64 * fetch["inst_entered"] + stream_bytes


# bytes a shaded vertex moves through k_shade: 72 in (ray origin / direction, hit record, weight,
# instance word), 48 out (next ray, weight), 48 for its shadow-queue entry
SHADE_BYTES_PER_VERTEX = 168


def traversal_roofline(bvh, peaks, ext_fetch, sh_fetch, stage, iters, ext_rays, sh_rays, vertices):
    """`roofline` of the dominant kernel (k_extend) + the side figures of k_shadow and k_shade, from per-stage CUDA-event
    times on the render stream.  Bound: L2 when the BVH fits the 126 MB L2 with room to spare (configs[1], [2]), HBM
    otherwise (configs[3]) — SURVEY 8(d)."""
    # a ray that leaves the scene also reads its path's sample (16 B) and adds it to the image (16 B read-modify-write) in k_extend
    miss_frac = max(0.0, 1.0 - vertices / ext_rays) if ext_rays else 0.0
    ext_b, sh_b = ray_bytes(ext_fetch, 32 + 24 + miss_frac * 48), ray_bytes(sh_fetch, 32 + 4)
    ext_ms, sh_ms, shade_ms = stage["extend"] / max(iters, 1), stage["shadow"] / max(iters, 1), stage["shade"] / max(iters, 1)
    achieved = ext_b * (ext_rays / max(iters, 1)) / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else 0.0
    sh_achieved = sh_b * (sh_rays / max(iters, 1)) / (sh_ms * 1e-3) / 1e9 if sh_ms > 0 else 0.0
    in_l2 = bvh.get("bvh_bytes", 0) < 64 * 1024 * 1024 and peaks["l2"]
    peak = peaks["l2"] if in_l2 else peaks["hbm"]
    total = sum(stage.values())
    shade_gbs = SHADE_BYTES_PER_VERTEX * vertices / (stage["shade"] * 1e-3) / 1e9 if stage["shade"] > 0 else 0.0
    ops = None
    opath = os.path.join(ROOT, "profiles", "shade_ops.json")
    if os.path.exists(opath):
        ops = json.load(open(opath))
    flops_v = ops.get("flops_per_vertex") if ops else None
    roof = {"kernel": "k_extend (closest-hit traversal)", "bound": "l2" if in_l2 else "hbm", "achieved": achieved, "peak": peak,
            "unit": "GB/s", "frac": achieved / peak if peak else None,
            "peak_source": "profiles/peaks_b200.json l2_read_gbs (lib/gsb_peaks on a B200 of this pool)" if in_l2 else peaks["hbm_src"],
            "frac_of_hbm_peak": achieved / peaks["hbm"], "hbm_peak": peaks["hbm"], "l2_peak": peaks["l2"],
            "frac_of_l2_peak": (achieved / peaks["l2"]) if peaks["l2"] else None,
            "bytes_per_ray": ext_b, "rays_per_launch": ext_rays / max(iters, 1), "ms_per_launch": ext_ms,
            "fetch_per_ray": ext_fetch, "miss_fraction": miss_frac, "launches": iters,
            "share_of_step": {k: v / total for k, v in stage.items()} if total > 0 else None,
            "shadow": {"bytes_per_ray": sh_b, "fetch_per_ray": sh_fetch, "achieved": sh_achieved, "ms_per_launch": sh_ms,
                       "frac": sh_achieved / peak if peak else None},
            "note": "achieved = algorithmic bytes (instrumented fetch counts x record sizes + the ray's wavefront streams) per second "
                    "of k_extend; the kernel is issue-bound, its DRAM traffic is `traffic`"}
    shade = {"kernel": "k_shade<TYPE> (all BSDF types of the scene)", "bytes_per_vertex": SHADE_BYTES_PER_VERTEX,
             "vertices_per_launch": vertices / max(iters, 1), "ms_per_launch": shade_ms, "achieved_gbs": shade_gbs,
             "frac_hbm": shade_gbs / peaks["hbm"], "flops_per_vertex": flops_v,
             "achieved_tflops": (flops_v * vertices / (stage["shade"] * 1e-3) / 1e12) if flops_v and stage["shade"] > 0 else None,
             "fp32_peak_tflops": peaks["fp32"],
             "frac_fp32": (flops_v * vertices / (stage["shade"] * 1e-3) / 1e12 / peaks["fp32"]) if flops_v and peaks["fp32"] and stage["shade"] > 0 else None,
             "flops_source": ops.get("source") if ops else None}
    return roof, shade


def build_roofline(bvh, peaks):
    """BVH build against the HBM roofline: per unique triangle 48 B of vertices in, 144 B of leaf records
    out, ~10 B of 8-wide nodes, and the (key, index) radix sort's four passes over 12 B in + 12 B out; build_ms is the
    device-event time of the batched BLAS build + world level + TLAS of the scene (queueing by the host included)."""
    tris = bvh.get("unique_triangles", 0) + bvh.get("baked_triangles", 0)
    ms = bvh.get("build_ms", 0.0)
    if not tris or ms <= 0:
        return None
    nbytes = tris * (48 + 144 + 10 + 4 * 24)
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"bound": "hbm", "bytes_per_triangle": 298, "achieved": gbs, "peak": peaks["hbm"], "unit": "GB/s", "frac": gbs / peaks["hbm"],
            "mtris_per_s": tris / (ms * 1e-3) / 1e6}


def short_run(g, torch, dev, stream, name, scene, W, H, S, max_depth, peaks, steps=2, warmup=2):
    """A short driver-run measurement of another BASELINE.json configuration on this GPU: value, stage times, fetch counts,
    roofline fractions (same method as the main line, fewer steps; no e2e / CPU legs)."""
    t_build = time.perf_counter()
    ctx = g.Context(dev.index)
    ctx.set_stream(stream.cuda_stream)
    ctx.upload(scene)
    first_build_ms = ctx.bvh_stats()["build_ms"]
    ctx.upload(scene)           # built twice: the second build does not pay for the growth of the allocation pools
    ctx.set_params(max_depth=max_depth)
    bvh = ctx.bvh_stats()
    bvh["first_build_ms"] = first_build_ms
    ctx.render(W, H, 0, 0)
    ctx.clear()
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t_build
    ext_fetch, sh_fetch = fetch_counts(ctx, W, H)
    for k in range(warmup):
        ctx.render(W, H, k * S, S, 1)
    ctx.clear()
    ctx.reset_counters()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for k in range(steps):
        ctx.render(W, H, (warmup + k) * S, S, 1)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    c = ctx.counters()
    ctx.set_profiling(True)
    c0 = ctx.counters()
    ctx.render(W, H, (warmup + steps) * S, S, 1)
    stage = {k: v for k, v in ctx.last_render_ms().items() if k != "total"}
    iters = ctx.launch_info()[1]
    c1 = ctx.counters()
    ctx.set_profiling(False)
    img = ctx.read_hdr()
    ok = bool(np.isfinite(img).all() and img[..., :3].mean() > 1e-3)
    ctx.close()
    d_ext, d_sh = c1["extension_rays"] - c0["extension_rays"], c1["shadow_rays"] - c0["shadow_rays"]
    roof, shade = traversal_roofline(bvh, peaks, ext_fetch, sh_fetch, stage, iters, d_ext, d_sh, c1["shaded_vertices"] - c0["shaded_vertices"])
    return {"workload": f"{name} {W}x{H}, NEE+MIS, max_depth {max_depth}, {S} spp per step", "value": c["paths"] / (ms * 1e-3) / 1e6,
            "unit": UNIT, "mrays_per_s": (c["extension_rays"] + c["shadow_rays"]) / (ms * 1e-3) / 1e6, "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "stage_ms_per_step": stage, "image_ok": ok, "setup_s": t_build,
            "rays_per_path": {"extension": c["extension_rays"] / c["paths"], "shadow": c["shadow_rays"] / c["paths"]},
            "roofline": {k: roof[k] for k in ("bound", "achieved", "peak", "frac", "frac_of_hbm_peak", "frac_of_l2_peak", "bytes_per_ray", "fetch_per_ray", "ms_per_launch")},
            "shade_gbs": shade["achieved_gbs"], "bvh": bvh, "bvh_roofline": build_roofline(bvh, peaks)}


def run_ours(a, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    import gpuspectral_b200 as g

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W, H, S = a.width, a.height, a.spp_per_step
    npix = W * H
    scene = g.Scene.builtin(a.scene, *scene_params(a))
    ctx = g.Context(local_rank)
    stream = torch.cuda.Stream(device=dev)     # one stream for the render, the NCCL reduce and the timing events
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ctx.upload(scene)
    first_build_ms = ctx.bvh_stats()["build_ms"]
    ctx.upload(scene)           # built twice: the second build does not pay for the growth of the allocation pools
    ctx.set_params(max_depth=a.max_depth)
    bvh = ctx.bvh_stats()       # device time of the BLAS + TLAS builds, node / triangle counts, SAH before / after refinement
    bvh["first_build_ms"] = first_build_ms
    ctx.render(W, H, 0, 0)      # allocate the frame
    ctx.clear()
    ptr, nbytes = ctx.device_sum_ptr()
    accum = torch.as_tensor(CudaBuffer(ptr, nbytes // 4), device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(k):
        """S samples per pixel on this GPU, then the frame's single reduce."""
        ctx.render(W, H, k * S * world + rank, S, world)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
            if rank != 0:
                accum.zero_()       # rank 0 keeps the running global sum; the others restart their partial sums

    # ---- instrumented pass (untimed): fetch counts of the rays this workload really traces ------------------
    peaks = load_peaks()
    zero = {"nodes_visited": 0.0, "tris_tested": 0.0, "inst_entered": 0.0}
    ext_fetch, sh_fetch = (zero, zero) if a.no_instrumented else fetch_counts(ctx, W, H)
    ctx.clear()
    accum.zero_()

    # ---- timed region ------------------------------------------------------------------------------------------
    for k in range(a.warmup):
        step(k)
    ctx.clear()
    accum.zero_()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for k in range(a.steps):
        step(a.warmup + k)      # asynchronous: the K steps queue up on the stream, the host never gates the device
    ev1.record(stream)
    barrier()
    device_ms = ev0.elapsed_time(ev1)
    c = ctx.counters()
    launches = ctx.launch_info()[0]
    # ---- stage breakdown: the same step with CUDA events around every stage of every wavefront iteration (events on
    # the render stream; this mode drives the iterations from the host instead of the device-side graph loop)
    ctx.set_profiling(True)
    stage = {"extend": 0.0, "shade": 0.0, "shadow": 0.0, "finish": 0.0}
    iters = 0
    prof_steps = max(1, min(a.steps, 2))
    prof_ext = prof_sh = prof_vert = 0
    for k in range(prof_steps):
        c0 = ctx.counters()
        ctx.render(W, H, (a.warmup + a.steps + k) * S * world + rank, S, world)
        ms = ctx.last_render_ms()
        for name in stage:
            stage[name] += ms[name]
        iters += ctx.launch_info()[1]
        c1 = ctx.counters()
        prof_ext += c1["extension_rays"] - c0["extension_rays"]
        prof_sh += c1["shadow_rays"] - c0["shadow_rays"]
        prof_vert += c1["shaded_vertices"] - c0["shaded_vertices"]
    ctx.set_profiling(False)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    elapsed_ms = ev0.elapsed_time(ev1)
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    totals = torch.tensor([c["paths"], c["extension_rays"], c["shadow_rays"], launches], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(totals, op=dist.ReduceOp.SUM)
    elapsed_ms = float(t.item())
    paths, ext, sh, launches = [float(x) for x in totals.tolist()]
    value = paths / (elapsed_ms * 1e-3) / 1e6
    mrays = (ext + sh) / (elapsed_ms * 1e-3) / 1e6

    # image sanity on rank 0: the running sum holds steps * S * world samples per pixel
    if rank == 0:
        img = ctx.read_hdr()
        assert np.isfinite(img).all() and img[..., :3].mean() > 0.01, "bench image is empty"
        samples = float(accum.view(-1, 4)[:, 3].mean().item())
        assert abs(samples - (a.steps * world + prof_steps) * S) < 0.5, f"sum buffer holds {samples} samples per pixel"

    # ---- roofline of the dominant kernel (k_extend: closest-hit traversal), from this rank's profiled steps ----------
    roofline, roofline_shade = traversal_roofline(bvh, peaks, ext_fetch, sh_fetch, stage, iters, prof_ext, prof_sh, prof_vert)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "extend_traffic.json")
    if os.path.exists(tpath) and a.scene == "cornell":
        traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
    roofline["traffic"] = traffic

    # ---- e2e: the per-frame API with host buffers ----------------------------------------------------------------------
    e2e = None
    if not a.no_e2e:
        mesh_map = list(ctx._mesh_map)
        host_img = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
        host_np = host_img.numpy()
        h2d = scene_h2d_bytes(scene)
        d2h = npix * 16 if rank == 0 else 0
        ctx.clear()
        accum.zero_()

        def e2e_step(k):
            per_frame_upload(ctx, scene, mesh_map)            # host arrays -> device, TLAS rebuilt (every frame, as the reference)
            step(k)
            if rank == 0:
                ctx.read_hdr(out=host_np)                     # resolve + device -> pinned host

        e2e_step(0)
        ctx.reset_counters()
        barrier()
        t0 = time.perf_counter()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for k in range(a.e2e_steps):
            e2e_step(1 + k)
        f1.record(stream)
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        dev_ms = f0.elapsed_time(f1)
        tt = torch.tensor([max(wall_ms, dev_ms)], dtype=torch.float64, device=dev)
        pp = torch.tensor([float(ctx.counters()["paths"])], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(pp, op=dist.ReduceOp.SUM)
        e2e = {"value": float(pp.item()) / (float(tt.item()) * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": npix * 16, "steps": a.e2e_steps, "ms_per_step": float(tt.item()) / a.e2e_steps,
               "what": "per step: BSDF/light/instance arrays + camera from host memory, TLAS rebuild, render, NCCL reduce (N>1), "
                       "HDR frame resolved and copied to pinned host memory on rank 0"}
        if rank == 0:
            assert np.isfinite(host_np).all() and host_np[..., :3].mean() > 0.01
        # the same per-frame loop for a scene whose instances MOVE: transforms from host memory through
        # gsb_scene_update_transforms (re-bake + refit over the kept topology) instead of the full instance upload + rebuild
        inst, ni = scene.instances
        xf = np.array([list(inst[i].transform) for i in range(ni)], np.float32)

        def refit_step(k):
            ctx.update_transforms(0, xf)
            ctx.set_camera(*scene.camera)
            step(k)
            if rank == 0:
                ctx.read_hdr(out=host_np)

        refit_step(100)
        ctx.reset_counters()
        barrier()
        t0 = time.perf_counter()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        for k in range(a.e2e_steps):
            refit_step(101 + k)
        g1.record(stream)
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        tt = torch.tensor([max(wall_ms, g0.elapsed_time(g1))], dtype=torch.float64, device=dev)
        pp = torch.tensor([float(ctx.counters()["paths"])], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(pp, op=dist.ReduceOp.SUM)
        st = ctx.bvh_stats()
        e2e["moving_instances"] = {"value": float(pp.item()) / (float(tt.item()) * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": float(tt.item()) / a.e2e_steps,
                                   "h2d_bytes_per_step": int(xf.nbytes) + 80, "refit_ms": st["refit_ms"], "rebuild_ms": bvh["build_ms"],
                                   "what": "per step: every instance transform from host memory (gsb_scene_update_transforms: re-bake + "
                                           "bottom-up refit), camera, render, reduce (N>1), HDR frame to pinned host memory on rank 0"}

    # ---- CPU baseline (rank 0, N = 1 only) ---------------------------------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        from oracle import oracle as orc
        orc.build()
        o = orc.OracleScene.from_host_scene(scene)
        o.set_params(max_depth=a.max_depth)
        o.render(W, H, 0, 1, sum_mode=True)
        spp = 6
        t0 = time.perf_counter()
        _, oc, threads = o.render(W, H, 100, spp, sum_mode=True, threads=host_threads())
        dt = time.perf_counter() - t0
        cpu = {"value": oc["paths"] / dt / 1e6, "unit": UNIT, "cores": threads, "kind": "port",
               "mrays_per_s": (oc["extension_rays"] + oc["shadow_rays"]) / dt / 1e6,
               "sample": f"{W}x{H} x {spp} spp of the same scene/params ({oc['paths']} paths, {dt:.1f} s)"}

    # ---- the other BASELINE.json configurations, short runs on this GPU (N = 1 only) ------------------------------------
    extra = None
    if rank == 0 and world == 1 and not a.no_extra and a.scene == "cornell" and not a.scene_params:
        extra = {}
        for name, scn, (ew, eh, es) in (("gallery", "gallery", (1920, 1080, 64)), ("field", "field", (1920, 1080, 16)),
                                        ("cornell_4k", "cornell", (4096, 2160, 32))):
            try:
                extra[name] = short_run(g, torch, dev, stream, scn, g.Scene.builtin(scn), ew, eh, es, a.max_depth, peaks)
            except Exception as e:      # a failed side run must not take the headline line with it
                extra[name] = {"error": repr(e)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": elapsed_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "mrays_per_s": mrays,
                "rays_per_path": {"extension": ext / paths, "shadow": sh / paths},
                "config": {"workload": workload_name(a), "width": W, "height": H, "spp_per_step": S, "max_depth": a.max_depth,
                           "scene": a.scene, "parallelism": f"sample-index sharding x{world}" + (", 1 NCCL reduce per step" if world > 1 else ""),
                           "l2": "path state streams of a step (up to 4.2 M slots x ~230 B) exceed the 126 MB L2; scene data is meant to be cache resident"},
                "e2e": e2e, "roofline": roofline, "roofline_shade": roofline_shade, "cpu_baseline": cpu, "clocks": clocks, "gpu_launches": int(launches),
                "stage_ms_per_step": {k: v / prof_steps for k, v in stage.items()},
                "device_ms_per_step": device_ms / a.steps, "bvh": bvh, "bvh_roofline": build_roofline(bvh, peaks), "extra_workloads": extra}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        run_reference(a, rank, world)
        return
    if world == 1 and a.gpus > 1:
        # launched without torchrun: re-launch under it, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    run_ours(a, rank, world, local_rank)


if __name__ == "__main__":
    main()
