#!/usr/bin/env python
"""bench.py — throughput of the per-pixel trace path (ray-gen → BVH8 traverse → shade → film) on B200.

    python bench.py --gpus N --steps K --warmup W                  # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path (oracle port) on host cores

One step = one full frame of the workload (all screen tiles, all samples).  Metric (BASELINE.json): Mrays/s (all ray
segments: camera + bounce + shadow) with samples/s beside it.  Prints ONE JSON line on rank 0.

Default workload = the north_star target, BASELINE configs[2]: the Sponza STAND-IN ("atrium", sponza.obj is absent from the
reference) at 1920x1080, 256 spp, 5 bounces.  configs[1] (teapot 1920x1080, 64 spp, 5 bounces) rides along as `extra` at N=1.

  value        device-resident: scene, camera and tiles already in HBM; CUDA events around each step on the launching stream,
               the pipeline MP_MODE_AUTO picks (what render() runs)
  e2e          the same frame through the reference-facing API with HOST buffers: N=1 `minipath_b200.render()` + wait() + image()
               (mp_render); N>1 the mp_group_* C ABI (scene upload from host arrays, dynamic tile queue, NCCL film reduce inside
               the library, RGBA8 image back on the host of rank 0)
  roofline     whole step against the measured FP32 / L2 / L1-gather / HBM denominators, SURVEY §8(d) constants
  cpu_baseline the oracle (CPU restatement of the reference path; `kind: port`) on the box's host cores, bounded sample

Multi-GPU (torchrun, one rank per GPU): tile sharding — the ranks pull batches of tiles from ONE shared counter (the
reference's scheduler, machinery.rs:205-209), total work fixed ("strong"), films combined by one ncclReduce inside the
library.  `--shard spp` splits the samples instead (configs[3], "weak": per-GPU work fixed).
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
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

# Algorithmic work per visit — SURVEY §8(d): bytes/ray = 128 n_inner + 144 n_pack + 36 (ray in) + 16 (hit out) + 24 (n_inner +
# n_leaf) for the stored per-node boxes; flops/ray = 390 n_inner + 620 n_pack.
BYTES_INNER = 128        # 6 x u16x8 child bounds + 8 links
BYTES_PACKET = 144       # 9 x u16x8
BYTES_BOX = 24           # enclosing box (min, size) of an inner node / a leaf, stored on the device
BYTES_RAY = 36 + 16      # ray in, hit out
FLOPS_INNER = 390
FLOPS_PACKET = 620
# what the device records actually hold per visit (round 1's constants; printed beside the above as `alt`): 160-byte node record,
# 144-byte packet + 32-byte leaf box; 8 x 48 flop per node, 8 x 87 per packet counted from the SASS
ALT = dict(BYTES_INNER=160, BYTES_PACKET=144, BYTES_LEAF=32, BYTES_RAY=52, FLOPS_INNER=384, FLOPS_PACKET=696)

# name -> (scene, width, height, spp, max_bounces); BASELINE.json configs
WORKLOADS = {
    "atrium-1080p": ("atrium", 1920, 1080, 256, 5),    # configs[2] on the procedural Sponza stand-in: the north_star target
    "teapot-1080p": ("teapot", 1920, 1080, 64, 5),     # configs[1]: teapot 1920x1080, 64 spp, 5 bounces
    "cube-1080p": ("cube", 1920, 1080, 64, 5),         # configs[1], the cube half
    "teapot-bench": ("teapot", 2048, 1536, 10, 0),     # configs[0]: benches/render_teapot.rs:12-24 (reference shader)
    "atrium-4k": ("atrium", 3840, 2160, 1024, 8),      # configs[3] on the stand-in: spp split over the GPUs, NCCL film reduce
    "atrium-1080p-materials": ("atrium+materials", 1920, 1080, 256, 5),  # the same frame with the nine per-part materials (RGB path state)
}
DEFAULT_WORKLOAD = "atrium-1080p"
PIPELINES = ["megakernel", "wavefront-octet", "wavefront-thread", "wavefront-persist", "wavefront-hybrid", "wavefront-pool",
             "wavefront-hybrid-pool", "wavefront-hybrid-pool-s0"]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="auto", choices=["auto"] + PIPELINES,
                    help="auto: the pipeline MP_MODE_AUTO / render() picks; otherwise force one (A/B runs)")
    ap.add_argument("--ab", action="store_true", help="also time every pipeline once (1 warm-up + 1 step each) and report it in `pipeline_ab_ms`")
    ap.add_argument("--order", default="reference", choices=["reference", "near_first"])
    ap.add_argument("--layout", default="auto", choices=["auto", "decoded", "quantised"])
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel")
    ap.add_argument("--bounces", type=int, default=-1, help="override the workload's max_bounces")
    ap.add_argument("--shard", default="tiles", choices=["tiles", "spp"], help="N>1: dynamic tile queue (strong scaling) or sample split (weak)")
    ap.add_argument("--min-batch", type=int, default=0, help="N>1 tile sharding: smallest batch of tiles a GPU claims (0 = the library's choice: about 2^20 paths)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the riders (teapot-1080p = configs[1]; the default workload on the device-built tree)")
    ap.add_argument("--builder", default="reference", choices=["reference", "device"],
                    help="who builds the stand-in's BVH: the reference split on the host (default; the CPU arm's tree too) or mp_bvh_build_device")
    ap.add_argument("--ref-seconds", type=float, default=12.0, help="CPU arms: target seconds of oracle work per bounded step")
    ap.add_argument("--timeline", default="", help="N>1: write a per-rank timeline of one step (JSON) to this path")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.1)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def workload_params(name, args):
    scene, w, h, spp, bounces = WORKLOADS[name]
    if name == args.workload:
        if args.spp > 0:
            spp = args.spp
        if args.bounces >= 0:
            bounces = args.bounces
    return scene, w, h, spp, bounces


def load_workload(name, args, builder=None):
    """Scene (host BVH), camera and frame parameters of a workload.  Built on the host, outside every timed region
    (benches/render_teapot.rs:25-27 excludes the scene build too)."""
    import minipath_b200 as mp

    scene, w, h, spp, bounces = workload_params(name, args)
    data = os.path.join(ROOT, "data")
    if scene == "teapot":  # camera of benches/render_teapot.rs:12-19
        bvh = mp.TriangleBvh.with_obj(os.path.join(data, "teapot.obj"))
        cam = mp.Camera.default().look_at((0, 2, 10), (0, 1.5, 0), (0, 1, 0)).f_number(4.8).focus_distance(10.0)
    elif scene == "cube":  # fan-triangulated quads (documented deviation), SURVEY §8(d) gate camera
        bvh = mp.TriangleBvh.with_obj(os.path.join(data, "cube.obj"), quad_mode=1)
        cam = mp.Camera.default().look_at((1.5, 1.2, 2.5), (0, 0, 0), (0, 1, 0)).f_number(64.0)
    else:
        from minipath_b200 import scenes

        bvh, cam = scenes.atrium_scene(builder=2 if (builder or args.builder) == "device" else 0, materials=scene.endswith("+materials"),
                                       device=int(os.environ.get("LOCAL_RANK", 0)))
    return bvh, cam, w, h, spp, bounces


def workload_config(name, args):
    """The `config` object both arms print: identical keys and values for the CUDA arm and the reference arm."""
    scene, w, h, spp, bounces = workload_params(name, args)
    return {"workload": name, "scene": scene + (" (procedural Sponza STAND-IN; sponza.obj is absent from the reference)" if scene.startswith("atrium") else ".obj of the reference"),
            "resolution": [w, h], "spp": spp, "max_bounces": bounces, "tile_size": 64,
            "l2": "flushed between timed steps (256 MB fill outside the event pairs); inputs resident in HBM",
            "note": "max_bounces>0 is the extension integrator (the reference traces primary rays only)"}


def data_note(name):
    scene = WORKLOADS[name][0]
    return "synthetic camera samples (counter-based RNG) on " + ("the reference's data/%s.obj" % scene if not scene.startswith("atrium") else
                                                                 "a procedural Sponza STAND-IN (sponza.obj is absent from the reference)")


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------------------------------
# CPU arms (the oracle is test infrastructure: it runs ONLY here, as the thing beside the measurement, never inside it)
# ------------------------------------------------------------------------------------------------------------------
def oracle_workload(oracle, name, args):
    """The workload on the CHECKER's side only: its own OBJ loader / mesh builder (oracle/oracle_build.inc), its own camera sampler
    and tile list — nothing from libminipath_b200.so, which a CPU arm therefore never maps."""
    from oracle_lib import scenes

    scene, w, h, spp, bounces = workload_params(name, args)
    data = os.path.join(ROOT, "data")
    if scene == "teapot":  # camera of benches/render_teapot.rs:12-19
        bvh = oracle.build_obj(os.path.join(data, "teapot.obj"))
        sampler = oracle.camera_sampler((0, 2, 10), (0, 1.5, 0), (0, 1, 0), (w, h), f_number=4.8, focus_distance=10.0)
    elif scene == "cube":
        bvh = oracle.build_obj(os.path.join(data, "cube.obj"), quad_mode=1)
        sampler = oracle.camera_sampler((1.5, 1.2, 2.5), (0, 0, 0), (0, 1, 0), (w, h), f_number=64.0, focus_distance=float(np.linalg.norm([1.5, 1.2, 2.5])))
    else:
        if scene.endswith("+materials"):
            from oracle_lib import materials_table

            p, n, t, uv, tm, table = scenes.atrium_mesh_materials()
            bvh = oracle.build_mesh(p, t, n, texcoords=uv, tri_material=tm,
                                    materials=materials_table(table[:, :3], table[:, 3] > 0, np.where(table[:, 3] > 0, table[:, 3], 1.0)))
        else:
            p, n, t = scenes.atrium_mesh()
            bvh = oracle.build_mesh(p, t, n)
        c = scenes.atrium_camera_params(*bvh.get_bounding_box())
        sampler = oracle.camera_sampler(c["eye"], c["target"], c["up"], (w, h), sensor_size=c["sensor_width"], sensor_is_width=True,
                                        focal_length=c["focal_length"], f_number=c["f_number"], focus_distance=c["focus"])
    return bvh, sampler, oracle.tiles(w, h, 64), w, h, spp, bounces


def render_params(spp, bounces, seed=7):
    from oracle_lib import _ffi

    return _ffi.MpRenderParams(spp, 0, bounces, 0, 0, 64, seed)


def oracle_frame(oracle, bvh, sampler, w, h, tiles, params, threads):
    """One bounded frame on the CPU oracle; returns (seconds, rays, camera samples)."""
    from oracle_lib import _ffi

    c = _ffi.MpCounters()
    film, secs, c = oracle.render_tiles(bvh, sampler, w, h, tiles, params, threads=threads, counters=c)
    return secs, int(c.rays), int(c.camera_rays)


def bounded_spp(oracle, bvh, sampler, w, h, tiles, bounces, spp, threads, seconds):
    """How many of the workload's spp fit `seconds` of oracle work (calibrated with one 1-spp frame)."""
    s1, _, _ = oracle_frame(oracle, bvh, sampler, w, h, tiles, render_params(1, bounces), threads)
    return int(max(1, min(spp, round(seconds / max(s1, 1e-3)))))


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port — the Rust crate cannot be built here) on all host cores, same
    workload, metric and config object; each step is a bounded sample: the full frame at `sample_spp` of the workload's spp."""
    if rank != 0:
        return
    from oracle_lib import Oracle

    oracle = Oracle()
    oracle.o.orc_set_simd(1)
    threads = host_cores()
    bvh, sampler, tiles, w, h, spp, bounces = oracle_workload(oracle, args.workload, args)
    n = bounded_spp(oracle, bvh, sampler, w, h, tiles, bounces, spp, threads, args.ref_seconds)
    for _ in range(min(args.warmup, 1)):  # the calibration frame has already warmed caches and threads
        oracle_frame(oracle, bvh, sampler, w, h, tiles, render_params(n, bounces), threads)
    secs = rays = samples = 0
    for _ in range(args.steps):
        s, r, c = oracle_frame(oracle, bvh, sampler, w, h, tiles, render_params(n, bounces), threads)
        secs, rays, samples = secs + s, rays + r, samples + c
    mrays = rays / secs / 1e6
    sample = f"{args.workload}: full {w}x{h} frame at {n} of {spp} spp per step, {bounces} bounces, AVX2+FMA oracle, {threads} threads"
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.shard == "tiles" else "weak", "vs_baseline": None, "dtype": "f32", "data": data_note(args.workload),
        "samples_per_s": samples / secs, "config": workload_config(args.workload, args),
        "sample": sample, "sample_spp": n,
        "reference_kind": "CPU restatement of the reference path (oracle port; cargo/rustc absent), thread per core, atomic tile queue; "
                          "scene from the oracle's own loader + builder, libminipath_b200.so not loaded",
        "product_library_loaded": "minipath_b200" in sys.modules,
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline(args, name):
    """Oracle on the host cores, bounded to roughly 10-20 s: 1 spp to calibrate, then as many spp as fit."""
    from oracle_lib import Oracle

    oracle = Oracle()
    oracle.o.orc_set_simd(1)
    threads = host_cores()
    bvh, sampler, tiles, w, h, spp, bounces = oracle_workload(oracle, name, args)
    n = bounded_spp(oracle, bvh, sampler, w, h, tiles, bounces, spp, threads, args.ref_seconds)
    secs, rays, samples = oracle_frame(oracle, bvh, sampler, w, h, tiles, render_params(n, bounces), threads)
    return {"value": rays / secs / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port", "samples_per_s": samples / secs,
            "seconds": secs, "sample": f"{name}: full {w}x{h} frame at {n} of {spp} spp, {bounces} bounces, AVX2+FMA oracle, {threads} threads"}


# ------------------------------------------------------------------------------------------------------------------
# CUDA arm
# ------------------------------------------------------------------------------------------------------------------
def pipeline_tables(mp):
    modes = {k: (mp.MODE_MEGAKERNEL if k == "megakernel" else mp.MODE_WAVEFRONT) for k in PIPELINES}
    variants = {"megakernel": mp.TRACE_THREAD, "wavefront-thread": mp.TRACE_THREAD, "wavefront-octet": mp.TRACE_OCTET,
                "wavefront-persist": mp.TRACE_PERSIST, "wavefront-hybrid": mp.TRACE_HYBRID, "wavefront-pool": mp.TRACE_POOL,
                "wavefront-hybrid-pool": mp.TRACE_HYBRID_POOL, "wavefront-hybrid-pool-s0": mp.TRACE_HYBRID_POOL_S0}
    return modes, variants


def auto_pipeline(bvh, bounces):
    """The pick of MP_MODE_AUTO (minipath_b200/csrc/host_render.cpp: mp_render), so that `value` times what render() runs."""
    st = bvh.stats()
    deep = st["max_stack"] > 128
    if not deep and (bounces == 0 or st["n_inner"] == 0):
        return "megakernel"
    return "wavefront-pool" if deep else "wavefront-hybrid-pool-s0"


class Timer:
    """Per-step CUDA events on the launching stream, L2 flushed between steps (outside the event pairs)."""

    def __init__(self, torch, stream):
        self.torch, self.stream = torch, stream
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def run(self, step, n_warm, n_steps):
        torch, stream = self.torch, self.stream
        with torch.cuda.stream(stream):
            for _ in range(n_warm):
                step()
            evs = []
            for _ in range(n_steps):
                self.flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                step()
                e1.record(stream)
                evs.append((e0, e1))
            stream.synchronize()
        return [a.elapsed_time(b) for a, b in evs]


def measure_single(args, name, torch, local, full=True, builder=None):
    """N = 1: device-resident `value` through a Context, `e2e` through minipath_b200.render()."""
    import minipath_b200 as mp

    bvh, cam, w, h, spp, bounces = load_workload(name, args, builder)
    order = mp.ORDER_REFERENCE if args.order == "reference" else mp.ORDER_NEAR_FIRST
    sampler = cam.build_sampler((w, h))
    tiles = mp.full_screen_tiles(w, h, 64)
    modes, variants = pipeline_tables(mp)
    if bvh.stats()["max_stack"] > 128:  # deeper than the local-memory stack of the thread-per-ray kernels
        modes = {k: v for k, v in modes.items() if k in ("wavefront-octet", "wavefront-pool")}
    stream = torch.cuda.Stream()
    ctx = mp.Context(local)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_node_layout({"auto": mp.LAYOUT_AUTO, "decoded": mp.LAYOUT_DECODED, "quantised": mp.LAYOUT_QUANTISED}[args.layout])
    ctx.upload_scene(bvh)
    ctx.set_camera(sampler)
    ctx.create_film(w, h)
    rgba = torch.empty((h, w), dtype=torch.int32, device="cuda")
    timer = Timer(torch, stream)

    def make_step(key):
        st = mp.RenderSettings(64, spp, (w, h), max_bounces=bounces, seed=7, mode=modes[key], order=order)

        def step():
            ctx.set_trace_variant(variants[key])
            ctx.clear_film()
            ctx.render_tiles(tiles, st.params())
            ctx.resolve_device(spp, rgba.data_ptr())
        return step

    mode_name = auto_pipeline(bvh, bounces) if (args.mode == "auto" or name != args.workload) else args.mode
    ab = {}
    if args.ab and full:
        for key in modes:
            ab[key] = sum(timer.run(make_step(key), 1, 1))
    step = make_step(mode_name)

    # counted pass (untimed): rays and node visits of one step, deterministic for a given seed
    ctx.enable_counters(True)
    ctx.reset_counters()
    with torch.cuda.stream(stream):
        step()
    stream.synchronize()
    cnt = ctx.counters()
    ctx.enable_counters(False)

    sampler_thread = ClockSampler(local) if full else None
    timer.run(step, args.warmup, 0)
    ctx.reset_counters()
    torch.cuda.synchronize()
    if sampler_thread:
        sampler_thread.start()
    per_step = timer.run(step, 0, args.steps)
    torch.cuda.synchronize()
    clocks = sampler_thread.stop() if sampler_thread else None
    launches = ctx.counters()["launches"]
    ms_per_step = sum(per_step) / args.steps
    res = dict(name=name, ctx=ctx, bvh=bvh, cam=cam, w=w, h=h, spp=spp, bounces=bounces, tiles=tiles, cnt=cnt, ab=ab, mode_name=mode_name,
               ms_per_step=ms_per_step, per_step_ms=per_step, value=cnt["rays"] / (ms_per_step * 1e-3) / 1e6, world=1,
               samples_per_s=cnt["camera_rays"] / (ms_per_step * 1e-3), clocks=clocks, launches=launches, rays_job=float(cnt["rays"]),
               n_gpus=1, sharding="none", scaling="strong" if args.shard == "tiles" else "weak")

    if not args.no_e2e:
        raw = bvh.raw
        scene_bytes = raw.n_inner * (128 + 32) + raw.n_packets * (144 + 32 + 8 * 16) + raw.n_vertices * 32
        st = mp.RenderSettings(64, spp, (w, h), max_bounces=bounces, seed=7, order=order,
                               **({} if mode_name == auto_pipeline(bvh, bounces) else {"mode": modes[mode_name], "trace_kernel": variants[mode_name]}))

        def once():
            prog = mp.render(mp.Scene(bvh), cam, st, devices=[local])
            prog.wait()
            return prog.image()

        for _ in range(max(1, min(args.warmup, 2))):
            once()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            once()
        torch.cuda.synchronize()
        secs = time.perf_counter() - t0
        res["e2e"] = {"value": res["rays_job"] * args.steps / secs / 1e6, "unit": "Mrays/s", "ms_per_step": secs / args.steps * 1e3,
                      "h2d_bytes_per_step": int(scene_bytes + len(tiles) * 16 + 52), "d2h_bytes_per_step": int(w * h * 4),
                      "api": "minipath_b200.render()+wait()+image() [mp_render, MP_MODE_AUTO]: host scene arrays in, host RGBA8 image out"}
    return res


def measure_group(args, name, torch, dist, rank, local, world):
    """N > 1 under torchrun: every rank joins ONE mp_group (NCCL communicator + shared-memory tile counter inside the library)."""
    import minipath_b200 as mp

    bvh, cam, w, h, spp, bounces = load_workload(name, args)
    order = mp.ORDER_REFERENCE if args.order == "reference" else mp.ORDER_NEAR_FIRST
    sampler = cam.build_sampler((w, h))
    tiles = mp.full_screen_tiles(w, h, 64)
    modes, variants = pipeline_tables(mp)
    mode_name = auto_pipeline(bvh, bounces) if args.mode == "auto" else args.mode
    shard = mp.SHARD_TILES if args.shard == "tiles" else mp.SHARD_SPP
    spp_total = spp if args.shard == "tiles" else spp * world  # weak scaling: every GPU keeps the workload's spp

    boot = [None, None]
    if rank == 0:
        boot = [mp.Group.unique_id(), f"/mp_b200_{os.getpid()}_{int(time.time() * 1e3) & 0xffffff}"]
    dist.broadcast_object_list(boot, src=0)
    g = mp.Group.for_rank(local, rank, world, boot[0], boot[1])
    stream = torch.cuda.Stream()
    ctx = g.context(0)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_node_layout({"auto": mp.LAYOUT_AUTO, "decoded": mp.LAYOUT_DECODED, "quantised": mp.LAYOUT_QUANTISED}[args.layout])
    g.set_trace_variant(variants[mode_name])
    g.upload_scene(bvh)
    g.set_camera(sampler)
    g.create_film(w, h)
    ctx.resolution = (w, h)
    rgba = torch.empty((h, w), dtype=torch.int32, device="cuda")
    timer = Timer(torch, stream)
    st = mp.RenderSettings(64, spp_total, (w, h), max_bounces=bounces, seed=7, mode=modes[mode_name], order=order)

    def step():
        g.clear_film()
        g.render(tiles, st.params(), shard=shard, min_batch=args.min_batch)
        g.reduce_film()  # ncclReduce(sum) to rank 0 on the same stream
        if rank == 0:
            ctx.resolve_device(spp_total, rgba.data_ptr())

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    def allsum(vals):
        t = torch.tensor(vals, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(x) for x in t]

    def allmax(vals):
        t = torch.tensor(vals, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t]

    ctx.enable_counters(True)
    ctx.reset_counters()
    with torch.cuda.stream(stream):
        step()
    stream.synchronize()
    cnt_local = ctx.counters()
    ctx.enable_counters(False)
    keys = ["rays", "camera_rays", "shadow_rays", "inner_visits", "packet_visits", "leaf_visits"]
    cnt = dict(zip(keys, allsum([cnt_local[k] for k in keys])))

    sampler_thread = ClockSampler(local) if rank == 0 else None
    timer.run(step, args.warmup, 0)
    ctx.reset_counters()
    barrier()
    if sampler_thread:
        sampler_thread.start()
    per_step = timer.run(step, 0, args.steps)
    barrier()
    clocks = sampler_thread.stop() if sampler_thread else None
    launches = ctx.counters()["launches"]
    total_ms = allmax([sum(per_step)])[0]
    ms_per_step = total_ms / args.steps
    stats = g.member_stats(0)
    res = dict(name=name, ctx=ctx, group=g, bvh=bvh, cam=cam, w=w, h=h, spp=spp, bounces=bounces, tiles=tiles, cnt=cnt, ab={}, mode_name=mode_name,
               ms_per_step=ms_per_step, per_step_ms=per_step, value=cnt["rays"] / (ms_per_step * 1e-3) / 1e6, world=world,
               samples_per_s=cnt["camera_rays"] / (ms_per_step * 1e-3), clocks=clocks, launches=launches, rays_job=cnt["rays"],
               n_gpus=world, sharding=args.shard, scaling="strong" if args.shard == "tiles" else "weak", spp_total=spp_total)

    if args.timeline:  # where one step's time goes on every rank: device time of the render, of the reduce, tiles and batches taken
        with torch.cuda.stream(stream):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            barrier()
            g.clear_film()
            ev[0].record(stream)
            g.render(tiles, st.params(), shard=shard, min_batch=args.min_batch)
            ev[1].record(stream)
            g.reduce_film()
            ev[2].record(stream)
            if rank == 0:
                ctx.resolve_device(spp_total, rgba.data_ptr())
            ev[3].record(stream)
            stream.synchronize()
        s = g.member_stats(0)
        mine = {"rank": rank, "render_ms": ev[0].elapsed_time(ev[1]), "reduce_wait_ms": ev[1].elapsed_time(ev[2]), "resolve_ms": ev[2].elapsed_time(ev[3]),
                "tiles": s["tiles"], "batches": s["batches"]}
        rows = [None] * world
        dist.all_gather_object(rows, mine)
        if rank == 0:
            json.dump({"workload": name, "n_gpus": world, "sharding": args.shard, "min_batch": args.min_batch, "ranks": rows}, open(args.timeline, "w"), indent=1)

    if not args.no_e2e:
        raw = bvh.raw
        scene_bytes = raw.n_inner * (128 + 32) + raw.n_packets * (144 + 32 + 8 * 16) + raw.n_vertices * 32
        host_img = np.empty((h, w, 4), np.uint8)
        phases = np.zeros(4)

        def once(record=False):
            t = [time.perf_counter()]
            g.upload_scene(bvh)                       # host arrays -> every device
            g.set_camera(cam.build_sampler((w, h)))
            g.clear_film()
            t.append(time.perf_counter())
            g.render(tiles, st.params(), shard=shard, min_batch=args.min_batch)
            t.append(time.perf_counter())
            g.reduce_film()
            g.resolve_rgba8(spp_total, host_img)       # rank 0: resolve + D2H into host memory; other ranks: drain their stream
            t.append(time.perf_counter())
            if record:
                phases[:3] += np.diff(t)
            return host_img

        for _ in range(max(1, min(args.warmup, 2))):
            once()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            once(record=True)
        barrier()
        secs = allmax([time.perf_counter() - t0])[0]
        res["e2e"] = {"value": res["rays_job"] * args.steps / secs / 1e6, "unit": "Mrays/s", "ms_per_step": secs / args.steps * 1e3,
                      "h2d_bytes_per_step": int(world * (scene_bytes + len(tiles) * 16 + 52)), "d2h_bytes_per_step": int(w * h * 4),
                      "rank0_phase_ms": {"upload+camera+clear": phases[0] / args.steps * 1e3, "claim loop (launches)": phases[1] / args.steps * 1e3,
                                         "reduce+resolve+D2H (waits for the render)": phases[2] / args.steps * 1e3},
                      "api": "mp_group_* C ABI per rank: scene_upload + camera_set + film_create + render (shared tile counter) + reduce_film (NCCL) + resolve_rgba8"}
    res["member_stats"] = stats
    res["barrier"] = barrier
    return res


def roofline(args, res, probes):
    """Whole step against the measured denominators.  Algorithmic work from the visit counters of the counted pass (reference
    visiting order) with SURVEY §8(d)'s constants; `bound` = the roof that needs the longest time."""
    cnt, n = res["cnt"], res["n_gpus"]
    secs = res["ms_per_step"] * 1e-3
    ni, npk, nl, nr = cnt["inner_visits"], cnt["packet_visits"], cnt["leaf_visits"], cnt["rays"]
    bytes_ = ni * (BYTES_INNER + BYTES_BOX) + npk * BYTES_PACKET + nl * BYTES_BOX + nr * BYTES_RAY
    flops = ni * FLOPS_INNER + npk * FLOPS_PACKET
    sectors = ni * 5 + npk * 6  # 32-byte sectors a lane gathers per visit: 152 -> 160-byte node record, 168 -> 192-byte packet record
    raw = res["bvh"].raw
    scene_mb = (raw.n_inner * 160 + raw.n_packets * 192) / 1e6
    fp32, l2, l1g, hbm = probes["fp32_tflops"], probes["l2_gbs"], probes["l1_gather_gsectors"], probes["hbm_gbs"]
    times = {"fp32": flops / (fp32 * 1e12 * n), "l1_gather": sectors / (l1g * 1e9 * n),
             ("hbm" if scene_mb > 100 else "l2"): bytes_ / ((hbm if scene_mb > 100 else l2) * 1e9 * n)}
    bound = max(times, key=times.get)
    if bound == "fp32":
        out = {"bound": "fp32", "achieved": flops / secs / 1e12 / n, "peak": fp32, "unit": "TFLOP/s"}
    elif bound == "l1_gather":
        out = {"bound": "l1_gather", "achieved": sectors * 32 / secs / 1e9 / n, "peak": l1g * 32, "unit": "GB/s"}
    else:
        out = {"bound": bound, "achieved": bytes_ / secs / 1e9 / n, "peak": hbm if bound == "hbm" else l2, "unit": "GB/s"}
    out["frac"] = times[bound] / secs
    a_bytes = ni * ALT["BYTES_INNER"] + npk * ALT["BYTES_PACKET"] + nl * ALT["BYTES_LEAF"] + nr * ALT["BYTES_RAY"]
    a_flops = ni * ALT["FLOPS_INNER"] + npk * ALT["FLOPS_PACKET"]
    traffic, traffic_note = None, "no ncu capture recorded for this workload/pipeline (scripts/ncu_traffic.py)"
    try:  # DRAM bytes of the traversal launches of one step, from an ncu capture of this very command line (profiles/traffic.json)
        rec = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"{res['name']}/{res['mode_name']}")
        if rec and res["world"] == 1:
            traffic = rec["dram_bytes_per_step"] * (res["spp"] / rec["spp"] if "spp" in rec else 1.0)  # captured at a reduced spp: linear in spp
            traffic_note = f"dram__bytes_read.sum + dram__bytes_write.sum over the {rec['launches']} traversal launches of one step ({rec['source']})"
    except Exception:
        pass
    kernels = {"megakernel": "render_mega_kernel", "wavefront-octet": "wf_trace_octet_kernel+wf_shadow_octet_kernel",
               "wavefront-thread": "wf_trace_kernel+wf_shadow_kernel", "wavefront-persist": "wf_trace_persist_kernel+wf_shadow_persist_kernel",
               "wavefront-hybrid": "wf_primary_kernel + wf_trace_persist_kernel+wf_shadow_persist_kernel",
               "wavefront-pool": "wf_trace_pool_kernel+wf_shadow_pool_kernel",
               "wavefront-hybrid-pool": "wf_primary_kernel + wf_trace_pool_kernel+wf_shadow_pool_kernel",
               "wavefront-hybrid-pool-s0": "wf_primary_kernel<SHADOW0> (camera rays + depth-0 sun rays) + wf_trace_pool_kernel + wf_shadow_pool_kernel"}
    out.update({
        "traffic": traffic, "traffic_note": traffic_note, "algorithmic_bytes_per_step": bytes_,
        "constants": "SURVEY §8(d): bytes = 152 n_inner + 144 n_packet + 24 n_leaf + 52 rays; flops = 390 n_inner + 620 n_packet; sectors = 5 n_inner + 6 n_packet",
        "scope": "whole step (ray-gen, shade and film included), all launches; per GPU at N>1",
        "kernel": kernels[res["mode_name"]] + " (share of step: see profiles/)",
        "peak_source": probes["source"],
        "fp32": {"achieved_tflops": flops / secs / 1e12 / n, "peak_tflops": fp32, "frac": times["fp32"] / secs},
        "l2": {"achieved_gbs": bytes_ / secs / 1e9 / n, "peak_gbs": l2, "frac": bytes_ / (l2 * 1e9 * n) / secs, "sweep": probes["l2_sweep"]},
        "l1_gather": {"achieved_gsectors": sectors / secs / 1e9 / n, "peak_gsectors": l1g, "frac": times["l1_gather"] / secs,
                      "note": "32-byte sectors per second an incoherent per-lane record gather gets from the L1 pipe (mp_probe_l1_gather)"},
        "hbm": {"achieved_gbs": bytes_ / secs / 1e9 / n, "peak_gbs": hbm, "frac": bytes_ / (hbm * 1e9 * n) / secs,
                "note": "roof only when the BVH outgrows the 126 MB L2 (soups); cache-resident here"},
        "alt": {"constants": "round-1 device-record constants: bytes = 160 n_inner + 144 n_packet + 32 n_leaf + 52 rays; flops = 384 n_inner + 696 n_packet",
                "l2_frac": a_bytes / (l2 * 1e9 * n) / secs, "fp32_frac": a_flops / (fp32 * 1e12 * n) / secs},
        "per_ray": {"inner_visits": ni / max(1, nr), "packet_visits": npk / max(1, nr), "leaf_visits": nl / max(1, nr),
                    "bytes": bytes_ / max(1, nr), "flops": flops / max(1, nr)},
        "scene_mb": scene_mb,
    })
    return out


def run_probes(ctx, bvh):
    """Roofline denominators measured on this device in this run."""
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    sweep, best = {}, 0.0
    for mb in (8, 16, 32, 64, 96):          # all far below the 126 MB L2
        for bps in (1, 2, 4):                # x 512 threads per SM
            for vec in (16, 32):             # LDG.128 / LDG.256
                v = ctx.probe_l2_gbs_shape(mb << 20, bps, vec)
                sweep[f"{mb}MB/{bps * 512}thr/{vec}B"] = round(v, 1)
                best = max(best, v)
    raw = bvh.raw
    n_rec = max(1024, int(raw.n_inner + raw.n_packets))
    return {"fp32_tflops": ctx.probe_fp32_tflops(), "l2_gbs": best, "l2_sweep": sweep, "l1_gather_gsectors": ctx.probe_l1_gather(n_rec, 5),
            "hbm_gbs": float(peaks.get("hbm_gbs", 6650.0)),
            "source": "FP32 FMA rate, L2 read bandwidth (maximum of the sweep) and L1 gather rate measured by mp_probe_* in this run; HBM "
                      + ("measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)")}


def rider_probes(probes, ctx, bvh):
    """A rider's own L1-gather roof: the sector rate depends on whether the record table fits L1 (profiles/r2c_l1_gather_sweep.txt), so the main
    workload's probe must not be reused for a scene of another size; the other roofs do not depend on the scene."""
    raw = bvh.raw
    return dict(probes, l1_gather_gsectors=ctx.probe_l1_gather(max(1024, int(raw.n_inner + raw.n_packets)), 5))


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        args.gpus = world
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    import minipath_b200 as mp

    if not torch.cuda.is_available() or mp.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device — the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        res = measure_group(args, args.workload, torch, dist, rank, local, world)
    else:
        res = measure_single(args, args.workload, torch, local)

    if rank == 0:
        probes = run_probes(res["ctx"], res["bvh"])
        rl = roofline(args, res, probes)
        cb = None
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_baseline(args, args.workload)
        extra = None
        if world == 1 and not args.no_extra and args.workload == DEFAULT_WORKLOAD:
            r2 = measure_single(args, "teapot-1080p", torch, local, full=False)
            rl2 = roofline(args, r2, rider_probes(probes, r2["ctx"], r2["bvh"]))
            extra = {"teapot-1080p": {"config": workload_config("teapot-1080p", args), "value": r2["value"], "unit": "Mrays/s", "ms_per_step": r2["ms_per_step"],
                                      "pipeline": r2["mode_name"], "samples_per_s": r2["samples_per_s"], "e2e": r2.get("e2e"),
                                      "roofline": {k: rl2[k] for k in ("bound", "achieved", "peak", "unit", "frac", "per_ray")},
                                      "note": "BASELINE configs[1] (teapot 1920x1080, 64 spp, 5 bounces), same run, same timing rules"}}
            r2["ctx"].close()
            if args.builder == "reference":  # the same frame on the tree mp_bvh_build_device makes: fewer visits per ray
                r3 = measure_single(args, args.workload, torch, local, full=False, builder="device")
                rl3 = roofline(args, r3, rider_probes(probes, r3["ctx"], r3["bvh"]))
                st3 = r3["bvh"].stats()
                extra[args.workload + "@device-built-tree"] = {
                    "value": r3["value"], "unit": "Mrays/s", "ms_per_step": r3["ms_per_step"], "pipeline": r3["mode_name"], "samples_per_s": r3["samples_per_s"],
                    "e2e": r3.get("e2e"), "bvh": {"inner_nodes": st3["n_inner"], "packets": st3["n_packets"], "device_build_ms": getattr(r3["bvh"], "last_build_ms", None)},
                    "roofline": {k: rl3[k] for k in ("bound", "achieved", "peak", "unit", "frac", "per_ray")},
                    "note": "same scene, camera, spp and bounces; BVH from mp_bvh_build_device (reference node format, Morton-order splits) instead of the reference split — "
                            "samples_per_s is the comparable figure (the rays per sample are the same, the visits per ray are not)"}
                r3["ctx"].close()
        st = res["bvh"].stats()
        line = {
            "metric": "Mrays/s", "value": res["value"], "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": res["scaling"],
            "vs_baseline": None, "dtype": "f32", "data": data_note(args.workload), "samples_per_s": res["samples_per_s"],
            "config": workload_config(args.workload, args),
            "pipeline": {"picked": res["mode_name"], "how": "MP_MODE_AUTO (what render() runs)" if args.mode == "auto" else "forced by --mode",
                         "traversal_order": args.order, "node_layout": args.layout, "pipeline_ab_ms": res["ab"],
                         "sharding": res["sharding"], "spp_total": res.get("spp_total", res["spp"]),
                         "bvh": {"triangles": st["n_triangles"], "inner_nodes": st["n_inner"], "packets": st["n_packets"], "builder": args.builder},
                         "per_step_ms_rank0": res["per_step_ms"], "member_stats_rank0": res.get("member_stats")},
            "rays_per_step": res["rays_job"], "gpu_launches": int(res["launches"]), "clocks": res["clocks"],
            "roofline": rl, "cpu_baseline": cb, "e2e": res.get("e2e"), "extra": extra,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        res["barrier"]()
        res["group"].close()
        dist.barrier()
        dist.destroy_process_group()
    else:
        res["ctx"].close()


if __name__ == "__main__":
    main()
