#!/usr/bin/env python
"""bench.py — throughput of the per-pixel trace path (ray-gen → BVH8 traverse → shade → film) on B200.

    python bench.py --gpus N --steps K --warmup W                  # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path (oracle port) on host cores

One step = one full frame of the workload (all screen tiles, all samples).  Metric (BASELINE.json): Mrays/s
(all ray segments: camera + bounce + shadow) with samples/s beside it.  Prints ONE JSON line on rank 0.

  value        device-resident: scene, camera and tiles already in HBM; CUDA events around each step on the launching stream
  e2e          the same frame through the reference-facing call `minipath_b200.render()` (mp_render: host scene arrays →
               device, tiles rendered, RGBA8 image back in host memory), host<->device copies inside the timed region
  roofline     dominant kernel against the measured FP32 / L2 / HBM denominators
  cpu_baseline the oracle (CPU restatement of the reference path; `kind: port`) on the box's host cores, bounded sample

Multi-GPU (torchrun, one rank per GPU): samples are split over ranks (`sample_base`), films summed by one NCCL reduce
to rank 0; per-GPU work is fixed as N grows ("weak").
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

# algorithmic work per visit: bytes fetched and f32 arithmetic of the reference algorithm
BYTES_INNER = 128 + 32   # node (6 x u16x8 + 8 links) + its enclosing box (min,size)
BYTES_PACKET = 144       # 9 x u16x8
BYTES_LEAF = 32          # enclosing box of the leaf
BYTES_RAY = 36 + 16      # ray in, hit out
FLOPS_INNER = 8 * 48     # per child: 6 dequant (cvt+mul+fma=4) + 6 sub + 6 mul + 12 min/max
FLOPS_PACKET = 8 * 87    # per triangle: 9 dequant (36) + Moller-Trumbore (51: 6 sub, 2 cross, 4 dot, 3 mul, 3 sub, 1 div)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="teapot-1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="auto", choices=["auto", "megakernel", "wavefront-octet", "wavefront-thread", "wavefront-persist", "wavefront-hybrid", "wavefront-pool",
                                                     "wavefront-hybrid-pool", "wavefront-hybrid-pool-s0"],
                    help="megakernel: one thread per pixel; wavefront-*: staged kernels with the thread-per-ray, the 8-lanes-per-ray "
                         "(octet) or the persistent phased thread-per-ray traversal kernel")
    ap.add_argument("--order", default="reference", choices=["reference", "near_first"])
    ap.add_argument("--layout", default="auto", choices=["auto", "decoded", "quantised"],
                    help="node layout read by the traversal kernels: f32 boxes/edges decoded at upload (default) or the reference's u16 records")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel")
    ap.add_argument("--bounces", type=int, default=-1, help="override the workload's max_bounces")
    ap.add_argument("--shard", default="spp", choices=["spp", "tiles"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--ref-spp", type=int, default=16, help="--impl reference: samples per pixel of one bounded step")
    return ap.parse_args()


# name -> (scene, width, height, spp, max_bounces); BASELINE.json configs
WORKLOADS = {
    "teapot-1080p": ("teapot", 1920, 1080, 64, 5),     # configs[1]: teapot 1920x1080, 64 spp, 5 bounces
    "cube-1080p": ("cube", 1920, 1080, 64, 5),         # configs[1], the cube half
    "teapot-bench": ("teapot", 2048, 1536, 10, 0),     # configs[0]: benches/render_teapot.rs:12-24 (reference shader)
    "atrium-1080p": ("atrium", 1920, 1080, 256, 5),    # configs[2] on the procedural Sponza stand-in (sponza.obj is absent)
    "atrium-4k": ("atrium", 3840, 2160, 1024, 8),      # configs[3] on the stand-in: spp split over the GPUs, NCCL film reduce
}


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


def load_workload(name, args):
    """Scene (host BVH), camera and frame parameters of a workload.  Built on the host, outside every timed region
    (benches/render_teapot.rs:25-27 excludes the scene build too)."""
    import minipath_b200 as mp

    scene, w, h, spp, bounces = WORKLOADS[name]
    if args.spp > 0:
        spp = args.spp
    if args.bounces >= 0:
        bounces = args.bounces
    data = os.path.join(ROOT, "data")
    if scene == "teapot":  # camera of benches/render_teapot.rs:12-19
        bvh = mp.TriangleBvh.with_obj(os.path.join(data, "teapot.obj"))
        cam = mp.Camera.default().look_at((0, 2, 10), (0, 1.5, 0), (0, 1, 0)).f_number(4.8).focus_distance(10.0)
    elif scene == "cube":  # fan-triangulated quads (documented deviation), SURVEY §8(d) gate camera
        bvh = mp.TriangleBvh.with_obj(os.path.join(data, "cube.obj"), quad_mode=1)
        cam = mp.Camera.default().look_at((1.5, 1.2, 2.5), (0, 0, 0), (0, 1, 0)).f_number(64.0)
    else:
        from minipath_b200 import scenes

        bvh, cam = scenes.atrium_scene()
    return bvh, cam, w, h, spp, bounces


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def oracle_frame(oracle, bvh, sampler, w, h, tiles, params, threads):
    """One bounded frame on the CPU oracle; returns (seconds, rays, camera samples)."""
    from minipath_b200 import _ffi

    c = _ffi.MpCounters()
    film, secs, c = oracle.render_tiles(bvh, sampler, w, h, tiles, params, threads=threads, counters=c)
    return secs, int(c.rays), int(c.camera_rays)


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port — the Rust crate cannot be built here) on all host
    cores, same workload/metric; each step is a bounded sample (ref_spp samples per pixel of the full frame)."""
    if rank != 0:
        return
    import minipath_b200 as mp
    from oracle_lib import Oracle

    bvh, cam, w, h, spp, bounces = load_workload(args.workload, args)
    oracle = Oracle()
    oracle.o.orc_set_simd(1)
    threads = host_cores()
    sampler = cam.build_sampler((w, h))
    tiles = mp.full_screen_tiles(w, h, 64)
    st = mp.RenderSettings(64, args.ref_spp, (w, h), max_bounces=bounces, seed=7)
    for _ in range(args.warmup):
        oracle_frame(oracle, bvh, sampler, w, h, tiles, st.params(), threads)
    secs = rays = samples = 0
    for _ in range(args.steps):
        s, r, c = oracle_frame(oracle, bvh, sampler, w, h, tiles, st.params(), threads)
        secs, rays, samples = secs + s, rays + r, samples + c
    mrays = rays / secs / 1e6
    sample = f"{args.workload}: full {w}x{h} frame at {args.ref_spp} of {spp} spp per step, {bounces} bounces"
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic camera samples on data/teapot.obj" if "teapot" in args.workload else "synthetic",
        "samples_per_s": samples / secs,
        "config": {"workload": args.workload, "resolution": [w, h], "spp": spp, "max_bounces": bounces, "tile_size": 64,
                   "note": "CPU restatement of the reference path (oracle port; cargo/rustc absent), AVX2+FMA lanes, thread per core"},
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline(args, bvh, cam, w, h, spp, bounces):
    """Oracle on the host cores, bounded to roughly 10-20 s: 1 spp to calibrate, then as many spp as fit."""
    import minipath_b200 as mp
    from oracle_lib import Oracle

    oracle = Oracle()
    oracle.o.orc_set_simd(1)
    threads = host_cores()
    sampler = cam.build_sampler((w, h))
    tiles = mp.full_screen_tiles(w, h, 64)
    st1 = mp.RenderSettings(64, 1, (w, h), max_bounces=bounces, seed=7)
    s1, _, _ = oracle_frame(oracle, bvh, sampler, w, h, tiles, st1.params(), threads)
    n = int(max(1, min(spp, round(12.0 / max(s1, 1e-3)))))
    st = mp.RenderSettings(64, n, (w, h), max_bounces=bounces, seed=7)
    secs, rays, samples = oracle_frame(oracle, bvh, sampler, w, h, tiles, st.params(), threads)
    return {"value": rays / secs / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port", "samples_per_s": samples / secs,
            "seconds": secs, "sample": f"{args.workload}: full {w}x{h} frame at {n} of {spp} spp, {bounces} bounces, AVX2 oracle"}


def init_dist(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, local, world


def run_b200(args, rank, local, world):
    import torch
    import torch.distributed as dist

    import minipath_b200 as mp
    from minipath_b200 import dist as mpdist

    if not torch.cuda.is_available() or mp.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device — the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    bvh, cam, w, h, spp, bounces = load_workload(args.workload, args)
    order = mp.ORDER_REFERENCE if args.order == "reference" else mp.ORDER_NEAR_FIRST
    sampler = cam.build_sampler((w, h))
    tiles = mp.full_screen_tiles(w, h, 64)

    stream = torch.cuda.Stream()
    ctx = mp.Context(local)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_node_layout({"auto": mp.LAYOUT_AUTO, "decoded": mp.LAYOUT_DECODED, "quantised": mp.LAYOUT_QUANTISED}[args.layout])
    ctx.upload_scene(bvh)
    ctx.set_camera(sampler)
    film = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
    ctx.bind_film(film.data_ptr(), w, h)
    rgba = torch.empty((h, w), dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def reduce_fn(t):
        dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)

    modes = {"megakernel": mp.MODE_MEGAKERNEL, "wavefront-thread": mp.MODE_WAVEFRONT, "wavefront-octet": mp.MODE_WAVEFRONT,
             "wavefront-persist": mp.MODE_WAVEFRONT, "wavefront-hybrid": mp.MODE_WAVEFRONT,
             "wavefront-pool": mp.MODE_WAVEFRONT, "wavefront-hybrid-pool": mp.MODE_WAVEFRONT,
             "wavefront-hybrid-pool-s0": mp.MODE_WAVEFRONT}
    variants = {"megakernel": mp.TRACE_THREAD, "wavefront-thread": mp.TRACE_THREAD, "wavefront-octet": mp.TRACE_OCTET,
                "wavefront-persist": mp.TRACE_PERSIST, "wavefront-hybrid": mp.TRACE_HYBRID,
                "wavefront-pool": mp.TRACE_POOL, "wavefront-hybrid-pool": mp.TRACE_HYBRID_POOL,
                "wavefront-hybrid-pool-s0": mp.TRACE_HYBRID_POOL_S0}
    if bvh.stats()["max_stack"] > 128:  # deeper than the local-memory stack of the thread-per-ray kernels
        modes = {"wavefront-octet": mp.MODE_WAVEFRONT, "wavefront-pool": mp.MODE_WAVEFRONT}

    def make_step(mode_key):
        mode = modes[mode_key]

        def render_fn(sub_tiles, base, count):
            st = mp.RenderSettings(64, count, (w, h), max_bounces=bounces, seed=7, mode=mode, order=order)
            ctx.render_tiles(sub_tiles, st.params(sample_base=base))

        def step():
            ctx.set_trace_variant(variants[mode_key])
            film.zero_()
            # weak scaling: every rank renders the full frame with its own `spp` sample indices; total spp = spp * world
            mpdist.render_sharded(render_fn, film, tiles, spp * world if args.shard == "spp" else spp, rank, world, args.shard,
                                  reduce_fn=reduce_fn if world > 1 else None)
            if rank == 0:
                ctx.resolve_device(spp * world if args.shard == "spp" else spp, rgba.data_ptr())
        return step

    def timed(step, n_warm, n_steps):
        """Per-step CUDA events on the launching stream, L2 flushed between steps (outside the event pairs)."""
        with torch.cuda.stream(stream):
            for _ in range(n_warm):
                step()
            evs = []
            for _ in range(n_steps):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                step()
                e1.record(stream)
                evs.append((e0, e1))
            stream.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- pick the pipeline: the A/B the north_star asks for, measured here, reported in config ----
    ab = {}
    if args.mode == "auto":
        for name in modes:
            barrier()
            ab[name] = timed(make_step(name), 1, 1)
        if world > 1:
            t = torch.tensor([ab[k] for k in modes], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ab = {k: float(t[i]) for i, k in enumerate(modes)}
        mode_name = min(ab, key=ab.get)
    else:
        mode_name = args.mode
    step = make_step(mode_name)

    # ---- counted pass (untimed): rays and node visits of one step, deterministic for a given seed ----
    ctx.enable_counters(True)
    ctx.reset_counters()
    with torch.cuda.stream(stream):
        step()
    stream.synchronize()
    cnt = ctx.counters()
    ctx.enable_counters(False)

    # ---- timed region ----
    sampler_thread = ClockSampler(local) if rank == 0 else None
    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
    ctx.reset_counters()
    barrier()
    if sampler_thread:
        sampler_thread.start()
    total_ms = timed(step, 0, args.steps)
    barrier()
    clocks = sampler_thread.stop() if sampler_thread else None
    launches = ctx.counters()["launches"]
    if world > 1:
        t = torch.tensor([total_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t[0])
        r = torch.tensor([cnt["rays"], cnt["camera_rays"]], device="cuda", dtype=torch.float64)
        dist.all_reduce(r, op=dist.ReduceOp.SUM)
        rays_job, samples_job = float(r[0]), float(r[1])
    else:
        rays_job, samples_job = float(cnt["rays"]), float(cnt["camera_rays"])
    ms_per_step = total_ms / args.steps
    value = rays_job / (ms_per_step * 1e-3) / 1e6

    res = dict(ctx=ctx, bvh=bvh, cam=cam, w=w, h=h, spp=spp, bounces=bounces, tiles=tiles, cnt=cnt, ab=ab, mode_name=mode_name,
               ms_per_step=ms_per_step, value=value, world=world, samples_per_s=samples_job / (ms_per_step * 1e-3), clocks=clocks,
               launches=launches, rays_job=rays_job, stream=stream, torch=torch, dist=dist, modes=modes, variants=variants, order=order, barrier=barrier)
    return res


def run_e2e(args, res, rank, local, world):
    """The same frame through the public API with host buffers.  N=1: `minipath_b200.render()` + wait() + image()
    (mp_render: context, scene upload from host arrays, tile batches, RGBA8 tiles copied back under the image lock).
    N>1: every rank uploads the host scene, renders its sample slice, one NCCL reduce, rank 0 resolves to host."""
    import minipath_b200 as mp

    torch, dist = res["torch"], res["dist"]
    w, h, spp, bounces, bvh, cam = res["w"], res["h"], res["spp"], res["bounces"], res["bvh"], res["cam"]
    raw = bvh.raw
    scene_bytes = (raw.n_inner * (128 + 32) + raw.n_packets * (144 + 32 + 8 * 16) + raw.n_vertices * 32)
    n_tiles = len(res["tiles"])
    mode = res["modes"][res["mode_name"]]
    if world == 1:
        st = mp.RenderSettings(64, spp, (w, h), max_bounces=bounces, seed=7, mode=mode, order=res["order"], trace_kernel=res["variants"][res["mode_name"]])

        def once():
            prog = mp.render(mp.Scene(bvh), cam, st, devices=[local])
            prog.wait()
            return prog.image()
    else:
        ctx, stream = res["ctx"], res["stream"]
        film = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
        host_img = np.empty((h, w, 4), np.uint8)

        def once():
            with torch.cuda.stream(stream):
                ctx.upload_scene(bvh)
                ctx.set_camera(cam.build_sampler((w, h)))
                ctx.bind_film(film.data_ptr(), w, h)
                film.zero_()
                base = rank * spp
                stp = mp.RenderSettings(64, spp, (w, h), max_bounces=bounces, seed=7, mode=mode, order=res["order"])
                ctx.render_tiles(res["tiles"], stp.params(sample_base=base))
                dist.reduce(film, dst=0, op=dist.ReduceOp.SUM)
                if rank == 0:
                    ctx.resolve_rgba8(spp * world, out=host_img)
                stream.synchronize()
            return host_img

    for _ in range(max(1, min(args.warmup, 2))):
        once()
    res["barrier"]()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        once()
    res["barrier"]()
    secs = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([secs], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = float(t[0])
    return {"value": res["rays_job"] * args.steps / secs / 1e6, "unit": "Mrays/s", "ms_per_step": secs / args.steps * 1e3,
            "h2d_bytes_per_step": int(scene_bytes + n_tiles * 16 + 52), "d2h_bytes_per_step": int(w * h * 4),
            "api": "minipath_b200.render()+wait()+image() [mp_render]" if world == 1 else "Context upload_scene+render_tiles+NCCL reduce+resolve_rgba8"}


def roofline(args, res):
    """Dominant kernel = the traversal (megakernel: the whole step is one launch; wavefront: wf_trace + wf_shadow).
    Algorithmic work from the visit counters of the counted pass (reference visiting order), DESIGN.md §5."""
    ctx, cnt = res["ctx"], res["cnt"]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    fp32 = ctx.probe_fp32_tflops()
    l2 = ctx.probe_l2_gbs(32 << 20)
    secs = res["ms_per_step"] * 1e-3
    bytes_ = cnt["inner_visits"] * BYTES_INNER + cnt["packet_visits"] * BYTES_PACKET + cnt["leaf_visits"] * BYTES_LEAF + cnt["rays"] * BYTES_RAY
    flops = cnt["inner_visits"] * FLOPS_INNER + cnt["packet_visits"] * FLOPS_PACKET
    raw = res["bvh"].raw
    scene_mb = (raw.n_inner * 160 + raw.n_packets * 176) / 1e6
    t_fp32, t_l2, t_hbm = flops / (fp32 * 1e12), bytes_ / (l2 * 1e9), bytes_ / (hbm * 1e9)
    # the BVH of every workload here is far smaller than the 126 MB L2, so the memory-side roof is L2, not HBM
    mem_bound = ("hbm", t_hbm, hbm) if scene_mb > 100 else ("l2", t_l2, l2)
    if t_fp32 >= mem_bound[1]:
        out = {"bound": "fp32", "achieved": flops / secs / 1e12, "peak": fp32, "unit": "TFLOP/s", "frac": t_fp32 / secs}
    else:
        out = {"bound": mem_bound[0], "achieved": bytes_ / secs / 1e9, "peak": mem_bound[2], "unit": "GB/s", "frac": mem_bound[1] / secs}
    traffic, traffic_note = None, "no ncu capture recorded for this workload/pipeline (scripts/ncu_traffic.py)"
    try:  # DRAM bytes of the traversal launches of one step, from an ncu capture of this very command line (profiles/traffic.json)
        rec = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"{args.workload}/{res['mode_name']}")
        if rec and args.spp == 0 and args.bounces < 0 and res["world"] == 1:
            traffic = rec["dram_bytes_per_step"] * (res["spp"] / rec["spp"] if "spp" in rec else 1.0)  # captured at a reduced spp: linear in spp
            traffic_note = f"dram__bytes_read.sum + dram__bytes_write.sum over the {rec['launches']} traversal launches of one step ({rec['source']})"
        elif rec:
            traffic_note = "recorded for the workload's own spp/bounces on 1 GPU only"
    except Exception:
        pass
    out.update({
        "traffic": traffic, "traffic_note": traffic_note, "algorithmic_bytes_per_step": bytes_,
        "kernel": {"megakernel": "render_mega_kernel", "wavefront-octet": "wf_trace_octet_kernel+wf_shadow_octet_kernel",
                   "wavefront-thread": "wf_trace_kernel+wf_shadow_kernel",
                   "wavefront-persist": "wf_trace_persist_kernel+wf_shadow_persist_kernel",
                   "wavefront-hybrid": "wf_primary_kernel (camera rays: ray-gen + trace + shade) + wf_trace_persist_kernel+wf_shadow_persist_kernel",
                   "wavefront-pool": "wf_trace_pool_kernel+wf_shadow_pool_kernel",
                   "wavefront-hybrid-pool": "wf_primary_kernel (camera rays: ray-gen + trace + shade) + wf_trace_pool_kernel+wf_shadow_pool_kernel",
                   "wavefront-hybrid-pool-s0": "wf_primary_kernel<SHADOW0> (camera rays + depth-0 sun rays: ray-gen + trace + shade + any-hit) + wf_trace_pool_kernel+wf_shadow_pool_kernel"}[res["mode_name"]] + " (share of step: see profiles/)",
        "peak_source": "FP32 FMA and L2 read bandwidth measured by mp_probe_* in this run; HBM " + hbm_src,
        "fp32": {"achieved_tflops": flops / secs / 1e12, "peak_tflops": fp32, "frac": t_fp32 / secs},
        "l2": {"achieved_gbs": bytes_ / secs / 1e9, "peak_gbs": l2, "frac": t_l2 / secs},
        "hbm": {"achieved_gbs": bytes_ / secs / 1e9, "peak_gbs": hbm, "frac": t_hbm / secs, "note": "BVH is cache-resident; HBM is not the roof"},
        "per_ray": {"inner_visits": cnt["inner_visits"] / max(1, cnt["rays"]), "packet_visits": cnt["packet_visits"] / max(1, cnt["rays"]),
                    "bytes": bytes_ / max(1, cnt["rays"]), "flops": flops / max(1, cnt["rays"])},
        "scene_mb": scene_mb,
    })
    return out


def main():
    args = parse_args()
    rank, local, world = init_dist(args)
    if world > 1 and args.gpus != world:
        args.gpus = world
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    res = run_b200(args, rank, local, world)
    e2e = None if args.no_e2e else run_e2e(args, res, rank, local, world)
    if rank == 0:
        rl = roofline(args, res)
        cb = None
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_baseline(args, res["bvh"], res["cam"], res["w"], res["h"], res["spp"], res["bounces"])
        w, h, spp, bounces = res["w"], res["h"], res["spp"], res["bounces"]
        st = res["bvh"].stats()
        line = {
            "metric": "Mrays/s", "value": res["value"], "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak" if args.shard == "spp" else "strong",
            "vs_baseline": None, "dtype": "f32",
            "data": "synthetic camera samples (counter-based RNG) on " + ("the reference's data/%s.obj" % WORKLOADS[args.workload][0]
                                                                         if WORKLOADS[args.workload][0] != "atrium" else
                                                                         "a procedural Sponza STAND-IN (sponza.obj is absent from the reference)"),
            "samples_per_s": res["samples_per_s"],
            "config": {"workload": args.workload, "resolution": [w, h], "spp_per_gpu": spp, "spp_total": spp * world if args.shard == "spp" else spp,
                       "max_bounces": bounces, "tile_size": 64, "pipeline": res["mode_name"], "traversal_order": args.order, "node_layout": args.layout,
                       "pipeline_ab_ms": res["ab"], "sharding": args.shard if world > 1 else "none",
                       "bvh": {"triangles": st["n_triangles"], "inner_nodes": st["n_inner"], "packets": st["n_packets"]},
                       "l2": "flushed between timed steps (256 MB fill outside the event pairs)",
                       "note": "max_bounces>0 is the extension integrator (the reference traces primary rays only)"},
            "rays_per_step": res["rays_job"], "gpu_launches": int(res["launches"]), "clocks": res["clocks"],
            "roofline": rl, "cpu_baseline": cb, "e2e": e2e,
        }
        print(json.dumps(line), flush=True)
    res["ctx"].close()
    if world > 1:
        res["dist"].barrier()
        res["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
