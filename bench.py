#!/usr/bin/env python
"""bench.py — one JSON line for BASELINE.json's metric on BASELINE config 2.

A "step" is ONE FRAME of the hot path (RayTracer::draw, src/kernel.cu:259-362) on the 1 047 552-triangle procedural
terrain at 1920x1080, 1 spp:
    LBVH rebuild (rtrt_build_bvh) -> path trace (rtrt_pathtrace: blue-noise sampler, BSDF + one-sample MIS, sky/sun
    CDF sampling, <= 5 TraverseBvh calls per pixel) -> temporal + a-trous denoise (rtrt_denoise).
`value`  : Mrays/s = TraverseBvh calls executed (device counter) / whole-step device time, inputs resident in HBM,
           CUDA events on the library's stream, L2 flushed between steps, max over ranks.
`e2e`    : same metric through the C ABI with HOST buffers: every step uploads the animated vertex + normal buffers
           from pinned host memory (rtrt_update_vertices), renders the frame (build / pathtrace / denoise) and reads the
           denoised frame back to pinned host memory (rtrt_read_buffer_async) inside the timed region.  Frames are kept in
           flight: the library runs both copies on its copy stream, ordered by events, so they overlap the kernels of the
           neighbouring frames; all K steps, their copies and the L2 flushes lie between ONE pair of events.
N > 1    : one process per GPU, the frame is partitioned BY SAMPLE (north_star: "by tile or by sample"): every rank
           renders the same view with its own sample index (frame number), BVH rebuilt redundantly per GPU, and the
           denoised frames are averaged with one NCCL all-reduce (the framebuffer combine) inside the timed region
           -> per-GPU work fixed -> "scaling": "weak".
           `--mode rows` shards ONE frame by row bands instead (strong scaling; parallel.RowShardedRayTracer: band path
           trace, NCCL all-gather of the G-buffers, band + halo denoise, NCCL gather of the frame); `--bvh broadcast` builds
           the BVH on rank 0 and broadcasts it instead of rebuilding it on every GPU.
`--workload C5` : BASELINE config 5 — 3840x2160, 64 spp progressive (no denoiser), 16 777 216-triangle terrain, row bands,
           one gather of the fp32 accumulation buffer per step (always `--mode rows`).
`--impl reference` : the reference has no CPU path and does not build on Linux; this tier's reference arm times the CPU
           oracle (oracle/, multithreaded C++ restatement) with all host threads; every step is one WHOLE frame of the workload
           (about 1.3 s on 16 threads), `ms_per_step` is that step's own wall time.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "real-time-ray-tracing_b200"

WIDTH, HEIGHT = 1920, 1080
SCENE = "C2ref"
WORKLOAD = ("C2: 1080p 1 spp frame, 1 047 552-tri procedural terrain (1023 batches), per-frame two-level LBVH rebuild + "
            "path trace (blue noise, Lambert BSDF + MIS, sky/sun) + temporal/a-trous denoise")
C5_SPP = 64
CAMERA = dict(pos=(0, 30, -60), yaw=0.0, pitch=-0.4)
BUILD_BYTES_PER_TRI = 248          # SURVEY.md §8(d) / BASELINE.md §4 / DESIGN.md
DENOISE_BYTES_PER_PX = 252         # reference pass sequence; fused lower bound 74
GBUF_BYTES_PER_PX = 30             # path-trace outputs: colour 8 + normal 8 + albedo 8 + depth 2 + motion 4
NODE_BYTES, TRI_BYTES = 64, 128    # reference layout bytes per node visit / leaf test (SURVEY §8d "secondary")
WIDE_NODE_BYTES, WIDE_TRI_BYTES = 80, 48   # wide layout (csrc/wide_bvh.cuh): 80 used bytes of one 128-byte line per visit, 48-byte triangles
CPU_ROWS = 1080                    # rows of the frame the CPU baseline path-traces and denoises: the whole frame (no extrapolation)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.stop_flag, self.idx = [], set(), False, gpu_index
        self.max_mhz = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.idx)],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0])); self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if "Active" in v and "Not" not in v:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def make_scene(pkg):
    gx, gz = pkg.scenes.TERRAIN_SIZES[SCENE]
    pos, nrm, idx = pkg.scenes.terrain(gx, gz, seed=1)
    cam = pkg.scenes.make_camera(CAMERA["pos"], CAMERA["yaw"], CAMERA["pitch"], WIDTH, HEIGHT)
    return pos, nrm, idx, cam


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (oracle/ — test infrastructure; executed here only as the timed CPU baseline / reference arm)
# ------------------------------------------------------------------------------------------------------------------
class CpuArm:
    """Bounded sample of the step on the host cores: FULL LBVH build + path trace + denoise of CPU_ROWS rows
    (5 bands of 27 rows spread over the frame).  Whole-frame time is estimated as build + (H / CPU_ROWS) * (pt + dn)."""

    def __init__(self, pkg):
        from oracle import orc
        orc.build()
        self.orc, self.pkg = orc, pkg
        self.cores = orc.default_threads()
        self.pos, self.nrm, self.idx, self.cam = make_scene(pkg)
        self.sky = orc.sky_generate(orc.sun_dir())
        nb = 5
        h = CPU_ROWS // nb
        self.bands = [(int((k + 0.5) * HEIGHT / nb) - h // 2, int((k + 0.5) * HEIGHT / nb) - h // 2 + h) for k in range(nb)]
        self.state = orc.DenoiseState(WIDTH, h * nb)
        self.sample = (f"one whole step per sample: LBVH build of {self.idx.shape[0]} tris + path trace + denoise of all {h * nb} rows of the {WIDTH}x{HEIGHT} frame "
                       f"(no extrapolation), {self.cores} host threads")

    def step(self, frame):
        orc = self.orc
        t0 = time.perf_counter()
        ob = orc.build_bvh(self.pos, self.nrm, self.idx, threads=self.cores)
        t1 = time.perf_counter()
        rays = 0
        parts = []
        for (b, e) in self.bands:
            f = orc.pathtrace(ob, self.cam, WIDTH, HEIGHT, frame, self.sky, threads=self.cores, rows=(b, e))
            rays += int(f.counters[0])
            parts.append((f, b, e))
        t2 = time.perf_counter()
        sl = np.concatenate([np.arange(b * WIDTH, e * WIDTH) for (_, b, e) in parts])
        f = parts[0][0]
        def cat(name):
            return np.concatenate([getattr(p, name)[b * WIDTH:e * WIDTH] for (p, b, e) in parts])
        n = sl.shape[0]
        color = np.zeros(n, orc.COLOR_DTYPE)
        color["rgb"] = cat("color").astype(np.float16); color["mask"] = cat("mask")
        nrm4 = np.zeros((n, 4), np.float16); nrm4[:, :3] = cat("normal")
        alb4 = np.zeros((n, 4), np.float16); alb4[:, :3] = cat("albedo")
        with np.errstate(over="ignore"):
            depth = cat("depth").astype(np.float16)
        motion = cat("motion").astype(np.float16)
        t3 = time.perf_counter()
        orc.denoise(self.state, frame, color, nrm4, depth, motion, alb4, threads=self.cores)
        t4 = time.perf_counter()
        scale = HEIGHT / float(n // WIDTH)
        build_s, pt_s, dn_s = t1 - t0, t2 - t1, t4 - t3
        frame_s = build_s + scale * (pt_s + dn_s)
        return {"rays": rays, "sample_s": build_s + pt_s + dn_s, "frame_s": frame_s, "frame_rays": rays * scale,
                "build_ms": 1e3 * build_s, "pathtrace_ms_frame": 1e3 * scale * pt_s, "denoise_ms_frame": 1e3 * scale * dn_s}

    def baseline(self, steps=1, warmup=0):
        rs = [self.step(2 + s) for s in range(warmup + steps)][warmup:]
        frame_s = float(np.mean([r["frame_s"] for r in rs])); frame_rays = float(np.mean([r["frame_rays"] for r in rs]))
        val = frame_rays / frame_s / 1e6
        return val, 1e3 * frame_s, {"value": val, "unit": "Mrays/s", "cores": self.cores, "kind": "port", "sample": self.sample,
                                    "build_ms": float(np.mean([r["build_ms"] for r in rs])),
                                    "pathtrace_ms_per_frame": float(np.mean([r["pathtrace_ms_frame"] for r in rs])),
                                    "denoise_ms_per_frame": float(np.mean([r["denoise_ms_frame"] for r in rs])),
                                    "pathtrace_mrays_s": float(np.mean([r["frame_rays"] / (r["pathtrace_ms_frame"] * 1e-3) / 1e6 for r in rs]))}


def run_reference(args, rank, world):
    """Reference arm (this tier): CPU oracle, all host threads, bounded sample of the same workload; rank 0 only."""
    if rank != 0:
        return
    pkg = importlib.import_module(PKG)
    arm = CpuArm(pkg)
    val, ms, cb = arm.baseline(steps=args.steps, warmup=args.warmup)
    line = {"impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": WORKLOAD, "sample": arm.sample},
            "cpu_baseline": cb, "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit_line(line)


class DevView:
    """A borrowed device buffer as a __cuda_array_interface__ object (so torch / NCCL can address it)."""

    def __init__(self, ptr, nbytes, typestr="<f2", itemsize=2):
        self.__cuda_array_interface__ = {"shape": (nbytes // itemsize,), "typestr": typestr, "data": (ptr, False), "version": 3,
                                         "strides": None}


# stdout carries exactly ONE JSON line: everything else that writes to fd 1 while the program runs (NCCL's version banner, the
# reference kernels' device printf) is sent to stderr; the line itself goes to the saved descriptor.
_REAL_STDOUT = None


def divert_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def ncu_traffic(key, field="dram_bytes"):
    """dram__bytes_read.sum + dram__bytes_write.sum (or another counter) of one launch, from the committed ncu --set full capture of this
    command (profiles/ncu_traffic.json, written by tools/ncu_to_profiles.py)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)[key][field]
    except Exception:
        return None


class quiet_stdout:
    """The reference's traversal kernel printf()s "push to full stack" from the device for a few rays per frame; stdout of this
    program must stay ONE JSON line, so fd 1 points at /dev/null while those kernels run (device printf is flushed by the sync)."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)
        return self

    def __exit__(self, *exc):
        try:
            import torch
            torch.cuda.synchronize()
        finally:
            os.dup2(self.saved, 1); os.close(self.saved); os.close(self.null)
        return False


def reference_kernels(pkg, rt, pos, nrm, idx, cam, rays_per_frame, warm=10, timed=50):
    """The reference's own CUDA kernels (oracle/_ref/librefcuda{,_fmad}.so, built for sm_100 from /root/reference by
    oracle/refcuda/build_ref.py) timed on this GPU on the same scene — the baseline north_star's ">= 3x" target is about.  SURVEY 8(d):
    10 warm-up + 50 timed launches per stage, CUDA events around each launch sequence, min and median; both the --fmad=false build (the
    one the parity tests compare with) and the default-flags (--fmad=true) build.  Baseline leg only, outside every timed region."""
    try:
        from oracle.refcuda import ref
        from oracle import orc
        if not ref.available() or not hasattr(ref.load(), "refcuda_frame_create") or idx.shape[0] >= 1023 * 1024 + 1:
            return {"unavailable": "oracle/_ref/librefcuda.so not present (or scene above the reference's 1023-batch limit)"}
        orc.build()
        out = {"what": "UpdateSceneGeometry+RadixSort+BuildLBVH+UpdateTLAS+RadixSort+BuildLBVH | PathTrace | TemporalSpatialDenoising kernel "
                       "sequence; CUDA events around each sequence, no per-stage device sync", "warmup": warm, "timed": timed}
        sky = orc.Sky(); sky.sky = rt.read_buffer(pkg.BUF_SKY); sky.sky_cdf = rt.read_buffer(pkg.BUF_SKY_CDF)
        sky.sun = rt.read_buffer(pkg.BUF_SUN); sky.sun_cdf = rt.read_buffer(pkg.BUF_SUN_CDF); sky.sun_dir, sky.cos_theta_max = rt.get_sun()
        for fmad in (False, True):
            if not ref.available(fmad):
                continue
            rb = ref.build_bvh(pos, nrm, idx, iterations=1, fmad=fmad)
            b = [rb.rebuild(1) for _ in range(warm + timed)][warm:]
            with ref.RefFrame(WIDTH, HEIGHT, fmad=fmad) as rf:
                rf.set_tables(orc.load_tables()[0]); rf.set_scene(rb)
                rf.set_materials(orc.default_materials(), np.full(rb.num_batches * 1024, 3, np.int32))
                rf.set_texture(0, orc.mip_chain(orc.DEFAULT_TEX[0])); rf.set_texture(1, orc.mip_chain(orc.DEFAULT_TEX[1])); rf.set_sky(sky)
                pt, dn = [], []
                for fr in range(1, warm + timed + 1):
                    pt.append(rf.pathtrace(cam, fr))
                    dn.append(rf.denoise(fr))
                pt, dn = pt[warm:], dn[warm:]
            rec = {"flags": "-arch=sm_100 --fmad=" + ("true" if fmad else "false"),
                   "build_ms": float(np.median(b)), "build_ms_min": float(np.min(b)),
                   "pathtrace_ms": float(np.median(pt)), "pathtrace_ms_min": float(np.min(pt)),
                   "denoise_ms": float(np.median(dn)), "denoise_ms_min": float(np.min(dn))}
            rec["frame_ms"] = rec["build_ms"] + rec["pathtrace_ms"] + rec["denoise_ms"]
            rec["frame_ms_min"] = rec["build_ms_min"] + rec["pathtrace_ms_min"] + rec["denoise_ms_min"]
            rec["mrays_s"] = rays_per_frame / (rec["frame_ms"] * 1e-3) / 1e6
            rec["pathtrace_mrays_s"] = rays_per_frame / (rec["pathtrace_ms"] * 1e-3) / 1e6
            if fmad:
                out["fmad_true"] = rec
            else:
                out.update(rec)
        return out
    except Exception as e:                                     # baseline leg: never fail the bench line
        return {"unavailable": repr(e)}


def _gpu_time(torch, rt, flush, fn, n=8, warm=2):
    """median CUDA-event time of fn(i) on the library's stream, L2 flushed before every call"""
    stream = torch.cuda.ExternalStream(rt.stream)
    ts = []
    for i in range(n + warm):
        with torch.cuda.stream(stream):
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); fn(i); b.record(stream); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts[warm:]))


def other_configs(pkg, torch, local, flush, hbm_peak):
    """BASELINE configs 1, 3 and 4 (the ones that are not the bench line) as sub-records of the N = 1 line, so that every config has a
    driver-run number: product times with CUDA events (L2 flushed), the reference kernels beside them where they can run the size."""
    out = {}
    try:
        from oracle.refcuda import ref
        have_ref = ref.available() and hasattr(ref.load(), "refcuda_frame_create")
    except Exception:
        ref, have_ref = None, False
    # C1: 10 240-triangle mesh, LBVH build + 256 x 256 primary closest hit
    W = H = 256
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    cam = pkg.scenes.make_camera(CAMERA["pos"], CAMERA["yaw"], CAMERA["pitch"], W, H)
    rays = pkg.scenes.primary_rays(cam, W, H)
    with pkg.RayTracer(W, H, device=local) as rt:
        rt.set_mesh(pos, idx, nrm)
        d_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda(); d_hits = torch.empty(rays.shape[0] * 16, dtype=torch.uint8, device="cuda")
        c1 = {"workload": "C1: 10 240-tri mesh, LBVH build + 256x256 primary closest hit", "build_ms": _gpu_time(torch, rt, flush, lambda i: rt.build_bvh()),
              "traverse_ms": _gpu_time(torch, rt, flush, lambda i: rt.traverse_device(d_rays.data_ptr(), rays.shape[0], d_hits.data_ptr()))}
    if have_ref:
        rb = ref.build_bvh(pos, nrm, idx, iterations=11); c1["ref_kernels_build_ms"] = rb.ms
        _, c1["ref_kernels_traverse_ms"] = ref.traverse(rb, rays, iterations=11)
    out["c1"] = c1
    # C3: 4 194 304 triangles rebuilt per frame (vertex upload excluded, Morton-ordered input: per-batch sort + global top level)
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C3"], seed=1)
    with pkg.RayTracer(64, 64, device=local) as rt:
        rt.set_mesh(pos, idx, nrm)
        ms = _gpu_time(torch, rt, flush, lambda i: rt.build_bvh())
    gbs = BUILD_BYTES_PER_TRI * idx.shape[0] / (ms * 1e-3) / 1e9
    out["c3"] = {"workload": "C3: per-frame LBVH rebuild of 4 194 304 triangles (4096 batches, global top level)", "triangles": int(idx.shape[0]), "build_ms": ms,
                 "mtris_s": idx.shape[0] / ms / 1e3, "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak},
                 "reference_kernels": "not runnable (its top level is limited to 1023 batches, src/init.cu:126)"}
    # C4: temporal accumulation + a-trous denoise at 3840 x 2160 on path-traced G-buffers: the reference's schedule (7x7 + 5x5 with strides 3, 6, 12)
    # and the 5-iteration schedule the config names (5x5 with strides 1, 2, 4, 8, 16); 4K path trace beside it
    W, H = 3840, 2160
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES[SCENE], seed=1)
    cam = pkg.scenes.make_camera(CAMERA["pos"], CAMERA["yaw"], CAMERA["pitch"], W, H)
    c4 = {"workload": "C4: temporal + a-trous denoise at 3840x2160 on path-traced G-buffers of the C2 scene", "width": W, "height": H}
    with pkg.RayTracer(W, H, device=local) as rt:
        rt.set_mesh(pos, idx, nrm); rt.build_bvh(); rt.set_camera(cam, reset_history=True)
        def pt(i): rt.set_frame_number(2 + i); rt.pathtrace()
        c4["pathtrace_ms"] = _gpu_time(torch, rt, flush, pt, n=5); c4["pathtrace_mrays_s"] = rt.timings()["rays_traced"] / c4["pathtrace_ms"] / 1e3
        for name, wide_pass in (("denoise_ms", 1), ("denoise_5_iterations_ms", 2)):
            rt.set_denoise_params(None, pkg.PassSettings(1, 1, wide_pass, 1))
            stream = torch.cuda.ExternalStream(rt.stream); ts = []
            for i in range(8):
                rt.set_frame_number(20 + i); rt.pathtrace()
                with torch.cuda.stream(stream):
                    flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream); rt.denoise(); b.record(stream); torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            c4[name] = float(np.median(ts[2:]))
        gbs = DENOISE_BYTES_PER_PX * W * H / (c4["denoise_ms"] * 1e-3) / 1e9
        c4["roofline"] = {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "bytes_per_px": DENOISE_BYTES_PER_PX}
        if have_ref:
            from oracle import orc
            sky = orc.Sky(); sky.sky = rt.read_buffer(pkg.BUF_SKY); sky.sky_cdf = rt.read_buffer(pkg.BUF_SKY_CDF)
            sky.sun = rt.read_buffer(pkg.BUF_SUN); sky.sun_cdf = rt.read_buffer(pkg.BUF_SUN_CDF); sky.sun_dir, sky.cos_theta_max = rt.get_sun()
    if have_ref:
        rb = ref.build_bvh(pos, nrm, idx)
        with ref.RefFrame(W, H) as rf:
            rf.set_tables(orc.load_tables()[0]); rf.set_scene(rb); rf.set_materials(orc.default_materials(), np.full(rb.num_batches * 1024, 3, np.int32))
            rf.set_texture(0, orc.mip_chain(orc.DEFAULT_TEX[0])); rf.set_texture(1, orc.mip_chain(orc.DEFAULT_TEX[1])); rf.set_sky(sky)
            c4["ref_kernels_pathtrace_ms"] = rf.pathtrace(cam, 2, iterations=3)
            for name, wide_pass in (("ref_kernels_denoise_ms", 1), ("ref_kernels_denoise_5_iterations_ms", 2)):
                ts = []
                for fr in range(1, 6):
                    rf.pathtrace(cam, fr); ts.append(rf.denoise(fr, passes=(1, 1, wide_pass, 1)))
                c4[name] = float(np.median(ts[2:]))
    out["c4"] = c4
    return out


def strong_c5(args, pkg, torch, dist, rank, world, local, make_group):
    """BASELINE config 5 as a sub-record of every --gpus N line: 3840x2160, 64 spp progressive (no denoiser), 16 777 216-triangle terrain,
    partitioned BY SAMPLE through the library's own multi-GPU entry point (rtrt_render_progressive_multi: rank r renders samples r, r + N, ...
    of the whole frame, BVH rebuilt per GPU, one NCCL all-reduce of the fp32 accumulation image).  Fixed total work -> strong scaling."""
    W, H, spp = 3840, 2160, C5_SPP
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C5"], seed=1)
    cam = pkg.scenes.make_camera(CAMERA["pos"], CAMERA["yaw"], CAMERA["pitch"], W, H)
    rt = pkg.RayTracer(W, H, device=local).init()
    rt.set_option("wide_bvh", int(args.layout == "wide"))
    rt.set_mesh(pos, idx, nrm); rt.set_camera(cam, reset_history=True)
    mg = make_group(rt, "samples")
    stream = torch.cuda.ExternalStream(rt.stream, device=torch.device("cuda", local))
    def step():
        if mg is None:
            rt.build_bvh()
            for s in range(spp):
                rt.set_frame_number(1 + s); rt.pathtrace(); rt.accumulate(reset=(s == 0))
        else:
            mg.render_progressive(spp, first_frame=1)
    step(); rt.sync()
    r0 = rt.timings()["rays_traced_total"]
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    steps = 2
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(steps): step()
    b.record(stream)
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
    r = torch.tensor([float(rt.timings()["rays_traced_total"] - r0)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(r, op=dist.ReduceOp.SUM)
    rt.cleanup()
    ms = t.item() / steps
    return {"workload": f"C5: 3840x2160, {spp} spp progressive path trace (no denoiser), {idx.shape[0]}-tri terrain (16 384 batches, global top level), "
                        f"sample-partitioned x{world}, BVH rebuilt per GPU and step, one NCCL all-reduce of the fp32 accumulation image",
            "scaling": "strong", "steps": steps, "ms_per_step": ms, "ms_per_sample": ms / spp, "rays_per_step": r.item() / steps,
            "mrays_s": r.item() / steps / (ms * 1e-3) / 1e6}


def run_c5(args, rank, world, local, torch, dist, pkg):
    """BASELINE config 5: 3840x2160, 64 spp progressive (no denoiser), 16 777 216-triangle terrain, row bands over the GPUs,
    BVH rebuilt per GPU or built once and broadcast (--bvh), one NCCL gather of the fp32 accumulation buffer per step."""
    W, H, spp = 3840, 2160, args.spp
    gx, gz = pkg.scenes.TERRAIN_SIZES["C5"]
    pos, nrm, idx = pkg.scenes.terrain(gx, gz, seed=1)
    cam = pkg.scenes.make_camera(CAMERA["pos"], CAMERA["yaw"], CAMERA["pitch"], W, H)
    rt = pkg.RayTracer(W, H, device=local).init()
    rt.set_mesh(pos, idx, nrm)
    rt.set_camera(cam, reset_history=True)
    sh = pkg.RowShardedRayTracer(rt, W, H, rank, world, bvh=args.bvh)
    stream = torch.cuda.ExternalStream(rt.stream, device=torch.device("cuda", local))
    h_out = torch.empty(W * H * 16, dtype=torch.uint8).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # rays per step are deterministic (same frame numbers every step): count them once, untimed
    sh.build_bvh()
    rays_rank = 0
    by_samples = args.partition == "samples" and world > 1
    if by_samples:
        rt.set_row_range(0, H); rt.set_row_strips(0, 1, 0)
    for smp in range(spp):
        if by_samples and smp % world != rank:
            continue
        rt.set_frame_number(1 + smp); rt.pathtrace(); rays_rank += rt.timings()["rays_traced"]
    if by_samples:
        rt.set_row_range(*sh.band); rt.set_row_strips(sh.strip_height, world, rank)
    sampler = ClockSampler(local); sampler.start()
    res = {}
    for name, e2e in (("device", False), ("e2e", True)):
        steps = args.steps if not e2e else max(2, args.steps // 2)
        for _ in range(args.warmup if not e2e else 1):
            sh.render_progressive(spp, partition=args.partition)
        barrier()
        l0 = rt.timings()["kernel_launches"]
        evs = []
        for _ in range(steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            sh.render_progressive(spp, partition=args.partition)
            if e2e:
                rt.read_buffer_into(pkg.BUF_ACCUM_RADIANCE, h_out.data_ptr(), h_out.numel())
            b.record(stream)
            evs.append((a, b))
        barrier()
        res[name] = (float(np.sum([a.elapsed_time(b) for a, b in evs])), steps, rt.timings()["kernel_launches"] - l0)
        if not e2e:
            sampler.stop_flag = True; sampler.join()
    t = torch.tensor([res["device"][0], res["e2e"][0]], dtype=torch.float64, device="cuda")
    r = torch.tensor([float(rays_rank)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(r, op=dist.ReduceOp.SUM)
    dev_ms, e2e_ms = t.tolist(); rays_step = r.item()
    if rank == 0:
        ms_step = dev_ms / res["device"][1]
        e2e_step = e2e_ms / res["e2e"][1]
        line = {"metric": "Mrays/s", "value": rays_step / (ms_step * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": f"C5: 3840x2160, {spp} spp progressive path trace (no denoiser), {idx.shape[0]}-tri procedural terrain "
                                       f"({(idx.shape[0] + 1023) // 1024} batches, global top level), {args.partition}-partitioned x{world}",
                           "width": W, "height": H, "triangles": int(idx.shape[0]), "spp": spp, "rays_per_step": rays_step,
                           "parallelism": (f"row strips x{world}, BVH {args.bvh} per step, NCCL all-gather of the fp32 accumulation buffer" if not by_samples else
                                           f"samples x{world} (rank r renders samples r, r+N, ...), BVH {args.bvh} per step, NCCL all-reduce of the fp32 accumulation buffer"),
                           "l2": "inputs larger than L2 (3.2 GB BVH)", "ms_per_sample": ms_step / spp},
                "e2e": {"value": rays_step / (e2e_step * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_step, "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": int(h_out.numel()), "note": "static scene: no per-step upload; the accumulated frame is read back"},
                "gpu_launches": int(res["device"][2]), "clocks": sampler.summary()}
        emit_line(line)
    rt.cleanup()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--mode", default="samples", choices=["samples", "rows"])
    ap.add_argument("--bvh", default="rebuild", choices=["rebuild", "broadcast"])
    ap.add_argument("--workload", default="C2", choices=["C2", "C5"])
    ap.add_argument("--spp", type=int, default=C5_SPP)
    ap.add_argument("--partition", default="samples", choices=["samples", "rows"], help="C5: how the 64 spp are split over the GPUs")
    ap.add_argument("--layout", default="wide", choices=["wide", "reference"],
                    help="traversal layout of the timed steps: the 8-wide compressed BVH (option wide_bvh, derived from the reference-layout arrays "
                         "inside every rtrt_build_bvh) or the reference's own node / triangle arrays; the other one is timed as a sub-record")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling sub-records (C5 at 64 spp, C2 rows mode)")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C3 / C4 sub-records of the N = 1 line")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-kernels", action="store_true")
    args = ap.parse_args()
    divert_stdout()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module(PKG)        # raises if librtrt.so is missing: no fallback
    if args.workload == "C5":
        run_c5(args, rank, world, local, torch, dist, pkg)
        if world > 1:
            dist.destroy_process_group()
        return
    pos, nrm, idx, cam = make_scene(pkg)
    n_tris, n_px = int(idx.shape[0]), WIDTH * HEIGHT
    hbm_peak, peak_kind = peaks()

    rt = pkg.RayTracer(WIDTH, HEIGHT, device=local).init()
    rt.set_mesh(pos, idx, nrm)
    rt.set_camera(cam, reset_history=True)
    wide = args.layout == "wide"
    rt.set_option("wide_bvh", int(wide))
    node_bytes, tri_bytes = (WIDE_NODE_BYTES, WIDE_TRI_BYTES) if wide else (NODE_BYTES, TRI_BYTES)
    by_rows = args.mode == "rows"
    # N > 1: the library's own multi-GPU entry points (csrc/multi.cu: NCCL inside librtrt.so); torch.distributed only hands out the NCCL id,
    # synchronises the ranks around the timed regions and reduces the timing scalars
    def make_group(ctx, mode):
        return pkg.parallel.MultiGpuRayTracer.bootstrap(ctx, mode=mode) if world > 1 else None
    mg = make_group(rt, "rows" if by_rows else "samples")
    rows = mg.band()[:2] if (mg is not None and by_rows) else (0, HEIGHT)
    stream = torch.cuda.ExternalStream(rt.stream, device=torch.device("cuda", local))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2
    h_pos = torch.from_numpy(pos.reshape(-1).copy()).pin_memory()
    h_nrm = torch.from_numpy(nrm.reshape(-1).copy()).pin_memory()
    h_out = [torch.empty(n_px * 8, dtype=torch.uint8).pin_memory() for _ in range(2)]     # frames in flight: two host frame buffers
    torch.cuda.synchronize()
    frame_no = [1]
    lib_stage_ms = []                           # N > 1: per step, the library's own CUDA-event stage times (build, path trace, denoise)
    wide_ms = []                                # per step: the collapse to the wide layout inside rtrt_build_bvh (library timing wide_build_ms)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def frame(e):
        """build | path trace | (G-buffer redistribution) | denoise with halo exchange | (frame gather / sample combine)"""
        rt.set_frame_number(frame_no[0])
        if mg is not None:
            frame_no[0] += 1 if by_rows else world       # samples mode: rank r renders sample index frame + r
            mg.draw()                                    # rtrt_render_frame_multi: every stage and every NCCL exchange inside the library
            return
        frame_no[0] += 1
        rt.build_bvh()
        if e: e[1].record(stream)
        rt.pathtrace()
        if e: e[2].record(stream); e[3].record(stream)
        rt.denoise()
        if e: e[4].record(stream)

    def step_device(e):
        if e: e[0].record(stream)
        frame(e)
        if e: e[5].record(stream)

    def step_e2e(i):
        """One frame through the public API with HOST buffers, nothing blocks the host: vertex upload (copy stream, overlaps the
        previous frame's path trace), frame, read-back of the denoised frame (snapshot + copy stream, overlaps the next frame)."""
        if mg is None:
            rt.update_vertices(h_pos.data_ptr(), h_nrm.data_ptr())             # H2D, pinned
        else:                                                                  # rank 0 uploads once, NCCL broadcast over NVLink to the other GPUs
            mg.update_vertices(h_pos.data_ptr() if rank == 0 else None, h_nrm.data_ptr() if rank == 0 else None)
        frame(None)
        if rank == 0:                                                          # the combined frame is needed on one host only
            rt.read_buffer_async(pkg.BUF_RENDER_COLOR, h_out[i & 1].data_ptr(), h_out[i & 1].numel())   # D2H, pinned

    def timed_e2e(steps, warmup):
        """K frames in flight, timed as ONE region on the library's stream: start event, K x (L2 flush, step), join of the copy
        stream, end event.  The flushes are inside the region (they cannot be subtracted from overlapped work)."""
        for i in range(warmup):
            step_e2e(i)
        rt.sync(); barrier()
        r0 = rt.timings()["rays_traced_total"]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(steps):
            with torch.cuda.stream(stream):
                flush.zero_()
            step_e2e(i)
        rt.join()
        e1.record(stream)
        rt.sync(); barrier()
        return e0.elapsed_time(e1), int(rt.timings()["rays_traced_total"] - r0)

    def timed(step_fn, steps, warmup):
        for _ in range(warmup):
            step_fn(None)
        barrier()
        lbefore = rt.timings()["kernel_launches"]
        evs, rays = [], 0
        kms.clear()
        for _ in range(steps):
            with torch.cuda.stream(stream):
                flush.zero_()                                            # L2 flush, outside the event pair
            e = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
            step_fn(e)
            evs.append(e)
            tmg = rt.timings()                                           # synchronises the stream (outside the event pair)
            rays += tmg["rays_traced"]
            kms.append(tmg["pt_kernel_ms"])
            lib_stage_ms.append([tmg["build_bvh_ms"], tmg["pathtrace_ms"], tmg["denoise_ms"]])
            wide_ms.append(tmg["wide_build_ms"])
        barrier()
        nl = rt.timings()["kernel_launches"] - lbefore
        tot = float(np.sum([e[0].elapsed_time(e[5]) for e in evs]))
        stages = None
        if step_fn is step_device and mg is None:
            stages = [float(np.mean([e[k].elapsed_time(e[k + 1]) for e in evs])) for k in range(5)]
        elif step_fn is step_device:
            b_ms, p_ms, d_ms = np.mean(np.asarray(lib_stage_ms[-steps:], np.float64), axis=0)    # denoise_ms includes its halo exchanges
            w_ms = float(np.mean(wide_ms[-steps:])) if wide_ms else 0.0                           # the wide collapse is its own entry of stages_ms
            stages = [float(b_ms), float(p_ms), 0.0, float(d_ms), max(tot / steps - float(b_ms + p_ms + d_ms) - w_ms, 0.0)]
        return tot, stages, int(nl), int(rays)

    kms = []                                    # per step: CUDA-event times of the path-trace kernels (library option kernel_timing)
    rt.set_option("kernel_timing", 1)
    sampler = ClockSampler(local); sampler.start()
    total_ms, stages, launches, rays = timed(step_device, args.steps, args.warmup)
    kernel_ms = np.mean(np.asarray(kms, np.float64), axis=0) if kms and not by_rows else np.zeros(6)
    rt.set_option("kernel_timing", 0)
    sampler.stop_flag = True; sampler.join()
    frame_sha1 = None
    if world == 1:                              # the frame the timed region ended on: two runs of one commit can be compared
        import hashlib
        frame_sha1 = hashlib.sha1(rt.read_buffer(pkg.BUF_RENDER_COLOR).tobytes()).hexdigest()
    e2e_steps = max(3, args.steps // 2)
    e2e_ms, e2e_rays = timed_e2e(e2e_steps, 3)

    # untimed: one frame with the visit counters on -> algorithmic bytes of the traversal
    rt.set_option("count_visits", 1)
    rt.set_frame_number(2); rt.pathtrace()
    tm = rt.timings()
    node_visits, leaf_tests, rays_cnt = tm["node_visits"], tm["leaf_tests"], tm["rays_traced"]
    r1_nodes, r1_leaves, r1_rays = tm["round_node_visits"][0], tm["round_leaf_tests"][0], tm["round_rays"][0]
    rt.set_option("count_visits", 0)

    # untimed: the ceilings a traversal without RT cores runs against on this GPU (csrc/microbench.cu)
    mb = rt.microbench() if rank == 0 else None

    # ---- the other traversal layout on the same workload (N = 1): both modes in one line ----
    other_layout = None
    if world == 1:
        rt.set_option("wide_bvh", 0 if wide else 1)
        for _ in range(3):
            step_device(None)
        torch.cuda.synchronize()
        ev, r0 = [], rt.timings()["rays_traced_total"]
        n_other = max(5, args.steps // 3)
        for _ in range(n_other):
            with torch.cuda.stream(stream):
                flush.zero_()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
            step_device(e); ev.append(e)
        torch.cuda.synchronize()
        o_ms = float(np.sum([e[0].elapsed_time(e[5]) for e in ev])) / n_other
        o_rays = (rt.timings()["rays_traced_total"] - r0) / n_other
        other_layout = {"layout": "reference" if wide else "wide", "steps": n_other, "ms_per_step": o_ms, "value": o_rays / (o_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                        "stages_ms": {k: float(np.mean([e[i].elapsed_time(e[i + 1]) for e in ev])) for i, k in ((0, "build_bvh"), (1, "pathtrace"), (3, "denoise"))},
                        "note": "reference = the reference's node / triangle arrays walked with its loop, caps and quirks (per-ray hits bit-identical to it); "
                                "wide = 8-wide fp16 layout derived from them inside every build"}
        rt.set_option("wide_bvh", int(wide))
        rt.build_bvh()

    # ---- strong-scaling sub-records (fixed total work): ONE C2 frame sharded by rows, and BASELINE config 5 at 64 spp ----
    strong = None
    if not args.no_strong:
        strong = {}
        if mg is not None and not by_rows:
          try:                                                     # sub-records never fail the bench line (an error every rank sees alike)
            mg.set_mode("rows")
            k = 0
            def rows_step():
                nonlocal k
                rt.set_frame_number(100 + k); k += 1
                mg.draw()
            for _ in range(3): rows_step()
            barrier()
            r0 = rt.timings()["rays_traced_total"]
            n_rows_steps, ev = 10, []
            for _ in range(n_rows_steps):
                with torch.cuda.stream(stream):
                    flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream); rows_step(); b.record(stream); ev.append((a, b))
            barrier()
            tt = torch.tensor([float(np.sum([a.elapsed_time(b) for a, b in ev]))], dtype=torch.float64, device="cuda")
            rr = torch.tensor([float(rt.timings()["rays_traced_total"] - r0)], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX); dist.all_reduce(rr, op=dist.ReduceOp.SUM)
            band = mg.band()
            strong["c2_rows"] = {"workload": "C2, ONE 1080p frame sharded over the GPUs: path trace in interleaved 8-row strips, G-buffer redistribution, denoise in "
                                             "16-row-aligned bands with an NCCL halo exchange before every filter pass, frame gather (rtrt_render_frame_multi, rows mode)",
                                 "scaling": "strong", "steps": n_rows_steps, "ms_per_step": tt.item() / n_rows_steps,
                                 "mrays_s": rr.item() / (tt.item() * 1e-3) / 1e6, "rank0_band": list(band[:2])}
          except Exception as e:
            strong["c2_rows"] = {"unavailable": repr(e)}
          finally:
            mg.set_mode("samples")
        try:
            strong["c5_64spp"] = strong_c5(args, pkg, torch, dist, rank, world, local, make_group)
        except Exception as e:
            strong["c5_64spp"] = {"unavailable": repr(e)}
    configs = None
    if world == 1 and not args.no_configs:
        try:
            with quiet_stdout():
                configs = other_configs(pkg, torch, local, flush, hbm_peak)
        except Exception as e:                                 # sub-records never fail the bench line
            configs = {"unavailable": repr(e)}

    t = torch.tensor([total_ms, e2e_ms] + stages, dtype=torch.float64, device="cuda")
    r = torch.tensor([rays, e2e_rays], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(r, op=dist.ReduceOp.SUM)
    total_ms, e2e_ms, build_ms, pt_ms, xch_ms, dn_ms, comb_ms = t.tolist()
    # the build stage = LBVH build (the reference's BuildBvhLevel1 + 2: what BUILD_BYTES_PER_TRI and the reference kernels' build time are about)
    # + the collapse to the wide layout, timed inside the library; N > 1 stages already carry the LBVH build alone
    wide_build_ms = float(np.mean(wide_ms[-args.steps:])) if wide_ms else 0.0
    if mg is None:
        build_ms = max(build_ms - wide_build_ms, 0.0)
    rays, e2e_rays = r.tolist()
    ms_step = total_ms / args.steps
    value = rays / (total_ms * 1e-3) / 1e6
    e2e_val = e2e_rays / (e2e_ms * 1e-3) / 1e6

    if rank == 0:
        rows_n = rows[1] - rows[0]
        build_gbs = BUILD_BYTES_PER_TRI * n_tris / (build_ms * 1e-3) / 1e9
        dn_gbs = DENOISE_BYTES_PER_PX * WIDTH * rows_n / (dn_ms * 1e-3) / 1e9
        pt_bytes = node_visits * node_bytes + leaf_tests * tri_bytes + GBUF_BYTES_PER_PX * WIDTH * rows_n
        pt_gbs = pt_bytes / (pt_ms * 1e-3) / 1e9
        # the dominant kernel: k_pt_trace of bounce round 1 (incoherent bounce / shadow rays).  Algorithmic bytes per launch = its own node
        # visits x 64 B + leaf tests x 128 B (counted on the device in an untimed frame) + ray fetch 32 B + hit record 16 B + 2 queue words
        # per ray; duration = CUDA events recorded by the library around that launch, averaged over the timed steps.
        tr_bytes = r1_nodes * node_bytes + r1_leaves * tri_bytes + r1_rays * (32 + 16 + 8)
        tr_ms = float(kernel_ms[2])
        tr_gbs = tr_bytes / (tr_ms * 1e-3) / 1e9 if tr_ms > 0 else 0.0
        # Ceilings of the dominant kernel (node fetches are L1/L2 hits, the HBM peak is not a ceiling this kernel can reach):
        #   bandwidth: algorithmic fetch bytes / time against the MEASURED L2 read bandwidth (rtrt_microbench, all SMs streaming an
        #              L2-resident buffer);
        #   latency:   every ray is a chain of (node visits + leaf fetches) DEPENDENT loads; with `lanes` rays resident at once the
        #              kernel cannot finish before  ceil(rays / lanes) x chain length x load latency  (L1-hit and L2-hit variants).
        props = torch.cuda.get_device_properties(local)
        lanes = props.multi_processor_count * 6 * 128          # RTRT_TRACE_MINB = 6 CTAs of 128 threads per SM
        chain = (r1_nodes + r1_leaves) / max(r1_rays, 1)
        waves = -(-int(r1_rays) // lanes)
        floor_l1 = waves * chain * mb["l1_latency_ns"] * 1e-6
        floor_l2 = waves * chain * mb["l2_latency_ns"] * 1e-6
        l2_peak = mb["l2_read_gbs"]
        # The binding ceiling (ncu: neither DRAM nor L2 nor the load chains come close, the SM issue slots do): warp instructions of one launch
        # (counted by ncu on this very command, profiles/ncu_traffic.json — the workload is deterministic) per second of the LIVE launch time,
        # against one warp instruction per SM sub-partition and clock at the SM clock sampled during the timed region.
        tkey = "k_pt_trace_round1_wide" if wide else "k_pt_trace_round1"
        winst = ncu_traffic(tkey, "warp_instructions")
        clk = sampler.summary()["sm_mhz"] or sampler.summary()["sm_max_mhz"] or 1965.0
        issue_peak = props.multi_processor_count * 4 * clk * 1e6 / 1e9                 # G warp-instructions / s
        issue_ach = (winst / (tr_ms * 1e-3) / 1e9) if (winst and tr_ms > 0) else None
        have_inst = issue_ach is not None      # no committed ncu instruction count for this layout: fall back to the L2-bandwidth ceiling
        trace_roofline = {"kernel": "k_pt_trace, bounce round 1 (%.0f%% of the step)" % (100 * tr_ms / max(ms_step, 1e-9)), "bound": "issue" if have_inst else "l2",
                          "achieved": issue_ach if have_inst else tr_gbs, "peak": issue_peak if have_inst else l2_peak, "unit": "Gwarp-inst/s" if have_inst else "GB/s",
                          "frac": (issue_ach / issue_peak) if have_inst else (tr_gbs / l2_peak if l2_peak else None),
                          "peak_kind": "1 warp instruction per SM sub-partition and clock: %d SMs x 4 x %.0f MHz (clock sampled under load); instruction count per launch "
                                       "from the committed ncu capture of this command (profiles/ncu_traffic.json)" % (props.multi_processor_count, clk),
                          "warp_instructions_per_launch": winst, "lanes_per_instruction": ncu_traffic(tkey, "lanes_per_instruction"),
                          "traffic": ncu_traffic(tkey),
                          "l2": {"achieved": tr_gbs, "peak": l2_peak, "unit": "GB/s", "frac": tr_gbs / l2_peak if l2_peak else None,
                                 "peak_kind": "measured L2 read bandwidth (rtrt_microbench: 32 MiB L2-resident buffer, ld.global.cg, all SMs)"},
                          "hbm_frac": tr_gbs / hbm_peak, "hbm_peak": hbm_peak,
                          "algorithmic_bytes_per_launch": int(tr_bytes), "ms_per_launch": tr_ms,
                          "rays_per_launch": int(r1_rays), "node_visits_per_ray": r1_nodes / max(r1_rays, 1), "leaf_tests_per_ray": r1_leaves / max(r1_rays, 1),
                          "latency_floor": {"resident_lanes": lanes, "dependent_loads_per_ray": chain, "waves": waves,
                                            "l1_hit_ns": mb["l1_latency_ns"], "l2_hit_ns": mb["l2_latency_ns"], "dram_ns": mb["dram_latency_ns"],
                                            "floor_ms_all_l1_hits": floor_l1, "floor_ms_all_l2_hits": floor_l2,
                                            "frac_of_l2_floor": floor_l2 / tr_ms if tr_ms > 0 else None},
                          "note": "BVH traversal without RT cores is instruction-bound on B200: `frac` = issued warp instructions per second against the SM issue "
                                  "rate.  The other candidate ceilings are far away: `l2` = node / leaf fetch bytes per second against the measured L2 read bandwidth, "
                                  "`latency_floor` = the time the dependent-load chains alone would take at the measured L1 / L2 hit latencies with every resident "
                                  "lane busy, `traffic` = DRAM bytes of one launch (ncu --set full).  What is left of the issue rate goes to SIMT divergence: "
                                  "`lanes_per_instruction` of 32 (profiles/*_ncu_summary.md)"}
        par = ("1 GPU" if world == 1 else
               f"sample-partitioned x{world} (weak scaling: N samples per pixel and frame; rank r renders the whole frame with sample index frame + r), BVH rebuilt "
               f"per GPU, one ncclAllReduce (mean of the rgb lanes) of the denoised frame; all of it inside rtrt_render_frame_multi" if not by_rows else
               f"ONE frame over {world} GPUs: path trace in interleaved 8-row strips, grouped ncclSend/Recv redistribution of the G-buffers to 16-row-aligned bands "
               f"(+- 32 rows), denoise per band with a halo exchange before every filter pass, frame gather (16 B/px); rtrt_render_frame_multi")
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak" if not by_rows else "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "layout": args.layout, "width": WIDTH, "height": HEIGHT, "triangles": n_tris, "spp": 1,
                       "rays_per_step": rays / args.steps, "parallelism": par,
                       "l2": "flushed between steps (256 MiB memset, outside the event pairs)",
                       "stages_ms": {"build_bvh": build_ms, "wide_collapse": wide_build_ms, "pathtrace": pt_ms, "exchange_gbuffers": xch_ms, "denoise": dn_ms,
                                     "combine": comb_ms},
                       "pathtrace_mrays_s": rays / args.steps / (pt_ms * 1e-3) / 1e6,
                       "build_mtris_s": n_tris / (build_ms * 1e-3) / 1e6},
            "roofline": trace_roofline,
            "roofline_pathtrace_stage": {"kernel": "k_pt_primary + k_pt_logic_first + 4 x (k_pt_trace, k_pt_logic): the path-trace stage, %.0f%% of the step"
                                   % (100 * pt_ms / max(ms_step, 1e-9)), "bound": "hbm",
                         "achieved": pt_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": pt_gbs / hbm_peak, "traffic": None,
                         "peak_kind": peak_kind, "algorithmic_bytes_per_launch": int(pt_bytes),
                         "note": "latency/divergence-bound traversal (no RT cores); bytes = node visits x 64 + leaf tests x 128 + 30 B/px G-buffer "
                                 "(reference layout, SURVEY 8d 'secondary'); most of them are L1/L2 hits, see profiles/ for dram__bytes",
                         "node_visits_per_ray": node_visits / max(rays_cnt, 1), "leaf_tests_per_ray": leaf_tests / max(rays_cnt, 1),
                         "kernels_ms": {"k_pt_primary": kernel_ms[0], "k_pt_logic_first": kernel_ms[1], "k_pt_trace_round1": kernel_ms[2],
                                        "k_pt_logic_round1": kernel_ms[3], "k_pt_trace_round2": kernel_ms[4], "rest": kernel_ms[5]}},
            "roofline_build": {"kernel": "k_build_blas (gather+AABB+Morton+sort+Karras+refit+TLAS, one launch; the wide collapse is stages_ms.wide_collapse)", "bound": "hbm",
                               "achieved": build_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": build_gbs / hbm_peak,
                               "algorithmic_bytes_per_launch": BUILD_BYTES_PER_TRI * n_tris},
            "roofline_denoise": {"kernel": "k_dn_* (7 launches)", "bound": "hbm", "achieved": dn_gbs, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": dn_gbs / hbm_peak, "algorithmic_bytes_per_step": DENOISE_BYTES_PER_PX * WIDTH * rows_n,
                                 "note": "252 B/px = the reference's pass-wise traffic; fused lower bound 74 B/px"},
            "e2e": {"value": e2e_val, "unit": "Mrays/s", "ms_per_step": e2e_ms / e2e_steps,
                    "h2d_bytes_per_step": int(h_pos.numel() * 4 + h_nrm.numel() * 4), "d2h_bytes_per_step": int(h_out[0].numel()),
                    "multi_gpu": None if world == 1 else "rank 0 uploads the vertex + normal buffers once per step and they are broadcast over NVLink (rtrt_update_vertices_multi); "
                                                         "only rank 0 reads the combined frame back",
                    "note": "frames in flight: every step uploads the vertex + normal buffers from pinned host memory (copy stream), rebuilds the "
                            "BVH, path-traces, denoises and reads the denoised frame back into pinned host memory (snapshot + copy stream); the "
                            "copies overlap the neighbouring frames' kernels; one timed region over all steps, L2 flushes included"},
            "microbench": mb,
            "last_frame_sha1": frame_sha1,
            "other_layout": other_layout,
            "strong": strong,
            "other_configs": configs,
            "gpu_launches": int(launches),
            "clocks": sampler.summary(),
        }
        if world == 1 and not args.no_ref_kernels:
            with quiet_stdout():
                line["reference_kernels"] = reference_kernels(pkg, rt, pos, nrm, idx, cam, rays / args.steps)
        if not args.no_cpu_baseline and world == 1:
            _, _, line["cpu_baseline"] = CpuArm(pkg).baseline(steps=4, warmup=0)      # four whole frames, ~6 s of CPU work on 16 threads
        emit_line(line)
    rt.cleanup()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
