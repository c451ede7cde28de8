#!/usr/bin/env python
"""bench.py — one JSON line for BASELINE.json's metric on BASELINE config 2.

A "step" is one frame of the hot path on the 1 047 552-triangle procedural terrain at 1920x1080:
  LBVH rebuild (rtrt_build_bvh) -> ray traversal of the frame's rays (rtrt_traverse; the path-trace and
  denoise stages are added to the step as they land).
`value`  : Mrays/s, device-resident inputs, CUDA events per step on the library's stream, max over ranks.
`e2e`    : same metric through the host-buffer entry point (rays H2D + hits D2H inside the timed region).
N > 1    : the frame is partitioned BY SAMPLE — every rank renders the same view with its own sample index,
           BVH rebuilt redundantly per GPU (no data-path collective) -> "scaling": "weak".
`--impl reference` : the reference has no CPU path and does not build here; the reference arm times the CPU
           oracle (oracle/, the multithreaded C++ restatement) on a bounded sample of the same workload.
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
WORKLOAD = "C2: 1080p frame, 1 047 552-tri procedural terrain (1023 batches), two-level LBVH rebuild + traversal"
CAMERA = dict(pos=(0, 30, -60), yaw=0.0, pitch=-0.4)
BUILD_BYTES_PER_TRI = 248          # SURVEY.md §8(d) / BASELINE.md §4


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
            time.sleep(0.1)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def make_scene(pkg):
    gx, gz = pkg.scenes.TERRAIN_SIZES["C2ref"]
    pos, nrm, idx = pkg.scenes.terrain(gx, gz, seed=1)
    cam = pkg.scenes.make_camera(CAMERA["pos"], CAMERA["yaw"], CAMERA["pitch"], WIDTH, HEIGHT)
    return pos, nrm, idx, cam


def run_reference(args, rank, world):
    """Reference arm: CPU oracle, all host threads, bounded sample of the same workload."""
    if rank != 0:
        return
    pkg = importlib.import_module(PKG)
    from oracle import orc
    orc.build()
    pos, nrm, idx, cam = make_scene(pkg)
    rays = pkg.scenes.primary_rays(cam, WIDTH, HEIGHT)
    rows = 135                                            # 1/8 of the frame's rows, spread over the image
    sel = rays.reshape(HEIGHT, WIDTH)[::HEIGHT // rows][:rows].reshape(-1).copy()
    cores = orc.default_threads()
    times = []
    for s in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        ob = orc.build_bvh(pos, nrm, idx, threads=cores)
        orc.traverse(ob, sel, threads=cores)
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    val = sel.shape[0] / (ms * 1e-3) / 1e6
    sample = f"full LBVH build of 1 047 552 tris + {rows}/{HEIGHT} rows ({sel.shape[0]} rays) of the 1080p frame"
    line = {"impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": WORKLOAD, "sample": sample},
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_baseline(pkg, pos, nrm, idx, rays):
    from oracle import orc
    orc.build()
    cores = orc.default_threads()
    rows = 135
    sel = rays.reshape(HEIGHT, WIDTH)[::HEIGHT // rows][:rows].reshape(-1).copy()
    t0 = time.perf_counter()
    ob = orc.build_bvh(pos, nrm, idx, threads=cores)
    t1 = time.perf_counter()
    orc.traverse(ob, sel, threads=cores)
    t2 = time.perf_counter()
    return {"value": sel.shape[0] / (t2 - t0) / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "port",
            "sample": f"full LBVH build of {idx.shape[0]} tris + {rows}/{HEIGHT} rows ({sel.shape[0]} rays) of the 1080p frame",
            "build_ms": 1e3 * (t1 - t0), "traverse_mrays_s": sel.shape[0] / (t2 - t1) / 1e6}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
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
    pos, nrm, idx, cam = make_scene(pkg)
    rays = pkg.scenes.primary_rays(cam, WIDTH, HEIGHT)
    n_rays = rays.shape[0]
    hbm_peak, peak_kind = peaks()

    rt = pkg.RayTracer(WIDTH, HEIGHT, device=local).init()
    rt.set_mesh(pos, idx, nrm)
    stream = torch.cuda.ExternalStream(rt.stream, device=torch.device("cuda", local))
    d_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda()
    d_hits = torch.empty(n_rays * 16, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2
    h_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1).copy()).pin_memory()
    h_hits = torch.empty(n_rays * 16, dtype=torch.uint8).pin_memory()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps, warmup, sub_events=False):
        for _ in range(warmup):
            step_fn(None)
        barrier()
        lbefore = rt.timings()["kernel_launches"]
        evs = []
        for _ in range(steps):
            with torch.cuda.stream(stream):
                flush.zero_()                                            # L2 flush, outside the event pair
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            step_fn(e)
            evs.append(e)
        barrier()
        tot = [e[0].elapsed_time(e[2]) for e in evs]
        first = [e[0].elapsed_time(e[1]) for e in evs]
        nl = rt.timings()["kernel_launches"] - lbefore
        return float(np.sum(tot)), float(np.mean(first)), float(np.mean(tot) - np.mean(first)), int(nl)

    def step_device(e):
        if e: e[0].record(stream)
        rt.build_bvh()
        if e: e[1].record(stream)
        rt.traverse_device(d_rays.data_ptr(), n_rays, d_hits.data_ptr())
        if e: e[2].record(stream)

    def step_e2e(e):
        if e: e[0].record(stream)
        rt.build_bvh()
        if e: e[1].record(stream)
        rt._check(rt._lib.rtrt_traverse_host(rt._h, h_rays.data_ptr(), n_rays, h_hits.data_ptr(), 0))
        if e: e[2].record(stream)

    sampler = ClockSampler(local); sampler.start()
    total_ms, build_ms, trav_ms, launches = timed(step_device, args.steps, args.warmup)
    sampler.stop_flag = True; sampler.join()
    e2e_ms, _, _, _ = timed(step_e2e, max(3, args.steps // 2), 3)
    e2e_steps = max(3, args.steps // 2)

    t = torch.tensor([total_ms, e2e_ms, build_ms, trav_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, build_ms, trav_ms = t.tolist()
    ms_step = total_ms / args.steps
    value = world * n_rays / (ms_step * 1e-3) / 1e6
    e2e_val = world * n_rays / (e2e_ms / e2e_steps * 1e-3) / 1e6

    if rank == 0:
        n_tris = idx.shape[0]
        build_gbs = BUILD_BYTES_PER_TRI * n_tris / (build_ms * 1e-3) / 1e9
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "width": WIDTH, "height": HEIGHT, "triangles": int(n_tris), "rays_per_step": int(n_rays),
                       "parallelism": f"sample-partitioned x{world}, BVH rebuilt per GPU", "l2": "flushed between steps (256 MiB memset, outside the event pairs)",
                       "stages": {"build_bvh_ms": build_ms, "traverse_ms": trav_ms}},
            "roofline": {"kernel": "k_build_blas (LBVH build, fused gather+sort+emit+refit+TLAS)", "bound": "hbm",
                         "achieved": build_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": build_gbs / hbm_peak, "traffic": None,
                         "peak_kind": peak_kind, "algorithmic_bytes_per_launch": BUILD_BYTES_PER_TRI * int(n_tris)},
            "e2e": {"value": e2e_val, "unit": "Mrays/s", "h2d_bytes_per_step": int(n_rays * 32), "d2h_bytes_per_step": int(n_rays * 16)},
            "gpu_launches": int(launches),
            "clocks": sampler.summary(),
        }
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(pkg, pos, nrm, idx, rays)
        print(json.dumps(line), flush=True)
    rt.cleanup()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
