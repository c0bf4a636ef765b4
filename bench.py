#!/usr/bin/env python
"""bench.py -- path samples/s of the B200 `src/spheres` render loop (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1..c5] [--spp-step S] [--impl reference]

One "step" = one pass of the hot path: `spp_per_step` x SpheresApp::DrawBitmap (spheres-app.cpp:163-184) over
every pixel of the workload, i.e. one rt_render() call per GPU (+ the sum-reduce onto rank 0 when N > 1).
Default workload c3 = BASELINE.json configs[2] (default 488-sphere scene, 1920x1080, depth 50: the full path with
indirect light, Fresnel glass and the reference's hemisphere sampler); configs[1] (direct only) is `--workload c2`.
N > 1: launched by torchrun, one rank per GPU; every rank renders `spp_per_step` frames of a disjoint sample
range (weak scaling), buffers are combined with ncclReduce.  Rank 0 prints ONE JSON line.

`--impl reference` times the reference's own CPU algorithm (per-material atomic Halton counters, BVH, recursive
integrator, all host threads): oracle/_ref (the reference's own common-lib sources, compiled headless) when built,
else the oracle port.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# name -> (scene, width, height, max_depth, config spp, description)
WORKLOADS = {
    "c1": ("d488", 1280, 720, 50, 256, "configs[0]: default scene 1280x720 depth 50 (reference defaults; 256 spp chosen)"),
    "c2": ("d488", 1920, 1080, 0, 256, "configs[1]: default scene, direct lighting only (1 vertex), 1920x1080, 256 spp"),
    "c3": ("d488", 1920, 1080, 50, 1024, "configs[2]: default scene, indirect + Fresnel glass, depth 50, 1920x1080, 1024 spp"),
    "c4": ("d488", 3840, 2160, 50, 1024, "configs[3]: default scene, thin lens aperture 0.4, 3840x2160, 1024 spp"),
    "c5": ("s4096", 3840, 2160, 50, 4096, "configs[4]: 4096-sphere scene, 3840x2160, 4096 spp, sample-split + ncclReduce"),
}
FLOP_PER_TEST = 17.0            # irreducible FP32 flops of one ray-sphere discriminant test
FP32_PEAK_MEASURED_TFLOPS = 72.5   # FFMA-chain microbenchmark on this pool's B200 (profiles/r01_fp32_peak.txt)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--spp-step", type=int, default=0, help="frames per step per GPU (0 = 256, or 32 for c5)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default, the driver's contract): every rank renders spp-step frames per step; strong: one step = the "
                         "workload's whole config spp, split over the ranks")
    ap.add_argument("--schedule", default="two-stage", choices=["two-stage", "one-kernel"], help="one-kernel = RT_FLAG_DEBUG_ONE_KERNEL (round-1 schedule)")
    ap.add_argument("--no-extras", action="store_true", help="skip the short runs of the other BASELINE configs and the RMSE-vs-golden figures")
    ap.add_argument("--query", default="brute", choices=["brute", "accel"],
                    help="closest-hit / occlusion query of the timed steps: brute force over every sphere (north_star, default) or the "
                         "box hierarchy RT_FLAG_ACCEL (same bits); the default run also reports the other one under 'accel'")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ CPU arm (oracle)
# This arm never imports cpuraytracer_b200 and never maps librt_b200.so: its scenes are the committed oracle/scenes/*.rtsc
# (frozen copies of the named synthetic scenes; tests/test_host_api.py checks them against the product's recorder).
class CpuArm:
    """The reference's CPU implementation of the path on all host threads.  kind "reference": oracle/_ref -- the reference's
    OWN common-lib sources compiled unmodified (g++ -O2, SSE2 baseline) + the headless DrawBitmap driver, ONE persistent world
    (scene, BVH, per-material Halton counters, worker pool) across frames as in the reference.  kind "port": the oracle's
    restatement in the reference's configuration (atomic Halton counters + BVH + recursive integrator)."""

    def __init__(self, workload):
        import numpy as np
        from oracle import orc
        self.orc, self.np = orc, np
        scene_name, self.w, self.h, self.depth, _, _ = WORKLOADS[workload]
        self.osc = orc.frozen_scene(scene_name)
        self.cp = orc.default_camera_params(self.w / self.h)
        self.hdr = np.zeros((self.h, self.w, 4), np.float32)
        self.frame = 0
        self.world = orc.RefWorld(self.osc, self.cp) if orc.load_ref() is not None else None
        self.kind = "reference" if self.world is not None else "port"
        self.threads = self.world.threads if self.world is not None else (os.cpu_count() or 1)
        self.build = (orc.load_ref().ref_build_info() if self.world is not None else orc.load().orc_build_info()).decode()

    def frames(self, n=1, flags=0):
        """n x DrawBitmap (1 spp over every pixel each), continuing the sample count and the sampler state."""
        orc = self.orc
        first = self.frame + 1
        self.frame += n
        if self.world is not None:
            _, st = self.world.render(orc.config(self.w, self.h, spp=n, max_depth=self.depth, first_sample=first, flags=flags), self.hdr)
        else:
            _, st = orc.render(self.osc, self.cp, orc.config(self.w, self.h, spp=n, max_depth=self.depth, first_sample=first, rng_mode=orc.RNG_HALTON,
                                                             accel=orc.ACCEL_BVH, integrator=orc.INTEGRATOR_RECURSIVE, threads=0), self.hdr)
        return st

    def describe(self):
        return {"cores": self.threads, "kind": self.kind, "threads": self.threads, "host_cpus": os.cpu_count(), "build": self.build,
                "compiler_flags": "g++ -O2 -std=c++17 -ffp-contract=off -fno-fast-math (x86-64 SSE2 baseline = /arch:SSE2, common-lib.vcxproj:72)",
                "counters": "per-thread (no shared cache line per ray)", "world": "persistent across frames (scene, BVH, Halton indices, worker pool)",
                "schedule": "rows handed to a persistent pool, primary ray generated by the tracing worker (kinder to the CPU than the "
                            "reference's serial GenerateRays; that structure is timed under 'drawbitmap_structure')"}

    def drawbitmap_structure(self, n=2):
        """Second, stated figure: the reference's own frame structure (serial GenerateRays :132-161, parallel trace :177-184,
        parallel resolve :186-214)."""
        if self.world is None:
            return None
        t0 = time.perf_counter()
        st = self.frames(n, self.orc.FLAG_REF_DRAWBITMAP)
        el = time.perf_counter() - t0
        return {"value": st.paths / el, "unit": "path samples/s", "frames": n}


def cpu_baseline(workload, budget_s):
    """Bounded sample: whole frames (1 spp over every pixel of the workload) until ~budget_s of CPU time."""
    arm = CpuArm(workload)
    arm.frames(1)                                            # first touch: page in, spin the pool up
    t0 = time.perf_counter()
    frames, paths = 0, 0
    while True:
        st = arm.frames(1)
        frames += 1
        paths += st.paths
        el = time.perf_counter() - t0
        if el >= budget_s or el + el / frames > 1.5 * budget_s:
            break
    out = {"value": paths / el, "unit": "path samples/s"}
    out.update(arm.describe())
    out["sample"] = (f"{frames} full frame(s) (1 spp each, {arm.w}x{arm.h}, depth {arm.depth}) of the workload in {el:.1f} s on {arm.threads} threads "
                     f"after 1 untimed frame; scene from oracle/scenes/*.rtsc")
    out["drawbitmap_structure"] = arm.drawbitmap_structure(2)
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    scene_name, w, h, depth, cfg_spp, desc = WORKLOADS[args.workload]
    arm = CpuArm(args.workload)
    for _ in range(args.warmup):
        arm.frames(1)
    t0 = time.perf_counter()
    paths = 0
    for _ in range(args.steps):
        paths += arm.frames(1).paths
    el = time.perf_counter() - t0
    value = paths / el
    sample = f"each step = 1 full frame (1 spp, {w}x{h}, depth {depth}) on {arm.threads} host threads; frames continue one persistent world"
    cb = {"value": value, "unit": "path samples/s", "sample": sample}
    cb.update(arm.describe())
    cb["drawbitmap_structure"] = arm.drawbitmap_structure(2)
    line = {"impl": "reference", "metric": "path samples/s", "value": value, "unit": "path samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * el / max(args.steps, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload} {desc}", "scene": scene_name, "width": w, "height": h, "max_depth": depth,
                       "spp_per_step": 1, "note": "the reference's Win32/Direct2D shell cannot build here; its common-lib sources can (oracle/_ref)"},
            "cpu_baseline": cb,
            "e2e": {"value": value, "unit": "path samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _emit(line)
    return 0


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        try:
            rows = [r.strip().split(", ") for r in open(self.path) if r.strip()]
            os.unlink(self.path)
        except Exception:
            return out
        sm, mx, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); power.append(float(r[3]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[5 + k].strip().lower().startswith("active"):
                    reasons.add(nm)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm),
                       power_w_max=max(power))
        return out


# ------------------------------------------------------------------------------------------------ B200 arm
def _fp32_roofline(tests, kern_s, world, traffic=None):
    achieved = tests * FLOP_PER_TEST / kern_s / 1e12 / world            # per-GPU TFLOP/s of the render kernels
    return {"bound": "fp32", "achieved": achieved, "peak": FP32_PEAK_MEASURED_TFLOPS, "unit": "TFLOP/s",
            "frac": achieved / FP32_PEAK_MEASURED_TFLOPS, "traffic": traffic,
            "note": "first-hit + path kernels: sphere tests x 17 flop / kernel time (cudaEvents on their stream); peak = measured "
                    "FFMA-chain FP32 rate (MEASURED_PEAKS.json has no FP32 figure; nominal 148x128x2x1.965 GHz = 74.5)",
            "sphere_tests_per_s": tests / kern_s / world}


def short_config_runs(rt, local, flush, skip):
    """Driver-observed figures for the BASELINE configs the timed region does not run (N = 1 only): 1 warm-up + 2 timed steps each,
    brute force and RT_FLAG_ACCEL, CUDA events on the launch stream, L2 flushed between steps.  Short steps: seconds in total."""
    import torch
    out = {}
    for wl, spp_b, spp_a in (("c1", 128, 256), ("c2", 128, 256), ("c3", 64, 128), ("c4", 32, 64), ("c5", 8, 32)):
        if wl == skip:
            continue
        scene_name, w, h, depth, cfg_spp, desc = WORKLOADS[wl]
        r = rt.Renderer(device=local, width=w, height=h, stream=torch.cuda.current_stream(local).cuda_stream)
        r.set_scene(rt.scenes.NAMED[scene_name]())
        r.set_camera(rt.scenes.default_camera(aspect=w / h))
        entry = {"workload": desc, "width": w, "height": h, "max_depth": depth}
        for name, flags, spp in (("brute", 0, spp_b), ("accel", rt.RT_FLAG_ACCEL, spp_a)):
            r.render(spp, depth, flags, first_sample=1)
            torch.cuda.synchronize()
            r.clear(); r.reset_stats()
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(2)]
            for i in range(2):
                flush.fill_(i)
                ev[i][0].record()
                r.render(spp, depth, flags, first_sample=1 + i * spp)
                ev[i][1].record()
            torch.cuda.synchronize()
            ms = sum(a.elapsed_time(b) for a, b in ev)
            st = r.stats()
            assert st.paths == 2 * spp * w * h
            entry[name] = {"value": st.paths / (ms * 1e-3), "unit": "path samples/s", "spp_per_step": spp, "steps": 2, "ms_per_step": ms / 2,
                           "rays_per_path": (st.closest_rays + st.shadow_rays) / st.paths,
                           "roofline_frac": st.sphere_tests * FLOP_PER_TEST / (st.render_ms * 1e-3) / 1e12 / FP32_PEAK_MEASURED_TFLOPS}
        r.close()
        out[wl] = entry
    return out


def static_entry_run(rt, local, workload, flush):
    """rt_render_static (the entry point that allocates nothing: scratch in the library's static device array) on the bench workload:
    1 warm-up + 2 timed calls of 64 frames, checked against rt_render on the same frames.  None where the job does not fit the array."""
    import numpy as np
    import torch
    scene_name, w, h, depth, _, _ = WORKLOADS[workload]
    spp = 64
    if rt.load().rt_render_static_scratch_bytes(w, h, spp) > rt._lib.RT_STATIC_SCRATCH_BYTES:
        return None
    r = rt.Renderer(device=local, width=w, height=h, stream=torch.cuda.current_stream(local).cuda_stream)
    r.set_scene(rt.scenes.NAMED[scene_name]())
    r.set_camera(rt.scenes.default_camera(aspect=w / h))
    r.render_static(spp, depth, 0, first_sample=1)
    torch.cuda.synchronize()
    r.clear()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(2)]
    for i in range(2):
        flush.fill_(i)
        ev[i][0].record()
        r.render_static(spp, depth, 0, first_sample=1 + i * spp)
        ev[i][1].record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    got = r.read_hdr_sum()
    r.clear()
    r.set_chunk_spp(1 << 20)                                  # rt_render with one work item per pixel and batch: the same sequential sums
    r.render(2 * spp, depth, 0, first_sample=1)
    ref = r.read_hdr_sum()
    r.close()
    a, b = got[..., :3].astype(np.float64), ref[..., :3].astype(np.float64)
    rel = float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-3 * b.mean())))
    assert np.array_equal(got[..., 3], ref[..., 3]) and rel < 2e-4, rel
    return {"entry_point": "rt_render_static (no allocation per call; scratch = 64 MiB static device array; batches of 16 frames)",
            "value": 2 * spp * w * h / (ms * 1e-3), "unit": "path samples/s", "spp_per_call": spp, "calls": 2,
            "max_rel_diff_vs_rt_render": rel, "bitwise_equal_to_rt_render": bool(np.array_equal(got, ref))}


def rmse_vs_golden(rt, local):
    """BASELINE.json's metric names "RMSE vs CPU ref": per config the GPU's converged linear radiance against the committed golden
    of the reference's own renderer (tests/golden/golden_converged.npz: oracle/_ref at 65,536 spp, 240x135), same spp on the GPU.
    Bars (north_star): per-pixel relRMSE < 1e-2, image-mean relative error < 0.5 %."""
    import numpy as np
    import torch
    path = os.path.join(ROOT, "tests", "golden", "golden_converged.npz")
    if not os.path.exists(path):
        return None
    gold = np.load(path)
    if "spp" not in gold.files:
        return None
    spp = int(gold["spp"])
    out = {"golden": "oracle/_ref (the reference's own sources) at %d spp, 240x135; tests/golden/make_golden_converged.py" % spp,
           "bars": {"rel_rmse": 1e-2, "mean_rel_err": 5e-3}}
    for key, scene_name, depth in (("c2", "d488", 0), ("c3", "d488", 50), ("c5", "s4096", 50)):
        if key + "_mean" not in gold.files:
            continue
        ref = gold[key + "_mean"].astype(np.float64) * 1024.0
        h, w = ref.shape[:2]
        r = rt.Renderer(device=local, width=w, height=h, stream=torch.cuda.current_stream(local).cuda_stream)
        r.set_scene(rt.scenes.NAMED[scene_name]())
        r.set_camera(rt.scenes.default_camera(aspect=w / h))
        r.render(spp, depth, rt.RT_FLAG_ACCEL)              # same bits as brute force (tests), several times faster on s4096
        g = r.read_hdr_sum()
        r.close()
        mean = g[..., :3].astype(np.float64) / spp * 2.0 ** 15
        rel = float(np.sqrt(np.mean(((mean - ref) / (ref + 1e-3 * ref.mean())) ** 2)))
        merr = float(abs(mean.mean() - ref.mean()) / ref.mean())
        out[key] = {"rel_rmse": rel, "mean_rel_err": merr, "golden_noise_estimate": float(gold[key + "_floor"]), "spp": spp, "scale": f"{w}x{h}",
                    "pass": bool(rel < 1e-2 and merr < 5e-3)}
    return out


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import cpuraytracer_b200 as rt
    from cpuraytracer_b200.distributed import ShardedRenderer

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; cpuraytracer_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; launch N > 1 with torch.distributed.run")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))

    scene_name, w, h, depth, cfg_spp, desc = WORKLOADS[args.workload]
    strong = args.scaling == "strong"
    # weak: every rank renders spp_step frames per step; strong: one step = the config's whole spp, split over the ranks
    spp_step = args.spp_step or (32 if args.workload == "c5" else 256)
    job_spp = cfg_spp if strong else world * spp_step        # frames of ONE step, all ranks together
    n_pix = w * h
    scene = rt.scenes.NAMED[scene_name]()
    flat = scene.export()
    cam = rt.scenes.default_camera(aspect=w / h)
    sr = ShardedRenderer(w, h, device=local)
    r = sr.renderer
    sr.set_scene(scene)
    sr.set_camera(cam)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")      # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    base_flags = {"brute": 0, "accel": rt.RT_FLAG_ACCEL}[args.query] | (rt.RT_FLAG_DEBUG_ONE_KERNEL if args.schedule == "one-kernel" else 0)

    def step(i, flags=None):
        # global frame range of step i: [1 + i*job_spp, ...); each rank takes its disjoint share
        sr.render(job_spp, depth, base_flags if flags is None else flags, first_sample=1 + i * job_spp, reduce=True)

    def timed(flags, steps):
        """steps x step() with CUDA events on the launching stream, L2 flushed before each; returns max-over-ranks seconds."""
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        for i in range(steps):
            flush.fill_(i & 0xFF)
            ev[i][0].record()
            step(i, flags)
            ev[i][1].record()
        barrier()
        ms = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=f"cuda:{local}")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) * 1e-3

    # ---- warm-up ----
    for i in range(args.warmup):
        step(i)
    barrier()
    sr.clear()
    r.reset_stats()

    # ---- timed region: exactly K steps ----
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    t_wall0 = time.perf_counter()
    total_s = timed(None, args.steps)
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    st = r.stats()
    counts = torch.tensor([st.paths, st.closest_rays, st.shadow_rays, st.sphere_tests, st.launches], dtype=torch.float64, device=f"cuda:{local}")
    kern_ms = torch.tensor([st.render_ms], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(kern_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    paths = job_spp * n_pix * args.steps
    assert int(counts[0].item()) == paths, (int(counts[0].item()), paths)
    value = paths / total_s
    if rank == 0:      # rank 0 holds the whole job's sum -> every pixel accumulated K * job_spp samples, exactly
        wch = sr.hdr[..., 3]
        assert float(wch.min().item()) == float(wch.max().item()) == float(args.steps * job_spp), (float(wch.min().item()), float(wch.max().item()))

    # ---- K5, the only HBM-bound kernel of the path (north_star: "HBM GB/s is reported only for the accumulation pass"):
    #      resolve = 16 B read + 4 B write per pixel, L2 flushed before every launch, CUDA events inside rt_resolve_ldr ----
    accumulation = None
    if rank == 0:
        r.reset_stats()
        n_res = 10
        for i in range(n_res):
            flush.fill_(i & 0xFF)
            r.resolve_ldr(sample_count=args.steps * job_spp, copy=False)
        rs = r.stats()
        res_s = rs.resolve_ms * 1e-3 / n_res
        hbm_peak = 6550.7
        try:
            hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except Exception:
            pass
        accumulation = {"kernel": "rt_resolve_kernel (mean, ACES, gamma, BGRA8)", "bytes_per_launch": 20 * n_pix, "us_per_launch": res_s * 1e6,
                        "achieved_gbs": 20 * n_pix / res_s / 1e9, "peak_gbs": hbm_peak, "frac": 20 * n_pix / res_s / 1e9 / hbm_peak,
                        "note": "peak = MEASURED_PEAKS.json hbm_gbs (driver-measured copy bandwidth)"}
        r.reset_stats()

    # ---- the same steps with the other query mode and with the round-1 schedule (informative; same bits, see tests) ----
    others = {}
    if args.query == "brute" and args.schedule == "two-stage" and not strong:
        for name, flags in (("accel", rt.RT_FLAG_ACCEL), ("one_kernel_schedule", rt.RT_FLAG_DEBUG_ONE_KERNEL)):
            k = min(args.steps, 3)
            step(0, flags)
            barrier()
            sr.clear()
            s_ = timed(flags, k)
            others[name] = {"value": job_spp * n_pix * k / s_, "unit": "path samples/s", "ms_per_step": 1e3 * s_ / k, "steps": k}
            sr.clear()
        others["accel"]["query"] = "RT_FLAG_ACCEL box hierarchy (bit-identical results)"
        others["one_kernel_schedule"]["note"] = "RT_FLAG_DEBUG_ONE_KERNEL: primary rays inside the path megakernel (round-1 schedule), brute force"

    # ---- end to end through the public API with HOST buffers: scene arrays -> rt_scene_import/rt_set_scene (H2D),
    #      render (+ reduce), running-sum buffer -> pinned host memory (D2H) ----
    host_hdr = torch.empty((h, w, 4), dtype=torch.float32).pin_memory()
    h2d = flat.spheres.nbytes + flat.materials.nbytes + flat.textures.nbytes + flat.lights.nbytes + 40
    d2h = host_hdr.numel() * 4
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        sc = rt.Scene.from_flat(flat)               # host arrays in, as a caller of the C ABI would hand them over
        sr.set_scene(sc)
        sr.set_camera(cam)
        sr.clear()
        step(i)
        if rank == 0:
            host_hdr.copy_(sr.hdr, non_blocking=True)
        torch.cuda.synchronize()
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = paths / float(e2e_s.item())

    parity_n = parity_across_gpus(rt, ShardedRenderer, scene, depth, world, rank, local) if world > 1 else None
    extra_configs = short_config_runs(rt, local, flush, args.workload) if (world == 1 and rank == 0 and not args.no_extras) else None
    rmse = rmse_vs_golden(rt, local) if (rank == 0 and not args.no_extras) else None
    static_entry = static_entry_run(rt, local, args.workload, flush) if (world == 1 and rank == 0 and not args.no_extras) else None

    if rank == 0:
        kern_s = float(kern_ms.item()) * 1e-3
        tests = float(counts[3].item())
        roof = _fp32_roofline(tests, kern_s, world, _ncu_traffic(args.workload, args.query))
        roof["kernel_ms_per_step"] = 1e3 * kern_s / args.steps
        line = {
            "metric": "path samples/s", "value": value, "unit": "path samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total_s / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload} {desc}", "scene": scene_name, "n_spheres": int(len(flat.spheres)), "width": w,
                       "height": h, "max_depth": depth, "spp_per_step_per_gpu": job_spp / world, "config_spp": cfg_spp,
                       "partition": "disjoint sample ranges per GPU + ncclReduce(sum) to rank 0" if world > 1 else "single GPU",
                       "query": "RT_FLAG_ACCEL box hierarchy in front of the cull + exact tests" if args.query == "accel" else "brute force over every sphere (north_star)",
                       "schedule": "two-stage: first-hit kernel + path megakernel" if args.schedule == "two-stage" else "one kernel (round-1 schedule)",
                       "l2": "flushed between timed steps (256 MiB fill); scene lives in shared memory"},
            "rays_per_s": (float(counts[1].item()) + float(counts[2].item())) / total_s,
            "rays_per_path": (float(counts[1].item()) + float(counts[2].item())) / paths,
            "gpu_launches": int(counts[4].item()),
            "e2e": {"value": e2e_value, "unit": "path samples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "roofline": roof,
            "clocks": clocks, "wall_s": t_wall,
        }
        if parity_n is not None:
            line["parity_n"] = parity_n
        for k_, v_ in others.items():
            v_["speedup_vs_value"] = v_["value"] / value
            line[k_] = v_
        if accumulation is not None:
            line["accumulation"] = accumulation
        if extra_configs:
            line["configs"] = extra_configs
        if rmse:
            line["rmse"] = rmse
        if static_entry:
            line["static_entry"] = static_entry
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args.workload, args.cpu_seconds)
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def parity_across_gpus(rt, ShardedRenderer, scene, depth, world, rank, local, w=480, h=270, spp_per_gpu=8):
    """N > 1 must equal N = 1 (one running sum per pixel, spheres-app.cpp:177-184): a small job of the same
    scene rendered (a) sharded over the N ranks by sample range and sum-reduced onto rank 0, in TWO calls so that the
    multi-call protocol (ranks >= 1 start each call from zero) is covered, and (b) on rank 0's GPU alone over the same frame
    range.  Asserted: sample-count channel exactly equal, xyz within 1e-5 relative (the reduction order differs), and
    rt_primary_hits buffers identical on every rank.  Runs after the timed region; returns the figures (rank 0)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    cam = rt.scenes.default_camera(aspect=w / h)
    spp = world * spp_per_gpu
    out = {"width": w, "height": h, "spp": spp, "max_depth": depth, "calls": 2}
    for name, flags in (("brute", 0), ("accel", rt.RT_FLAG_ACCEL)):
        sr = ShardedRenderer(w, h, device=local)
        sr.set_scene(scene)
        sr.set_camera(cam)
        sr.clear()
        half = spp // 2 + 1                                   # uneven split: ranks get shares that differ by one frame
        sr.render(half, depth, flags, first_sample=1, reduce=True)
        sr.render(spp - half, depth, flags, first_sample=1 + half, reduce=True)
        torch.cuda.synchronize()
        ids, mids = sr.renderer.primary_hits(1, flags)        # jitter + lens off (the bit-exact sub-test of configs[1])
        mine = torch.from_numpy(np.stack([ids, mids]).astype(np.int32)).to(f"cuda:{local}")
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        if rank == 0:
            sharded = sr.hdr.cpu().numpy()
            one = rt.Renderer(device=local, width=w, height=h)
            one.set_scene(scene)
            one.set_camera(cam)
            one.render(spp, depth, flags, first_sample=1)
            single = one.read_hdr_sum()
            ids1, mids1 = one.primary_hits(1, flags)
            one.close()
            count_equal = bool(np.array_equal(sharded[..., 3], single[..., 3]) and np.all(single[..., 3] == spp))
            a, b = sharded[..., :3].astype(np.float64), single[..., :3].astype(np.float64)
            rel = float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-3 * b.mean())))
            ref_ids = torch.from_numpy(np.stack([ids1, mids1]).astype(np.int32)).to(f"cuda:{local}")
            hits_equal = all(bool(torch.equal(g, ref_ids)) for g in gathered)
            out[name] = {"max_rel_diff_xyz": rel, "sample_count_equal": count_equal, "primary_hits_identical_on_all_ranks": hits_equal,
                         "bitwise_equal_xyz": bool(np.array_equal(sharded[..., :3], single[..., :3]))}
            assert count_equal, f"parity_n[{name}]: sample-count channel differs between {world} GPUs and 1 GPU"
            assert rel < 1e-5, f"parity_n[{name}]: N={world} vs N=1 max relative difference {rel}"
            assert hits_equal, f"parity_n[{name}]: primary-hit buffers differ across ranks"
        dist.barrier()
    return out if rank == 0 else None


def _ncu_traffic(workload: str, query: str):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE rt_render_kernel launch of this workload's step (same scene, size,
    spp per step and depth), from the committed ncu capture (profiles/traffic.json, written from tools/gpu_round.sh's
    `traffic_*.csv`); None when no capture of that launch exists.  A profiler number, never measured inside bench.py."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            return json.load(f).get(f"{workload}:{query}", {}).get("bytes")
    except (OSError, ValueError):
        return None


def _emit(line: dict) -> None:
    """The ONE JSON line goes to the process's original stdout; everything else (NCCL's version banner, warnings of
    libraries that print to fd 1) was redirected to stderr by main()."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)                       # native libraries (NCCL) write banners to fd 1: keep stdout clean for the JSON line
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
