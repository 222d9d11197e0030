#!/usr/bin/env python
"""bench.py -- headline benchmark: path Msamples/s on the bedroom (stand-in) 720p, 64 spp per GPU, path-mis loop.

One *step* = one full render of the workload: every rank renders its image bands (all samples of those pixels) into a
private RGBW film, the films are summed with one NCCL all-reduce, the result is developed.  Weak scaling: the job at
N GPUs is the 1280x720 image at 64*N spp, i.e. 58 982 400 paths per GPU (BASELINE.json configs[1] at N = 1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b2|reference] [--workload bedroom|cornell]

Prints ONE JSON line (rank 0).  `value` = device-timed throughput with the scene resident in HBM; `e2e` = the same
metric through the host-buffer C-ABI call (film H2D + render + develop + film/image D2H inside the timed region);
`roofline` = the dominant kernel, k_wf_trace (closest + shadow rays): algorithmic bytes = rays x B_ray (SURVEY.md 8d,
exact ray counters) / the CUDA-event time of its launches inside the timed region (kernel history of the C-ABI),
against the measured HBM copy bandwidth; `cpu_baseline` = the CPU oracle (a port of the reference's
llvm_ad_rgb path; Mitsuba itself cannot be installed here) on a bounded sample of the same workload.
"""
import argparse
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

HBM_FALLBACK_GBS = 6650.0


def measured_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback"


def build_workload(name, n_gpus):
    from mitsuba3_experiments_b200 import scenes
    if name == "bedroom":
        spp = 64 * n_gpus
        d = scenes.bedroom_standin(1280, 720, spp)
        return d, dict(variant=1, max_depth=8, rr_depth=2), spp, "bedroom stand-in 1280x720 (1.8M tris, 2 area lights + envmap), path-mis.py loop"
    if name == "bedroom4k":      # BASELINE.json configs[4]: fixed total work (strong scaling), 2 passes x 512 spp (integrator.cpp:251-257)
        spp = 1024
        d = scenes.bedroom_standin(3840, 2160, spp)
        return d, dict(variant=1, max_depth=8, rr_depth=2), spp, "bedroom stand-in 3840x2160 1024 spp (1.8M tris, 2 area lights + envmap), path-mis.py loop, lanes sharded over the GPUs"
    if name == "cornell":
        spp = 16 * n_gpus
        return scenes.cornell_box(256, 256, spp), dict(variant=0, max_depth=8, rr_depth=4), spp, "mi.cornell_box() 256x256, path.py loop"
    raise SystemExit(f"unknown workload {name}")


def algorithmic_bytes(cnt, n_faces, tex_bytes_per_eval=32):
    """bytes any implementation must move from global memory for these events."""
    D = max(1, math.ceil(math.log(max(n_faces / 3.0, 1.0), 8)))
    b_ray = 80 * D + 48 * 3
    return ((cnt["closest_rays"] + cnt["shadow_rays"]) * b_ray + cnt["vertices"] * 140 + (cnt["tex_evals"] // 2) * tex_bytes_per_eval
            + cnt["nee_area"] * 64 + cnt["nee_env"] * 224 + cnt["env_escapes"] * 48 + cnt["splat_taps"] * 16), b_ray


def trace_bytes(cnt, n_faces):
    """Algorithmic bytes of the trace launches alone: every ray fetches D wide nodes and 3 triangles."""
    D = max(1, math.ceil(math.log(max(n_faces / 3.0, 1.0), 8)))
    return (cnt["closest_rays"] + cnt["shadow_rays"]) * (80 * D + 48 * 3)


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        self.stop_flag = True
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def cpu_oracle_rate(scene_dict, kw, label_spp=1, threads=None):
    """Times the CPU oracle on `label_spp` samples per pixel of the workload (BVH build excluded, like the GPU arm)."""
    from mitsuba3_experiments_b200.scene import SceneDesc
    from oracle.oracle_py import OracleScene, lib
    desc = SceneDesc(scene_dict)
    orc = OracleScene(desc)
    # all host threads this process may run on: torchrun exports OMP_NUM_THREADS=1 to every rank, which would otherwise
    # silently turn the "all host threads" baseline into a single-thread one at N > 1
    lib().orc_set_num_threads(threads or len(os.sched_getaffinity(0)))
    t0 = time.perf_counter()
    _, cnt = orc.render_film(seed=0, spp=label_spp, **kw)
    dt = time.perf_counter() - t0
    return cnt["samples"] / dt / 1e6, dt, lib().orc_num_threads(), cnt["samples"]


def run_reference(args, rank):
    """--impl reference: the reference's CPU path.  Mitsuba/Dr.Jit (llvm_ad_rgb) cannot be installed on this box, so the
    timed implementation is the oracle port (oracle/), with all host threads, on a bounded sample per step."""
    if rank != 0:
        return
    d, kw, spp, wl = build_workload(args.workload, 1)
    from mitsuba3_experiments_b200.scene import SceneDesc
    from oracle.oracle_py import OracleScene, lib
    orc = OracleScene(SceneDesc(d))
    lib().orc_set_num_threads(len(os.sched_getaffinity(0)))     # torchrun sets OMP_NUM_THREADS=1 per rank; use every host thread
    sample_spp = 2
    times, n = [], 0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        _, cnt = orc.render_film(seed=i, spp=sample_spp, **kw)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt); n = cnt["samples"]
    ms = 1e3 * float(np.mean(times)); value = n / (ms * 1e-3) / 1e6
    sample = f"{sample_spp} spp of the {spp} spp workload per step ({n} paths)"
    line = {"impl": "reference", "metric": "path Msamples/s bedroom 720p", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if args.workload == "bedroom4k" else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": wl, "sample": sample, "note": "CPU oracle port of the reference's llvm_ad_rgb path (Mitsuba not installable offline)"},
            "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": lib().orc_num_threads(), "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b2", choices=["b2", "reference"])
    ap.add_argument("--workload", default="bedroom", choices=["bedroom", "cornell", "bedroom4k"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-prb", action="store_true")
    ap.add_argument("--shard", default="samples", choices=["samples", "bands"], help="multi-GPU split: sample slices of every pixel (balanced) or image bands")
    ap.add_argument("--bands", type=int, default=2, help="image bands per rank (N > 1); bands are interleaved across ranks")
    args = ap.parse_args()
    rank_env = int(os.environ.get("RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank_env)

    import torch
    import torch.distributed as tdist
    from mitsuba3_experiments_b200 import dist as b2dist
    from mitsuba3_experiments_b200.integrator import make_integrator
    from mitsuba3_experiments_b200.runtime import DeviceScene

    rank, world, local = b2dist.init_from_env()
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    scene_dict, kw, spp, wl = build_workload(args.workload, world)
    ds = DeviceScene(scene_dict, device=local)
    H, W = ds.film_shape
    n_total = ds.wavefront_size(spp)
    spp_pp = n_total // (H * W)                              # samples per pixel per pass (< spp when H W spp exceeds 2^32 - 1 lanes)
    paths_total = H * W * spp
    plan = b2dist.shard_plan(n_total, spp_pp, W, world, rank, bands_per_rank=args.bands if world > 1 else 1)
    n_local = sum(b - a for a, b in plan)
    shard = b2dist.sample_shard(spp_pp, world, rank) if args.shard == "samples" else None
    shard_kw = {}
    if shard is not None:                                   # every rank: all pixels, its slice of the samples of each pass
        shard_kw = dict(sample_begin=shard[0], sample_count=shard[1])
        plan = [(0, H * W * shard[1])]; n_local = plan[0][1]
    film = ds.new_film()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def render_bands(counters=False):
        tot = None
        for a, b in plan:
            _, c = ds.render_film(film=film, counters=counters, seed=0, spp=spp, lane_begin=a, lane_end=b, **shard_kw, **kw)
            if counters:
                tot = c if tot is None else {k: tot[k] + c[k] for k in c}
        return tot

    def step():
        film.zero_()
        render_bands()
        b2dist.merge_sum(film)
        return ds.develop(film)

    def barrier():
        if world > 1:
            tdist.barrier()
        torch.cuda.synchronize(dev)

    # exact event counters of this rank's shard (one untimed, counted run) -> algorithmic bytes
    film.zero_()
    cnt = render_bands(counters=True)
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    clocks = ClockSampler(local); clocks.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    ds.set_kernel_timing(True); ds.kernel_times()          # kernel history on for the timed region (events per launch)
    for e0, e1, e2 in ev:
        flush.zero_()                       # L2 flush between timed iterations (outside the event pair)
        film.zero_()
        e0.record(); render_bands(); e1.record()
        b2dist.merge_sum(film); img = ds.develop(film); e2.record()
    barrier()
    ktimes = ds.kernel_times(); ds.set_kernel_timing(False)
    clk = clocks.result()
    step_ms = [e0.elapsed_time(e2) for e0, _, e2 in ev]
    kern_ms = [e0.elapsed_time(e1) for e0, e1, _ in ev]
    t = torch.tensor([float(np.mean(step_ms)), float(np.mean(kern_ms))], device=dev, dtype=torch.float64)
    if world > 1:
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
    ms_per_step, kernel_ms = float(t[0]), float(t[1])
    value = paths_total / (ms_per_step * 1e-3) / 1e6

    # end-to-end through host buffers
    film_h = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory()
    img_h = torch.zeros((H, W, 3), dtype=torch.float32).pin_memory()
    e2e_times = []
    e2e_warm = 3
    for i in range(e2e_warm + args.steps):
        film_h.zero_()
        barrier(); t0 = time.perf_counter()
        if world == 1:
            ds.render_host(film_h.numpy(), img_h.numpy(), seed=0, spp=spp, **kw)      # b2_render_host: H2D film, render, develop, D2H film + image
        else:
            film.copy_(film_h, non_blocking=True)
            render_bands(); b2dist.merge_sum(film)
            img_h.copy_(ds.develop(film), non_blocking=True); film_h.copy_(film, non_blocking=True)
        barrier(); dt = time.perf_counter() - t0
        if i >= e2e_warm:
            e2e_times.append(dt)
    t = torch.tensor([float(np.mean(e2e_times))], device=dev, dtype=torch.float64)
    if world > 1:
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
    e2e_value = paths_total / float(t[0]) / 1e6

    if world > 1:
        tdist.barrier(); tdist.destroy_process_group()
    if rank != 0:
        return
    peak, peak_kind = measured_peak()
    alg_bytes, b_ray = algorithmic_bytes(cnt, ds.desc.c.n_faces)
    tr_ms, tr_n = ktimes["trace"]                           # summed over the timed steps
    tr_bytes_step = trace_bytes(cnt, ds.desc.c.n_faces)
    tr_ms_launch = tr_ms / max(tr_n, 1)
    achieved = tr_bytes_step * args.steps / (tr_ms * 1e-3) / 1e9 if tr_ms > 0 else 0.0
    step_gbs = alg_bytes / (kernel_ms * 1e-3) / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload)
    except Exception:
        pass
    launches_per_step = cnt["launches"] + 1                 # wavefront kernels of all bands (counted run) + k_develop
    line = {
        "metric": {"bedroom": "path Msamples/s bedroom 720p", "bedroom4k": "path Msamples/s bedroom 2160p 1024 spp", "cornell": "path Msamples/s cornell 256x256"}[args.workload],
        "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if args.workload == "bedroom4k" else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl, "spp_per_gpu": spp // world, "spp_total": spp, "paths_per_step": paths_total, "passes": spp // spp_pp, "max_depth": kw["max_depth"],
                   "rr_depth": kw["rr_depth"], "parallelism": (f"all pixels x {shard[1]} of {spp_pp} samples per pass on each of {world} GPU(s)" if shard is not None else f"image bands x all samples over {world} GPU(s)") + ", NCCL film all-reduce",
                   "l2": "flushed between timed iterations (256 MiB memset)", "bvh_build_ms": ds.build_stats.ms_total,
                   "bvh_nodes": ds.build_stats.n_nodes8, "triangles": ds.build_stats.n_tris,
                   "rays_per_step_rank0": cnt["closest_rays"] + cnt["shadow_rays"], "Mrays_per_s_rank0": (cnt["closest_rays"] + cnt["shadow_rays"]) / (kernel_ms * 1e-3) / 1e6},
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": H * W * 16, "d2h_bytes_per_step": H * W * 16 + H * W * 12,
                "ms_each_rank0": [round(x * 1e3, 2) for x in e2e_times]},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "kernel": "k_wf_trace", "kernel_ms_per_launch": tr_ms_launch, "launches_per_step": tr_n // max(args.steps, 1),
                     "kernel_share_of_step": tr_ms / max(args.steps, 1) / kernel_ms,
                     "algorithmic_bytes_per_launch": tr_bytes_step / max(tr_n // max(args.steps, 1), 1), "bytes_per_ray": b_ray, "peak_kind": peak_kind,
                     "whole_step": {"achieved": step_gbs, "frac": step_gbs / peak, "algorithmic_bytes": alg_bytes, "render_ms": kernel_ms,
                                    "ms_per_step_by_kind": {k: v[0] / max(args.steps, 1) for k, v in ktimes.items()}},
                     "note": "irregular gathers: nodes + triangles (100 MB) stay L2-resident on B200, so DRAM traffic is below the algorithmic bytes; ncu: issue slots 66 % busy, ALU pipe 55-72 %, 44 % of stall cycles long-scoreboard (profiles/r01_v4_*), see DESIGN.md 4.3"},
    }
    if world == 1 and not args.no_cpu_baseline:
        v1, dt1, cores, n1 = cpu_oracle_rate(scene_dict, kw, 1)              # calibrate, then ~15 s of CPU work
        s_spp = int(min(spp, max(1, round(15.0 / max(dt1, 1e-3)))))
        v, dt, cores, n = cpu_oracle_rate(scene_dict, kw, s_spp) if s_spp > 1 else (v1, dt1, cores, n1)
        line["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": cores, "kind": "port",
                                "sample": f"{s_spp} spp of the {spp} spp workload ({n} paths, {dt:.1f} s on {cores} threads; BVH build excluded)"}
    else:
        line["cpu_baseline"] = None
    if world == 1 and not args.no_prb:
        try:
            line["prb"] = prb_bench(local)
        except Exception as e:  # the headline line must still print
            line["prb"] = {"error": str(e)}
    print(json.dumps(line), flush=True)


def prb_bench(local, iters=5):
    """BASELINE metric 2: PRB fwd + adjoint ms/iter -- Cornell box 512x512, 32 spp, textured `white` albedo 960x870 fp32."""
    import torch
    from mitsuba3_experiments_b200 import scenes
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, traverse
    from mitsuba3_experiments_b200.runtime import DeviceScene
    d = scenes.cornell_prb_scene(512, 512, 32)
    ds = DeviceScene(d, device=local)
    prb = PRBIntegrator({"max_depth": 8, "rr_depth": 5})
    params = traverse(ds); params.keep(["white.reflectance.data"])
    target = prb.render(ds, seed=1, spp=32) * 0.5
    times = []
    for i in range(3 + iters):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        img = prb.render(ds, seed=i, spp=32, evaluate=False)
        grad_in = 2.0 * (img - target) / img.numel()
        params.zero_grad()
        prb.render_backward(ds, params, grad_in, seed=1000 + i, spp=32)
        e1.record(); torch.cuda.synchronize()
        if i >= 3:
            times.append(e0.elapsed_time(e1))
    return {"metric": "PRB fwd+adjoint ms/iter", "value": float(np.mean(times)), "unit": "ms", "higher_is_better": False,
            "config": "Cornell box 512x512, 32 spp, max_depth 8, rr_depth 5, white.reflectance = WoodFloor.jpg 960x870 as an fp32 bitmap parameter",
            "paths_per_iter": 512 * 512 * 32}


if __name__ == "__main__":
    main()
