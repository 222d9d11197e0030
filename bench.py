#!/usr/bin/env python
"""bench.py -- headline benchmark: path Msamples/s on the bedroom (stand-in) 720p, 64 spp per GPU, path-mis loop, and the
PRB fwd + adjoint iteration time (BASELINE.json metric, both halves).

One *step* = one full render of the workload: every rank renders every pixel with its slice of the samples into a private
RGBW film, the films are summed with one NCCL all-reduce, the result is developed.  Weak scaling: the job at N GPUs is the
1280x720 image at 64*N spp, i.e. 58 982 400 paths per GPU (BASELINE.json configs[1] at N = 1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b2|reference] [--workload bedroom|cornell|bedroom4k|bedroom_xml|soup]

Prints ONE JSON line (rank 0):
  value        device-timed throughput with the scene resident in HBM (CUDA events, max over ranks)
  e2e          the same metric through the host-buffer C-ABI call (film H2D + render + develop + film / image D2H timed)
  roofline     the dominant kernel, k_wf_trace on the rays of bounces >= 1: algorithmic bytes = rays x B_ray (exact ray
               counters) over the CUDA-event time of its launches inside the timed region, against the measured HBM copy bandwidth;
               bounce 0 (k_wf_trace_packet, one traversal per 64 lanes) is timed alone and reported beside it
  cpu_baseline the CPU oracle (a port of the reference's llvm_ad_rgb path; Mitsuba itself cannot be installed here) on a
               bounded sample of the same workload; the same oracle image is the checker of the `parity` record
  parity       CUDA vs oracle on the benchmark scene itself (same seed, the cpu_baseline sample) + converged relMSE of C1
  prb          PRB fwd + adjoint ms/iter (C3) with its own roofline / e2e / cpu_baseline; at N GPUs the fixed 32 spp are
               split by sample slices, film and gradients merged by NCCL all-reduce
  strong       (N > 1) the bedroom at a FIXED 64 spp total, with the same-run 1-GPU time and the merged-image difference
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
    if name == "bedroom_xml":   # the material palette of data/bedroom/scene.xml (roughplastic, conductor, roughdielectric, mask included)
        spp = 64 * n_gpus
        d = scenes.bedroom_standin(1280, 720, spp, palette="scene_xml")
        return d, dict(variant=1, max_depth=8, rr_depth=2), spp, "bedroom stand-in 1280x720, scene.xml material palette (1.8M tris), path-mis.py loop"
    if name == "bedroom4k":      # BASELINE.json configs[4]: fixed total work (strong scaling), 2 passes x 512 spp (integrator.cpp:251-257)
        spp = 1024
        d = scenes.bedroom_standin(3840, 2160, spp)
        return d, dict(variant=1, max_depth=8, rr_depth=2), spp, "bedroom stand-in 3840x2160 1024 spp (1.8M tris, 2 area lights + envmap), path-mis.py loop, lanes sharded over the GPUs"
    if name == "soup":           # BASELINE.json configs[3]: BVH build + traversal stress
        spp = 128 * n_gpus
        return scenes.random_triangles(10_000_000, 1920, 1080, spp), dict(variant=0, max_depth=8, rr_depth=4), spp, "10M random triangles 1920x1080, path.py loop"
    if name == "cornell":
        spp = 16 * n_gpus
        return scenes.cornell_box(256, 256, spp), dict(variant=0, max_depth=8, rr_depth=4), spp, "mi.cornell_box() 256x256, path.py loop"
    raise SystemExit(f"unknown workload {name}")


def ray_bytes(n_faces):
    D = max(1, math.ceil(math.log(max(n_faces / 3.0, 1.0), 8)))
    return 80 * D + 48 * 3


def algorithmic_bytes(cnt, n_faces, tex_bytes_per_eval=32, splat=True):
    """bytes any implementation must move from global memory for these events."""
    b_ray = ray_bytes(n_faces)
    return ((cnt["closest_rays"] + cnt["shadow_rays"]) * b_ray + cnt["vertices"] * 140 + (cnt["tex_evals"] // 2) * tex_bytes_per_eval
            + cnt["nee_area"] * 64 + cnt["nee_env"] * 224 + cnt["env_escapes"] * 48 + (cnt["splat_taps"] * 16 if splat else 0)), b_ray


def trace_bytes(cnt, n_faces):
    """Algorithmic bytes of the trace launches alone: every ray fetches D wide nodes and 3 triangles."""
    return (cnt["closest_rays"] + cnt["shadow_rays"]) * ray_bytes(n_faces)


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


def host_threads():
    # all host threads this process may run on: torchrun exports OMP_NUM_THREADS=1 to every rank, which would otherwise
    # silently turn the "all host threads" baseline into a single-thread one at N > 1
    return len(os.sched_getaffinity(0))


def oracle_scene(scene_dict):
    from mitsuba3_experiments_b200.scene import SceneDesc
    from oracle.oracle_py import OracleScene, lib
    lib().orc_set_num_threads(host_threads())
    return OracleScene(SceneDesc(scene_dict))


def cpu_oracle_render(orc, kw, spp, seed=0):
    """Times the CPU oracle on `spp` samples per pixel of the workload (BVH build excluded, like the GPU arm)."""
    from oracle.oracle_py import lib
    t0 = time.perf_counter()
    film, cnt = orc.render_film(seed=seed, spp=spp, **kw)
    dt = time.perf_counter() - t0
    return orc.develop(film), cnt["samples"] / dt / 1e6, dt, lib().orc_num_threads(), cnt["samples"]


def run_reference(args, rank):
    """--impl reference: the reference's CPU path.  Mitsuba/Dr.Jit (llvm_ad_rgb) cannot be installed on this box, so the
    timed implementation is the oracle port (oracle/), with all host threads, on a bounded sample per step."""
    if rank != 0:
        return
    d, kw, spp, wl = build_workload(args.workload, 1)
    from oracle.oracle_py import lib
    orc = oracle_scene(d)
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
            "dtype": "f32", "data": "synthetic", "config": {"workload": wl, "sample": sample, "note": "CPU oracle port of the reference's llvm_ad_rgb path (scalar C++ restatement, -O3, OpenMP over lanes; Mitsuba not installable offline)"},
            "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": lib().orc_num_threads(), "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def rel_err(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 2e-2)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b2", choices=["b2", "reference"])
    ap.add_argument("--workload", default="bedroom", choices=["bedroom", "cornell", "bedroom4k", "bedroom_xml", "soup"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-prb", action="store_true")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--env-warp", default="hierarchical", choices=["hierarchical", "alias"], help="envmap importance sampling: the reference's Hierarchical2D warp or the alias table (in-expectation parity only)")
    ap.add_argument("--shard", default="samples", choices=["samples", "bands"], help="multi-GPU split: sample slices of every pixel (balanced) or image bands")
    ap.add_argument("--bands", type=int, default=2, help="image bands per rank (N > 1); bands are interleaved across ranks")
    args = ap.parse_args()
    rank_env = int(os.environ.get("RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank_env)

    import torch
    import torch.distributed as tdist
    from mitsuba3_experiments_b200 import capi
    from mitsuba3_experiments_b200 import dist as b2dist
    from mitsuba3_experiments_b200.runtime import DeviceScene

    rank, world, local = b2dist.init_from_env()
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    scene_dict, kw, spp, wl = build_workload(args.workload, world)
    kw = dict(kw, env_warp=capi.ENV_WARP_ALIAS if args.env_warp == "alias" else capi.ENV_WARP_HIERARCHICAL)
    ds = DeviceScene(scene_dict, device=local)
    H, W = ds.film_shape
    n_total = ds.wavefront_size(spp)
    spp_pp = n_total // (H * W)                              # samples per pixel per pass (< spp when H W spp exceeds 2^32 - 1 lanes)
    paths_total = H * W * spp
    plan = b2dist.shard_plan(n_total, spp_pp, W, world, rank, bands_per_rank=args.bands if world > 1 else 1)
    shard = b2dist.sample_shard(spp_pp, world, rank) if args.shard == "samples" else None
    shard_kw = {}
    if shard is not None:                                   # every rank: all pixels, its slice of the samples of each pass
        shard_kw = dict(sample_begin=shard[0], sample_count=shard[1])
        plan = [(0, H * W * shard[1])]
    film = ds.new_film()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def render_bands(counters=False, plan=plan, shard_kw=shard_kw, spp=spp, seed=0):
        tot = None
        for a, b in plan:
            _, c = ds.render_film(film=film, counters=counters, seed=seed, spp=spp, lane_begin=a, lane_end=b, **shard_kw, **kw)
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

    def max_over_ranks(vals):
        t = torch.tensor([float(v) for v in vals], device=dev, dtype=torch.float64)
        if world > 1:
            tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
        return [float(x) for x in t]

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
    # bounce 0 goes through the packet variant of the trace kernel (k_wf_trace_packet): time that launch alone (same lanes, max_depth 1,
    # untimed region, L2 flushed) so that the roofline of the dominant kernel -- k_wf_trace on the incoherent rays of bounces >= 1 -- excludes it
    kw_primary = dict(kw, max_depth=1)
    for i in range(2 + args.steps):
        if i == 2:
            ds.set_kernel_timing(True); ds.kernel_times()
        flush.zero_(); film.zero_()
        for a, b in plan:
            ds.render_film(film=film, seed=0, spp=spp, lane_begin=a, lane_end=b, **shard_kw, **kw_primary)
    primary_ms, primary_n = ds.kernel_times()["trace"]; ds.set_kernel_timing(False)
    step_ms = [e0.elapsed_time(e2) for e0, _, e2 in ev]
    kern_ms = [e0.elapsed_time(e1) for e0, e1, _ in ev]
    ms_per_step, kernel_ms = max_over_ranks([np.mean(step_ms), np.mean(kern_ms)])
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
    e2e_value = paths_total / max_over_ranks([np.mean(e2e_times)])[0] / 1e6

    strong = None
    if world > 1 and args.workload == "bedroom" and not args.no_strong:
        strong = strong_record(ds, film, kw, world, rank, dev, args.steps, flush, barrier, max_over_ranks)
    prb = None
    if not args.no_prb:
        try:
            prb = prb_bench(local, rank, world, dev, barrier, max_over_ranks, cpu=(world == 1 and not args.no_cpu_baseline))
        except Exception as e:  # the headline line must still print
            prb = {"error": repr(e)}

    if world > 1:
        tdist.barrier()
    if rank != 0:
        if world > 1:
            tdist.destroy_process_group()
        return
    peak, peak_kind = measured_peak()
    alg_bytes, b_ray = algorithmic_bytes(cnt, ds.desc.c.n_faces)
    tr_all_ms, tr_all_n = ktimes["trace"]                   # summed over the timed steps, bounce 0 (packet kernel) included
    primary_n = args.steps * len(plan)                      # (a max_depth 1 render also makes an empty second trace launch: its few us stay in primary_ms)
    tr_ms, tr_n = tr_all_ms - primary_ms, tr_all_n - primary_n
    primary_rays = cnt["samples"]                           # one primary ray per path
    tr_bytes_step = trace_bytes(cnt, ds.desc.c.n_faces) - primary_rays * ray_bytes(ds.desc.c.n_faces)
    tr_ms_launch = tr_ms / max(tr_n, 1)
    achieved = tr_bytes_step * args.steps / (tr_ms * 1e-3) / 1e9 if tr_ms > 0 else 0.0
    step_gbs = alg_bytes / (kernel_ms * 1e-3) / 1e9
    traffic = dram_by_kind = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload)
        if isinstance(tj, dict):
            traffic, dram_by_kind = tj.get("k_wf_trace_per_launch"), tj.get("dram_bytes_per_step_by_kind")
        else:
            traffic = tj
    except Exception:
        pass
    launches_per_step = cnt["launches"] + 1                 # wavefront kernels of all bands (counted run) + k_develop
    bs = ds.build_stats
    line = {
        "metric": {"bedroom": "path Msamples/s bedroom 720p", "bedroom_xml": "path Msamples/s bedroom 720p (scene.xml palette)", "bedroom4k": "path Msamples/s bedroom 2160p 1024 spp",
                   "cornell": "path Msamples/s cornell 256x256", "soup": "path Msamples/s 10M random triangles 1080p"}[args.workload],
        "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if args.workload == "bedroom4k" else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl, "spp_per_gpu": spp // world, "spp_total": spp, "paths_per_step": paths_total, "passes": spp // spp_pp, "max_depth": kw["max_depth"],
                   "rr_depth": kw["rr_depth"], "env_warp": args.env_warp, "parallelism": (f"all pixels x {shard[1]} of {spp_pp} samples per pass on each of {world} GPU(s)" if shard is not None else f"image bands x all samples over {world} GPU(s)") + ", NCCL film all-reduce",
                   "l2": "flushed between timed iterations (256 MiB memset)", "bvh_build_ms": bs.ms_total, "bvh_builder": {0: "lbvh", 1: "ploc"}.get(bs.builder, bs.builder),
                   "bvh_depth": bs.max_depth, "bvh_stack_capacity": 48, "bvh_sah_cost": bs.sah_cost,
                   "bvh_nodes": bs.n_nodes8, "triangles": bs.n_tris,
                   "rays_per_step_rank0": cnt["closest_rays"] + cnt["shadow_rays"], "Mrays_per_s_rank0": (cnt["closest_rays"] + cnt["shadow_rays"]) / (kernel_ms * 1e-3) / 1e6,
                   "node_fetches_per_ray": cnt["node_fetches"] / max(cnt["closest_rays"] + cnt["shadow_rays"], 1), "tri_tests_per_ray": cnt["tri_fetches"] / max(cnt["closest_rays"] + cnt["shadow_rays"], 1)},
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": H * W * 16, "d2h_bytes_per_step": H * W * 16 + H * W * 12,
                "ms_each_rank0": [round(x * 1e3, 2) for x in e2e_times]},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "kernel": "k_wf_trace", "kernel_ms_per_launch": tr_ms_launch, "launches_per_step": tr_n // max(args.steps, 1),
                     "kernel_share_of_step": tr_ms / max(args.steps, 1) / kernel_ms,
                     "rays_per_step": cnt["closest_rays"] + cnt["shadow_rays"] - primary_rays,
                     "bounce0_packet_kernel": {"kernel": "k_wf_trace_packet", "ms_per_launch": primary_ms / max(primary_n, 1), "launches_per_step": primary_n // max(args.steps, 1),
                                               "rays_per_step": primary_rays, "model_GBps_at_bytes_per_ray": primary_rays * ray_bytes(ds.desc.c.n_faces) / max(primary_ms / max(args.steps, 1) * 1e-3, 1e-12) / 1e9,
                                               "note": "timed alone (max_depth 1 run of the same lanes); one traversal per 64 consecutive lanes, so its real node traffic is ~1/64 of the per-ray model"},
                     "algorithmic_bytes_per_launch": tr_bytes_step / max(tr_n // max(args.steps, 1), 1), "bytes_per_ray": b_ray, "peak_kind": peak_kind,
                     "whole_step": {"achieved": step_gbs, "frac": step_gbs / peak, "algorithmic_bytes": alg_bytes, "render_ms": kernel_ms,
                                    "ms_per_step_by_kind": {k: v[0] / max(args.steps, 1) for k, v in ktimes.items()}, "trace_ms_bounce0": primary_ms / max(args.steps, 1),
                                    "dram_bytes_per_step_by_kind": dram_by_kind},
                     "note": "irregular gathers: nodes + triangles (100 MB) stay L2-resident on B200, so DRAM traffic is below the algorithmic bytes; the binding limit of k_wf_trace is the ALU pipe / SIMT efficiency, not DRAM (profiles/README.md, DESIGN.md 4.3)"},
    }
    if world == 1 and not args.no_cpu_baseline:
        orc = oracle_scene(scene_dict)
        _, v1, dt1, cores, n1 = cpu_oracle_render(orc, kw_oracle(kw), 1, seed=1)          # calibrate, then ~15 s of CPU work
        s_spp = int(min(spp, max(1, round(15.0 / max(dt1, 1e-3)))))
        ref_img, v, dt, cores, n = cpu_oracle_render(orc, kw_oracle(kw), s_spp)
        line["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": cores, "kind": "port",
                                "sample": f"{s_spp} spp of the {spp} spp workload ({n} paths, {dt:.1f} s on {cores} threads; BVH build excluded)"}
        # parity on the benchmark scene itself: the CUDA render of the same sample set (same seed, same spp) against the oracle image
        parity = {}
        if args.env_warp == "hierarchical":
            img_s, _ = ds.render(seed=0, spp=s_spp, **kw)
            e = rel_err(img_s.cpu().numpy(), ref_img)
            parity["benchmark_scene"] = {"spp": s_spp, "frac_within_1e-3": float((e <= 1e-3).mean()), "mean_rel_err": float(np.minimum(e, 1.0).mean()),
                                         "criterion": "|b2 - oracle| / max(|oracle|, 2e-2), same seeds (SURVEY.md 8d: >= 0.99 within 1e-3)"}
        try:
            parity["c1_converged"] = c1_converged_relmse(local)
        except Exception as e:
            parity["c1_converged"] = {"error": repr(e)}
        line["parity"] = parity
        del orc
    else:
        line["cpu_baseline"] = None
    if prb is not None:
        line["prb"] = prb
    if strong is not None:
        line["strong"] = strong
    print(json.dumps(line), flush=True)
    if world > 1:
        tdist.destroy_process_group()


def kw_oracle(kw):
    return {k: v for k, v in kw.items() if k in ("variant", "max_depth", "rr_depth")}


def c1_converged_relmse(local):
    """north_star: converged-image relative MSE at high spp -- CUDA C1 at 4 096 spp against the stored 16 384-spp oracle image."""
    from mitsuba3_experiments_b200 import scenes
    from mitsuba3_experiments_b200.runtime import DeviceScene
    g = np.load(os.path.join(ROOT, "tests", "golden", "c1_oracle_16384spp.npz"))
    ref = g["image"]
    ds = DeviceScene(scenes.cornell_box(256, 256, 4096), device=local)
    img, _ = ds.render(seed=0, spp=4096, max_depth=8, rr_depth=4, variant=0)
    img = img.cpu().numpy()
    return {"relmse": float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2))), "relmse_oracle_4096spp_same_reference": float(g["relmse_oracle_4096"]),
            "spp": 4096, "reference": "CPU oracle at 16384 spp (tests/golden/c1_oracle_16384spp.npz)", "mean_ratio": float(img.mean() / ref.mean())}


def strong_record(ds, film, kw, world, rank, dev, steps, flush, barrier, max_over_ranks):
    """Fixed total work: the 1280x720 bedroom at 64 spp TOTAL, sample slices over the ranks (64 / N spp each), film all-reduce.
    Rank 0 also renders the same job alone (same run, same clocks) for the 1-GPU time and the reference image."""
    import torch
    from mitsuba3_experiments_b200 import dist as b2dist
    spp = 64
    H, W = ds.film_shape
    sh = b2dist.sample_shard(spp, world, rank)
    if sh is None:
        return {"skipped": f"{spp} spp do not divide by {world}"}
    skw = dict(sample_begin=sh[0], sample_count=sh[1])

    def job():
        film.zero_()
        ds.render_film(film=film, seed=0, spp=spp, **skw, **kw)
        b2dist.merge_sum(film)
        return ds.develop(film)

    for _ in range(3):
        img = job()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    for e0, e1 in ev:
        flush.zero_()
        e0.record(); img = job(); e1.record()
    barrier()
    ms = max_over_ranks([np.mean([a.elapsed_time(b) for a, b in ev])])[0]
    merged = img.clone()
    n1_ms, diff = 0.0, 0.0
    if rank == 0:                       # the same job on one GPU, in the same run
        f1 = ds.new_film()
        for _ in range(2):
            f1.zero_(); ds.render_film(film=f1, seed=0, spp=spp, **kw)
        ev1 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        torch.cuda.synchronize(dev)
        for e0, e1 in ev1:
            flush.zero_(); f1.zero_()
            e0.record(); ds.render_film(film=f1, seed=0, spp=spp, **kw); img1 = ds.develop(f1); e1.record()
        torch.cuda.synchronize(dev)
        n1_ms = float(np.mean([a.elapsed_time(b) for a, b in ev1]))
        ref = img1.cpu().numpy(); got = merged.cpu().numpy()
        diff = float(np.max(np.abs(got - ref) / np.maximum(np.abs(ref), 1e-3)))
    barrier()
    paths = H * W * spp
    return {"workload": "bedroom stand-in 1280x720 at a fixed 64 spp total", "scaling": "strong", "n_gpus": world, "spp_per_gpu": sh[1], "ms_per_step": ms,
            "value": paths / (ms * 1e-3) / 1e6, "unit": "Msamples/s", "n1_ms_per_step_same_run": n1_ms,
            "efficiency_vs_n1": (n1_ms / ms / world) if ms > 0 else None,
            "merged_image_max_rel_diff_vs_1gpu": diff, "note": "max |merged - 1-GPU| / max(|1-GPU|, 1e-3): fp32 summation order only"}


def prb_bench(local, rank, world, dev, barrier, max_over_ranks, iters=5, cpu=True):
    """BASELINE metric 2: PRB fwd + adjoint ms/iter -- Cornell box 512x512, 32 spp, textured `white` albedo 960x870 fp32.
    One iteration = primal render (seed i) + loss gradient + render_backward (seed 1000 + i) [+ NCCL film / gradient merges]."""
    import torch
    from mitsuba3_experiments_b200 import dist as b2dist
    from mitsuba3_experiments_b200 import scenes
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, traverse
    from mitsuba3_experiments_b200.runtime import DeviceScene
    spp, res = 32, 512
    d = scenes.cornell_prb_scene(res, res, spp)
    ds = DeviceScene(d, device=local)
    prb = PRBIntegrator({"max_depth": 8, "rr_depth": 5})
    key = "white.reflectance.data"
    params = traverse(ds); params.keep([key])
    target = prb.render(ds, seed=1, spp=spp) * 0.5
    sh = b2dist.sample_shard(spp, world, rank)
    split = "samples" if (world == 1 or sh is not None) else "bands"
    film = ds.new_film()

    def iteration(i, counters=False):
        film.zero_()
        if world == 1:
            prb.render(ds, seed=i, spp=spp, develop=False, evaluate=False, film=film, counters=counters)
            c_primal = prb.last_counters
            img = ds.develop(film)
        else:
            img = b2dist.render_sharded_samples(prb, ds, seed=i, spp=spp, rank=rank, world=world, film=film)
            c_primal = None
        grad_in = 2.0 * (img - target) / img.numel()
        params.zero_grad()
        if world == 1:
            prb.render_backward(ds, params, grad_in, seed=1000 + i, spp=spp, counters=counters)
        else:
            b2dist.render_backward_sharded(prb, ds, params, grad_in, seed=1000 + i, spp=spp, rank=rank, world=world, split=split)
        return c_primal, prb.last_counters

    c_primal, c_back = iteration(0, counters=(world == 1))
    times = []
    ds.set_kernel_timing(False)
    for i in range(3 + iters):
        if i == 3:
            barrier(); ds.set_kernel_timing(True); ds.kernel_times()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(); iteration(i); e1.record()
        barrier()
        if i >= 3:
            times.append(e0.elapsed_time(e1))
    kt = ds.kernel_times(); ds.set_kernel_timing(False)
    ms = max_over_ranks([np.mean(times)])[0]
    out = {"metric": "PRB fwd+adjoint ms/iter", "value": ms, "unit": "ms", "higher_is_better": False, "n_gpus": world,
           "scaling": "strong" if world > 1 else "n/a", "dtype": "f32",
           "config": {"workload": "Cornell box 512x512, 32 spp total, max_depth 8, rr_depth 5, white.reflectance = WoodFloor.jpg 960x870 as an fp32 bitmap parameter",
                      "paths_per_iter": res * res * spp, "parallelism": (f"{sh[1]} of {spp} samples of every pixel per GPU" if sh else "image bands") + ", NCCL all-reduce of the film and of the gradient" if world > 1 else "1 GPU"},
           "ms_by_kind_rank0": {k: v[0] / iters for k, v in kt.items()}, "launches_per_iter_rank0": sum(v[1] for v in kt.values()) // iters}
    # end to end: host target image in, host gradient + loss out, every iteration
    if world == 1:
        tgt_h = target.cpu().pin_memory(); grad_h = torch.empty((870, 960, 3), dtype=torch.float32).pin_memory(); loss_h = torch.empty((), dtype=torch.float32).pin_memory()
        e2e = []
        for i in range(2 + iters):
            torch.cuda.synchronize(dev); t0 = time.perf_counter()
            tgt = tgt_h.to(dev, non_blocking=True)
            film.zero_(); prb.render(ds, seed=i, spp=spp, develop=False, evaluate=False, film=film); img = ds.develop(film)
            diff = img - tgt
            params.zero_grad()
            prb.render_backward(ds, params, 2.0 * diff / img.numel(), seed=1000 + i, spp=spp)
            grad_h.copy_(params.grad[key], non_blocking=True); loss_h.copy_((diff * diff).mean(), non_blocking=True)
            torch.cuda.synchronize(dev)
            if i >= 2:
                e2e.append((time.perf_counter() - t0) * 1e3)
        out["e2e"] = {"value": float(np.mean(e2e)), "unit": "ms", "h2d_bytes_per_step": res * res * 12, "d2h_bytes_per_step": 870 * 960 * 12 + 4}
        # roofline: SURVEY.md 8(d) "PRB iter" = bytes(primal render) + 2 x bytes(sample without splat) + textured vertices x B_grad + splat read
        peak, peak_kind = measured_peak()
        b_primal, b_ray = algorithmic_bytes(c_primal, 36, tex_bytes_per_eval=64)
        b_back, _ = algorithmic_bytes(c_back, 36, tex_bytes_per_eval=64, splat=False)      # primal replay + adjoint pass
        b_grad = c_back["grad_atomics"] * 8                                                   # fp32 read-modify-write per atomic
        b_gather = c_primal["splat_taps"] * 16                                                # adjoint gather of grad_in / W over the footprint
        total = b_primal + b_back + b_grad + b_gather
        gbs = total / (ms * 1e-3) / 1e9
        out["roofline"] = {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": None, "peak_kind": peak_kind,
                           "algorithmic_bytes_per_iter": total, "bytes_per_ray": b_ray, "grad_atomics_per_iter": c_back["grad_atomics"],
                           "note": "36 triangles: nothing here is HBM-bound by construction (SURVEY.md 8d); the iteration is bound by issue slots / latency of the "
                                   "trace + shade kernels and by L2 reductions (film splat, texel scatter); see profiles/README.md for the ncu counters"}
        if cpu:
            orc = oracle_scene(d)
            s_spp = 4
            t0 = time.perf_counter()
            img_o, _ = orc.render(seed=0, spp=s_spp, max_depth=8, rr_depth=5, variant=2)
            gi = (2.0 * (img_o - target.cpu().numpy()) / img_o.size).astype(np.float32)
            gref = np.zeros((870, 960, 3), np.float32)
            orc.render_backward(gi, [(params._index[key][0], 1, gref)], seed=1000, spp=s_spp, max_depth=8, rr_depth=5)
            dt = time.perf_counter() - t0
            out["cpu_baseline"] = {"value": dt * 1e3 * spp / s_spp, "unit": "ms", "cores": host_threads(), "kind": "port",
                                   "sample": f"{s_spp} of 32 spp (render + render_backward took {dt:.2f} s), scaled linearly to 32 spp"}
    return out


if __name__ == "__main__":
    main()
