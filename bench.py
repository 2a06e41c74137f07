#!/usr/bin/env python
"""bench.py -- PLT Msamples/s + Mrays/s on the cbox-grating configuration (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--spp S] [--res R] [--impl ours|reference]

A *step* is one complete forward render of the cbox-grating scene (1024x1024, S spp in total, default 1024 = the
configuration the metric is quoted on) through the plt path: seeding, ray generation, sample phase, NEE, film splat and
-- for N > 1 -- one NCCL reduce of the RGBW accumulator.  The spp range is sharded over the N ranks (strong scaling on
the named configuration; scene replicated, sample streams offset by rank).

  value : Msamples/s with everything resident in HBM (no host copies in the timed region), device time from CUDA
          events on the launching stream (pltb_stats.ms_total), max over ranks.
  e2e   : the same metric through the public API `PLTIntegrator.render(scene, spp=..)` with HOST buffers: every step
          uploads the sensor block (camera update, H2D) and downloads the developed H x W x 3 image (D2H) into
          pinned host memory.
  --impl reference : the CPU oracle (the restated reference path; the fork itself cannot be built on this image,
          DESIGN.md) timed on all host cores on a bounded sample of the same workload.
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

METRIC = "PLT Msamples/s (cbox-grating, spectral, plt max_depth=7)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--spp", type=int, default=1024, help="samples per pixel of one step (whole job)")
    ap.add_argument("--res", type=int, default=1024)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target size of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].startswith("Active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


DATA = ("the reference's own scene file scenes/cbox-grating/cbox.xml (+ meshes/*.ply, textures/UVtest.png; shipped under "
        "tests/golden/scenes/), -Dresx/-Dresy as configured, fixed seed; no dataset or weights are involved")


def load_scene(args):
    from mitsuba3_plt_b200 import scenes
    return scenes.load("cbox-grating", resx=args.res, resy=args.res, spp=args.spp)


def cpu_oracle_run(args, scene, seconds, threads=0):
    """The oracle (kind "port") on a bounded sample of the same workload: full frame, as many spp as fit `seconds`."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    probe = oracle_lib.render(scene.flat, args.spp, spp_begin=0, spp_count=1, crop=(0, 0, args.res, max(1, args.res // 16)),
                              film=False, n_threads=threads)
    rate = probe["stats"]["samples"] / max(probe["seconds"], 1e-9)
    n_spp = int(max(1, min(args.spp, seconds * rate / (args.res * args.res))))
    r = oracle_lib.render(scene.flat, args.spp, spp_begin=0, spp_count=n_spp, n_threads=threads)
    st = r["stats"]
    return {"seconds": r["seconds"], "samples": st["samples"], "rays": st["closest_rays"] + st["shadow_rays"],
            "sample": f"full {args.res}x{args.res} frame, spp indices [0,{n_spp}) of {args.spp}, film splat included",
            "cores": threads or os.cpu_count()}


def algorithmic_bytes(st):
    """Algorithmic bytes per kernel over the counted work of one step (DESIGN.md "Rooflines": per-unit figures x counted units).
    S samples, R closest-hit rays (R - S of them secondary), V recorded vertices (Vg on a grating), Q NEE items, Sh shadow rays.
      k_raygen            per sample: 16 B wavelengths + 16 B sensor weight + 8 B film position + 16 B RNG state + 4 B record count,
                          32 B ray, 16 B hit + 4 B queue entry                                                   = 112 B
      k_shade<c>          per vertex: in 4 B entry + 16 B hit + 32 B ray + 16 B RNG + 16 B throughput + 20 B prev; out 16 B weight +
                          16 B replayed throughput + 16 B emissive + 16 B zeroed NEE slot + 88 B NEE record + 4 B NEE entry +
                          32 B next ray + 16 B RNG + 36 B path state + 4 B live entry                                = 364 B
      k_trace             per secondary ray: 4 B entry + 32 B ray in, 16 B hit out                                  = 52 B
      k_classify          per secondary ray: 4 B entry + 16 B hit in, 4 B queue entry out                           = 24 B
      k_nee_light         per NEE item: 4 B entry + 88 B record + 16 B wavelengths + 16 B RNG in; 48 B light sample + 4 B out = 176 B
      k_trace<shadow>     per shadow ray: 4 B entry + 32 B ray in, 4 B visibility out                               = 40 B
      k_nee_post<c>       per NEE item: 4 B entry + 88 B record + 48 B light sample + 16 B wavelengths + 16 B throughput in, 16 B out = 188 B
      k_nee_eval          per evaluated item: 4 B entry + 88 B record + 48 B + 16 B in, 16 B out                    = 172 B
      k_gather            per sample 4 B + 16 B out (+ 16 B dump); per vertex 32 B                                  = 20 B + 32 B/vertex
      k_splat             per sample 16 B L + 16 B + 16 B + 8 B in; film RMW 2 x 16 B x 180 strip pixels per 256 samples = 78.5 B"""
    S, R, V, Vg = st["samples"], st["closest_rays"], st["vertices"], st["grating_vertices"]
    Q0, Q1, Sh, E = st["nee_other"], st["nee_grating"], st["shadow_rays"], st["nee_grating_eval"]
    return {"k_raygen": 112 * S, "k_shade<grating>": 364 * Vg, "k_shade<other>": 364 * (V - Vg), "k_trace": 52 * (R - S),
            "k_classify": 24 * (R - S), "k_nee_light": 176 * (Q0 + Q1), "k_trace<shadow>": 40 * Sh,
            "k_nee_post<other>": 188 * Q0, "k_nee_post<grating>": 188 * Q1, "k_nee_eval": 172 * E,
            "k_gather": 20 * S + 32 * V, "k_splat": 78.5 * S}


def run_ours(args):
    import numpy as np
    import torch
    import mitsuba3_plt_b200 as mi

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun --nproc-per-node N for --gpus N")
    torch.cuda.set_device(local)
    ctx = mi.init_distributed() if world > 1 else mi.context()
    import torch.distributed as dist

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    scene = load_scene(args)
    integ = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
    W, H = scene.film_size()
    cam = np.array(list(scene.flat.sensors[0].to_world), np.float64).reshape(4, 4)
    host_img = torch.empty((H, W, 3), dtype=torch.float32, pin_memory=True)

    def step_resident():
        integ.render(scene, spp=args.spp, seed=0, ctx=ctx, download=False, develop=False)
        return integ.last_stats

    def step_e2e():
        scene.update_sensor(0, cam)                                     # H2D: the step's input (camera block)
        integ.render(scene, spp=args.spp, seed=0, ctx=ctx, out=host_img.numpy())   # D2H: developed image
        return integ.last_stats

    def timed(fn, k):
        barrier()
        t0 = time.perf_counter()
        dev_ms, last = 0.0, None
        for _ in range(k):
            last = fn()
            dev_ms += last["ms_total"]
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t[0].item(), t[1].item(), last

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dev_ms, wall_ms, st = timed(step_resident, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    e_dev_ms, e_wall_ms, _ = timed(step_e2e, args.steps)

    # per-kernel profile of one step (events around every kernel; not part of the timed numbers)
    integ.render(scene, spp=args.spp, seed=0, ctx=ctx, download=False, develop=False, profile=True)
    pst = integ.last_stats
    # whole-job counters: every rank does the same amount of work up to the spp split
    cnt = torch.tensor([st[k] for k in ("samples", "closest_rays", "shadow_rays", "vertices", "grating_vertices",
                                        "kernel_launches")], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(cnt)
        dist.barrier()
        ctx.close()
        dist.destroy_process_group()
    if rank != 0:
        return
    samples, closest, shadow, verts, gverts, launches = (int(x) for x in cnt.tolist())
    ms_step = dev_ms / args.steps
    value = samples / (ms_step * 1e-3) / 1e6
    e2e_ms = e_wall_ms / args.steps
    kern = {"k_sample": pst["ms_sample"], "k_nee": pst["ms_nee_grating"] + pst["ms_nee_other"], "k_splat": pst["ms_splat"]}
    per_kernel = {k: v for k, v in pst["kernels"].items() if v["launches"]}
    top = max(per_kernel, key=lambda k: per_kernel[k]["ms"])          # the dominant CUDA kernel of the step
    top_ms, n_launch = per_kernel[top]["ms"], per_kernel[top]["launches"]
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = peaks.get("hbm_gbs", 6650.0)
    alg = algorithmic_bytes(pst)
    ach = alg[top] / (top_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")   # dram bytes per launch from the committed ncu --set full capture
    traffic_note = None
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        t = tj.get(top)
        if t:   # the capture is one pass of 2^23 samples (ncu replays every launch ~40x); a bench launch covers a larger pass
            scale = (pst["samples"] / max(1, pst["passes"])) / tj.get("_samples_per_pass_in_capture", 1 << 23)
            traffic = round(t["dram_bytes_per_launch"] * scale)
            traffic_note = (f"dram__bytes_read+write per launch from profiles/ncu_traffic.json ({t['dram_bytes_per_launch']} B at "
                            f"{tj.get('_samples_per_pass_in_capture', 1 << 23)} samples per pass) scaled by {scale:.2f} to this run's pass size")
    # the limiter of this path is the SM issue rate at low occupancy, not HBM: next to the contract's HBM roofline the
    # line carries the issue-slot utilisation of the dominant kernel from the committed ncu capture (share must agree, not absolutes)
    issue = None
    if os.path.exists(tpath) and tj.get(top, {}).get("issue_slots_busy_pct") is not None:
        t = tj[top]
        issue = {"kernel": top, "bound": "issue", "achieved": t["issue_slots_busy_pct"], "peak": 100.0, "unit": "% of issue slots busy",
                 "frac": round(t["issue_slots_busy_pct"] / 100.0, 3), "active_lanes_of_32": t.get("active_lanes"),
                 "occupancy_pct": t.get("occupancy_pct"), "registers": t.get("registers"),
                 "source": "profiles/ncu_traffic.json (ncu --set full, one pass of the same scene)"}
    out = {
        "metric": METRIC, "value": round(value, 3), "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": DATA,
        "config": {"workload": f"cbox-grating {args.res}x{args.res}, {args.spp} spp per step sharded over {world} GPU(s), "
                   "plt max_depth=7 rr_depth=50, gaussian rfilter", "samples_per_step": samples,
                   "l2": "per-pass wavefront state (up to 2^25 samples x ~1.2 KB = 41 GB) exceeds L2; no explicit flush",
                   "timing": "CUDA events on the render stream inside pltb_render (pltb_stats.ms_total), max over ranks"},
        "mrays_per_s": round((closest + shadow) / (ms_step * 1e-3) / 1e6, 3),
        "rays_per_sample": round((closest + shadow) / samples, 4), "mean_depth": round(verts / samples, 4),
        "mean_grating_vertices": round(gverts / samples, 4),
        "wall_ms_per_step": round(wall_ms / args.steps, 3),
        "gpu_launches": launches * args.steps // 1,
        "clocks": clocks,
        "e2e": {"value": round(samples / (e2e_ms * 1e-3) / 1e6, 3), "unit": "Msamples/s",
                "h2d_bytes_per_step": 156, "d2h_bytes_per_step": int(H * W * 3 * 4), "ms_per_step": round(e2e_ms, 3),
                "note": "wall clock around PLTIntegrator.render incl. sensor upload and image download to pinned host memory"},
        "kernel_ms_per_step": {k: round(v, 3) for k, v in kern.items()} | {
            "reduce": round(pst["ms_reduce"], 3), "develop": round(pst["ms_develop"], 3), "copy": round(pst["ms_copy"], 3)},
        "kernels": {k: {"ms": round(v["ms"], 3), "launches": v["launches"],
                        "algorithmic_GB_per_s": round(alg[k] / max(v["ms"], 1e-6) / 1e6, 1)} for k, v in per_kernel.items()},
        "roofline": {"kernel": top, "bound": "hbm", "achieved": round(ach, 2), "peak": peak, "unit": "GB/s",
                     "frac": round(ach / peak, 4), "traffic": traffic, "traffic_note": traffic_note, "launches_per_step": n_launch,
                     "avg_launch_ms": round(top_ms / n_launch, 4), "algorithmic_bytes_per_launch": round(alg[top] / n_launch),
                     "share_of_step": round(top_ms / sum(v["ms"] for v in per_kernel.values()), 4),
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (sustained copy)" if peaks else "fallback 6650 GB/s",
                     "note": "the path is SM-issue (FP32+MUFU) bound, not HBM bound (SURVEY.md 8d): profiles/ holds issue-slot "
                             "utilisation, active lanes and stall mix per kernel from ncu; kernel times here come from CUDA "
                             "events after every launch in a separate profiled step"},
    }
    if issue:
        out["roofline_issue"] = issue
    if not args.no_cpu_baseline:
        c = cpu_oracle_run(args, scene, args.cpu_seconds)
        out["cpu_baseline"] = {"value": round(c["samples"] / c["seconds"] / 1e6, 4), "unit": "Msamples/s",
                               "cores": c["cores"], "kind": "port", "sample": c["sample"],
                               "mrays_per_s": round(c["rays"] / c["seconds"] / 1e6, 4)}
    emit(out)


def run_reference(args):
    """Reference arm: the reference's own implementation cannot be built on this image (no Dr.Jit/LLVM, missing
    submodules), so this times the oracle port of it on all host cores.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    scene = load_scene(args)
    t_each = max(2.0, min(args.cpu_seconds, 60.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_oracle_run(args, scene, min(t_each, 2.0))
    tot_s, tot_n, tot_r, last = 0.0, 0, 0, None
    for _ in range(args.steps):
        last = cpu_oracle_run(args, scene, t_each)
        tot_s += last["seconds"]; tot_n += last["samples"]; tot_r += last["rays"]
    v = tot_n / tot_s / 1e6
    out = {"impl": "reference", "metric": METRIC, "value": round(v, 4), "unit": "Msamples/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(tot_s / args.steps * 1e3, 3),
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
           "data": DATA,
           "config": {"workload": f"cbox-grating {args.res}x{args.res}, {args.spp} spp, plt max_depth=7 rr_depth=50 "
                                  f"(each step a bounded sample: {last['sample']})"},
           "mrays_per_s": round(tot_r / tot_s / 1e6, 4),
           "cpu_baseline": {"value": round(v, 4), "unit": "Msamples/s", "cores": last["cores"], "kind": "port",
                            "sample": last["sample"]},
           "e2e": {"value": round(v, 4), "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(out)


def claim_stdout():
    """Only the JSON line may reach stdout: native libraries (NCCL prints its version banner from C) write to fd 1 directly, so
    fd 1 is pointed at stderr for the whole run and the real stdout is kept aside for emit()."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(obj):
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()


if __name__ == "__main__":
    a = parse()
    claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
