#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 ray/path back end (driver contract: see README/DESIGN.md).

Workload (BASELINE.json configs[2], "C3"): the 1M-triangle procedurally displaced sphere and, per
GPU and per step, one batch each of 16 Mi coherent primary closest-hit rays, 16 Mi any-hit shadow
rays and 16 Mi incoherent cosine-diffuse closest-hit rays, traced through the C-ABI
(include/y4gpu.h) by the hand-written sm_100a BVH8 kernels.  metric = Mrays/s, whole job.

  python bench.py [--gpus N --steps K --warmup W]          our arm (torchrun for N > 1)
  python bench.py --impl reference [...]                    the reference's CPU kd-tree, host cores

Only the CPU legs (cpu_baseline, --impl reference) touch oracle/ (the compiled, unmodified
reference); the measured GPU path never does and fails loudly without liby4gpu.so + a GPU.
"""
from __future__ import annotations

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

METRIC = "path-traced Msamples/s (Cornell box 1920x1080, 8 bounces + NEE, reference sample sequences); ray-batch Mrays/s in `raybatch`"
UNIT = "Msamples/s"
RB_METRIC = "Mrays/s (closest-hit primary + any-hit shadow + cosine-diffuse closest-hit, 1M-triangle displaced sphere)"
RB_UNIT = "Mrays/s"
RB_METRIC_C5 = "Mrays/s (incoherent closest-hit + any-hit segments, 50M-triangle random soup)"
BATCHES = ("primary", "shadow", "diffuse")
LIGHT = (2.0, 3.0, 4.0)
SHADOW_BIAS, MIN_RAYDIST = 5.0e-4, 5.0e-5


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--res", type=int, default=4096, help="primary grid is res x res rays (16 Mi at 4096)")
    ap.add_argument("--nu", type=int, default=1000, help="sphere longitude segments (nu * nu/2 * 2 triangles)")
    ap.add_argument("--cpu-sample-res", type=int, default=512, help="CPU legs trace a (res/k)^2 strided subsample")
    ap.add_argument("--pt-width", type=int, default=1920)
    ap.add_argument("--pt-height", type=int, default=1080)
    ap.add_argument("--pt-spp", type=int, default=16, help="samples per pixel per GPU per step (one AA pass); the config's 1024 spp = 64 such steps")
    ap.add_argument("--pt-bounces", type=int, default=8)
    ap.add_argument("--cpu-pt-width", type=int, default=0, help="reference arm film width (0 = the GPU arm's: 1920)")
    ap.add_argument("--cpu-pt-height", type=int, default=0)
    ap.add_argument("--cpu-pt-spp", type=int, default=1, help="reference arm samples per pixel per step (the rate is per pixel-sample)")
    ap.add_argument("--workload", default="c2", choices=["c2", "c4", "c5"],
                    help="c2 (default, the headline): Cornell path tracer + C3 ray batches; c4: 10 M-triangle instanced city at 3840x2160; "
                         "c5: 50 M-triangle soup, incoherent closest-hit + any-hit segment batches")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: per-GPU work fixed (spp / rays per GPU); strong: the job is fixed (spp per pass / rays per batch split over the GPUs)")
    ap.add_argument("--city-tris", type=int, default=10_000_000)
    ap.add_argument("--soup-tris", type=int, default=50_000_000)
    ap.add_argument("--no-raybatch", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--only-raybatch", action="store_true", help="(profiling runs) skip the path tracer part")
    ap.add_argument("--no-profile", action="store_true", help="(ncu runs) skip the extra profiled pass; the line then carries no roofline")
    args = ap.parse_args()
    if args.workload == "c4" and args.pt_width == 1920 and args.pt_height == 1080:
        args.pt_width, args.pt_height = 3840, 2160
    if not args.cpu_pt_width:
        args.cpu_pt_width, args.cpu_pt_height = args.pt_width, args.pt_height
    return args


def pt_config_dict(args, world=1):
    strong = args.scaling == "strong"
    spp_gpu = args.pt_spp / world if strong else args.pt_spp
    if args.workload == "c4":
        what = ("C4 (BASELINE configs[3]): procedural instanced city (64 base meshes on a jittered grid through yafaray_addInstance, %d instanced "
                "triangles + ground, flattened to world space), path tracer at %dx%d, %d bounces + NEE, one sun-like area light, gauss filter; "
                "one step = one AA pass of %g spp per GPU" % (args.city_tris, args.pt_width, args.pt_height, args.pt_bounces, spp_gpu))
    else:
        what = ("C2 (BASELINE configs[1]): procedural Cornell box built through the C-API mirror, path tracer 1024 spp at "
                "%dx%d, %d bounces, 1 area light with MIS NEE, box filter; one step = one AA pass of %g spp per GPU "
                "(1024 spp = %d steps)" % (args.pt_width, args.pt_height, args.pt_bounces, spp_gpu, max(1, int(1024 // max(spp_gpu, 1e-9)))))
    return {
        "workload": what,
        "width": args.pt_width, "height": args.pt_height, "spp_per_step_per_gpu": spp_gpu, "bounces": args.pt_bounces,
        "path_samples": 1, "russian_roulette": "off (min_bounces = bounces: the reference's deterministic mode)",
        "samples_per_step_per_gpu": args.pt_width * args.pt_height * spp_gpu,
        "scaling": args.scaling,
        "l2": "no flush needed: the path state of the wavefronts in flight (2 x 8 Mi pixel-samples x ~1 KB = 17 GB) is far larger than the 126 MB L2",
        "sharding": "AA-pass samples s mod N per GPU, scene + BVH replicated, one NCCL all-reduce of the pass film (5 floats/pixel) per step",
        "secondary": "`raybatch`: " + ("C5 (configs[4]) 50M-triangle soup, 2 x 16 Mi ray batches" if args.workload == "c5" else "C3 (configs[2]) 1M-triangle sphere, 3 x 16 Mi ray batches per GPU per step"),
    }


def config_dict(args, world=1):
    n = args.res * args.res
    n_gpu = n // world if args.scaling == "strong" else n
    if args.workload == "c5":
        return {
            "workload": "C5 (BASELINE configs[4]): %d random triangles (centres uniform in the unit cube, edges log-uniform 1e-3..1e-2); per step 2 ray batches: "
                        "incoherent closest-hit (uniform origins and directions) and any-hit segments between random point pairs" % args.soup_tris,
            "n_tri": args.soup_tris, "rays_per_batch": n, "rays_per_batch_per_gpu": n_gpu, "rays_per_step_per_gpu": 2 * n_gpu, "scaling": args.scaling,
            "l2": "no flush needed: the BVH8 (%.1f GB) and each batch (%d MiB in, %d MiB out) are far larger than the 126 MB L2" % (args.soup_tris * 72e-9, n_gpu * 32 >> 20, n_gpu * 16 >> 20),
            "sharding": "each GPU holds the whole scene + BVH8 and traces a contiguous slice of every batch, no data-path collective",
        }
    return {
        "workload": "C3 (BASELINE configs[2]): 1M-triangle displaced sphere; per GPU per step 3 ray batches "
                    "(coherent primary closest-hit, any-hit shadow, incoherent cosine-diffuse closest-hit)",
        "n_tri": args.nu * (args.nu // 2) * 2,
        "rays_per_batch": n, "rays_per_batch_per_gpu": n_gpu,
        "rays_per_step_per_gpu": 3 * n_gpu, "scaling": args.scaling,
        "l2": "no flush needed: each batch streams %d MiB of rays in and %d MiB of hits out (> 126 MB L2)" % (n_gpu * 32 >> 20, n_gpu * 16 >> 20),
        "sharding": "independent ray batches per GPU (scene + BVH replicated), no data-path collective",
    }


# ------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi in the background during the timed regions)
class ClockSampler:
    Q = "timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.rows, self.marks = [], []
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.p = None
            return
        self.t = threading.Thread(target=self._read, daemon=True)
        self.t.start()

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self, t0: float, t1: float):
        self.marks.append((t0, t1))

    def finish(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        inside = [r for (t, r) in self.rows if any(a - 0.05 <= t <= b + 0.15 for a, b in self.marks)]
        scope = "timed regions"
        if len(inside) < 3:
            inside, scope = [r for _, r in self.rows], "whole run (timed regions shorter than the sampling period)"
        sm, mx, pw, reasons = [], [], [], set()
        for r in inside:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "scope": scope, "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# synthetic ray batches, generated on the device with torch (set-up plumbing, never timed)
def camera_basis(rank: int):
    """Pinhole at distance 3 from the sphere centre; rank r looks from a yaw of 45*r degrees."""
    a = np.radians(45.0 * rank)
    eye = np.array([3.0 * np.sin(a), 0.0, 3.0 * np.cos(a)])
    fwd = -eye / np.linalg.norm(eye)
    right = np.cross(fwd, [0.0, 1.0, 0.0])
    right /= np.linalg.norm(right)
    up = np.cross(right, fwd)
    return eye, fwd, right, up


def gen_primary(torch, res: int, rank: int, device, stride: int = 1):
    eye, fwd, right, up = camera_basis(rank)
    half = np.tan(np.radians(43.6) * 0.5)
    idx = torch.arange(0, res, stride, device=device, dtype=torch.float64) + 0.5 * stride
    x = (idx / res * 2 - 1) * half
    y = (1 - idx / res * 2) * half
    f, r, u = (torch.tensor(v, device=device, dtype=torch.float64) for v in (fwd, right, up))
    d = f[None, None, :] + x[None, :, None] * r + y[:, None, None] * u
    d = d / d.norm(dim=-1, keepdim=True)
    n = d.shape[0] * d.shape[1]
    rays = torch.empty((n, 8), device=device, dtype=torch.float32)
    rays[:, 0:3] = torch.tensor(eye, device=device, dtype=torch.float32)
    rays[:, 3] = 0.0
    rays[:, 4:7] = d.reshape(-1, 3).float()
    rays[:, 7] = -1.0
    return rays


def gen_secondary(torch, tris_d, rays, hits, seed: int):
    """Shadow rays toward LIGHT and cosine-diffuse rays from the primary hit points (misses are
    recycled from the hit list), reference bias conventions (integrator_montecarlo.cc:105-114,
    integrator_path_tracer.cc:189)."""
    n = rays.shape[0]
    prim = hits[:, 3].view(torch.int32).long()
    ok = prim >= 0
    src_idx = torch.nonzero(ok).squeeze(1)
    if src_idx.numel() == 0:
        raise RuntimeError("no primary ray hit the scene")
    pick = src_idx[torch.arange(n, device=rays.device) % src_idx.numel()]
    o = rays[pick, 0:3].double() + rays[pick, 4:7].double() * hits[pick, 0:1].double()
    tr = tris_d[prim[pick]].double().reshape(-1, 3, 3)
    nn = torch.linalg.cross(tr[:, 1] - tr[:, 0], tr[:, 2] - tr[:, 0])
    nn = nn / nn.norm(dim=-1, keepdim=True).clamp_min(1e-300)
    flip = (nn * rays[pick, 4:7].double()).sum(-1) > 0
    nn = torch.where(flip[:, None], -nn, nn)
    del tr
    # shadow
    light = torch.tensor(LIGHT, device=rays.device, dtype=torch.float64)
    d = light[None, :] - o
    dist = d.norm(dim=-1)
    sh = torch.empty((n, 8), device=rays.device, dtype=torch.float32)
    sh[:, 0:3] = o.float()
    sh[:, 3] = (SHADOW_BIAS * o.norm(dim=-1).clamp_min(1.0)).float()
    sh[:, 4:7] = (d / dist[:, None]).float()
    sh[:, 7] = dist.float()
    del d, dist
    # diffuse
    g = torch.Generator(device=rays.device)
    g.manual_seed(seed)
    s1 = torch.rand(n, device=rays.device, dtype=torch.float64, generator=g)
    s2 = torch.rand(n, device=rays.device, dtype=torch.float64, generator=g)
    ax = torch.where(nn[:, 0:1].abs() > 0.5, torch.tensor([[0.0, 1.0, 0.0]], device=rays.device, dtype=torch.float64),
                     torch.tensor([[1.0, 0.0, 0.0]], device=rays.device, dtype=torch.float64))
    tu = torch.linalg.cross(nn, ax)
    tu = tu / tu.norm(dim=-1, keepdim=True)
    tv = torch.linalg.cross(nn, tu)
    dd = (tu * torch.cos(2 * np.pi * s2)[:, None] + tv * torch.sin(2 * np.pi * s2)[:, None]) * torch.sqrt(1 - s1)[:, None] + nn * torch.sqrt(s1)[:, None]
    dd = dd / dd.norm(dim=-1, keepdim=True)
    df = torch.empty((n, 8), device=rays.device, dtype=torch.float32)
    df[:, 0:3] = o.float()
    df[:, 3] = MIN_RAYDIST
    df[:, 4:7] = dd.float()
    df[:, 7] = -1.0
    return sh, df


# ------------------------------------------------------------------------------------------------
# CPU legs: the UNMODIFIED reference kd-tree (oracle/_ref, built by oracle/Makefile) on the host cores
def cpu_sample_batches_numpy(tris, res: int, server):
    """All-CPU construction of the sample batches (used by --impl reference): primary grid, then
    shadow / diffuse rays from the reference's own hits."""
    from libyafaray_b200 import scenes

    prim = scenes.primary_rays(res, res)
    h, _ = server.trace(prim, "closest")
    p, n, ok = scenes.hit_points_normals(tris, prim, h["t"], h["prim"])
    return {"primary": prim, "shadow": scenes.shadow_rays(p, ok, LIGHT), "diffuse": scenes.diffuse_rays(p, n, ok)}


def reference_raybatch(args):
    """C3 ray batches through the UNMODIFIED reference kd-tree on the host cores -> dict."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refserver
    from libyafaray_b200 import scenes

    line = {"impl": "reference", "metric": RB_METRIC, "unit": RB_UNIT, "steps": args.steps, "warmup": args.warmup, "config": config_dict(args)}
    if not refserver.have_ref():
        line["unavailable"] = "oracle/_ref/ref_raybatch is not built (run __graft_entry__.build() where /root/reference exists)"
        return line
    tris = scenes.displaced_sphere(args.nu, args.nu // 2)
    with refserver.RefServer(tris, threads=0) as srv:
        batches = cpu_sample_batches_numpy(tris, args.cpu_sample_res, srv)
        kinds = {"primary": "closest", "shadow": "shadow", "diffuse": "closest"}
        for _ in range(args.warmup):
            for b in BATCHES:
                srv.trace(batches[b], kinds[b])
        per = {b: 0.0 for b in BATCHES}
        for _ in range(args.steps):
            for b in BATCHES:
                per[b] += srv.trace(batches[b], kinds[b])[1]["trace_s"]
        n = batches["primary"].shape[0]
        total = sum(per.values())
        value = 3 * n * args.steps / total * 1e-6
        threads = srv.info["threads"]
        line.update({
            "value": value, "ms_per_step": total / args.steps * 1e3,
            "breakdown_mrays_s": {b: n * args.steps / per[b] * 1e-6 for b in BATCHES},
            "kd_build_s": srv.info["build_s"],
            "cpu_baseline": {"value": value, "unit": RB_UNIT, "cores": threads, "kind": "reference",
                             "sample": "%dx%d primary grid + shadow + diffuse batches (%d rays each) per step; yafaray-kdtree-original, build (%.1f s, serial) not timed"
                                       % (args.cpu_sample_res, args.cpu_sample_res, n, srv.info["build_s"])},
        })
    return line


# ------------------------------------------------------------------------------------------------
def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md; MEASURED_PEAKS.json absent)"


def issue_rate(env):
    """warp-instructions the GPU can issue per second: SMs x 4 schedulers x SM clock (the clock bench.py samples under load)."""
    torch = env["torch"]
    sms = torch.cuda.get_device_properties(env["dev"]).multi_processor_count
    mhz = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            mhz = float(json.load(f)["sm_max_mhz"])
    except Exception:
        pass
    mhz = mhz or 1965.0
    return sms * 4 * mhz * 1e6, "%d SMs x 4 schedulers x %.0f MHz (1 warp-instruction per scheduler and clock)" % (sms, mhz)


def load_profile_json(name, want):
    path = os.path.join(ROOT, "profiles", name)
    try:
        t = json.load(open(path))
        if t.get("workload") != want:
            return None, "profiles/%s is for '%s', this run is '%s'" % (name, t.get("workload"), want)
        return t, "profiles/%s (%s)" % (name, t.get("source"))
    except (OSError, KeyError, ValueError) as e:
        return None, "no capture: %s" % e


def raybatch_arm(args, env):
    """Ray batches (device-resident + end-to-end with host buffers) -> dict on rank 0, None elsewhere.
    workload c2 / c4: C3 (1M-triangle sphere: primary, shadow, diffuse); c5: the 50M-triangle soup (incoherent closest, any-hit segments)."""
    torch, dist = env["torch"], env["dist"]
    world, rank, local, dev, barrier, sampler = env["world"], env["rank"], env["local"], env["dev"], env["barrier"], env["sampler"]
    from libyafaray_b200 import scenes
    from libyafaray_b200.accelerator import AcceleratorB200

    soup = args.workload == "c5"
    BATCHES = ("incoherent", "segment") if soup else ("primary", "shadow", "diffuse")      # noqa: N806 (shadows the module constant on purpose)
    n_total = args.res * args.res
    n = n_total // world if args.scaling == "strong" else n_total      # rays per batch on THIS GPU
    first = rank * n                                                    # strong: a contiguous slice of the batch; weak: distinct rays per GPU
    if soup:
        tris = np.concatenate([scenes.triangle_soup(min(5_000_000, args.soup_tris - f), first=f) for f in range(0, args.soup_tris, 5_000_000)])
    else:
        tris = scenes.displaced_sphere(args.nu, args.nu // 2)
    acc = AcceleratorB200(tris, device=local)
    stream = torch.cuda.Stream(device=dev)
    acc.set_stream(stream.cuda_stream)
    st0 = acc.stats()
    tris_d = None if soup else torch.from_numpy(tris).to(dev)
    if soup:
        kinds = {"incoherent": acc.intersect, "segment": acc.isShadowed}
        chunks = lambda fn: np.concatenate([fn(min(1 << 22, n - f), first=first + f) for f in range(0, n, 1 << 22)])
        rays = {"incoherent": torch.from_numpy(chunks(scenes.incoherent_rays)).to(dev), "segment": torch.from_numpy(chunks(scenes.segment_rays)).to(dev)}
        hits = {b: torch.empty((n, 4), device=dev, dtype=torch.float32) for b in BATCHES}
    else:
        kinds = {"primary": acc.intersect, "shadow": acc.isShadowed, "diffuse": acc.intersect}
        # ---- set-up: batches resident in HBM ---------------------------------------------------
        full = gen_primary(torch, args.res, rank if args.scaling == "weak" else 0, dev)
        rays = {"primary": full[first:first + n].contiguous() if args.scaling == "strong" else full}
        del full
        hits = {b: torch.empty((n, 4), device=dev, dtype=torch.float32) for b in BATCHES}
        torch.cuda.synchronize()
        acc.intersect(rays["primary"], hits["primary"])
        acc.synchronize()
        rays["shadow"], rays["diffuse"] = gen_secondary(torch, tris_d, rays["primary"], hits["primary"], seed=0xD1F + rank)
    torch.cuda.synchronize()

    # visit counters (deterministic per batch) -> algorithmic bytes per ray, SURVEY.md 8(d)
    alg = {}
    for b in BATCHES:
        kinds[b](rays[b], hits[b], counted=True)
        s = acc.stats()
        npr, tpr = s["nodes_visited"] / n, s["tris_tested"] / n
        out_b = 16 if b != "shadow" else 16   # the shadow kernel also writes the 16 B blocker record
        alg[b] = {"nodes_per_ray": npr, "tris_per_ray": tpr, "bytes_per_ray": 32 + out_b + 80 * npr + 48 * tpr}

    def step():
        for b in BATCHES:
            kinds[b](rays[b], hits[b])

    for _ in range(args.warmup):
        step()
    barrier()
    # ---- timed region 1: device-resident throughput ------------------------------------------
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.steps)]
    launches0 = acc.stats()["kernel_launches"]
    barrier()
    w0 = time.time()
    for k in range(args.steps):
        ev[k][0].record(stream)
        for j, b in enumerate(BATCHES):
            kinds[b](rays[b], hits[b])
            ev[k][j + 1].record(stream)
    stream.synchronize()
    barrier()
    w1 = time.time()
    if sampler:
        sampler.mark(w0, w1)
    total_ms = ev[0][0].elapsed_time(ev[-1][len(BATCHES)])
    per_ms = {b: float(np.mean([ev[k][j].elapsed_time(ev[k][j + 1]) for k in range(args.steps)])) for j, b in enumerate(BATCHES)}
    launches = acc.stats()["kernel_launches"] - launches0
    hit_frac = {b: float((hits[b][:, 3].contiguous().view(torch.int32) >= 0).float().mean()) for b in BATCHES}

    # ---- timed region 2: end to end through the plug-in call with HOST buffers ------------------
    e2e = None
    if not args.no_e2e:
        h_rays = {b: torch.empty((n, 8), dtype=torch.float32, pin_memory=True) for b in BATCHES}
        h_hits = {b: torch.empty((n, 4), dtype=torch.float32, pin_memory=True) for b in BATCHES}
        for b in BATCHES:
            h_rays[b].copy_(rays[b])
        torch.cuda.synchronize()
        e_steps = max(1, min(args.steps, 10))
        for _ in range(min(args.warmup, 2)):
            for b in BATCHES:
                kinds[b](h_rays[b], h_hits[b])
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w2 = time.time()
        e0.record(stream)
        for _ in range(e_steps):
            for b in BATCHES:
                kinds[b](h_rays[b], h_hits[b])   # H2D of the rays, kernel, D2H of the hit records, sync
        e1.record(stream)
        stream.synchronize()
        barrier()
        w3 = time.time()
        if sampler:
            sampler.mark(w2, w3)
        e2e_ms = e0.elapsed_time(e1)
        # closest-hit batches: the host-buffer path must reproduce the resident run bit for bit; any-hit batches return whichever
        # blocker a warp meets first (schedule-dependent by design), so they are compared by occlusion
        same = {}
        for b in BATCHES:
            hh, hd = h_hits[b].view(torch.int32), hits[b].cpu().view(torch.int32)
            same[b] = ("same occlusion: %s" % bool(torch.equal(hh[:, 3] >= 0, hd[:, 3] >= 0))) if kinds[b] == acc.isShadowed else ("bit-identical: %s" % bool(torch.equal(hh, hd)))
        e2e = {"ms": e2e_ms, "steps": e_steps, "identical_to_resident_run": same}

    # ---- max over ranks ------------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e["ms"] if e2e else 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(t[0]), float(t[1])

    if rank == 0:
        peak, peak_src = load_peaks()
        dom = max(BATCHES, key=lambda b: per_ms[b])
        achieved = alg[dom]["bytes_per_ray"] * n / (per_ms[dom] * 1e-3) * 1e-9
        value = float(len(BATCHES)) * n * world * args.steps / (total_ms * 1e-3) * 1e-6
        want = ("soup %d tris %d rays" % (args.soup_tris, n_total)) if soup else ("sphere nu%d res%d" % (args.nu, args.res))
        prof, prof_src = load_profile_json("rb_classes_c5.json" if soup else "rb_classes.json", want)
        irate, irate_src = issue_rate(env)
        per_batch_roof = {}
        for b in BATCHES:
            r_hbm = peak * 1e9 / alg[b]["bytes_per_ray"] * 1e-6                      # Mrays/s if every algorithmic byte came from HBM at the measured peak
            ent = {"achieved_mrays_s": n / per_ms[b] * 1e-3, "r_hbm_mrays_s": r_hbm}
            if prof and b in prof["batches"]:
                pb = prof["batches"][b]
                wpr = pb["warp_inst"] / n_total                                       # warp-instructions per ray (ncu: smsp__inst_executed.sum)
                ent.update({"r_issue_mrays_s": irate / wpr * 1e-6, "warp_inst_per_ray": wpr, "active_lanes_per_inst": pb["active_lanes_per_inst"],
                            "dram_bytes_per_ray_measured": pb["dram_bytes"] / n_total})
                ent["binding"] = "issue" if ent["r_issue_mrays_s"] < r_hbm else "hbm"
                ent["frac_of_binding"] = ent["achieved_mrays_s"] / min(ent["r_issue_mrays_s"], r_hbm)
            per_batch_roof[b] = ent
        line = {
            "metric": RB_METRIC_C5 if soup else RB_METRIC, "value": value, "unit": RB_UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "scaling": args.scaling, "config": config_dict(args, world),
            "breakdown_mrays_s": {b: n / per_ms[b] * 1e-3 for b in BATCHES},
            "hit_fraction": hit_frac,
            "bvh": {k: acc.stats()[k] for k in ("n_tris", "n_nodes", "node_bytes", "tri_bytes", "build_ms")},
            "gpu_launches": int(launches),
            "roofline": {
                "bound": "hbm", "kernel": "k_trace<%s> (%s batch)" % ("SHADOW" if kinds[dom] == acc.isShadowed else "CLOSEST", dom),
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": (prof["batches"][dom]["dram_bytes"] if prof and dom in prof["batches"] else None), "traffic_source": prof_src,
                "peak_source": peak_src, "launch_ms": per_ms[dom],
                # north_star: the roofline is the slower of (algorithmic bytes per ray / HBM bandwidth) and the instruction-issue bound
                "min_of_hbm_and_issue": per_batch_roof, "issue_rate_source": irate_src,
                "algorithmic_bytes_per_ray": alg[dom]["bytes_per_ray"], "per_batch": alg,
                "note": "algorithmic bytes = 32 B ray + 16 B hit + 80 B x BVH8 nodes visited + 48 B x triangles tested (SURVEY.md 8(d)); "
                        "most node/triangle fetches are served by L2, DRAM traffic is far lower (see profiles/)",
            },
        }
        if e2e:
            line["e2e"] = {"value": float(len(BATCHES)) * n * world * e2e["steps"] / (e2e_ms * 1e-3) * 1e-6, "unit": RB_UNIT,
                           "h2d_bytes_per_step": len(BATCHES) * n * 32, "d2h_bytes_per_step": len(BATCHES) * n * 16, "steps": e2e["steps"],
                           "identical_to_resident_run": e2e["identical_to_resident_run"],
                           "api": "AcceleratorB200.intersect/isShadowed -> y4gpu_trace with pinned host ray/hit buffers"}
        if world == 1 and not args.no_cpu_baseline and not soup:
            line.update(cpu_baseline_leg(args, torch, acc, tris, tris_d, rays, hits, kinds))
        if soup:
            line["cpu_baseline"] = {"value": None, "unit": RB_UNIT, "cores": 0, "kind": "reference",
                                    "sample": "not timed in this run: yafaray-kdtree-original needs 593 s to build its tree on the 50 M triangles (tests/golden/make_golden_large.py); "
                                              "parity at full size is pinned by tests/golden/rays_soup50m.npz"}
        acc.close()
        return line
    acc.close()
    return None


def cpu_baseline_leg(args, torch, acc, tris, tris_d, rays, hits, kinds):
    """cpu_baseline: the unmodified reference kd-tree on the host cores, on a strided subsample of
    the very batches the GPU just traced; doubles as a full-size parity spot check."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refserver

    if not refserver.have_ref():
        return {"cpu_baseline": {"value": None, "unit": RB_UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}}
    stride = max(1, args.res // max(1, args.cpu_sample_res))
    g = torch.arange(stride // 2, args.res, stride, device=tris_d.device)
    idx = (g[:, None] * args.res + g[None, :]).reshape(-1)
    m = int(idx.numel())
    per, parity = {}, {}
    with refserver.RefServer(tris, threads=0) as srv:
        for b in BATCHES:
            r = rays[b][idx].cpu().numpy()
            got = hits[b][idx].cpu().numpy().view(refserver.HIT_DTYPE).reshape(-1)
            ref, info = srv.trace(r, "shadow" if b == "shadow" else "closest", repeats=2)
            per[b] = info["trace_s"]
            if b == "shadow":
                parity[b] = {"n": m, "occlusion_mismatch": int(((got["prim"] >= 0) != (ref["prim"] != -1)).sum())}
            else:
                same = got["prim"] == ref["prim"]
                both = (got["prim"] >= 0) & (ref["prim"] >= 0)
                dt = np.abs(got["t"].astype(np.float64) - ref["t"].astype(np.float64))
                near = ~same & both & (dt <= 1e-5 * np.abs(ref["t"]))
                ok = same & both
                parity[b] = {"n": m, "same_prim": int(same.sum()), "near_tie": int(near.sum()), "bad": int((~same & ~near).sum()),
                             "t_bit_exact": int((got["t"][ok] == ref["t"][ok]).sum()), "n_hit_both": int(ok.sum()),
                             "uv_bit_exact": int(((got["bu"][ok] == ref["bu"][ok]) & (got["bv"][ok] == ref["bv"][ok])).sum())}
        threads, build_s = srv.info["threads"], srv.info["build_s"]
    total = sum(per.values())
    return {
        "cpu_baseline": {"value": 3 * m / total * 1e-6, "unit": RB_UNIT, "cores": threads, "kind": "reference",
                         "breakdown_mrays_s": {b: m / per[b] * 1e-6 for b in BATCHES}, "kd_build_s": build_s,
                         "sample": "every %dth ray in x and y of each 16 Mi batch (%d rays per batch, mean of 2 repeats); "
                                   "yafaray-kdtree-original from oracle/_ref on all host threads, serial kd build not timed" % (stride, m)},
        "parity_sample_vs_reference": parity,
    }


# ------------------------------------------------------------------------------------------------
# path tracer (BASELINE configs[1])
def reference_pt(args):
    """Reference PathIntegrator + kd-tree through the reference's own C API on all host threads, bounded sample."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refserver
    from libyafaray_b200 import scenes_pt

    if not refserver.have_ref("ref_render"):
        return {"unavailable": "oracle/_ref/ref_render is not built"}
    rec = refserver.ScriptRecorder()
    scenes_pt.cornell_box(rec, width=args.cpu_pt_width, height=args.cpu_pt_height, spp=args.cpu_pt_spp, bounces=args.pt_bounces, threads=-1)
    n = args.cpu_pt_width * args.cpu_pt_height * args.cpu_pt_spp
    for _ in range(max(0, min(args.warmup, 1))):
        refserver.ref_render(rec)
    t = [refserver.ref_render(rec)[1]["render_s"] for _ in range(max(1, args.steps))]
    return {"value": n * len(t) / sum(t) * 1e-6, "ms_per_step": sum(t) / len(t) * 1e3, "cores": os.cpu_count(),
            "sample": "%dx%d at %d spp per step (%d pixel-samples), same scene / bounces / light as the GPU step; reference PathIntegrator + "
                      "yafaray-kdtree-original through its C API, threads = all (%d)" % (args.cpu_pt_width, args.cpu_pt_height, args.cpu_pt_spp, n, os.cpu_count())}


def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    # the configuration this arm REALLY runs: the same scene, film size, bounces and light as the GPU arm, fewer samples per pixel
    # per step (the metric is a rate per pixel-sample)
    cfg = pt_config_dict(args)
    cfg.update({"width": args.cpu_pt_width, "height": args.cpu_pt_height, "spp_per_step_per_gpu": args.cpu_pt_spp,
                "samples_per_step_per_gpu": args.cpu_pt_width * args.cpu_pt_height * args.cpu_pt_spp,
                "reference_arm": "unmodified libYafaRay PathIntegrator + yafaray-kdtree-original through its own C API on all host threads; "
                                 "%dx%d film, %d spp per step" % (args.cpu_pt_width, args.cpu_pt_height, args.cpu_pt_spp)})
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg}
    r = reference_pt(args)
    if "unavailable" in r:
        line["unavailable"] = r["unavailable"]
        print_json(line)
        return
    line.update({"value": r["value"], "ms_per_step": r["ms_per_step"],
                 "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "reference", "sample": r["sample"]},
                 "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})
    if not args.no_raybatch:
        line["raybatch"] = reference_raybatch(args)
    print_json(line)


# Algorithmic bytes moved by the shading stages: SoA path state read + written, 16-byte fields.
SHADE_BYTES = {
    # per pixel-sample: pix/samp written, camera ray written
    "gen": 8 + 32,
    # per pixel-sample: hit + ray + pix/samp read; col written; for camera hits also hit0, pcol, lsamp, flags, scontrib x2, 1 shadow ray, 2 queue entries
    "primary_miss": 16 + 32 + 8 + 16 + 4, "primary_hit": 16 + 32 + 8 + 16 + 16 + 16 + 16 + 4 + 32 + 32 + 8,
    # per camera hit: queue entry, flags, pix/samp, hit0, shadow hits, scontrib x2, col read; col, thr, flags, next ray, queue entry written
    "start": 4 + 4 + 8 + 16 + 8 + 32 + 16 + 16 + 16 + 4 + 32 + 4,
    # per path-stage entry: queue entry, flags, hit, ray, thr, pix/samp, lsamp + previous estimate (shadow hits, scontrib, pthr, pcol) read;
    # flags, pthr, thr, scontrib, 1 shadow ray, next ray, pcol, 2 queue entries written.  The second contribution slot and
    # lcol move only when non-zero (flag bits) and are not counted: this is the lower bound the stage can reach.
    "path": (4 + 4 + 16 + 32 + 16 + 8 + 16) + (8 + 16 + 16 + 16) + (4 + 16 + 16 + 16 + 32 + 32 + 16 + 8),
    # per pixel-sample: col, pcol, flags, pix/samp read; 5 float atomics to the film (box filter: one pixel)
    "splat": 16 + 16 + 4 + 8 + 5 * 4 * 2,
}


def pt_workload_key(args):
    return "%s %dx%d spp%d bounces%d" % ("city%d" % args.city_tris if args.workload == "c4" else "cornell", args.pt_width, args.pt_height, args.pt_spp, args.pt_bounces)


def measured_traffic(class_name, args):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of this kernel class, from the committed ncu capture of one
    pass of the same workload (tools/ncu_traffic.py -> profiles/pt_traffic.json); None when the workload differs."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "pt_traffic.json")
    want = pt_workload_key(args)
    if args.workload == "c4":
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "pt_traffic_c4.json")
    try:
        t = json.load(open(path))
        if t.get("workload") != want:
            return None, "profiles/pt_traffic.json is for '%s', this run is '%s'" % (t.get("workload"), want)
        return t["classes"][class_name]["dram_bytes_per_launch"], "profiles/pt_traffic.json (%s)" % t.get("source")
    except (OSError, KeyError, ValueError) as e:
        return None, "no capture: %s" % e


def pt_arm(args, env):
    torch, dist = env["torch"], env["dist"]
    world, rank, local, dev, barrier, sampler = env["world"], env["rank"], env["local"], env["dev"], env["barrier"], env["sampler"]
    from libyafaray_b200 import interface, scenes_pt
    from libyafaray_b200.accelerator import ACCELERATOR_TYPE

    W, H, spp, bounces = args.pt_width, args.pt_height, args.pt_spp, args.pt_bounces
    stream = torch.cuda.Stream(device=dev)
    # samples per pixel of one step over ALL GPUs: weak scaling keeps `spp` per GPU, strong scaling keeps the pass (spp in total)
    spp_step = spp if args.scaling == "strong" else spp * world
    if args.scaling == "strong" and spp % world:
        raise SystemExit("bench.py: --scaling strong needs --pt-spp divisible by the number of GPUs")

    def make(profile=0, comm=None):
        yi = interface.Interface(device=local, shard_index=rank, shard_count=world, comm=comm)
        if args.workload == "c4":
            scenes_pt.city(yi, n_tri_target=args.city_tris, width=W, height=H, spp=spp_step, bounces=bounces, accelerator=ACCELERATOR_TYPE, filter_type="gauss")
        else:
            scenes_pt.cornell_box(yi, width=W, height=H, spp=spp_step, bounces=bounces, accelerator=ACCELERATOR_TYPE)
        yi.render_params["b200_profile"] = profile
        if os.environ.get("Y4_BENCH_WAVE"):          # (tuning runs: pixel-samples per wavefront, default 2^23)
            yi.render_params["b200_wave_size"] = int(os.environ["Y4_BENCH_WAVE"])
        return yi

    # N > 1: the product's sharded mode (Interface with a comm): every rank splats into its pass film, ONE NCCL all-reduce of
    # [weights | rgba] per pass merges the ranks (the film merge is linear), the sum goes into the total film every rank keeps
    torch.cuda.set_stream(stream)
    yi = make(comm=dist if world > 1 else None)
    yi.upload()
    acc_stats = yi.accelerator.stats()
    yi.accelerator.set_stream(stream.cuda_stream)

    def step(k):
        yi.render_pass_reduced(spp_step, k * spp_step, 0)      # this GPU renders the samples s with s mod N == rank

    for k in range(args.warmup):
        step(k)
    stream.synchronize()
    barrier()
    launches0 = yi.accelerator.stats()["kernel_launches"]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    w0 = time.time()
    e0.record(stream)
    for k in range(args.steps):
        step(args.warmup + k)
    e1.record(stream)
    stream.synchronize()
    barrier()
    w1 = time.time()
    if sampler:
        sampler.mark(w0, w1)
    total_ms = e0.elapsed_time(e1)
    launches = yi.accelerator.stats()["kernel_launches"] - launches0      # this repo's kernels only (the NCCL all-reduce and the two film adds per pass are library kernels)
    st = yi.pass_stats()
    fw_, frgba_ = yi.read_film()
    wsum = float(fw_.sum(dtype=np.float64))
    mean_rgb = float((frgba_.reshape(-1, 4)[:, :3].sum(0, dtype=np.float64) / max(wsum, 1e-30)).mean())

    # ---- end to end: the whole render call with host-side scene + host film (upload, build, pass, D2H film) ----
    e2e = None
    if not args.no_e2e:
        e_steps = max(1, min(args.steps, 5))
        ye = make()
        ye.upload()
        ye.render_pass(spp_step, 0, 0)
        ye.read_film(pinned=True)
        barrier()
        asm = ye._assembled
        if asm.get("two_level"):      # meshes once + one matrix per instance
            h2d = int(sum(m.nbytes for m in asm["meshes"]) + sum(f.nbytes for f in asm["mesh_flags"]) + 56 * len(asm["instances"]) + 4096)
        else:
            h2d = int(asm["tris"].nbytes + asm["flags"].nbytes + 4096)
        t0 = time.perf_counter()
        for k in range(e_steps):
            ye.upload()                                      # host triangles/materials/lights/camera -> device, BVH8 build
            ye.render_pass(spp_step, k * spp_step, 0)
            fw, frgba = ye.read_film(pinned=True)            # device -> host film (weights + rgba), page-locked staging
        barrier()
        e2e = {"ms": (time.perf_counter() - t0) * 1e3, "steps": e_steps, "h2d": h2d, "d2h": int(fw.nbytes + frgba.nbytes)}
        ye.clearAll()

    # ---- one profiled pass: device time per kernel class + BVH visit counters (for the roofline) ----
    if args.no_profile:
        if rank == 0:
            print_json({"metric": METRIC, "value": W * H * spp_step * args.steps / (total_ms * 1e-3) * 1e-6, "unit": UNIT, "ms_per_step": total_ms / args.steps,
                        "note": "--no-profile run (for ncu): no roofline"})
        yi.clearAll()
        raise SystemExit(0)
    yp = make(profile=1)
    yp.upload()
    yp.render_pass(spp_step, 0, 0)
    prof = yp.pass_stats()
    yp.clearAll()

    t = torch.tensor([total_ms, e2e["ms"] if e2e else 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(t[0]), float(t[1])
    if rank != 0:
        yi.clearAll()
        return None
    n = W * H * spp_step // world                             # pixel-samples per GPU per step
    value = n * world * args.steps / (total_ms * 1e-3) * 1e-6
    peak, peak_src = load_peaks()
    names = ["camera rays", "trace_closest", "trace_shadow", "path stages (camera hit, start, per-bounce)", "extra light-sample stages", "splat"]
    cms = prof["class_ms"][:6]
    dom = int(np.argmax(cms))
    hits0 = prof["camera_hits"]
    alg_bytes = {
        0: n * SHADE_BYTES["gen"],
        1: prof["rays_closest"] * 48 + (prof["rays_closest"] - n) * 4 + prof["nodes_closest"] * 80 + prof["tris_closest"] * 48,
        2: prof["rays_shadow"] * 52 + prof["nodes_shadow"] * 80 + prof["tris_shadow"] * 48,
        3: (n - hits0) * SHADE_BYTES["primary_miss"] + hits0 * (SHADE_BYTES["primary_hit"] + SHADE_BYTES["start"]) + prof["path_entries"] * SHADE_BYTES["path"],
        4: 0,
        5: n * SHADE_BYTES["splat"],
    }
    achieved = alg_bytes[dom] / (cms[dom] * 1e-3) * 1e-9 if cms[dom] > 0 else 0.0
    traffic, traffic_src = measured_traffic(names[dom], args)
    # north_star: roofline = the slower of (bytes / HBM bandwidth) and the instruction-issue bound.  Warp-instructions per class
    # come from the committed ncu launch list of this workload (smsp__inst_executed.sum, tools/ncu_traffic.py -> profiles/pt_traffic.json)
    irate, irate_src = issue_rate(env)
    tj, _ = load_profile_json("pt_traffic_c4.json" if args.workload == "c4" else "pt_traffic.json", pt_workload_key(args))
    per_class, tot_hbm_ms, tot_issue_ms = {}, 0.0, 0.0
    for k in range(6):
        if not prof["class_launches"][k]:
            continue
        ent = {"ms": cms[k], "hbm_bound_ms": alg_bytes[k] / (peak * 1e9) * 1e3}
        c = tj["classes"].get(names[k]) if tj else None
        if c and c.get("warp_inst"):
            winst = c["warp_inst_per_launch"] * prof["class_launches"][k]
            ent.update({"issue_bound_ms": winst / irate * 1e3, "warp_inst_per_pass": winst, "active_lanes_per_inst": c.get("active_lanes_per_inst"),
                        "dram_bytes_per_pass_measured": c["dram_bytes_per_launch"] * prof["class_launches"][k]})
            ent["binding"] = "issue" if ent["issue_bound_ms"] > ent["hbm_bound_ms"] else "hbm"
            ent["frac_of_binding"] = max(ent["issue_bound_ms"], ent["hbm_bound_ms"]) / cms[k] if cms[k] > 0 else None
            tot_issue_ms += ent["issue_bound_ms"]
        tot_hbm_ms += ent["hbm_bound_ms"]
        per_class[names[k]] = ent
    pass_ms = total_ms / args.steps
    pass_roof = {"hbm_bound_ms": tot_hbm_ms, "issue_bound_ms": tot_issue_ms or None, "pass_ms_timed": pass_ms,
                 "r_hbm_msamples_s": n / tot_hbm_ms * 1e-3, "r_issue_msamples_s": (n / tot_issue_ms * 1e-3) if tot_issue_ms else None,
                 "achieved_msamples_s": n / pass_ms * 1e-3}
    if tot_issue_ms:
        pass_roof["binding"] = "issue" if tot_issue_ms > tot_hbm_ms else "hbm"
        pass_roof["frac_of_binding"] = max(tot_issue_ms, tot_hbm_ms) / pass_ms
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": pt_config_dict(args, world),
        "rays_per_sample": {"closest": st["rays_closest"] / st["samples"], "shadow": st["rays_shadow"] / st["samples"]},
        "mrays_s": (st["rays_closest"] + st["rays_shadow"]) / st["samples"] * value,
        "film_check": {"weight_sum": wsum, "mean_rgb": mean_rgb},
        "accelerator": dict({k: acc_stats[k] for k in ("n_tris", "n_nodes", "node_bytes", "tri_bytes", "build_ms", "bvh_levels")},
                            two_level=bool(yi._assembled.get("two_level")), instances=len(yi._assembled["instances"]), meshes=len(yi._assembled["meshes"]),
                            note="node_bytes + tri_bytes = the resident accelerator (BVH8 nodes + leaf records [+ instance records]); a flat scene also "
                                 "keeps 40 B per triangle of vertices and flag words, a two-level scene 40 B per BASE-mesh triangle"),
        "gpu_launches": int(launches),
        "roofline": {
            "bound": "hbm", "kernel": "%s kernels of one pass" % names[dom], "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": alg_bytes[dom] / max(1, prof["class_launches"][dom]),
            "launch_ms": cms[dom] / max(1, prof["class_launches"][dom]),
            "class_ms_per_pass": dict(zip(names, cms)), "class_launches_per_pass": dict(zip(names, prof["class_launches"][:6])),
            "class_algorithmic_GB": {names[k]: alg_bytes[k] * 1e-9 for k in range(6)},
            "min_of_hbm_and_issue": {"per_class": per_class, "whole_pass": pass_roof, "issue_rate_source": irate_src,
                                     "note": "hbm_bound_ms = algorithmic bytes / measured HBM peak; issue_bound_ms = warp-instructions (ncu) / issue rate; "
                                             "binding = the slower bound; frac_of_binding = bound / measured time (whole pass: against the timed, lane-overlapped pass)"},
            "pass_ms_profiled": sum(cms),
            "pass_ms_timed": total_ms / args.steps,
            "note": "class times come from one profiled pass whose kernels run one by one on one stream; the timed passes overlap "
                    "wavefronts on two streams, so pass_ms_timed < pass_ms_profiled. Algorithmic bytes: traversal = 48 B per ray "
                    "(32 B ray in + 16 B hit out, + 4 B queue entry) + 80 B per BVH8 node visited + 48 B per triangle tested; "
                    "path stages = SoA path state read + written per queue entry (DESIGN.md section 5)",
        },
    }
    if e2e:
        line["e2e"] = {"value": n * world * e2e["steps"] / (e2e_ms * 1e-3) * 1e-6, "unit": UNIT, "h2d_bytes_per_step": e2e["h2d"],
                       "d2h_bytes_per_step": e2e["d2h"], "steps": e2e["steps"], "timing": "host wall clock around the calls (max over ranks)",
                       "api": "Interface.upload() [host geometry + scene -> device, BVH8 build] + render_pass() + read_film() [film -> host]"}
    if world == 1 and not args.no_cpu_baseline:
        r = reference_pt(argparse.Namespace(**{**vars(args), "steps": 2, "warmup": 1}))
        line["cpu_baseline"] = ({"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "reference", "sample": r["sample"]}
                                if "value" in r else {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": r["unavailable"]})
    yi.clearAll()
    return line


def main():
    args = parse_args()
    # stdout carries exactly ONE JSON line: anything libraries print to fd 1 (e.g. NCCL's version banner) goes to stderr
    out_fd = os.dup(1)
    os.dup2(2, 1)
    global print_json

    def print_json(obj):
        os.write(out_fd, (json.dumps(obj) + "\n").encode())

    if args.impl == "reference":
        run_reference_arm(args)
        return
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- this back end has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    env = {"torch": torch, "dist": dist, "world": world, "rank": rank, "local": local, "dev": dev, "barrier": barrier,
           "sampler": ClockSampler(local) if rank == 0 else None}
    if args.workload == "c5" or args.only_raybatch:
        # the soup has no path-traced part: its ray batches are the line
        line = raybatch_arm(args, env)
        if rank == 0:
            line.update({"higher_is_better": True, "vs_baseline": None, "dtype": "f32", "data": "synthetic"})
        rb = None
    else:
        line = pt_arm(args, env)
        rb = None if args.no_raybatch else raybatch_arm(args, env)
    if rank == 0:
        if rb is not None:
            line["raybatch"] = rb
            line["gpu_launches"] += rb.get("gpu_launches", 0)
        line["clocks"] = env["sampler"].finish()
        print_json(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
