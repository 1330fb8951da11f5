"""Ray-sort sweep on one GPU: resident batches of C5 (50 M-triangle soup: incoherent closest-hit, any-hit segments) and C3
(1 M-triangle sphere: primary, shadow, diffuse) traced unsorted and sorted with several key layouts.  Prints one JSON line per
measurement: Mrays/s from the context's own events (probe + key kernel + radix sort + traversal all inside the span), the
probe's coherence figure, and that the sorted hits are bit-identical to the unsorted ones.
usage: python tools/sweep_raysort.py [c5_tris] [n_rays]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from libyafaray_b200 import scenes
from libyafaray_b200.accelerator import AcceleratorB200


def chunked(fn, n, chunk=4_000_000, width=9):
    out = np.empty((n, width), np.float32)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        out[lo:hi] = fn(lo, hi)
    return out


def measure(acc, fn, d_rays, d_out, reps=3):
    ms = []
    for _ in range(reps):
        fn(d_rays, d_out)
        acc.synchronize()
        ms.append(acc.stats()["last_trace_ms"])
    return min(ms)


def sweep(tag, acc, batches, settings):
    ref = {}
    for name, rays, kind in batches:
        d_rays = torch.from_numpy(rays).cuda()
        d_out = torch.empty((len(rays), 4), dtype=torch.float32, device="cuda")
        fn = acc.isShadowed if kind else acc.intersect
        for setting in settings:
            mode, obits, dir_major = setting[:3]
            os.environ["Y4GPU_RAY_SORT_OBITS"] = str(obits)
            os.environ["Y4GPU_RAY_SORT_DIR_MAJOR"] = str(dir_major)
            acc.set_ray_sort(mode)
            ms = measure(acc, fn, d_rays, d_out)
            st = acc.stats()
            hits = d_out.clone()
            if mode == 0:
                ref[name] = hits
                same = True
            elif kind:
                same = bool(torch.equal(hits[:, 3].view(torch.int32) >= 0, ref[name][:, 3].view(torch.int32) >= 0))     # any-hit: same occlusion
            else:
                same = bool(torch.equal(hits.view(torch.int32), ref[name].view(torch.int32)))
            print(json.dumps({"scene": tag, "batch": name, "mode": mode, "obits": obits, "dir_major": dir_major, "ms": round(ms, 3),
                              "mrays_s": round(len(rays) / ms * 1e-3, 1), "sorted": st["last_trace_sorted"], "probe_distinct": round(st["last_probe_distinct"], 2),
                              "same_as_unsorted": same}), flush=True)
        del d_rays, d_out


def main():
    a = [int(x) for x in sys.argv[1:]]
    n_tri, n_rays = (a + [50_000_000, 1 << 24][len(a):])[:2]
    settings = [(0, 6, 0), (1, 4, 0), (1, 5, 0), (1, 6, 0), (1, 7, 0), (1, 8, 0), (1, 5, 1), (1, 6, 1), (1, 7, 1), (2, 6, 0)]
    # C3 first (small): the bench's own generators (camera rays, then shadow / cosine-diffuse rays from the primary hits)
    from bench import gen_primary, gen_secondary
    tris = scenes.displaced_sphere(1000, 500)
    acc = AcceleratorB200(tris)
    res = int(round(n_rays ** 0.5))
    prim = gen_primary(torch, res, 0, torch.device("cuda:0"))
    hits = torch.empty((len(prim), 4), device="cuda", dtype=torch.float32)
    acc.intersect(prim, hits)
    acc.synchronize()
    sh, di = gen_secondary(torch, torch.from_numpy(tris).cuda(), prim, hits, seed=0xD1F)
    batches = [("primary", prim.cpu().numpy(), 0), ("shadow", sh.cpu().numpy(), 1), ("diffuse", di.cpu().numpy(), 0)]
    del prim, hits, sh, di
    os.environ["Y4GPU_RAY_SORT_MIN_BVH_MB"] = "0"
    sweep("c3", acc, batches, settings)
    acc.close()
    os.environ.pop("Y4GPU_RAY_SORT_MIN_BVH_MB")
    tris = chunked(lambda lo, hi: scenes.triangle_soup(hi - lo, first=lo), n_tri)
    acc = AcceleratorB200(tris)
    del tris
    inc = chunked(lambda lo, hi: scenes.incoherent_rays(hi - lo, first=lo), n_rays, width=8)
    seg = chunked(lambda lo, hi: scenes.segment_rays(hi - lo, first=lo), n_rays, width=8)
    sweep("c5", acc, [("incoherent", inc, 0), ("segment", seg, 1)], settings)


if __name__ == "__main__":
    main()
