"""C5 (BASELINE configs[4]) at full size on one GPU: 50 M random triangles, 16 Mi incoherent closest-hit rays and
16 Mi any-hit segment rays.  Reports build time, Mrays/s, BVH visits per ray, and checks
  * closest-hit / any-hit consistency on every ray (a segment is occluded iff the closest-hit query over the same range
    finds a hit) -- a size-independent property;
  * a brute-force spot check (tests/emu exact triangle test over ALL triangles) on a few rays.
usage: python tools/c5bench.py [n_tri] [n_rays] [n_spot]"""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from libyafaray_b200 import scenes
from libyafaray_b200.accelerator import AcceleratorB200


def chunked(fn, n, chunk=4_000_000, width=9):
    out = np.empty((n, width), np.float32)
    # the generators hash the element index, so chunks are generated with an index offset through their `k` argument
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        out[lo:hi] = fn(lo, hi)
    return out


def main():
    a = [int(x) for x in sys.argv[1:]]
    n_tri, n_rays, n_spot = (a + [50_000_000, 1 << 24, 16][len(a):])[:3]
    t0 = time.time()
    tris = chunked(lambda lo, hi: scenes.triangle_soup(hi - lo, first=lo), n_tri)
    t_gen = time.time() - t0
    t0 = time.time()
    acc = AcceleratorB200(tris)
    st = acc.stats()
    out = {"config": "C5", "n_tri": n_tri, "gen_s": t_gen, "build_ms": st["build_ms"], "upload_build_wall_s": time.time() - t0, "n_nodes": st["n_nodes"],
           "bvh_GB": (st["n_nodes"] * 80 + n_tri * 48) * 1e-9}
    print(json.dumps(out), flush=True)
    inc = chunked(lambda lo, hi: scenes.incoherent_rays(hi - lo, first=lo), n_rays, width=8)
    seg = chunked(lambda lo, hi: scenes.segment_rays(hi - lo, first=lo), n_rays, width=8)
    res = {}
    for name, rays, kind in (("incoherent_closest", inc, 0), ("segment_anyhit", seg, 1)):
        d_rays = torch.from_numpy(rays).cuda()
        d_out = torch.empty((len(rays), 4), dtype=torch.float32, device="cuda")
        fn = acc.isShadowed if kind else acc.intersect
        fn(d_rays, d_out, counted=True)
        s = acc.stats()
        npr, tpr = s["nodes_visited"] / len(rays), s["tris_tested"] / len(rays)
        ms = []
        for _ in range(3):
            fn(d_rays, d_out)
            acc.synchronize()
            ms.append(acc.stats()["last_trace_ms"])
        best = min(ms)
        prim = d_out[:, 3].view(torch.int32)
        res[name] = {"n": len(rays), "ms": best, "mrays_s": len(rays) / best * 1e-3, "nodes_per_ray": npr, "tris_per_ray": tpr,
                     "alg_bytes_per_ray": 48 + npr * 80 + tpr * 48, "hit_frac": float((prim >= 0).float().mean())}
        print(json.dumps({name: res[name]}), flush=True)
        if kind == 1:
            # same segments through the closest-hit query with the wrapper's shadow range: origin advanced by tmin, t in [0, tmax - 2 tmin)
            adv = rays.copy()
            adv[:, 0:3] += rays[:, 4:7] * rays[:, 3:4]
            adv[:, 7] = rays[:, 7] - 2 * rays[:, 3]
            adv[:, 3] = 0.0
            c_out = torch.empty_like(d_out)
            d_adv = torch.from_numpy(adv).cuda()
            torch.cuda.synchronize()
            acc.intersect(d_adv, c_out)
            acc.synchronize()          # device-resident batches run asynchronously on the accelerator's own stream
            occl_any = (prim >= 0).cpu().numpy()
            occl_closest = (c_out[:, 3].view(torch.int32) >= 0).cpu().numpy()
            # the advanced origin is rounded once more in the closest query: allow disagreement only for hits within 1e-5 of the range ends
            dis = np.nonzero(occl_any != occl_closest)[0]
            res["anyhit_vs_closest_disagree"] = int(len(dis))
            print(json.dumps({"anyhit_vs_closest_disagree": int(len(dis)), "of": len(rays), "any_only": int((occl_any & ~occl_closest).sum()),
                              "closest_only": int((~occl_any & occl_closest).sum())}), flush=True)
            if len(dis) and n_spot:
                import helpers

                pick = dis[:6]
                b_any = helpers.emu_brute(tris, rays[pick], 1)
                b_cl = helpers.emu_brute(tris, adv[pick], 0)
                g_any = d_out.cpu().numpy()[pick]
                g_cl = c_out.cpu().numpy()[pick]
                for j in range(len(pick)):
                    print("ray", rays[pick[j]].tolist(), "gpu_any", g_any[j, 0], int(g_any[j, 3].view(np.int32)), "gpu_closest", g_cl[j, 0], int(g_cl[j, 3].view(np.int32)),
                          "brute_any", float(b_any["t"][j]), int(b_any["prim"][j]), "brute_closest", float(b_cl["t"][j]), int(b_cl["prim"][j]), flush=True)
        if n_spot:
            import helpers

            pick = np.linspace(0, len(rays) - 1, n_spot).astype(np.int64)
            got = d_out.cpu().numpy()[pick]
            exp = helpers.emu_brute(tris, rays[pick], kind)
            if kind == 0:
                ok = (got[:, 3].view(np.int32) == exp["prim"]) & (got[:, 0] == exp["t"])
            else:
                ok = (got[:, 3].view(np.int32) >= 0) == (exp["prim"] >= 0)
            res[name]["spot_check"] = {"n": int(n_spot), "agree": int(ok.sum())}
            print(json.dumps({name + "_spot_check": res[name]["spot_check"]}), flush=True)
    out.update(res)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
