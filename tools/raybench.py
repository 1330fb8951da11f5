"""Quick Mrays/s probe of the CUDA accelerator on the C3 scene (device-resident rays, CUDA-event kernel time)."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from libyafaray_b200 import scenes
from libyafaray_b200.accelerator import AcceleratorB200


def main():
    nu = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    res = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
    leaf_max = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    tris = scenes.displaced_sphere(nu, nu // 2)
    t0 = time.time()
    acc = AcceleratorB200(tris, leaf_max=leaf_max)
    st = acc.stats()
    print(json.dumps({"n_tri": len(tris), "n_nodes": st["n_nodes"], "build_ms": st["build_ms"], "wall_build_s": time.time() - t0}))
    prim = scenes.primary_rays(res, res)
    h = acc.intersect(prim)
    p, n, ok = scenes.hit_points_normals(tris, prim, h["t"], h["prim"])
    batches = {"primary": (prim, 0), "shadow": (scenes.shadow_rays(p, ok), 1), "diffuse": (scenes.diffuse_rays(p, n, ok), 0)}
    for name, (rays, kind) in batches.items():
        d_rays = torch.from_numpy(rays).cuda()
        d_out = torch.empty((len(rays), 4), dtype=torch.float32, device="cuda")
        fn = acc.isShadowed if kind else acc.intersect
        fn(d_rays, d_out, counted=True)
        s = acc.stats()
        npr, tpr = s["nodes_visited"] / len(rays), s["tris_tested"] / len(rays)
        ms = []
        for _ in range(5):
            fn(d_rays, d_out)
            acc.synchronize()
            ms.append(acc.stats()["last_trace_ms"])
        best = min(ms)
        bytes_per_ray = 32 + 16 + npr * 80 + tpr * 48
        print(json.dumps({"batch": name, "n": len(rays), "ms": best, "mrays_s": len(rays) / best * 1e-3, "nodes_per_ray": npr, "tris_per_ray": tpr,
                          "alg_bytes_per_ray": bytes_per_ray, "alg_GBs": bytes_per_ray * len(rays) / best * 1e-6, "hit_frac": float((d_out[:, 3].view(torch.int32) >= 0).float().mean())}))


if __name__ == "__main__":
    main()
