"""Knob sweep of the traversal kernels on the C3 batches (device-resident rays, CUDA-event kernel time).
usage: python tools/sweep.py [res] [nu] ["fetch,postpone,minb;..."]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from libyafaray_b200 import scenes  # noqa: E402
from libyafaray_b200.accelerator import AcceleratorB200  # noqa: E402


def main():
    res = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    nu = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    grid = sys.argv[3] if len(sys.argv) > 3 else "20,5,3;20,5,4;0,0,3;0,0,4;20,0,4;0,5,4;12,5,4;26,5,4;20,3,4;20,8,4"
    dev = torch.device("cuda", 0)
    tris = scenes.displaced_sphere(nu, nu // 2)
    tris_d = torch.from_numpy(tris).to(dev)
    n = res * res
    rays = {"primary": bench.gen_primary(torch, res, 0, dev)}
    hits = torch.empty((n, 4), device=dev, dtype=torch.float32)
    ref_hits = {}
    first = True
    for cfg in grid.split(";"):
        vals = [int(x) for x in cfg.split(",")]
        f, p, m = vals[:3]
        leaf = vals[3] if len(vals) > 3 else 0
        os.environ["Y4GPU_FETCH_THRESHOLD"], os.environ["Y4GPU_POSTPONE_DIV"], os.environ["Y4GPU_MIN_BLOCKS"] = str(f), str(p), str(m)
        acc = AcceleratorB200(tris, leaf_max=leaf)
        if first:
            acc.intersect(rays["primary"], hits)
            acc.synchronize()
            rays["shadow"], rays["diffuse"] = bench.gen_secondary(torch, tris_d, rays["primary"], hits, seed=0xD1F)
            first = False
        row = {"fetch": f, "postpone": p, "minb": m, "leaf": leaf, "n_nodes": acc.stats()["n_nodes"]}
        for b in bench.BATCHES:
            fn = acc.isShadowed if b == "shadow" else acc.intersect
            ms = []
            for _ in range(6):
                fn(rays[b], hits)
                acc.synchronize()
                ms.append(acc.stats()["last_trace_ms"])
            row[b] = round(n / min(ms[1:]) * 1e-3, 1)
            h = hits.clone()
            key = h[:, 3].contiguous().view(torch.int32)
            if b == "shadow":
                key = (key >= 0)
                cur = (key,)
            else:
                cur = (key, h[:, 0].clone())
            if b in ref_hits:
                row[b + "_same"] = all(bool(torch.equal(a, c)) for a, c in zip(cur, ref_hits[b]))
            else:
                ref_hits[b] = cur
        row["total"] = round(3 * n / sum(n / row[b] for b in bench.BATCHES), 1)
        print(json.dumps(row), flush=True)
        acc.close()


if __name__ == "__main__":
    main()
