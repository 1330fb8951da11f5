"""Differential fuzzing of the traversal (builder + BVH8 encoder + closest / any-hit traversal: the Y4_HD source the kernels compile, run by
tests/emu) against the reference's kd-tree (oracle/_ref/ref_raybatch; this container only).  Random triangle sets from 1 to 4000
triangles at scales 1e-3 .. 1e3 -- degenerate, coplanar / axis-aligned and single-plane cases included -- and rays aimed at them, random,
or along the axes, with random [tmin, tmax): closest hits must agree (t, u, v bit-exact or inside the documented near-tie class),
shadow rays must agree on occlusion.     usage: python tools/fuzz_rays_vs_reference.py <first seed> <count>"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers
bad_total = 0
for seed in range(int(sys.argv[1]), int(sys.argv[1]) + int(sys.argv[2])):
    rng = np.random.default_rng(seed)
    nt = int(rng.choice([1, 2, 7, 60, 500, 4000]))
    scale = float(rng.choice([1e-3, 1.0, 1e3]))
    c = rng.uniform(-1, 1, (nt, 1, 3)) * scale
    size = float(rng.choice([0.01, 0.2, 1.0])) * scale
    tris = (c + rng.uniform(-1, 1, (nt, 3, 3)) * size).astype(np.float32)
    k = rng.integers(0, 5)
    if k == 0 and nt > 4: tris[: nt // 4, 2] = tris[: nt // 4, 1]                      # degenerate (zero area)
    if k == 1: tris[:, :, 2] = np.round(tris[:, :, 2] / (0.25 * scale)) * (0.25 * scale)   # many coplanar / axis-aligned
    if k == 2: tris[:, :, 0] = tris[:, :1, 0]                                            # all in planes x = const
    tris = tris.reshape(nt, 9)
    nr = 4000
    o = (rng.uniform(-2, 2, (nr, 3)) * scale).astype(np.float32)
    tgt = tris.reshape(nt, 3, 3)[rng.integers(0, nt, nr)].mean(1) + rng.normal(0, 0.05 * size, (nr, 3))
    d = tgt - o
    m = rng.random(nr) < 0.2
    d[m] = rng.normal(0, 1, (int(m.sum()), 3))
    ax = rng.random(nr) < 0.1
    d[ax] = np.eye(3)[rng.integers(0, 3, int(ax.sum()))] * rng.choice([-1, 1], (int(ax.sum()), 1))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    tmin = np.where(rng.random(nr) < 0.5, 0.0, rng.uniform(0, 0.5 * scale, nr))
    tmax = np.where(rng.random(nr) < 0.5, -1.0, rng.uniform(0.1 * scale, 4 * scale, nr))
    rays = np.concatenate([o, tmin[:, None], d, tmax[:, None]], 1).astype(np.float32)
    for kind_name, kind in (("closest", 0), ("shadow", 1)):
        ref, _ = helpers.ref_raybatch(tris, rays, kind_name)
        got = helpers.emu_build_trace(tris, rays, kind)
        got = got[0] if isinstance(got, tuple) else got
        if kind == 0:
            r = helpers.compare_closest(got, ref)
            bad = r["bad"] + r["t_out_of_tol"] + r["uv_out_of_tol"]
            info = r
        else:
            bad = int(((got["prim"] != -1) != (ref["prim"] != -1)).sum()); info = {"occlusion mismatches": bad}
        if bad: print(seed, kind_name, "nt", nt, "scale", scale, "case", int(k), info)
        bad_total += bad
print("bad total", bad_total)
