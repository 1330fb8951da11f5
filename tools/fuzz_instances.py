"""Differential fuzzing of the two-level (instanced) traversal against the flattened scene, both run by the emulation of the device code
(tests/emu), with the reference's kd-tree (oracle/_ref/ref_raybatch, this container only) as the arbiter where they differ.
Random meshes, 1 .. 30 instances with rotations, mirrors, shears, translations up to 500 and axis scales of the chosen class:
  uniform   one scale from 0.01 .. 100 for all axes          aniso5 / aniso20   per-axis scales within 5 : 1 / 20 : 1
  extreme   per-axis scales drawn from {0.01, 0.3, 1, 5, 100} (needles, 100 : 1 and more)
usage: python tools/fuzz_instances.py <first seed> <count> [uniform|aniso5|aniso20|extreme]      (results: profiles/r2_v16_fuzz_vs_reference.log)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers
bad_total = 0
for seed in range(int(sys.argv[1]), int(sys.argv[1]) + int(sys.argv[2])):
    rng = np.random.default_rng(seed)
    n_mesh = int(rng.integers(1, 5))
    meshes = []
    for _ in range(n_mesh):
        nt = int(rng.choice([1, 5, 40, 300]))
        c = rng.uniform(-1, 1, (nt, 1, 3))
        meshes.append((c + rng.uniform(-1, 1, (nt, 3, 3)) * rng.choice([0.05, 0.3])).astype(np.float32).reshape(nt, 9))
    inst = []
    for _ in range(int(rng.integers(1, 30))):
        mesh = int(rng.integers(0, n_mesh))
        if rng.random() < 0.15: inst.append((mesh, None)); continue
        A = rng.normal(size=(3, 3))
        q, _ = np.linalg.qr(A)
        MODE = sys.argv[3] if len(sys.argv) > 3 else "extreme"
        if MODE == "extreme": sc = np.diag(rng.choice([0.01, 0.3, 1.0, 5.0, 100.0], 3) * rng.choice([1, 1, 1, -1], 3))
        elif MODE == "uniform": sc = np.eye(3) * rng.choice([0.01, 0.3, 1.0, 5.0, 100.0]) * rng.choice([1, -1])
        elif MODE == "aniso5": sc = np.diag(rng.uniform(0.2, 1.0, 3) * rng.choice([0.1, 1.0, 10.0]) * rng.choice([1, 1, 1, -1], 3))
        elif MODE == "aniso20": sc = np.diag(rng.uniform(0.05, 1.0, 3) * rng.choice([0.1, 1.0, 10.0]) * rng.choice([1, 1, 1, -1], 3))
        M = q @ sc
        if rng.random() < 0.2: M = M @ np.array([[1, 0.7, 0], [0, 1, 0.4], [0, 0, 1.0]])               # shear
        t = rng.uniform(-5, 5, 3) * rng.choice([1.0, 100.0])
        inst.append((mesh, np.concatenate([M, t[:, None]], 1).reshape(12).tolist()))
    flat = helpers.flatten_instances(meshes, inst)
    nr = 3000
    tgt = flat.reshape(-1, 3, 3)[rng.integers(0, len(flat), nr)].mean(1)
    ext = np.abs(flat).max()
    o = tgt + rng.normal(size=(nr, 3)) * rng.choice([0.5, 5.0, 50.0]) * max(1.0, ext * 0.05)
    d = tgt - o; d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((nr, 8), np.float32); rays[:, 0:3] = o; rays[:, 3] = 0.0; rays[:, 4:7] = d; rays[:, 7] = -1
    for kind in (0, 1):
        a = helpers.emu_build_trace(flat, rays, kind); a = a[0] if isinstance(a, tuple) else a
        b = helpers.emu_build_trace2(meshes, inst, rays, kind)[0]
        if kind == 0:
            same = (a["prim"] == b["prim"]) & (a["t"].view(np.uint32) == b["t"].view(np.uint32)) & (a["bu"].view(np.uint32) == b["bu"].view(np.uint32)) & (a["bv"].view(np.uint32) == b["bv"].view(np.uint32))
            # exact ties in t between different primitives may resolve differently: not a difference in the surface reached
            tie = (~same) & (a["t"] == b["t"]) & (a["prim"] != b["prim"])
            # the documented near-tie class (tests/helpers.compare_closest): two different primitives within 1e-5 of each other in t
            near = (~same) & (a["prim"] >= 0) & (b["prim"] >= 0) & (a["prim"] != b["prim"]) & (np.abs(a["t"] - b["t"]) <= 1e-5 * np.maximum(a["t"], b["t"]))
            n_near = int((near & ~tie).sum())
            if n_near: print("   ", seed, "closest:", n_near, "near-ties (two primitives within 1e-5 in t: the reported t of needle-like triangles is only that accurate)")
            tie = tie | near
            bad = int((~same & ~tie).sum())
            if bad:
                idx = np.nonzero(~same & ~tie)[0]
                ref, _ = helpers.ref_raybatch(flat, rays[idx], "closest")
                two_ok = int(((ref["prim"] == b["prim"][idx]) & (ref["t"] == b["t"][idx])).sum()); flat_ok = int(((ref["prim"] == a["prim"][idx]) & (ref["t"] == a["t"][idx])).sum())
                print("   ", seed, "closest: reference agrees with two-level on", two_ok, "and with flat on", flat_ok, "of", bad)
        else:
            bad = int(((a["prim"] != -1) != (b["prim"] != -1)).sum())
            if bad:
                idx = np.nonzero((a["prim"] != -1) != (b["prim"] != -1))[0]
                ref, _ = helpers.ref_raybatch(flat, rays[idx], "shadow")
                print("   ", seed, "shadow: reference agrees with two-level on", int(((ref["prim"] != -1) == (b["prim"][idx] != -1)).sum()), "and with flat on", int(((ref["prim"] != -1) == (a["prim"][idx] != -1)).sum()), "of", bad)
        if bad: print(seed, "kind", kind, "bad", bad, "n_inst", len(inst))
        bad_total += bad
print("bad total", bad_total)
