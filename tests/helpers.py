"""Test helpers: run the reference oracle binaries (oracle/_ref, built from /root/reference by
oracle/Makefile; shipped to the GPU box as built files) and the unit-test emulation."""
import ctypes
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
HIT_DTYPE = np.dtype([("t", "f4"), ("bu", "f4"), ("bv", "f4"), ("prim", "i4")])


def have_ref(name="ref_raybatch"):
    return os.path.exists(os.path.join(REF_DIR, name))


def ref_raybatch(tris, rays, kind, threads=0, accel="yafaray-kdtree-original"):
    """Run the UNMODIFIED reference kd-tree on a ray batch -> (hits, info json)."""
    with tempfile.TemporaryDirectory() as d:
        tp, rp, hp = (os.path.join(d, n) for n in ("t.f32", "r.f32", "h.bin"))
        np.ascontiguousarray(tris, np.float32).tofile(tp)
        np.ascontiguousarray(rays, np.float32).tofile(rp)
        cmd = [os.path.join(REF_DIR, "ref_raybatch"), tp, rp, kind, hp]
        if threads:
            cmd += [str(threads), accel]
        out = subprocess.run(cmd, capture_output=True, text=True, check=True)
        return np.fromfile(hp, HIT_DTYPE), json.loads(out.stdout.strip().splitlines()[-1])


_emu = None


def emu():
    global _emu
    if _emu is None:
        d = os.path.join(ROOT, "tests", "emu")
        so = os.path.join(d, "libemu.so")
        subprocess.run(["make", "-s", "-C", d], check=True)
        _emu = ctypes.CDLL(so)
    return _emu


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def emu_build_trace(tris, rays, kind, leaf_max=1, flags=None):
    tris = np.ascontiguousarray(tris, np.float32)
    rays = np.ascontiguousarray(rays, np.float32)
    hits = np.zeros(len(rays), HIT_DTYPE)
    stats = np.zeros(2, np.uint64)
    n_nodes = emu().emu_build_trace(_p(tris), _p(flags), len(tris), _p(rays), len(rays), kind, leaf_max, _p(hits), _p(stats))
    return hits, n_nodes, stats


def flatten_instances(meshes, instances):
    """World-space triangles of an instanced scene in primitive order, with the reference's Matrix4 * Point3 arithmetic
    (fp32, ((m0 x + m1 y) + m2 z) + m3; matrix4.h:85-90) -- what PrimitiveInstance applies per test.
    meshes: list of [n, 9] fp32 arrays; instances: list of (mesh index, 3x4 matrix or None for a plain object)."""
    out = []
    for mesh, mat in instances:
        v = np.ascontiguousarray(meshes[mesh], np.float32).reshape(-1, 3)
        if mat is not None:
            m = np.asarray(mat, np.float32).reshape(3, 4)
            v = np.stack([((m[r, 0] * v[:, 0] + m[r, 1] * v[:, 1]) + m[r, 2] * v[:, 2]) + m[r, 3] for r in range(3)], -1).astype(np.float32)
        out.append(v.reshape(-1, 9))
    return np.ascontiguousarray(np.concatenate(out), np.float32)


def mini_city(n_mesh=4, n_inst=40, tris_per_mesh=600, seed=5):
    """A small instanced scene for the two-level tests: `n_mesh` building meshes (scenes_pt._building) placed `n_inst` times with
    random yaw / scale / position like scenes_pt.city, plus a plain ground quad.  -> (meshes, instances)."""
    from libyafaray_b200 import scenes_pt
    rng = np.random.default_rng(seed)
    meshes = []
    for b in range(n_mesh):
        v, t = scenes_pt._building(0xC4, b, tris_per_mesh)
        meshes.append(np.asarray(v, np.float64).astype(np.float32)[np.asarray(t, np.int64)].reshape(-1, 9))
    meshes.append(np.asarray([[-2, 0, -2, -2, 0, 14, 14, 0, 14], [-2, 0, -2, 14, 0, 14, 14, 0, -2]], np.float32))
    inst = [(n_mesh, None)]
    for _ in range(n_inst):
        yaw, sc = rng.uniform(0, 2 * np.pi), rng.uniform(0.55, 0.9)
        cx, cz = rng.uniform(0, 12, 2)
        c, s = np.cos(yaw) * sc, np.sin(yaw) * sc
        inst.append((int(rng.integers(0, n_mesh)), [c, 0, s, cx, 0, sc, 0, 0, -s, 0, c, cz]))
    return meshes, inst


def city_rays(n, seed=6, origin_scale=1.0):
    """Rays aimed at points inside the mini city from origins `origin_scale` away (far origins stress the object-space padding)."""
    rng = np.random.default_rng(seed)
    tgt = np.stack([rng.uniform(0, 12, n), rng.uniform(0, 3, n), rng.uniform(0, 12, n)], -1)
    o = tgt + rng.normal(size=(n, 3)) * origin_scale
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3], rays[:, 3], rays[:, 4:7], rays[:, 7] = o, 1e-4, d, -1
    return rays


def emu_build_trace2(meshes, instances, rays, kind, mesh_flags=None):
    """Two-level emulation (tests/emu/emu_bvh.cpp: emu_build_trace2): bottom-level trees per mesh, top level over instances."""
    mv = np.ascontiguousarray(np.concatenate([np.asarray(m, np.float32).reshape(-1, 9) for m in meshes]), np.float32)
    mt = np.asarray([len(np.asarray(m).reshape(-1, 9)) for m in meshes], np.int32)
    mf = None if mesh_flags is None else np.ascontiguousarray(np.concatenate(mesh_flags), np.uint32)
    im = np.asarray([i[0] for i in instances], np.int32)
    ii = np.asarray([1 if i[1] is None else 0 for i in instances], np.int32)
    ident = np.asarray([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
    mat = np.ascontiguousarray(np.stack([ident if i[1] is None else np.asarray(i[1], np.float32).reshape(12) for i in instances]), np.float32)
    rays = np.ascontiguousarray(rays, np.float32)
    hits = np.zeros(len(rays), HIT_DTYPE)
    stats = np.zeros(2, np.uint64)
    n_nodes = emu().emu_build_trace2(_p(mv), _p(mf), _p(mt), len(mt), _p(im), _p(ii), _p(mat), len(im), _p(rays), len(rays), kind, _p(hits), _p(stats))
    return hits, n_nodes, stats


def emu_brute(tris, rays, kind, flags=None):
    tris = np.ascontiguousarray(tris, np.float32)
    rays = np.ascontiguousarray(rays, np.float32)
    hits = np.zeros(len(rays), HIT_DTYPE)
    emu().emu_brute(_p(tris), _p(flags), len(tris), _p(rays), len(rays), kind, _p(hits))
    return hits


def compare_closest(got, ref, rel=1e-5):
    """Parity rule of BASELINE.json: identical prim except documented near-ties (|dt| <= rel*t);
    t and barycentrics within rel where the prim agrees.  Returns a dict of counts."""
    same = got["prim"] == ref["prim"]
    both = (got["prim"] >= 0) & (ref["prim"] >= 0)
    dt = np.abs(got["t"].astype(np.float64) - ref["t"].astype(np.float64))
    near_tie = ~same & both & (dt <= rel * np.abs(ref["t"]))
    bad = ~same & ~near_tie
    ok = same & both
    tol_t = dt[ok] <= rel * np.abs(ref["t"][ok])
    tol_u = np.abs(got["bu"][ok] - ref["bu"][ok]) <= rel
    tol_v = np.abs(got["bv"][ok] - ref["bv"][ok]) <= rel
    return {
        "n": len(got), "same": int(same.sum()), "near_tie": int(near_tie.sum()), "bad": int(bad.sum()),
        "t_out_of_tol": int((~tol_t).sum()), "uv_out_of_tol": int((~tol_u).sum() + (~tol_v).sum()),
        "bit_exact_t": int((got["t"][ok] == ref["t"][ok]).sum()), "n_hit_both": int(ok.sum()),
    }


sys.path.insert(0, os.path.join(ROOT, "oracle"))
from refserver import RefServer  # noqa: E402,F401

from refserver import ScriptRecorder, ref_render  # noqa: E402,F401


def emu_film_flags(a, w, rgba, aa):
    """CPU emulation of y4gpu_pt_next_pass: -> flagged pixel indices (raster order)."""
    import ctypes as C

    pr = a["params"]
    flags = np.zeros(pr.height * pr.width, np.uint8)
    f = emu().emu_film_flags
    f.restype = C.c_uint64
    n = f(_p(np.ascontiguousarray(w, np.float32)), _p(np.ascontiguousarray(rgba, np.float32)), C.c_int(pr.width), C.c_int(pr.height), C.byref(aa), _p(flags))
    idx = np.nonzero(flags)[0].astype(np.uint32)
    assert len(idx) == n
    return idx


def emu_pt_render(a, n_samples=None, offset=0, aa_pass=0, shard_index=0, shard_count=1, pixel_list=None, film=None, ref_light_order=False):
    """CPU emulation of the wavefront path tracer stages on an Interface.assemble() result -> (weights, rgba).
    pixel_list: adaptive pass over these pixels only; film: (weights, rgba) to accumulate into.
    ref_light_order: pick the one-light-per-vertex lights with the reference's single-thread call counter (threads=1,
    linear tiles of 32) instead of the product's per-vertex index -- sample-exact comparisons of multi-light scenes."""
    import ctypes as C

    if ref_light_order:
        assert pixel_list is None and film is None and shard_count == 1
        emu().emu_light_choice_mode(1)
        emu_pt_render(a, n_samples, offset, aa_pass)
        emu().emu_light_choice_finalize(C.c_int(a["params"].width), C.c_int(32))
        emu().emu_light_choice_mode(2)
        try:
            return emu_pt_render(a, n_samples, offset, aa_pass)
        finally:
            emu().emu_light_choice_mode(0)

    pr = a["params"]
    n_samples = pr.aa_samples if n_samples is None else n_samples
    w = np.zeros((pr.height, pr.width), np.float32) if film is None else film[0]
    rgba = np.zeros((pr.height, pr.width, 4), np.float32) if film is None else film[1]
    pl = None if pixel_list is None else np.ascontiguousarray(pixel_list, np.uint32)
    st = emu().emu_pt_set_textures
    st.restype = C.c_int
    optr = a["tri_orco"].ctypes.data_as(C.c_void_p) if a.get("tri_orco") is not None else None
    uptr = a["tri_uv"].ctypes.data_as(C.c_void_p) if a.get("tri_uv") is not None else None
    assert st(a["textures"], C.c_uint32(a["n_tex"]), a["layers"], C.c_uint32(a["n_layers"]), optr, uptr, C.c_uint64(len(a["tris"]))) == 0
    f = emu().emu_pt_render
    f.restype = C.c_int
    nptr = a["normals"].ctypes.data_as(C.c_void_p) if a["normals"] is not None else None
    rc = f(C.byref(a["cam"]), a["mats"], C.c_uint32(a["n_mat"]), a["lights"], C.c_uint32(a["n_light"]), C.byref(pr), _p(a["tris"]), _p(a["flags"]),
           C.c_int(len(a["tris"])), nptr, C.c_int(n_samples), C.c_int(offset), C.c_int(aa_pass), C.c_int(shard_index), C.c_int(shard_count), _p(w), _p(rgba),
           _p(pl) if pl is not None else None, C.c_uint64(0 if pl is None else len(pl)))
    assert rc == 0
    return w, rgba


def psnr(img, ref, peak=None):
    """PSNR in dB over RGB, peak = max of the reference unless given."""
    a, b = np.asarray(img, np.float64)[..., :3], np.asarray(ref, np.float64)[..., :3]
    mse = np.mean((a - b) ** 2)
    peak = float(b.max()) if peak is None else peak
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)


def with_render_params(rec, **params):
    """Re-issue a recorded scene's render parameters with extra / overriding keys (before its final setupRender)."""
    i = max(k for k, l in enumerate(rec.lines) if l == "setupRender")
    rec.lines.pop(i)
    for k, v in params.items():
        rec.lines = [l for l in rec.lines if not (l.split()[:2] == ["paramsSetInt", k] or l.split()[:2] == ["paramsSetFloat", k])]
        if isinstance(v, bool):
            rec.lines.append(f"paramsSetBool {k} {1 if v else 0}")
        elif isinstance(v, int):
            rec.lines.append(f"paramsSetInt {k} {v}")
        elif isinstance(v, float):
            rec.lines.append(f"paramsSetFloat {k} {v!r}")
        else:
            rec.lines.append(f"paramsSetString {k} {v}")
    rec.lines.append("setupRender")
    return rec


def emu_adaptive_render(yi):
    """TiledIntegrator::render's pass loop (the one Interface.render runs on the GPU) over the CPU emulation:
    -> (image, resampled pixels per adaptive pass)."""
    from libyafaray_b200 import interface

    a = yi.assemble()
    pr, rp = a["params"], yi.render_params
    f32 = np.float32
    film = emu_pt_render(a, n_samples=pr.aa_samples, offset=0, aa_pass=0)
    threshold = f32(rp.get("AA_threshold", 0.05))
    floor_pixels = int(np.floor(f32(f32(rp.get("AA_resampled_floor", 0.0)) * f32(pr.width * pr.height)) / f32(100.0)))
    acum, mult, resampled, changed, hist = pr.aa_samples, f32(1.0), 0, True, []
    for i in range(1, pr.aa_passes):
        mult = f32(mult * f32(pr.aa_sample_multiplier_factor))
        plist, skip = None, False
        if resampled <= 0 and not changed:
            skip = True
        elif threshold > 0:
            plist = emu_film_flags(a, film[0], film[1], yi.aa_noise_params(threshold))
            resampled, changed = len(plist), False
        else:
            resampled, changed = pr.width * pr.height, False
        hist.append(0 if skip else resampled)
        n = int(np.ceil(f32(pr.aa_inc_samples) * mult))
        if not skip and resampled > 0:
            emu_pt_render(a, n_samples=n, offset=acum, aa_pass=i, pixel_list=plist, film=film)
        acum += n
        if resampled < floor_pixels:
            ratio = min(f32(8.0), f32(floor_pixels) / f32(resampled)) if resampled > 0 else f32(8.0)
            threshold = f32(threshold * f32(f32(1.0) - f32(0.1) * ratio))
            if threshold > 0:
                changed = True
    return interface.normalize_film(film[0], film[1]), hist


# ---- the compiled libYafaRay plug-ins (integration/: AcceleratorB200 + PathIntegratorB200 inside a patched copy of the reference) ----
B200_DIR = os.path.join(ROOT, "integration", "_build")


def have_plugin(name="b200_render"):
    return os.path.exists(os.path.join(B200_DIR, name))


def b200_render(recorder, gpu_integrator=True, cwd=None, env=None):
    """Replays a recorded C-API script through the REFERENCE'S OWN C API of the patched libYafaRay build.  The script must set
    scene_accelerator = yafaray-b200-bvh8; gpu_integrator selects PathIntegratorB200 (YAFARAY_B200_BACKEND=1) instead of the
    reference's CPU PathIntegrator over the GPU accelerator's single-ray virtuals.  -> (float32 [H, W, 4], info)."""
    with tempfile.TemporaryDirectory() as d:
        sp, op = os.path.join(d, "scene.script"), os.path.join(d, "out.f32")
        with open(sp, "w") as f:
            f.write("\n".join(recorder.lines) + f"\nrender {op}\n")
        e = dict(os.environ)
        e["YAFARAY_B200_BACKEND"] = "1" if gpu_integrator else "0"
        e.update(env or {})
        out = subprocess.run([os.path.join(B200_DIR, "b200_render"), sp], capture_output=True, text=True, env=e, cwd=cwd)
        if out.returncode != 0:
            raise RuntimeError(f"b200_render failed rc={out.returncode}: {out.stdout[-2000:]} {out.stderr[-2000:]}")
        info = json.loads(out.stdout.strip().splitlines()[-1])
        img = np.fromfile(op, np.float32).reshape(info["height"], info["width"], 4)
        return img, info


def b200_raybatch(tris, rays, kind, threads=8, accel="yafaray-b200-bvh8"):
    """Accelerator::factory(type) + intersect / isShadowed called from `threads` host threads of the patched libYafaRay build."""
    with tempfile.TemporaryDirectory() as d:
        tp, rp, hp = (os.path.join(d, n) for n in ("t.f32", "r.f32", "h.bin"))
        np.ascontiguousarray(tris, np.float32).tofile(tp)
        np.ascontiguousarray(rays, np.float32).tofile(rp)
        out = subprocess.run([os.path.join(B200_DIR, "b200_raybatch"), tp, rp, kind, hp, str(threads), accel], capture_output=True, text=True)
        if out.returncode != 0:
            raise RuntimeError(f"b200_raybatch failed rc={out.returncode}: {out.stdout[-2000:]} {out.stderr[-2000:]}")
        return np.fromfile(hp, HIT_DTYPE), json.loads(out.stdout.strip().splitlines()[-1])
