"""CPU-side checks (no GPU): the host emulation of the device logic (tests/emu, same Y4_HD source
the kernels compile) and the brute-force restatement of the reference triangle test against
(a) the committed golden hit records produced by the UNMODIFIED reference kd-tree and
(b) the live reference binary when oracle/_ref is built."""
import os

import numpy as np
import pytest

import helpers
from libyafaray_b200 import scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _ok_closest(got, ref):
    r = helpers.compare_closest(got, ref)
    assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0, r
    return r


@pytest.mark.parametrize("leaf_max", [1, 2, 3])
def test_sphere_golden(leaf_max):
    d = np.load(os.path.join(GOLD, "rays_sphere.npz"))
    tris = scenes.displaced_sphere(*d["scene"])
    for name in ("primary", "diffuse"):
        got, n_nodes, _ = helpers.emu_build_trace(tris, d[name], 0, leaf_max)
        r = _ok_closest(got, d["hit_" + name])
        # the triangle test is restated operation for operation: t and (u, v) are bit-exact
        assert r["bit_exact_t"] == r["n_hit_both"]
    got, _, _ = helpers.emu_build_trace(tris, d["shadow"], 1, leaf_max)
    assert np.array_equal(got["prim"] >= 0, d["hit_shadow"]["prim"] != -1)


def test_soup_golden():
    d = np.load(os.path.join(GOLD, "rays_soup.npz"))
    tris = scenes.triangle_soup(int(d["scene"][0]))
    got, _, _ = helpers.emu_build_trace(tris, d["incoherent"], 0)
    _ok_closest(got, d["hit_incoherent"])
    got, _, _ = helpers.emu_build_trace(tris, d["segment"], 1)
    assert np.array_equal(got["prim"] >= 0, d["hit_segment"]["prim"] != -1)
    # brute force over all triangles in primitive order == the reference's acceptance rules
    got = helpers.emu_brute(tris, d["incoherent"], 0)
    _ok_closest(got, d["hit_incoherent"])


def test_golden_bound():
    d = np.load(os.path.join(GOLD, "rays_sphere.npz"))
    tris = scenes.displaced_sphere(*d["scene"])
    lo, hi = tris.reshape(-1, 3).min(0), tris.reshape(-1, 3).max(0)
    foo = (hi - lo).astype(np.float64) * 0.001
    exp = np.concatenate([(lo.astype(np.float64) - foo).astype(np.float32), (hi.astype(np.float64) + foo).astype(np.float32)])
    assert np.allclose(exp, d["bound"], rtol=0, atol=1e-6)


@pytest.mark.skipif(not helpers.have_ref(), reason="oracle/_ref not built")
def test_live_reference_sphere_and_flags():
    tris = scenes.displaced_sphere(120, 60)
    prim = scenes.primary_rays(96, 96)
    with helpers.RefServer(tris) as srv:
        ref, _ = srv.trace(prim, "closest")
        p, n, ok = scenes.hit_points_normals(tris, prim, ref["t"], ref["prim"])
        sh, df = scenes.shadow_rays(p, ok), scenes.diffuse_rays(p, n, ok)
        ref_s, _ = srv.trace(sh, "shadow")
        ref_d, _ = srv.trace(df, "closest")
    got, _, _ = helpers.emu_build_trace(tris, prim, 0)
    _ok_closest(got, ref)
    got, _, _ = helpers.emu_build_trace(tris, df, 0)
    _ok_closest(got, ref_d)
    got, _, _ = helpers.emu_build_trace(tris, sh, 1)
    assert np.array_equal(got["prim"] >= 0, ref_s["prim"] != -1)


def test_edge_cases_emulation():
    rays = scenes.primary_rays(16, 16)
    got, n_nodes, _ = helpers.emu_build_trace(np.zeros((0, 9), np.float32), rays, 0)
    assert (got["prim"] == -1).all()
    one = np.array([[-1, -1, 0, 1, -1, 0, 0, 1, 0]], np.float32)
    for n in (1, 2, 3, 8, 9, 65):
        tris = np.repeat(one, n, 0) + np.arange(n, dtype=np.float32)[:, None] * np.array([0, 0, -0.1] * 3, np.float32)
        got, _, _ = helpers.emu_build_trace(tris, rays, 0)
        exp = helpers.emu_brute(tris, rays, 0)
        assert np.array_equal(got["prim"], exp["prim"]) and np.array_equal(got["t"], exp["t"])
    # degenerate (zero-area) and duplicated triangles, axis-parallel rays with zero direction components
    tris = np.concatenate([one, one, np.zeros((3, 9), np.float32), one + np.float32(2)])
    ax = np.zeros((6, 8), np.float32)
    ax[:, 0:3] = [0.0, -0.5, 5.0]
    ax[:, 7] = -1
    ax[:, 4:7] = [[0, 0, -1], [0, 0, 1], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0, -1, 0]]
    got, _, _ = helpers.emu_build_trace(tris, ax, 0)
    exp = helpers.emu_brute(tris, ax, 0)
    assert np.array_equal(got["t"], exp["t"]) and got["prim"][0] in (0, 1) and (got["prim"][1:] == -1).all()


def test_visibility_flags_and_ranges():
    """Visibility matrix of accelerator_kdtree.cc:733-737,869-873 and the wrapper ray ranges."""
    one = np.array([[-1, -1, 0, 1, -1, 0, 0, 1, 0]], np.float32)
    tris = np.concatenate([one, one - np.array([0, 0, 1] * 3, np.float32)])
    ray = np.array([[0, -0.5, 5, 0, 0, 0, -1, -1]], np.float32)
    flags = np.array([2, 1 | 2 | 4], np.uint32)   # triangle 0 = shadows only, triangle 1 = normal
    got = helpers.emu_brute(tris, ray, 0, flags)
    assert got["prim"][0] == 1 and got["t"][0] == 6.0
    got, _, _ = helpers.emu_build_trace(tris, ray, 0, 1, flags)
    assert got["prim"][0] == 1
    got, _, _ = helpers.emu_build_trace(tris, ray, 1, 1, flags)
    assert got["prim"][0] in (0, 1)
    # closest: tmin is inclusive, tmax exclusive (accelerator_kdtree.cc:731)
    r2 = np.array([[0, -0.5, 5, 5.0, 0, 0, -1, 6.0], [0, -0.5, 5, 5.5, 0, 0, -1, 6.0], [0, -0.5, 5, 6.0, 0, 0, -1, -1]], np.float32)
    got = helpers.emu_brute(tris, r2, 0)
    assert list(got["prim"]) == [0, -1, 1]
    # shadow: origin advanced by tmin, t_max = tmax - 2 tmin, lower bound 0 (accelerator.cc:69-80)
    r3 = np.array([[0, -0.5, 5, 1.0, 0, 0, -1, 5.5], [0, -0.5, 5, 1.0, 0, 0, -1, 6.5]], np.float32)
    got = helpers.emu_brute(tris, r3, 1)
    assert list(got["prim"] >= 0) == [False, True]


def test_two_level_traversal_equals_the_flattened_scene_bit_for_bit():
    """base meshes kept once in object space + instances (top-level BVH8 over instance boxes, bottom-level BVH8 per mesh,
    object-space node tests padded by the transform's rounding, exact WORLD-space triangle test with the reference's per-test
    vertex transform).  Every hit record equals the one the flattened scene gives, also for rays that start far away."""
    meshes, inst = helpers.mini_city()
    flat = helpers.flatten_instances(meshes, inst)
    for scale, n in ((1.0, 20000), (300.0, 8000)):
        rays = helpers.city_rays(n, origin_scale=scale)
        h1, n1, s1 = helpers.emu_build_trace(flat, rays, 0, leaf_max=3)
        h2, n2, s2 = helpers.emu_build_trace2(meshes, inst, rays, 0)
        assert np.array_equal(h1.view(np.uint32), h2.view(np.uint32)), int((h1["prim"] != h2["prim"]).sum())
        assert (h1["prim"] >= 0).mean() > 0.5
        assert n2 * 5 < n1                       # the instanced structure is a fraction of the flattened tree
        o1, _, _ = helpers.emu_build_trace(flat, rays, 1, leaf_max=3)
        o2, _, _ = helpers.emu_build_trace2(meshes, inst, rays, 1)
        assert np.array_equal(o1["prim"] >= 0, o2["prim"] >= 0)
    # transparent shadows (intersectTs): two of the meshes are transparent; same verdict and the same number of filtering layers
    from libyafaray_b200 import _lib
    mesh_flags = [np.full(len(m), _lib.TRI_VIS_CLOSEST | _lib.TRI_VIS_SHADOW | _lib.TRI_VIS_TSHADOW | (_lib.TRI_TRANSPARENT if i in (0, 2) else 0) | (i << _lib.TRI_MAT_SHIFT), np.uint32)
                  for i, m in enumerate(meshes)]
    flat_flags = np.concatenate([mesh_flags[mi] for mi, _ in inst])
    rays = helpers.city_rays(20000, seed=9)
    t1, _, _ = helpers.emu_build_trace(flat, rays, 2, leaf_max=3, flags=flat_flags)
    t2, _, _ = helpers.emu_build_trace2(meshes, inst, rays, 2, mesh_flags=mesh_flags)
    lit = t1["prim"] < 0           # (a shadowed ray stops at its opaque blocker: how many layers it met before depends on the walk order)
    assert np.array_equal(t1["prim"] >= 0, t2["prim"] >= 0) and np.array_equal(t1["bu"][lit], t2["bu"][lit])
    assert (t1["bu"][lit] > 0).mean() > 0.02 and (t1["bu"][lit] > 1).any()
