"""GPU parity: CUDA BVH8 accelerator (through the C-ABI) vs the reference kd-tree."""
import numpy as np
import pytest

import helpers
from libyafaray_b200 import scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sphere():
    from libyafaray_b200.accelerator import AcceleratorB200
    tris = scenes.displaced_sphere(200, 100)
    return tris, AcceleratorB200(tris)


def _batches(tris, ref_fn, n=256):
    prim = scenes.primary_rays(n, n)
    h = ref_fn(tris, prim, "closest")[0]
    p, nrm, ok = scenes.hit_points_normals(tris, prim, h["t"], h["prim"])
    return prim, h, scenes.shadow_rays(p, ok), scenes.diffuse_rays(p, nrm, ok)


@pytest.mark.skipif(not helpers.have_ref(), reason="oracle/_ref not built")
def test_sphere_vs_reference_kdtree(sphere):
    tris, acc = sphere
    prim, h_ref, sh, df = _batches(tris, helpers.ref_raybatch)
    got = acc.intersect(prim)
    r = helpers.compare_closest(got, h_ref)
    assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0, r
    got_d = acc.intersect(df)
    ref_d = helpers.ref_raybatch(tris, df, "closest")[0]
    r = helpers.compare_closest(got_d, ref_d)
    assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0, r
    got_s = acc.isShadowed(sh)
    ref_s = helpers.ref_raybatch(tris, sh, "shadow")[0]
    assert np.array_equal(got_s["prim"] >= 0, ref_s["prim"] != -1)


def test_sphere_vs_emulation_bit_exact(sphere):
    tris, acc = sphere
    prim = scenes.primary_rays(128, 128)
    got = acc.intersect(prim)
    exp, _, _ = helpers.emu_build_trace(tris, prim, 0)
    assert np.array_equal(got["prim"], exp["prim"])
    assert np.array_equal(got["t"], exp["t"]) and np.array_equal(got["bu"], exp["bu"]) and np.array_equal(got["bv"], exp["bv"])


def test_bound_matches_reference_inflation(sphere):
    tris, acc = sphere
    lo, hi = tris.reshape(-1, 3).min(0), tris.reshape(-1, 3).max(0)
    foo = (hi - lo).astype(np.float64) * 0.001
    exp = np.concatenate([(lo.astype(np.float64) - foo).astype(np.float32), (hi.astype(np.float64) + foo).astype(np.float32)])
    assert np.array_equal(acc.getBound(), exp)


def test_edge_cases():
    from libyafaray_b200.accelerator import AcceleratorB200
    rays = scenes.primary_rays(16, 16)
    empty = AcceleratorB200(np.zeros((0, 9), np.float32))
    assert (empty.intersect(rays)["prim"] == -1).all()
    one = np.array([[-1, -1, 0, 1, -1, 0, 0, 1, 0]], np.float32)
    for n in (1, 2, 3, 7, 9):
        tris = np.repeat(one, n, 0) + np.arange(n, dtype=np.float32)[:, None] * np.array([0, 0, -0.1] * 3, np.float32)
        acc = AcceleratorB200(tris)
        got = acc.intersect(rays)
        exp = helpers.emu_brute(tris, rays, 0)
        assert np.array_equal(got["prim"], exp["prim"]) and np.array_equal(got["t"], exp["t"])
        assert (acc.intersect(rays[:0])["prim"].shape[0]) == 0


def test_soup_golden_on_gpu():
    """Committed hit records of the reference kd-tree on a random triangle soup (incoherent closest + segment any-hit)."""
    import os
    d = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rays_soup.npz"))
    from libyafaray_b200.accelerator import AcceleratorB200
    tris = scenes.triangle_soup(int(d["scene"][0]))
    acc = AcceleratorB200(tris)
    r = helpers.compare_closest(acc.intersect(d["incoherent"]), d["hit_incoherent"])
    assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0 and r["bit_exact_t"] == r["n_hit_both"], r
    assert np.array_equal(acc.isShadowed(d["segment"])["prim"] >= 0, d["hit_segment"]["prim"] != -1)


def test_degenerate_inputs_and_ray_ranges():
    from libyafaray_b200.accelerator import AcceleratorB200
    one = np.array([[-1, -1, 0, 1, -1, 0, 0, 1, 0]], np.float32)
    # zero-area and duplicated triangles, axis-parallel rays (zero direction components)
    tris = np.concatenate([one, one, np.zeros((3, 9), np.float32), one + np.float32(2)])
    ax = np.zeros((6, 8), np.float32)
    ax[:, 0:3] = [0.0, -0.5, 5.0]
    ax[:, 7] = -1
    ax[:, 4:7] = [[0, 0, -1], [0, 0, 1], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0, -1, 0]]
    acc = AcceleratorB200(tris)
    got, exp = acc.intersect(ax), helpers.emu_brute(tris, ax, 0)
    assert np.array_equal(got["t"], exp["t"]) and got["prim"][0] == 0 and (got["prim"][1:] == -1).all()   # exact tie: lowest primitive index
    # visibility matrix + wrapper ranges (closest: tmin inclusive, tmax exclusive; shadow: advanced origin, t_max = tmax - 2 tmin)
    tris = np.concatenate([one, one - np.array([0, 0, 1] * 3, np.float32)])
    flags = np.array([2, 1 | 2 | 4], np.uint32)
    acc = AcceleratorB200(tris, flags=flags)
    ray = np.array([[0, -0.5, 5, 0, 0, 0, -1, -1]], np.float32)
    assert acc.intersect(ray)["prim"][0] == 1 and acc.intersect(ray)["t"][0] == 6.0
    acc = AcceleratorB200(tris)
    r2 = np.array([[0, -0.5, 5, 5.0, 0, 0, -1, 6.0], [0, -0.5, 5, 5.5, 0, 0, -1, 6.0], [0, -0.5, 5, 6.0, 0, 0, -1, -1], [0, -0.5, 5, 7.0, 0, 0, -1, 6.5]], np.float32)
    assert list(acc.intersect(r2)["prim"]) == [0, -1, 1, -1]
    r3 = np.array([[0, -0.5, 5, 1.0, 0, 0, -1, 5.5], [0, -0.5, 5, 1.0, 0, 0, -1, 6.5]], np.float32)
    assert list(acc.isShadowed(r3)["prim"] >= 0) == [False, True]
    # rays that can hit nothing must come back as misses, not hang: NaN / infinite origins, zero direction
    bad = np.array([[np.nan, 0, 5, 0, 0, 0, -1, -1], [np.inf, 0, 5, 0, 0, 0, -1, -1], [0, -0.5, 5, 0, 0, 0, 0, -1], [0, -0.5, 5, 0, np.nan, 0, -1, -1]], np.float32)
    assert (acc.intersect(bad)["prim"] == -1).all() and (acc.isShadowed(bad)["prim"] == -1).all()


def test_transparent_shadows_and_device_resident_batches(sphere):
    import torch
    tris, acc = sphere
    prim = scenes.primary_rays(96, 96)
    h = acc.intersect(prim)
    p, nrm, ok = scenes.hit_points_normals(tris, prim, h["t"], h["prim"])
    sh = scenes.shadow_rays(p, ok)
    # all-opaque materials: intersectTs reports "shadowed" exactly where it finds a blocker in [tmin, t_max) with material visibility
    ts = acc.isShadowedTransparent(sh, 4, np.ones((1, 3), np.float32))
    adv = sh.copy()
    adv[:, 0:3] += sh[:, 4:7] * sh[:, 3:4]
    adv[:, 7] = sh[:, 7] - 2 * sh[:, 3]
    exp = acc.intersect(adv)                                    # closest hit of the advanced ray with the same [tmin, t_max) interval
    assert np.array_equal(ts["prim"] >= 0, exp["prim"] >= 0)
    # device tensors in, device tensors out, on the caller's stream
    s = torch.cuda.Stream()
    acc.set_stream(s.cuda_stream)
    with torch.cuda.stream(s):
        d_rays = torch.from_numpy(prim).cuda()
        d_out = torch.empty((len(prim), 4), dtype=torch.float32, device="cuda")
        acc.intersect(d_rays, d_out)
    s.synchronize()
    acc.set_stream(0)
    got = d_out.cpu().numpy()
    assert np.array_equal(got[:, 0], h["t"]) and np.array_equal(got[:, 3].view(np.int32), h["prim"])


def test_sorted_ray_order_is_bit_identical_and_the_automatic_mode_leaves_camera_rays_alone(sphere, monkeypatch):
    """y4gpu_set_ray_sort: tracing a batch in key order (origin cell + direction class, through an index queue) changes only the
    order in which warps pick rays up -- every hit record equals the unsorted run's, resident and through host buffers."""
    import torch
    tris, acc = sphere
    n = 512
    prim = scenes.primary_rays(n, n)
    h = acc.intersect(prim)
    p, nrm, ok = scenes.hit_points_normals(tris, prim, h["t"], h["prim"])
    rng = np.random.default_rng(7)
    batches = {"diffuse": (scenes.diffuse_rays(p, nrm, ok)[rng.permutation(n * n)], acc.intersect),     # shuffled: as incoherent as it gets
               "shadow": (scenes.shadow_rays(p, ok)[rng.permutation(n * n)], acc.isShadowed), "primary": (prim, acc.intersect)}
    monkeypatch.setenv("Y4GPU_RAY_SORT_MIN_BVH_MB", "0")       # the automatic mode would skip a BVH that fits L2
    for name, (rays, fn) in batches.items():
        acc.set_ray_sort(0)
        d_rays = torch.from_numpy(rays).cuda()
        ref = torch.empty((len(rays), 4), dtype=torch.float32, device="cuda")
        fn(d_rays, ref)
        acc.synchronize()
        assert acc.stats()["last_trace_sorted"] == 0
        for mode in (1, 2):
            acc.set_ray_sort(mode)
            out = torch.zeros((len(rays), 4), dtype=torch.float32, device="cuda")
            fn(d_rays, out)
            acc.synchronize()
            st = acc.stats()
            if mode == 1 or name != "primary":
                assert st["last_trace_sorted"] == 1, (name, mode, st)
            else:                                   # camera rays: a handful of distinct keys per warp
                assert st["last_trace_sorted"] == 0 and st["last_probe_distinct"] < 6, (name, st)
            if fn == acc.isShadowed:                # any-hit: which blocker is found first may depend on the order, occlusion does not
                assert torch.equal(out[:, 3].view(torch.int32) >= 0, ref[:, 3].view(torch.int32) >= 0), (name, mode)
            else:
                assert torch.equal(out.view(torch.int32), ref.view(torch.int32)), (name, mode)
        # host buffers (the chunk-pipelined path sorts every chunk on its own stream)
        acc.set_ray_sort(1)
        big = np.ascontiguousarray(np.tile(rays, (3, 1)))      # 786 432 rays: several chunks
        got = fn(big)
        exp = ref.cpu().numpy()
        if fn == acc.isShadowed:
            assert np.array_equal(got["prim"] >= 0, np.tile(exp[:, 3].view(np.int32) >= 0, 3))
        else:
            assert np.array_equal(got["t"], np.tile(exp[:, 0], 3)) and np.array_equal(got["prim"], np.tile(exp[:, 3].view(np.int32), 3))
        assert acc.stats()["last_trace_sorted"] == 1
    acc.set_ray_sort(2)


def test_two_level_instanced_scene_equals_the_flattened_upload():
    """y4gpu_upload_instanced: meshes once + instances on the device; closest hits bit-identical to the flattened upload, any-hit
    same occlusion, resident and host batches, and a fraction of the memory."""
    import torch
    from libyafaray_b200.accelerator import AcceleratorB200
    meshes, inst = helpers.mini_city(n_mesh=6, n_inst=300, tris_per_mesh=1500)
    flat = helpers.flatten_instances(meshes, inst)
    a_flat = AcceleratorB200(flat, leaf_max=3)
    a_two = AcceleratorB200.from_instances(meshes, inst, leaf_max=3)
    assert a_two.n_primitives == len(flat)
    assert np.array_equal(a_flat.getBound(), a_two.getBound())
    s1, s2 = a_flat.stats(), a_two.stats()
    assert (s2["node_bytes"] + s2["tri_bytes"]) * 8 < s1["node_bytes"] + s1["tri_bytes"], (s1, s2)
    for scale, n in ((1.0, 200000), (300.0, 50000)):
        rays = helpers.city_rays(n, origin_scale=scale)
        h1, h2 = a_flat.intersect(rays), a_two.intersect(rays)
        assert np.array_equal(h1.view(np.uint32), h2.view(np.uint32)), int((h1["prim"] != h2["prim"]).sum())
        assert (h1["prim"] >= 0).mean() > 0.5
        assert np.array_equal(a_flat.isShadowed(rays)["prim"] >= 0, a_two.isShadowed(rays)["prim"] >= 0)
        d_rays = torch.from_numpy(rays).cuda()
        d_out = torch.empty((n, 4), dtype=torch.float32, device="cuda")
        a_two.intersect(d_rays, d_out)
        a_two.synchronize()
        assert np.array_equal(d_out.cpu().numpy().view(np.uint32).reshape(-1), h1.view(np.uint32).reshape(-1))
    a_two.intersect(rays, counted=True)
    st = a_two.stats()
    assert st["nodes_visited"] > len(rays) and st["tris_tested"] > 0           # visit counters (the roofline's algorithmic bytes)
    # filtered transparent shadows (intersectTs): two meshes transparent, one colour per material
    from libyafaray_b200 import _lib
    vis = _lib.TRI_VIS_CLOSEST | _lib.TRI_VIS_SHADOW | _lib.TRI_VIS_TSHADOW
    mesh_flags = [np.full(len(m), vis | (_lib.TRI_TRANSPARENT if i in (0, 2) else 0) | (i << _lib.TRI_MAT_SHIFT), np.uint32) for i, m in enumerate(meshes)]
    flat_flags = np.concatenate([mesh_flags[mi] for mi, _ in inst])
    mt = np.random.default_rng(3).uniform(0.2, 0.9, (len(meshes), 3)).astype(np.float32)
    b_flat, b_two = AcceleratorB200(flat, flags=flat_flags, leaf_max=3), AcceleratorB200.from_instances(meshes, inst, mesh_flags=mesh_flags, leaf_max=3)
    ts_rays = helpers.city_rays(100000, seed=9)
    ts_rays[:, 7] = 2.5                     # segments: many end inside the city after crossing only transparent walls
    for depth in (1, 3, 8):
        s1, s2 = b_flat.isShadowedTransparent(ts_rays, depth, mt), b_two.isShadowedTransparent(ts_rays, depth, mt)
        assert np.array_equal(s1["prim"] >= 0, s2["prim"] >= 0)
        lit = s1["prim"] < 0
        # the layers multiply in the order the walk meets them: the products agree to fp32 rounding
        assert np.allclose(s1["filter"][lit], s2["filter"][lit], rtol=1e-6, atol=0)
        assert (s1["filter"][lit] < 0.999).any()
    # a collapsed instance cannot be walked in object space: refused (the Interface then uploads the flattened scene); an instance
    # of an empty mesh owns no primitives and changes nothing
    with pytest.raises(Exception, match="singular"):
        AcceleratorB200.from_instances(meshes, inst + [(0, [1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 0])])
    a_e = AcceleratorB200.from_instances(meshes + [np.zeros((0, 9), np.float32)], inst + [(len(meshes), None)], leaf_max=3)
    assert np.array_equal(a_e.intersect(rays).view(np.uint32), h1.view(np.uint32))
    # back to a flat scene in the same context
    a_two.rebuild(flat[:1000])
    assert np.array_equal(a_two.intersect(rays[:1000]).view(np.uint32), AcceleratorB200(flat[:1000]).intersect(rays[:1000]).view(np.uint32))


def test_full_size_c3_properties_and_tree_independence(monkeypatch):
    """BASELINE config C3 at full size (1 M-triangle sphere, 4 Mi + 4 Mi rays here): size-independent properties.
    * a segment is occluded (any-hit) exactly when the closest-hit query over the same range finds something;
    * the hit records do not depend on the BVH: plain LBVH (no SAH treelets) and the default tree agree bit for bit;
    * a brute-force pass over ALL triangles with the exact triangle test agrees on sampled rays."""
    from libyafaray_b200.accelerator import AcceleratorB200
    tris = scenes.displaced_sphere(1000, 500)
    assert len(tris) == 1_000_000
    prim = scenes.primary_rays(2048, 2048)
    acc = AcceleratorB200(tris)
    h = acc.intersect(prim)
    p, nrm, ok = scenes.hit_points_normals(tris, prim, h["t"], h["prim"])
    df = scenes.diffuse_rays(p, nrm, ok)
    hd = acc.intersect(df)
    # any-hit over [tmin, t_closest + margin] must report an occluder iff the closest-hit query hit; over a range that ends
    # before the closest hit it must report none
    seg = df.copy()                      # rays are [n, 8] float32: origin, tmin, direction, tmax
    seg[:, 7] = np.where(hd["prim"] >= 0, hd["t"] * np.float32(0.5), np.float32(1e30))
    # (a blocker a rounding error below tmin is invisible to the closest-hit rule `tmin <= t` but can land at t' >= 0 after the
    # shadow rule's origin shift -- the reference's two wrappers differ the same way; grazing bounce rays only)
    early = acc.isShadowed(seg)["prim"][hd["prim"] >= 0] >= 0
    assert early.mean() <= 1e-4, (int(early.sum()), len(early))
    occ = acc.isShadowed(df)["prim"] >= 0
    # shadow rays shift the origin by dir * tmin and accept 0 <= t (accelerator.cc:71-80), closest-hit accepts tmin <= t:
    # the two agree wherever the closest hit is clear of tmin
    clear = (hd["prim"] < 0) | (hd["t"] > 4 * df[:, 3])
    differ = occ[clear] != (hd["prim"] >= 0)[clear]
    assert differ.mean() <= 1e-4, (int(differ.sum()), int(clear.sum()))
    st_default = acc.stats()
    monkeypatch.setenv("Y4GPU_TREELET", "0")
    plain = AcceleratorB200(tris)
    for rays, ref in ((prim, h), (df, hd)):
        g = plain.intersect(rays)
        assert np.array_equal(g["prim"], ref["prim"]) and np.array_equal(g["t"], ref["t"]) and np.array_equal(g["bu"], ref["bu"])
    assert plain.stats()["n_nodes"] != st_default["n_nodes"]          # the two trees really differ
    pick = np.linspace(0, len(df) - 1, 24).astype(np.int64)
    exp = helpers.emu_brute(tris, df[pick], 0)
    assert np.array_equal(hd["prim"][pick], exp["prim"]) and np.array_equal(hd["t"][pick], exp["t"])
    plain.close()
    acc.close()


def test_many_scale_chain_scene_deep_tree_stack_bound():
    """k_trace's 64-entry stack.  A scene whose triangles shrink geometrically towards one corner gives the
    deepest hierarchy the builder can produce in fp32 (every split peels a few octaves off); sibling groups (<= levels - 1)
    plus postponed triangle groups (pushed only while sp < 64 - levels) must fit, and the hits must equal brute force."""
    from libyafaray_b200.accelerator import AcceleratorB200
    rng = np.random.default_rng(7)
    k = np.arange(0, 240, dtype=np.float64)
    s = (0.5 ** (k / 2.0))                                   # 120 octaves of scale
    base = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float64)
    tris = []
    for rep in range(12):                                    # several triangles per scale, scattered around the diagonal
        jitter = 1.0 + 0.3 * rng.random((len(k), 1, 3))
        rot = rng.random((len(k), 1, 3)) * 0.2
        tris.append(((base[None] * jitter + rot) * s[:, None, None] + s[:, None, None]).reshape(len(k), 9))
    tris = np.concatenate(tris).astype(np.float32)
    acc = AcceleratorB200(tris)
    levels = acc.stats()["bvh_levels"]
    assert 1 <= levels <= 56
    # rays from outside towards the accumulation corner, from many directions, plus rays along the diagonal
    n = 4096
    o = rng.normal(size=(n, 3)); o = 3.0 * o / np.linalg.norm(o, axis=1, keepdims=True)
    target = (rng.random((n, 1)) ** 6) * np.array([[1.2, 1.2, 0.2]]) * rng.random((n, 3))
    d = target - o; d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = o; rays[:, 4:7] = d; rays[:, 3] = 0.0; rays[:, 7] = -1.0
    got, exp = acc.intersect(rays), helpers.emu_brute(tris, rays, 0)
    assert (exp["prim"] >= 0).sum() > n // 20
    # triangles 2^-100 across seen from distance 3 collide in fp32 t: ties / 1-ulp near-ties are the documented class
    r = helpers.compare_closest(got, exp)
    assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["near_tie"] <= 16 and r["same"] >= n - 16, r
    assert np.array_equal(acc.isShadowed(rays)["prim"] >= 0, helpers.emu_brute(tris, rays, 1)["prim"] >= 0)
    acc.close()


# ---- full-size configurations against the reference (pin C3 / C5 / intersectTs) ------------------------------------
@pytest.mark.skipif(not helpers.have_ref(), reason="oracle/_ref not built")
def test_full_size_c3_vs_live_reference_kdtree():
    """BASELINE config C3 at FULL size: the 1 000 000-triangle displaced sphere; 3 x 262 144 rays (primary / shadow / cosine-diffuse)
    against yafaray-kdtree-original built on the same triangles by the unmodified reference (oracle/_ref, live on this box)."""
    from libyafaray_b200.accelerator import AcceleratorB200
    tris = scenes.displaced_sphere(1000, 500)
    assert len(tris) == 1_000_000
    acc = AcceleratorB200(tris)
    with helpers.RefServer(tris, threads=0) as srv:
        prim = scenes.primary_rays(512, 512)
        h_ref, _ = srv.trace(prim, "closest")
        got = acc.intersect(prim)
        r = helpers.compare_closest(got, h_ref)
        assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0 and r["near_tie"] <= 8 and r["bit_exact_t"] == r["n_hit_both"], r
        assert r["n_hit_both"] > 100_000
        p, nrm, ok = scenes.hit_points_normals(tris, prim, h_ref["t"], h_ref["prim"])
        sh, df = scenes.shadow_rays(p, ok), scenes.diffuse_rays(p, nrm, ok)
        h_sh, _ = srv.trace(sh, "shadow")
        assert np.array_equal(acc.isShadowed(sh)["prim"] >= 0, h_sh["prim"] != -1)
        h_df, _ = srv.trace(df, "closest")
        r = helpers.compare_closest(acc.intersect(df), h_df)
        assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0 and r["near_tie"] <= 8 and r["bit_exact_t"] == r["n_hit_both"], r
        assert np.array_equal(acc.getBound(), np.array(srv.info["bound"], np.float32))
    acc.close()


@pytest.mark.parametrize("name", ["rays_soup5m.npz", "rays_soup50m.npz"])
def test_c5_soup_vs_reference_kdtree_golden(name):
    """BASELINE config C5: the random triangle soup -- its first 5 M triangles (the oracle subsample SURVEY.md 8(d) prescribes)
    and all 50 M (full size) -- 65 536 incoherent closest-hit rays and 65 536 any-hit segments against the committed hit records
    of yafaray-kdtree-original (tests/golden/make_golden_large.py)."""
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name)
    if not os.path.exists(path):
        pytest.skip(name + " not generated")
    from libyafaray_b200.accelerator import AcceleratorB200
    d = np.load(path)
    n_tri, n_rays = int(d["scene"][0]), int(d["scene"][1])
    tris = np.concatenate([scenes.triangle_soup(min(5_000_000, n_tri - f), first=f) for f in range(0, n_tri, 5_000_000)])
    acc = AcceleratorB200(tris)
    inc, seg = scenes.incoherent_rays(n_rays), scenes.segment_rays(n_rays)
    r = helpers.compare_closest(acc.intersect(inc), d["hit_incoherent"])
    assert r["bad"] == 0 and r["t_out_of_tol"] == 0 and r["uv_out_of_tol"] == 0 and r["near_tie"] <= 8 and r["bit_exact_t"] == r["n_hit_both"], r
    assert r["n_hit_both"] > n_rays // 2
    occ = np.unpackbits(d["occluded_segment"])[:n_rays].astype(bool)
    assert np.array_equal(acc.isShadowed(seg)["prim"] >= 0, occ)
    assert np.array_equal(acc.getBound(), d["bound"])
    acc.close()


def _sheet_scene(rng, n_sheets=12):
    """Stacked sheets (2 triangles each) across the unit square at increasing z; random materials: several transparent
    shinydiffuse filters and a few opaque ones."""
    tris, tri_mat = [], []
    mats = [(0.0, 0.8, 0.8, 0.8, 1.0)]                         # material 0: opaque
    for k in range(5):
        mats.append((float(np.float32(0.3 + 0.13 * k)), *(float(np.float32(c)) for c in rng.random(3)), float(np.float32(0.2 + 0.15 * k))))
    for s in range(n_sheets):
        z = 0.1 + 0.07 * s
        x0, y0 = rng.random(2) * 0.4 - 0.2
        a, b, c, e = (x0, y0, z), (x0 + 1.0, y0, z + 0.01), (x0 + 1.0, y0 + 1.0, z), (x0, y0 + 1.0, z + 0.01)
        tris += [[*a, *b, *c], [*a, *c, *e]]
        m = int(rng.integers(1, len(mats))) if s % 5 != 4 else 0
        tri_mat += [m, m]
    return np.array(tris, np.float32), np.array(tri_mat, np.int32), mats


@pytest.mark.skipif(not helpers.have_ref(), reason="oracle/_ref not built")
def test_transparent_shadows_vs_reference_intersectTs(tmp_path):
    """y4gpu_trace_tshadow against Accelerator::isShadowed(ray, depth, bias, camera) -> AcceleratorKdTree::intersectTs of the
    unmodified reference on a scene of transparent shinydiffuse sheets: shadowed flag, filter colour product, depth limit."""
    import os
    import subprocess
    from libyafaray_b200 import _lib
    from libyafaray_b200.accelerator import AcceleratorB200
    rng = np.random.default_rng(11)
    tris, tri_mat, mats = _sheet_scene(rng)
    n = 20000
    o = np.stack([rng.random(n) * 1.2 - 0.1, rng.random(n) * 1.2 - 0.1, np.zeros(n)], -1)
    e = np.stack([rng.random(n) * 1.2 - 0.1, rng.random(n) * 1.2 - 0.1, rng.random(n) * 1.1 + 0.05], -1)   # end points inside the stack
    d = e - o
    dist = np.linalg.norm(d, axis=1)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = o; rays[:, 4:7] = d / dist[:, None]; rays[:, 3] = 5e-4; rays[:, 7] = dist
    rays[: n // 10, 7] = -1.0                                   # some unbounded rays
    mf, tf = tmp_path / "mats.txt", tmp_path / "trimat.i32"
    mf.write_text("".join("%r %r %r %r %r\n" % m for m in mats))
    tri_mat.tofile(tf)
    f32 = np.float32
    # Material::getTransparency of an untextured shinydiffusemat (material_shiny_diffuse.cc:442-465)
    mt = np.array([[f32(f32(m[0]) * f32(f32(m[4]) * f32(m[1 + c]) + f32(f32(1.0) - f32(m[4])))) for c in range(3)] for m in mats], f32)
    flags = ((tri_mat.astype(np.uint32) << _lib.TRI_MAT_SHIFT) | 7 | np.where(np.array([mats[m][0] for m in tri_mat]) > 1e-5, _lib.TRI_TRANSPARENT, 0)).astype(np.uint32)
    acc = AcceleratorB200(tris, flags=flags)
    for depth in (2, 4, 16):
        with helpers.tempfile.TemporaryDirectory() as dd:
            tp, rp, hp = (os.path.join(dd, x) for x in ("t.f32", "r.f32", "h.bin"))
            tris.tofile(tp); rays.tofile(rp)
            env = dict(os.environ, REF_RAYBATCH_MATS=str(mf), REF_RAYBATCH_TRIMAT=str(tf), REF_RAYBATCH_TS_DEPTH=str(depth))
            subprocess.run([os.path.join(helpers.REF_DIR, "ref_raybatch"), tp, rp, "tshadow", hp, "4"], check=True, capture_output=True, env=env)
            ref = np.fromfile(hp, helpers.HIT_DTYPE)
        got = acc.isShadowedTransparent(rays, depth, mt)
        shadowed_ref = ref["prim"] != -1
        assert np.array_equal(got["prim"] >= 0, shadowed_ref), depth
        assert 0.05 < shadowed_ref.mean() < 0.95
        lit = ~shadowed_ref
        col_ref = np.stack([ref["t"], ref["bu"], ref["bv"]], -1)[lit]
        col = got["filter"][lit]
        assert (col_ref < 0.999).any(axis=1).mean() > 0.3           # many rays really cross transparent sheets
        assert np.allclose(col, col_ref, rtol=4e-6, atol=1e-9), float(np.abs(col - col_ref).max())   # products in another order: a few ulp
    # the compiled plug-in (AcceleratorB200::intersectTs inside libYafaRay), when built: same answers from the reference's own wrapper
    if helpers.have_plugin("b200_raybatch"):
        with helpers.tempfile.TemporaryDirectory() as dd:
            tp, rp, hp = (os.path.join(dd, x) for x in ("t.f32", "r.f32", "h.bin"))
            tris.tofile(tp); rays.tofile(rp)
            env = dict(os.environ, REF_RAYBATCH_MATS=str(mf), REF_RAYBATCH_TRIMAT=str(tf), REF_RAYBATCH_TS_DEPTH="4")
            subprocess.run([os.path.join(helpers.B200_DIR, "b200_raybatch"), tp, rp, "tshadow", hp, "8", "yafaray-b200-bvh8"], check=True, capture_output=True, env=env)
            plug = np.fromfile(hp, helpers.HIT_DTYPE)
            subprocess.run([os.path.join(helpers.REF_DIR, "ref_raybatch"), tp, rp, "tshadow", hp, "4"], check=True, capture_output=True, env=env)
            ref = np.fromfile(hp, helpers.HIT_DTYPE)
        assert np.array_equal(plug["prim"] != -1, ref["prim"] != -1)
        lit = ref["prim"] == -1
        assert np.allclose(np.stack([plug["t"], plug["bu"], plug["bv"]], -1)[lit], np.stack([ref["t"], ref["bu"], ref["bv"]], -1)[lit], rtol=4e-6, atol=1e-9)
    acc.close()
