"""Regenerates the committed golden fixtures from the UNMODIFIED reference (oracle/_ref, built
from /root/reference by `make -C oracle ref`).  Run in the build container only:

    python tests/golden/make_golden.py

  kat_sampling.json  <- oracle/_ref/ref_kat   (Halton/LDS/riVdC/riS/riLp/fnv, FAST_TRIG, cosHemisphere)
  rays_sphere.npz    <- oracle/_ref/ref_raybatch on scenes.displaced_sphere(40, 20): primary,
                        shadow and diffuse batches with the reference kd-tree's hit records
  rays_soup.npz      <- same on scenes.triangle_soup(3000) with incoherent + segment rays
  pt_dof.npz         <- the same harness on the Cornell box with a lens (PT_GOLDEN_DOF)
  pt_lights.npz      <- the same harness on the Cornell box with directional / sun / sphere lights (PT_GOLDEN_LIGHTS)
  pt_city.npz        <- the same harness on scenes_pt.city (instanced buildings) for PT_GOLDEN_CITY
  pt_cornell.npz     <- oracle/_ref/ref_render (reference PathIntegrator + kd-tree through the C API,
                        threads=1, tiles linear) on scenes_pt.cornell_box for the PT_GOLDEN configurations
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from libyafaray_b200 import scenes  # noqa: E402


# configurations of the committed path-tracer golden images (tests import this dict)
PT_GOLDEN = {
    "area_box": dict(width=96, height=54, spp=4, bounces=4),
    "point_gauss": dict(width=96, height=54, spp=4, bounces=4, point_light=True, filter_type="gauss"),
    "area3_oren_paths2": dict(width=64, height=36, spp=3, bounces=3, light_samples=3, oren_nayar=True, path_samples=2),
}
# C4 in miniature (instanced buildings through yafaray_addInstance, sky background, gauss filter) -> pt_city.npz
PT_GOLDEN_CITY = {
    "city_small": dict(scene="city", n_tri_target=6000, n_base=4, grid=6, tri_range=(300, 900), width=80, height=45, spp=3, bounces=4, filter_type="gauss"),
}


# depth of field (aperture != 0: lens samples from Halton(3)/(5), Shirley disk) -> pt_dof.npz
PT_GOLDEN_DOF = {
    "dof_disk": dict(scene="cornell", width=64, height=36, spp=4, bounces=3, aperture=0.06, dof_distance=1.9),
    # DirectLightIntegrator on the same stages (no path loop), 3 samples per light, gauss filter
    "direct_light3": dict(scene="cornell", width=64, height=36, spp=3, bounces=3, integrator="directlighting", light_samples=3, filter_type="gauss"),
}


# directional (infinite + finite beam), sunlight and spherelight next to the ceiling area light -> pt_lights.npz
PT_GOLDEN_LIGHTS = {
    "six_lights": dict(scene="cornell", width=64, height=36, spp=3, bounces=3, light_samples=2,
                        extra_lights=("sun", "directional", "directional_finite", "sphere", "spot")),
    "sun_sphere_spot_direct": dict(scene="cornell", width=64, height=36, spp=2, bounces=2, integrator="directlighting", light_samples=3,
                              point_light=True, extra_lights=("sun", "sphere", "spot"), filter_type="gauss"),
}


def make_extra(configs, filename):
    from libyafaray_b200 import scenes_pt

    out = {}
    for key, kw in configs.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.build(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
    np.savez_compressed(os.path.join(HERE, filename), **out)


def make_city():
    make_extra(PT_GOLDEN_CITY, "pt_city.npz")
    make_extra(PT_GOLDEN_DOF, "pt_dof.npz")
    make_extra(PT_GOLDEN_LIGHTS, "pt_lights.npz")


def main():
    out = subprocess.run([os.path.join(helpers.REF_DIR, "ref_kat")], capture_output=True, text=True, check=True).stdout
    with open(os.path.join(HERE, "kat_sampling.json"), "w") as f:
        f.write(out)

    tris = scenes.displaced_sphere(40, 20)
    prim = scenes.primary_rays(64, 64)
    with helpers.RefServer(tris, threads=1) as srv:
        h_prim, _ = srv.trace(prim, "closest")
        p, n, ok = scenes.hit_points_normals(tris, prim, h_prim["t"], h_prim["prim"])
        sh, df = scenes.shadow_rays(p, ok), scenes.diffuse_rays(p, n, ok)
        h_sh, _ = srv.trace(sh, "shadow")
        h_df, _ = srv.trace(df, "closest")
        bound = np.array(srv.info["bound"], np.float32)
    np.savez_compressed(os.path.join(HERE, "rays_sphere.npz"), scene=np.array([40, 20]), primary=prim, shadow=sh, diffuse=df,
                        hit_primary=h_prim, hit_shadow=h_sh, hit_diffuse=h_df, bound=bound)

    tris = scenes.triangle_soup(3000)
    inc, seg = scenes.incoherent_rays(4096), scenes.segment_rays(4096)
    with helpers.RefServer(tris, threads=1) as srv:
        h_inc, _ = srv.trace(inc, "closest")
        h_seg, _ = srv.trace(seg, "shadow")
        bound = np.array(srv.info["bound"], np.float32)
    np.savez_compressed(os.path.join(HERE, "rays_soup.npz"), scene=np.array([3000]), incoherent=inc, segment=seg,
                        hit_incoherent=h_inc, hit_segment=h_seg, bound=bound)
    # path tracer: reference PathIntegrator renders (deterministic mode) of the Cornell scene
    from libyafaray_b200 import scenes_pt

    out = {}
    for key, kw in PT_GOLDEN.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.cornell_box(rec, **kw)
        out[key] = helpers.ref_render(rec)[0]
    np.savez_compressed(os.path.join(HERE, "pt_cornell.npz"), **out)
    make_city()
    print("golden fixtures written")


if __name__ == "__main__":
    main()
