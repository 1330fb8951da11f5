"""Regenerates the committed golden fixtures from the UNMODIFIED reference (oracle/_ref, built
from /root/reference by `make -C oracle ref`).  Run in the build container only:

    python tests/golden/make_golden.py

  kat_sampling.json  <- oracle/_ref/ref_kat   (Halton/LDS/riVdC/riS/riLp/fnv, FAST_TRIG, cosHemisphere)
  rays_sphere.npz    <- oracle/_ref/ref_raybatch on scenes.displaced_sphere(40, 20): primary,
                        shadow and diffuse batches with the reference kd-tree's hit records
  rays_soup.npz      <- same on scenes.triangle_soup(3000) with incoherent + segment rays
  pt_dof.npz         <- the same harness on the Cornell box with a lens (PT_GOLDEN_DOF)
  pt_mirror.npz      <- the same harness on the Cornell box with `mirror` materials (PT_GOLDEN_MIRROR)
  pt_lights.npz      <- the same harness on the Cornell box with directional / sun / sphere lights (PT_GOLDEN_LIGHTS)
  pt_city.npz        <- the same harness on scenes_pt.city (instanced buildings) for PT_GOLDEN_CITY
  pt_test01.npz      <- the same harness on scenes_pt.test01 (the reference's tests/test01 scene: image textures) for PT_GOLDEN_TEST01
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
    # the other lens shapes of PerspectiveCamera::getLensUv: polygon corner table + edge bias, ring
    "dof_hexagon_edge": dict(scene="cornell", width=64, height=36, spp=4, bounces=3, aperture=0.08, dof_distance=1.9,
                             bokeh=dict(bokeh_type="hexagon", bokeh_bias="edge", bokeh_rotation=20.0)),
    # yafaray_smoothMesh (angle-limited vertex normals) on a sphere next to flat-shaded meshes
    "smooth_sphere_60": dict(scene="cornell", width=64, height=36, spp=3, bounces=3, smooth_sphere=60.0),
    # the other camera types: orthographic, architect (with a square lens), equirectangular from inside the box
    "cam_orthographic": dict(scene="cornell", width=64, height=36, spp=3, bounces=2, camera_extra=dict(type="orthographic", scale=1.3)),
    "cam_architect_dof": dict(scene="cornell", width=64, height=36, spp=3, bounces=2,
                              camera_extra=dict(type="architect", aperture=0.05, dof_distance=1.8, bokeh_type="square")),
    "cam_equirectangular": dict(scene="cornell", width=64, height=36, spp=3, bounces=2,
                                camera_extra={"type": "equirectangular", "from": (0.5, 0.5, 0.3), "to": (0.5, 0.45, 1.0), "up": (0.5, 1.5, 0.3)}),
    "dof_ring": dict(scene="cornell", width=64, height=36, spp=4, bounces=3, aperture=0.08, dof_distance=1.9, bokeh=dict(bokeh_type="ring")),
    # DirectLightIntegrator on the same stages (no path loop), 3 samples per light, gauss filter
    "direct_light3": dict(scene="cornell", width=64, height=36, spp=3, bounces=3, integrator="directlighting", light_samples=3, filter_type="gauss"),
}


# directional (infinite + finite beam), sunlight and spherelight next to the ceiling area light -> pt_lights.npz
PT_GOLDEN_LIGHTS = {
    # objectlight / meshlight: an emitting sphere mesh (70 triangles) sampled through the area cdf, single-light path tracing (sample-exact)
    "mesh_light_sphere": dict(scene="cornell", width=64, height=36, spp=3, bounces=3, light_samples=2, mesh_light=dict(shape="sphere", power=40.0)),
    "mesh_light_quad_double_sided_direct": dict(scene="cornell", width=64, height=36, spp=2, bounces=2, integrator="directlighting", light_samples=3,
                                                mesh_light=dict(double_sided=True)),
    "six_lights": dict(scene="cornell", width=64, height=36, spp=3, bounces=3, light_samples=2,
                        extra_lights=("sun", "directional", "directional_finite", "sphere", "spot")),
    "sun_sphere_spot_direct": dict(scene="cornell", width=64, height=36, spp=2, bounces=2, integrator="directlighting", light_samples=3,
                              point_light=True, extra_lights=("sun", "sphere", "spot"), filter_type="gauss"),
}


# `mirror` materials: the ray-level loop of recursiveRaytrace / specularReflect, specular bounces inside paths,
# emission seen through them (caustic_type path), transparent background behind reflections -> pt_mirror.npz
_MIRROR = dict(objects=("tall", "back"), color=(0.9, 0.95, 0.8), reflect=0.95)
PT_GOLDEN_MIRROR = {
    "mirror_box_and_wall": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, mirror=_MIRROR, integrator_extra=dict(raydepth=3, caustic_type="path")),
    "mirror_glow_caustic_paths": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, mirror=_MIRROR, glow=0.7, path_samples=2, light_samples=2,
                                      integrator_extra=dict(raydepth=2, caustic_type="path")),
    "mirror_hall_raydepth8": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, point_light=True, filter_type="gauss",
                                  mirror=dict(objects=("tall", "back", "left", "short")), integrator_extra=dict(raydepth=8, caustic_type="path")),
    # shinydiffusemat with a mirror component (specular_reflect, mirror_color, Fresnel): one estimate per ray level
    "shiny_walls_raydepth3": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, shiny=dict(specular_reflect=0.3, mirror_color=(0.9, 0.7, 0.5)),
                                  integrator_extra=dict(raydepth=3, caustic_type="path")),
    "shiny_fresnel_mirror_box_glow": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, shiny=dict(specular_reflect=0.6, fresnel_effect=True, IOR=1.8),
                                          mirror=dict(objects=("tall",), color=(0.9, 0.95, 0.8)), glow=0.5, oren_nayar=True, path_samples=2, light_samples=2,
                                          integrator_extra=dict(raydepth=2, caustic_type="path")),
    # shinydiffusemat transparency / transmit_filter: straight-through refraction rays (specularRefract), the Transmit | Filter
    # lobe inside paths, getAlpha in the sample weight; shadows stay opaque (transpShad off, the reference's default)
    "pane_filter_glow_caustic": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, pane=dict(transparency=0.8, transmit_filter=0.4, diffuse_reflect=0.7),
                                     glow=0.5, integrator_extra=dict(raydepth=3, caustic_type="path")),
    "pane_fresnel_shiny_walls_transp_bg_direct": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, integrator="directlighting",
                                                      pane=dict(transparency=0.5, fresnel_effect=True), shiny=dict(specular_reflect=0.3), background=(0.3, 0.2, 0.1),
                                                      integrator_extra=dict(raydepth=3, bg_transp=True, bg_transp_refract=True)),
    # transpShad: shadow rays filtered by the transparent pane (intersectTs colour product, shadowDepth)
    "pane_filtered_shadows": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, light_samples=2, path_samples=2,
                                  pane=dict(transparency=0.9, transmit_filter=1.0, color=(0.9, 0.3, 0.2)),
                                  integrator_extra=dict(raydepth=3, transpShad=True, caustic_type="path")),
    # shinydiffusemat translucency (diffuse transmission) with transparency and Oren-Nayar on the same sheet
    "pane_translucent_transparent": dict(scene="cornell", width=64, height=36, spp=3, bounces=4, oren_nayar=True,
                                         pane=dict(translucency=0.6, transparency=0.3, transmit_filter=0.5), integrator_extra=dict(raydepth=2, caustic_type="path")),
    "gradient_background_mirrors": dict(scene="cornell", width=64, height=36, spp=3, bounces=3, mirror=dict(objects=("tall", "back")), integrator_extra=dict(raydepth=3),
                                        background={"type": "gradientback", "horizon_color": (1.0, 0.8, 0.6, 1.0), "zenith_color": (0.2, 0.4, 0.9, 1.0),
                                                    "horizon_ground_color": (0.3, 0.25, 0.2, 1.0), "zenith_ground_color": (0.05, 0.1, 0.02, 1.0), "power": 0.8}),
    "mirror_sphere_transp_bg_direct": dict(scene="cornell", width=64, height=36, spp=2, bounces=2, integrator="directlighting", smooth_sphere=180.0,
                                           mirror=dict(objects=("sphere", "short")), background=(0.3, 0.2, 0.1),
                                           integrator_extra=dict(raydepth=2, bg_transp=True, bg_transp_refract=True)),
}


# C1 (BASELINE configs[0]): the reference's tests/test01 scene -- six cubes with image textures (layer + cube-mapped
# texture_mapper nodes on orco coordinates) -- with the path tracer, as shipped (directlighting), and variants that walk
# through the other texture / mapper / layer modes -> pt_test01.npz.  The reference build here has the TGA and HDR
# loaders only (oracle/Makefile), so four of the six textures fail to load and their materials fall back to the plain
# colour, in the reference and in the mirror alike.  The two image files it can load are copied next to the fixtures
# (tests/golden/test01/: test input data of the reference's own test, needed where /root/reference does not exist).
TEST01_DIR = os.path.join(HERE, "test01")
PT_GOLDEN_TEST01 = {
    "c1_test01_pathtracing": dict(scene="test01", width=480, height=270, spp=3, bounces=4),
    "c1_test01_as_shipped_directlighting": dict(scene="test01", width=480, height=270, spp=3, integrator="directlighting"),
    # tests/test00: the texture is filled texel by texel through yafaray_setImageColor (HDR values), un-optimised storage
    "test00_setimagecolor_directlighting": dict(scene="test00", width=200, height=200, spp=1, integrator="directlighting"),
    "test00_setimagecolor_pathtracing": dict(scene="test00", width=160, height=160, spp=2, integrator="pathtracing", bounces=3, tex=64),
    "tex_uv_repeat_mirror_adjust_multiply": dict(scene="test01", width=160, height=90, spp=2, bounces=3, with_uv=True,
        texture_extra={"xrepeat": 3, "yrepeat": 2, "mirror_x": True, "mirror_y": True, "adj_intensity": 1.2, "adj_contrast": 0.8,
                       "adj_mult_factor_red": 0.7, "adj_clamp": True},
        mapper_extra={"texco": "uv", "mapping": "plain", "offset": (0.3, 0.1, 0.0)}, layer_extra={"blend_mode": "multiply", "colfac": 0.7}),
    "tex_global_checker_nearest_stencil": dict(scene="test01", width=160, height=90, spp=2, bounces=3,
        texture_extra={"clipping": "checker", "checker_dist": 0.2, "even_tiles": True, "odd_tiles": False, "interpolate": "none"},
        mapper_extra={"texco": "global", "mapping": "cube", "scale": (0.7, 0.6, 0.5), "proj_x": 2, "proj_y": 3, "proj_z": 1},
        layer_extra={"stencil": True, "negative": True, "upper_color": (0.5, 0.6, 0.7, 0.6)}),
    "tex_orco_clip_crop_rot_screen_norgb": dict(scene="test01", width=160, height=90, spp=2, bounces=3, filter_type="box",
        texture_extra={"clipping": "clip", "rot90": True, "cropmin_x": 0.1, "cropmax_y": 0.8},
        mapper_extra={"scale": (1.7, 2.3, 1.0), "offset": (0.2, -0.3, 0.0)},
        layer_extra={"blend_mode": "screen", "noRGB": True, "def_col": (0.9, 0.2, 0.1, 1.0)}),
}


def make_extra(configs, filename, cwd=None):
    from libyafaray_b200 import scenes_pt

    out = {}
    for key, kw in configs.items():
        rec = helpers.ScriptRecorder()
        scenes_pt.build(rec, **kw)
        out[key] = helpers.ref_render(rec, cwd=cwd)[0]
    np.savez_compressed(os.path.join(HERE, filename), **out)


def make_test01():
    import shutil

    os.makedirs(TEST01_DIR, exist_ok=True)
    for f in ("tex.tga", "tex.hdr"):        # tests/test00/tex.tga is the same file as tests/test01/tex.tga
        shutil.copyfile(os.path.join("/root/reference/tests/test01", f), os.path.join(TEST01_DIR, f))
        os.chmod(os.path.join(TEST01_DIR, f), 0o644)
    make_extra(PT_GOLDEN_TEST01, "pt_test01.npz", cwd=TEST01_DIR)


def make_city():
    make_extra(PT_GOLDEN_CITY, "pt_city.npz")
    make_extra(PT_GOLDEN_DOF, "pt_dof.npz")
    make_extra(PT_GOLDEN_LIGHTS, "pt_lights.npz")
    make_extra(PT_GOLDEN_MIRROR, "pt_mirror.npz")
    make_test01()


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
