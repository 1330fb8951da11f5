"""C4 (BASELINE configs[3]) on one GPU: instanced city flattened to ~10 M world-space triangles, 3840x2160, 8 bounces + NEE,
16 spp per pass.  Reports scene assembly / upload+build times and Msamples/s of the passes.
usage: python tools/c4bench.py [n_tri] [width] [height] [spp] [passes] [profile]"""
import json
import sys
import time

sys.path.insert(0, ".")
from libyafaray_b200 import interface, scenes_pt


def main():
    a = [int(x) for x in sys.argv[1:]]
    n_tri, w, h, spp, passes, prof = (a + [10_000_000, 3840, 2160, 16, 3, 0][len(a):])[:6]
    yi = interface.Interface()
    t0 = time.time()
    n = scenes_pt.city(yi, n_tri_target=n_tri, width=w, height=h, spp=spp, bounces=8, accelerator="yafaray-b200-bvh8")
    t_scene = time.time() - t0
    if prof:
        yi.render_params["b200_profile"] = 1
    t0 = time.time()
    asm = yi.upload()
    t_upload = time.time() - t0
    st = yi.accelerator.stats()
    print(json.dumps({"config": "C4", "instanced_tris": int(n), "tris": int(len(asm["tris"])), "scene_calls_s": t_scene, "assemble_upload_build_s": t_upload,
                      "build_ms": st["build_ms"], "n_nodes": st["n_nodes"]}), flush=True)
    for r in range(passes):
        yi.render_pass(spp, r * spp, 0)
        s = yi.pass_stats()
        print(json.dumps({"pass": r, "samples": s["samples"], "pass_ms": s["pass_ms"], "msamples_s": s["samples"] / s["pass_ms"] * 1e-3,
                          "rays_closest_per_sample": s["rays_closest"] / s["samples"], "rays_shadow_per_sample": s["rays_shadow"] / s["samples"],
                          "mrays_s": (s["rays_closest"] + s["rays_shadow"]) / s["pass_ms"] * 1e-3, "kernel_launches": s["kernel_launches"],
                          **({"nodes_per_closest": s["nodes_closest"] / max(1, s["rays_closest"]), "tris_per_closest": s["tris_closest"] / max(1, s["rays_closest"]),
                              "nodes_per_shadow": s["nodes_shadow"] / max(1, s["rays_shadow"]), "tris_per_shadow": s["tris_shadow"] / max(1, s["rays_shadow"]),
                              "class_ms": s["class_ms"][:6]} if prof else {})}), flush=True)
    fw, rgba = yi.read_film()
    img = interface.normalize_film(fw, rgba)
    print(json.dumps({"film_mean_rgb": float(img[..., :3].mean()), "film_weight_sum": float(fw.sum())}))


if __name__ == "__main__":
    main()
