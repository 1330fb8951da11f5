"""Msamples/s of Cornell variants with specular materials at raydepth 5, caustic_type path:
  mirror: tall box, back and left wall are `mirror` materials (chains followed inside the camera-hit step)
  shiny : every wall material is shinydiffusemat with specular_reflect 0.3 (one estimate per ray level)
usage: python tools/mirrorbench.py [width height spp bounces reps]"""
import json
import sys

sys.path.insert(0, ".")
from libyafaray_b200 import interface, scenes_pt


def main():
    a = [int(x) for x in sys.argv[1:]]
    w, h, spp, bounces, reps = (a + [1920, 1080, 16, 8, 3][len(a):])[:5]
    for name, kw in (("mirror", dict(mirror=dict(objects=("tall", "back", "left"), color=(0.9, 0.95, 0.8)))), ("shiny", dict(shiny=dict(specular_reflect=0.3)))):
        yi = interface.Interface()
        scenes_pt.cornell_box(yi, width=w, height=h, spp=spp, bounces=bounces, accelerator="yafaray-b200-bvh8",
                              integrator_extra=dict(raydepth=5, caustic_type="path"), **kw)
        yi.upload()
        for r in range(reps):
            st = {"samples": 0, "rays_closest": 0, "rays_shadow": 0, "kernel_launches": 0, "pass_ms": 0.0}
            yi.render_pass(spp, r * spp, 0)
            yi._add_stats(st)
            st["msamples_s"] = st["samples"] / st["pass_ms"] * 1e-3
            st["mrays_s"] = (st["rays_closest"] + st["rays_shadow"]) / st["pass_ms"] * 1e-3
            print(json.dumps({"scene": name, **st}), flush=True)
        yi.clearAll()


if __name__ == "__main__":
    main()
