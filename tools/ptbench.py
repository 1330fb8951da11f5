"""Quick Msamples/s probe of the wavefront path tracer on the Cornell scene.
usage: python tools/ptbench.py [width height spp bounces wave_size reps profile]"""
import json
import sys
import time

sys.path.insert(0, ".")
from libyafaray_b200 import interface, scenes_pt


def main():
    a = [int(x) for x in sys.argv[1:]]
    w, h, spp, bounces, wave, reps, prof = (a + [1920, 1080, 16, 8, 0, 3, 0][len(a):])[:7]
    yi = interface.Interface()
    scenes_pt.cornell_box(yi, width=w, height=h, spp=spp, bounces=bounces, accelerator="yafaray-b200-bvh8")
    if wave:
        yi.render_params["b200_wave_size"] = wave
    if prof:
        yi.render_params["b200_profile"] = 1
    yi.upload()
    for r in range(reps):
        t0 = time.time()
        st = {"samples": 0, "rays_closest": 0, "rays_shadow": 0, "kernel_launches": 0, "pass_ms": 0.0}
        yi.render_pass(spp, r * spp, 0)
        yi._add_stats(st)
        st["wall_s"] = time.time() - t0
        st["msamples_s"] = st["samples"] / st["pass_ms"] * 1e-3
        st["mrays_s"] = (st["rays_closest"] + st["rays_shadow"]) / st["pass_ms"] * 1e-3
        if prof:
            st.update({k: v for k, v in yi.pass_stats().items() if k not in st})
        print(json.dumps(st), flush=True)


if __name__ == "__main__":
    main()
