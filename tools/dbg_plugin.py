import sys, os, subprocess, tempfile
sys.path.insert(0, "."); sys.path.insert(0, "tests"); sys.path.insert(0, "tests/golden")
import helpers
from libyafaray_b200 import scenes_pt
from make_golden_large import PT_GOLDEN_SHINY2
from make_golden import TEST01_DIR
rec = helpers.ScriptRecorder()
scenes_pt.build(rec, accelerator="yafaray-b200-bvh8", **PT_GOLDEN_SHINY2["pane_all_four_fresnel_gauss"])
with tempfile.TemporaryDirectory() as d:
    sp, op = os.path.join(d, "scene.script"), os.path.join(d, "out.f32")
    open(sp, "w").write("\n".join(rec.lines) + f"\nrender {op}\n")
    e = dict(os.environ); e["YAFARAY_B200_BACKEND"] = "1"; e["REF_RENDER_VERBOSITY"] = "debug"
    out = subprocess.run([os.path.join(helpers.B200_DIR, "b200_render"), sp], capture_output=True, text=True, env=e, cwd=TEST01_DIR)
    print("rc", out.returncode); print(out.stdout[-3000:]); print(out.stderr[-3000:])
