"""A few seeds of every differential fuzzer of tools/ (the campaigns of the round are logged in profiles/r2_v16_fuzz_vs_reference.log):
the emulation of the device code against the LIVE reference, so this needs oracle/_ref (this container)."""
import os
import subprocess
import sys

import pytest

import helpers

pytestmark = pytest.mark.skipif(not (helpers.have_ref("ref_render") and helpers.have_ref("ref_raybatch")), reason="oracle/_ref not built")
TOOLS = os.path.join(helpers.ROOT, "tools")


def _run(tool, *args):
    out = subprocess.run([sys.executable, os.path.join(TOOLS, tool), *map(str, args)], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout


@pytest.mark.parametrize("family,first,count", [("cornell", 7000, 10), ("textures", 7100, 6), ("lights", 7200, 8), ("materials", 7300, 8), ("adaptive", 7400, 6), ("city", 205, 3)])
def test_path_tracer_fuzz_families_agree_with_the_live_reference(family, first, count):
    out = _run("fuzz_vs_reference.py", first, count, family)
    assert out.strip().splitlines()[-1] == "bad: 0", out[-3000:]
    assert out.count(" ok ") + out.count("skipped") + out.count("refused") == count


def test_traversal_fuzz_agrees_with_the_reference_kdtree():
    assert _run("fuzz_rays_vs_reference.py", 500, 4).strip().splitlines()[-1] == "bad total 0"


def test_two_level_fuzz_agrees_with_the_flattened_scene():
    assert _run("fuzz_instances.py", 0, 6, "aniso5").strip().splitlines()[-1] == "bad total 0"
