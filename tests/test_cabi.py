"""The C-ABI shared library loads and exports every symbol include/y4gpu.h declares; without a
GPU every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from libyafaray_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "y4gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(y4gpu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    L = _lib.lib()
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/y4gpu.h but not exported by liby4gpu.so"
    assert sorted(_lib.EXPORTED) == names, "libyafaray_b200/_lib.py EXPORTED list out of sync with the header"


def test_hand_written_struct_mirrors_have_the_sizes_the_library_was_compiled_with():
    L = _lib.lib()          # (lib() itself refuses to load a library whose sizes differ)
    sizes = [L.y4gpu_struct_size(k) for k in range(16)]
    assert sizes[12:] == [32, 16, 32, 0] and all(s > 0 for s in sizes[:12])
    for which, cls in enumerate(_lib.MIRRORED_STRUCTS):
        if cls is not None:
            assert C.sizeof(cls) == sizes[which], cls.__name__
    assert L.y4gpu_struct_size(-1) == 0


def test_no_cpu_fallback_without_gpu():
    L = _lib.lib()
    if L.y4gpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    ctx = C.c_void_p()
    rc = L.y4gpu_create(0, C.byref(ctx))
    assert rc == _lib.ERR_CUDA and not ctx.value
    assert b"no CPU fallback" in L.y4gpu_last_error(None)
    from libyafaray_b200.accelerator import AcceleratorB200
    import numpy as np

    with pytest.raises(_lib.Y4Error):
        AcceleratorB200(np.zeros((1, 9), np.float32))


def test_product_does_not_import_oracle():
    """The product package must never import, link or execute anything under oracle/ or tests/."""
    pkg = os.path.join(ROOT, "libyafaray_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            src = open(os.path.join(dp, f), errors="ignore").read() if f.endswith((".py", ".cu", ".cuh", ".h", "Makefile")) else ""
            if f.endswith(".py"):
                assert not re.search(r"^\s*(from|import)\s+(refserver|helpers|oracle|tests)\b", src, re.M), f
                assert "oracle/_ref" not in src and "liboracle" not in src and "libemu" not in src, f
            else:
                assert not re.search(r"#include[^\n]*(oracle|tests/)", src), f
                assert "oracle/" not in src or f != "Makefile", f


def test_plugin_build_without_a_gpu_fails_the_render_and_does_not_fall_back_to_the_kdtree():
    """The patched libYafaRay (integration/) with scene_accelerator = yafaray-b200-bvh8 on a machine without a CUDA device:
    Accelerator::factory must NOT build a kd-tree silently (accelerator.cc:46-51 does for unknown types); yafaray_render fails."""
    import pytest
    import torch

    import helpers
    from libyafaray_b200 import scenes_pt

    if torch.cuda.is_available() or not helpers.have_plugin():
        pytest.skip("needs the plug-in build and no GPU")
    rec = helpers.ScriptRecorder()
    scenes_pt.cornell_box(rec, accelerator="yafaray-b200-bvh8", width=32, height=18, spp=1, bounces=4)
    img, info = helpers.b200_render(rec, gpu_integrator=False)
    assert info["width"] == 0 and img.size == 0
