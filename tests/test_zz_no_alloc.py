"""y4gpu_trace_noalloc (allocation-free ray batches; no counterpart in the reference)
against the existing entry point y4gpu_trace, through the C-ABI.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from libyafaray_b200 import _lib, scenes

pytestmark = pytest.mark.gpu


def _call(L, ctx, fn, kind, rays, out):
    rc = fn(ctx, kind, C.c_void_p(rays.ctypes.data), len(rays), C.c_void_p(out.ctypes.data))
    assert rc == _lib.OK, (rc, L.y4gpu_last_error(ctx))
    return out


def test_noalloc_entry_point_equals_y4gpu_trace_on_two_calls_in_a_row():
    from libyafaray_b200.accelerator import AcceleratorB200

    L = _lib.lib()
    tris = scenes.displaced_sphere(120, 60)
    acc = AcceleratorB200(tris)
    ctx = acc.ctx
    first = scenes.primary_rays(256, 256)                       # 65536 rays = Y4GPU_NOALLOC_MAX_RAYS exactly
    assert len(first) == _lib.NOALLOC_MAX_RAYS
    ref1 = _call(L, ctx, L.y4gpu_trace, _lib.TRACE_CLOSEST, first, np.empty(len(first), _lib.HIT_DTYPE))
    p, nrm, ok = scenes.hit_points_normals(tris, first, ref1["t"], ref1["prim"])
    second = scenes.diffuse_rays(p, nrm, ok)[: 40000]             # a different, shorter, incoherent batch: stale scratch must not leak
    shadow = scenes.shadow_rays(p, ok)
    for rays in (first, second):                                  # two calls in a row (and a third / fourth below)
        exp = _call(L, ctx, L.y4gpu_trace, _lib.TRACE_CLOSEST, rays, np.empty(len(rays), _lib.HIT_DTYPE))
        got = _call(L, ctx, L.y4gpu_trace_noalloc, _lib.TRACE_CLOSEST, rays, np.full(len(rays), -7, _lib.HIT_DTYPE))
        assert (exp["prim"] >= 0).sum() > len(rays) // 8
        # closest hits are schedule-independent: bit-exact, the tolerance the other closest-hit tests use
        for f in ("prim", "t", "bu", "bv"):
            assert np.array_equal(got[f], exp[f]), f
    for _ in range(2):
        exp = _call(L, ctx, L.y4gpu_trace, _lib.TRACE_SHADOW, shadow, np.empty(len(shadow), _lib.HIT_DTYPE))
        got = _call(L, ctx, L.y4gpu_trace_noalloc, _lib.TRACE_SHADOW, shadow, np.full(len(shadow), -7, _lib.HIT_DTYPE))
        assert np.array_equal(got["prim"] >= 0, exp["prim"] >= 0)  # any-hit: same occlusion (which blocker is found first is free)
    # device-resident buffers: queued on the context's stream, no staging
    import torch

    d_rays = torch.from_numpy(second).cuda()
    d_out = torch.full((len(second), 4), -7.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    for _ in range(2):
        rc = L.y4gpu_trace_noalloc(ctx, _lib.TRACE_CLOSEST, C.c_void_p(d_rays.data_ptr()), len(second), C.c_void_p(d_out.data_ptr()))
        assert rc == _lib.OK, L.y4gpu_last_error(ctx)
        acc.synchronize()
        g = d_out.cpu().numpy()
        exp = _call(L, ctx, L.y4gpu_trace, _lib.TRACE_CLOSEST, second, np.empty(len(second), _lib.HIT_DTYPE))
        assert np.array_equal(g[:, 0], exp["t"]) and np.array_equal(g[:, 3].view(np.int32), exp["prim"])
    # sizes beyond the static scratch are rejected with a clear error; nothing is written
    big = np.zeros((_lib.NOALLOC_MAX_RAYS + 1, 8), np.float32)
    out = np.full(len(big), -7, _lib.HIT_DTYPE)
    rc = L.y4gpu_trace_noalloc(ctx, _lib.TRACE_CLOSEST, C.c_void_p(big.ctypes.data), len(big), C.c_void_p(out.ctypes.data))
    assert rc == _lib.ERR_ARG and b"Y4GPU_NOALLOC_MAX_RAYS" in L.y4gpu_last_error(ctx)
    assert (out["prim"] == -7).all()
    assert L.y4gpu_trace_noalloc(ctx, _lib.TRACE_TSHADOW, C.c_void_p(big.ctypes.data), 16, C.c_void_p(out.ctypes.data)) == _lib.ERR_ARG
    acc.close()
