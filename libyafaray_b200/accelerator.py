"""Host-side mirror of the reference's Accelerator plugin interface for the B200 back end.

Reference interface (libYafaRay tree): include/accelerator/accelerator.h:53-69 --
``Accelerator::factory(logger, primitives, params)`` selected by ``params["type"]``, and the
virtuals ``intersect / intersectS / intersectTs / getBound`` plus the wrappers
``isShadowed`` (src/accelerator/accelerator.cc:32-94).  The C++ subclass a libYafaRay
maintainer would add is shown in INTEGRATION.md; this module mirrors the same names,
argument meaning and error behaviour over the same C-ABI (include/y4gpu.h) so the parity
tests read like tests of the reference accelerator.  Batched entry points take N rays at once
(the GPU-friendly form); single-ray calls are 1-ray batches through the same kernels.

There is no CPU fallback: construction fails when liby4gpu.so or a CUDA device is missing.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib
from ._lib import HIT_DTYPE, TSHIT_DTYPE, Y4Error

ACCELERATOR_TYPE = "yafaray-b200-bvh8"  # value of the `scene_accelerator` render parameter


def factory(primitives: np.ndarray, params: Optional[dict] = None) -> "AcceleratorB200":
    """Mirror of Accelerator::factory (accelerator.cc:32-53) for ``type == "yafaray-b200-bvh8"``.

    Unlike the reference factory there is no silent kd-tree fallback: an unknown type or a
    failed build raises."""
    params = dict(params or {})
    typ = params.get("type", ACCELERATOR_TYPE)
    if typ != ACCELERATOR_TYPE:
        raise Y4Error(f"accelerator type '{typ}' is not provided by this back end (only '{ACCELERATOR_TYPE}')")
    return AcceleratorB200(primitives, flags=params.get("flags"), leaf_max=int(params.get("leaf_max", 0)), device=int(params.get("device", 0)))


def _as_ptr(a) -> C.c_void_p:
    """numpy array -> host pointer; torch CUDA tensor / int -> device pointer."""
    if isinstance(a, np.ndarray):
        return C.c_void_p(a.ctypes.data)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(int(a))


class AcceleratorB200:
    """BVH8 accelerator resident in HBM; primitives = float32 [n_tri, 9] world-space triangles
    in the order of the vector handed to Accelerator::factory (primitive index == row)."""

    def __init__(self, primitives: Optional[np.ndarray], flags: Optional[np.ndarray] = None, leaf_max: int = 0, device: int = 0):
        self._L = _lib.lib()
        self._ctx = C.c_void_p()
        rc = self._L.y4gpu_create(device, C.byref(self._ctx))
        if rc != _lib.OK:
            msg = self._L.y4gpu_last_error(None)
            self._ctx = C.c_void_p()
            raise Y4Error(f"y4gpu_create failed (rc={rc}): {msg.decode() if msg else ''}")
        if primitives is not None:
            self.rebuild(primitives, flags, leaf_max)

    @classmethod
    def from_instances(cls, meshes, instances, mesh_flags=None, leaf_max: int = 0, device: int = 0) -> "AcceleratorB200":
        """Two-level accelerator (yafaray_addInstance: PrimitiveInstance / ObjectInstance).  meshes: [n_i, 9] fp32 arrays in object
        space; instances: (mesh index, 3x4 obj_to_world rows of the reference's Matrix4, or None for a plain world-space object),
        in primitive order.  Hits are bit-identical to the flattened scene's (include/y4gpu.h: y4gpu_upload_instanced)."""
        acc = cls(None, device=device)
        acc.rebuild_instances(meshes, instances, mesh_flags, leaf_max)
        return acc

    def rebuild_instances(self, meshes, instances, mesh_flags=None, leaf_max: int = 0) -> None:
        ms = [np.ascontiguousarray(m, dtype=np.float32).reshape(-1, 9) for m in meshes]
        fs = [None if (mesh_flags is None or mesh_flags[i] is None) else np.ascontiguousarray(mesh_flags[i], dtype=np.uint32).reshape(-1) for i in range(len(ms))]
        M = (_lib.Mesh * max(1, len(ms)))()
        for i, m in enumerate(ms):
            if fs[i] is not None and len(fs[i]) != len(m):
                raise Y4Error("mesh flags must have one word per triangle")
            M[i].v_xyz, M[i].flags, M[i].n_tri = m.ctypes.data, (fs[i].ctypes.data if fs[i] is not None else None), len(m)
        I = (_lib.Instance * max(1, len(instances)))()
        ident = (1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0)
        for i, (mesh, mat) in enumerate(instances):
            I[i].mesh, I[i].identity = int(mesh), 1 if mat is None else 0
            I[i].obj_to_world[:] = ident if mat is None else [float(x) for x in np.asarray(mat, np.float32).reshape(12)]
        self.n_primitives = int(sum(len(ms[int(mesh)]) for mesh, _ in instances))
        _lib.check(self._ctx, self._L.y4gpu_upload_instanced(self._ctx, M, len(ms), I, len(instances)), "y4gpu_upload_instanced")
        bp = _lib.BuildParams()
        bp.leaf_max = leaf_max
        _lib.check(self._ctx, self._L.y4gpu_build(self._ctx, C.byref(bp)), "y4gpu_build")

    def rebuild(self, primitives: np.ndarray, flags: Optional[np.ndarray] = None, leaf_max: int = 0) -> None:
        """Upload a new primitive list into the same context and build its BVH8 (device arrays are reused when the
        size is unchanged: the per-frame path of an animation or a repeated render)."""
        prims = np.ascontiguousarray(primitives, dtype=np.float32).reshape(-1, 9)
        self.n_primitives = int(prims.shape[0])
        fl = None
        if flags is not None:
            fl = np.ascontiguousarray(flags, dtype=np.uint32).reshape(-1)
            if fl.shape[0] != self.n_primitives:
                raise Y4Error("flags must have one word per primitive")
        _lib.check(self._ctx, self._L.y4gpu_upload_triangles(self._ctx, _as_ptr(prims), _as_ptr(fl) if fl is not None else None, self.n_primitives), "y4gpu_upload_triangles")
        bp = _lib.BuildParams()
        bp.leaf_max = leaf_max
        _lib.check(self._ctx, self._L.y4gpu_build(self._ctx, C.byref(bp)), "y4gpu_build")

    # -- lifetime -------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_ctx", None) and self._ctx.value:
            self._L.y4gpu_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def ctx(self) -> C.c_void_p:
        return self._ctx

    # -- Accelerator virtuals (batched) ------------------------------------------------------
    def getBound(self) -> np.ndarray:
        """Accelerator::getBound: tree bound incl. the reference's 0.1 % inflation -> [lo xyz, hi xyz]."""
        b = (C.c_float * 6)()
        _lib.check(self._ctx, self._L.y4gpu_get_bound(self._ctx, C.byref(b)), "y4gpu_get_bound")
        return np.array(list(b), dtype=np.float32)

    def _trace(self, kind: int, rays, out=None, counted: bool = False):
        n = int(rays.shape[0])
        if isinstance(rays, np.ndarray):
            rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
            if out is None:
                out = np.empty(n, dtype=HIT_DTYPE)
        elif out is None:
            raise Y4Error("device-resident rays need a device-resident `out` buffer")
        fn = self._L.y4gpu_trace_counted if counted else self._L.y4gpu_trace
        _lib.check(self._ctx, fn(self._ctx, kind, _as_ptr(rays), n, _as_ptr(out)), "y4gpu_trace")
        return out

    def intersect(self, rays, out=None, counted: bool = False):
        """Closest hit (Accelerator::intersect(ray, t_max) with the wrapper's t_max rule,
        accelerator.cc:55-67).  rays: [n, 8] (o, tmin, d, tmax); returns HIT_DTYPE records.
        Host arrays: synchronous.  Device tensors (rays and `out`): queued on this accelerator's stream and NOT waited
        for -- call synchronize(), or set_stream() to the stream that produces the rays / consumes the hits."""
        return self._trace(_lib.TRACE_CLOSEST, rays, out, counted)

    def isShadowed(self, rays, out=None, counted: bool = False):
        """Any-hit shadow query (Accelerator::isShadowed(ray, shadow_bias) -> intersectS,
        accelerator.cc:69-80); record.prim >= 0 means shadowed (prim = the blocker found)."""
        return self._trace(_lib.TRACE_SHADOW, rays, out, counted)

    intersectS = isShadowed

    def isShadowedTransparent(self, rays: np.ndarray, max_depth: int, mat_transparency: np.ndarray):
        """Transparent shadows (Accelerator::isShadowed(ray, max_depth, bias, camera) -> intersectTs,
        accelerator.cc:82-94).  mat_transparency: [n_mat, 3] Material::getTransparency colours."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        mt = np.ascontiguousarray(mat_transparency, dtype=np.float32).reshape(-1, 3)
        out = np.empty(rays.shape[0], dtype=TSHIT_DTYPE)
        _lib.check(self._ctx, self._L.y4gpu_trace_tshadow(self._ctx, _as_ptr(rays), rays.shape[0], int(max_depth), _as_ptr(mt), mt.shape[0], _as_ptr(out)), "y4gpu_trace_tshadow")
        return out

    intersectTs = isShadowedTransparent

    def set_stream(self, cuda_stream: int) -> None:
        """Run this accelerator's kernels/copies on the caller's CUDA stream (0/None = its own)."""
        _lib.check(self._ctx, self._L.y4gpu_set_stream(self._ctx, C.c_void_p(int(cuda_stream or 0))), "y4gpu_set_stream")

    def set_ray_sort(self, mode: int) -> None:
        """Ray order of large batches: 0 as given, 1 always sorted by origin cell + direction class, 2 automatic (default).
        Results are bit-identical in every mode (include/y4gpu.h: y4gpu_set_ray_sort)."""
        _lib.check(self._ctx, self._L.y4gpu_set_ray_sort(self._ctx, int(mode)), "y4gpu_set_ray_sort")

    # -- diagnostics -------------------------------------------------------------------------
    def synchronize(self) -> None:
        _lib.check(self._ctx, self._L.y4gpu_synchronize(self._ctx), "y4gpu_synchronize")

    def stats(self) -> dict:
        s = _lib.Stats()
        _lib.check(self._ctx, self._L.y4gpu_get_stats(self._ctx, C.byref(s)), "y4gpu_get_stats")
        return {k: getattr(s, k) for k, _ in s._fields_}
