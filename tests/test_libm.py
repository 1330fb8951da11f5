"""The fp32 inverse trigonometric functions the device evaluates (csrc/y4_libm.cuh: atanf, atan2f, asinf, acosf restated from
glibc's flt-32 routines) against the C library of this machine, bit for bit.  The reference reaches these through std::atan2 /
std::atan / std::asin / std::acos (camera_angular.cc:66-71, math.h:253-267), so "same bits as this libm" is the parity bar;
`python tools/libm_check.py 1` runs the exhaustive version (profiles/r2_v11_libm_exhaustive.log)."""
import ctypes as C
import math

import numpy as np
import pytest

import helpers


def _lib():
    lib = helpers.emu()
    lib.emu_libm_mismatch.restype = C.c_uint64
    lib.emu_libm_mismatch.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
    lib.emu_libm_atan2_mismatch.restype = C.c_uint64
    lib.emu_libm_atan2_mismatch.argtypes = [C.c_uint64, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.emu_libm_eval.restype = C.c_float
    lib.emu_libm_eval.argtypes = [C.c_int, C.c_float, C.c_float]
    return lib


def test_one_argument_functions_match_libm_on_a_strided_sweep_of_all_floats():
    lib = _lib()
    for fn, hi in ((0, 0x7F800001), (1, 0x3F800010), (2, 0x3F800010)):      # atanf over every finite float, asinf / acosf past |x| = 1
        for stride, lo in ((211, 0), (1, 0x3EDFFF00), (1, 0x3F2FFF00), (1, 0x3F97FF00), (1, 0x401BFF00), (1, 0x3EFFFF00), (1, 0x3F799900), (1, 0x3F7FFF00)):
            bad = C.c_uint32(0)
            top = hi if stride > 1 else min(hi, lo + 0x200)       # every float around the break points of the argument reductions
            assert lib.emu_libm_mismatch(fn, lo, top, stride, C.byref(bad)) == 0, (fn, hex(bad.value))


def test_atan2f_matches_libm_on_random_pairs_and_special_values():
    lib = _lib()
    by, bx = C.c_uint32(0), C.c_uint32(0)
    assert lib.emu_libm_atan2_mismatch(5_000_000, 12345, C.byref(by), C.byref(bx)) == 0, (hex(by.value), hex(bx.value))


def test_values_are_the_functions_they_claim_to_be():
    lib = _lib()
    rng = np.random.default_rng(5)
    for x in rng.uniform(-1.0, 1.0, 200):
        assert abs(lib.emu_libm_eval(1, x, 0.0) - math.asin(x)) < 3e-7 and abs(lib.emu_libm_eval(2, x, 0.0) - math.acos(x)) < 5e-7
    for y, x in rng.uniform(-4.0, 4.0, (200, 2)):
        assert abs(lib.emu_libm_eval(3, y, x) - math.atan2(np.float32(y), np.float32(x))) < 5e-7
        assert abs(lib.emu_libm_eval(0, y, 0.0) - math.atan(np.float32(y))) < 3e-7


@pytest.mark.gpu
def test_device_functions_have_the_bits_of_this_machines_libm():
    """The same source compiled by nvcc for sm_100a (no FMA contraction: every operation goes through __fmul_rn / __fadd_rn / ...)
    evaluated on the GPU through y4gpu_libm_eval, against atanf / asinf / acosf / atan2f of the host's C library."""
    from libyafaray_b200 import _lib as y4
    lib = _lib()
    lib.emu_libm_host.restype = None
    lib.emu_libm_host.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
    L = y4.lib()
    ctx = C.c_void_p()
    assert L.y4gpu_create(0, C.byref(ctx)) == 0
    try:
        rng = np.random.default_rng(11)
        for fn, hi in ((0, 0x7F800001), (1, 0x3F800010), (2, 0x3F800010), (3, 0)):
            if fn < 3:
                bits = np.arange(0, hi, 53, dtype=np.uint32)
                edges = np.concatenate([np.arange(c - 0x400, c + 0x400, dtype=np.uint32) for c in (0x3EE00000, 0x3F300000, 0x3F980000, 0x401C0000, 0x4C000000,
                                                                                                  0x3F000000, 0x3F79999A, 0x3F800000, 0x32000000, 0x31000000)])
                bits = np.concatenate([bits, edges[edges < hi]])
                a = np.concatenate([bits, bits | np.uint32(0x80000000)]).view(np.float32)
                b = np.zeros_like(a)
            else:
                n = 20_000_000
                a = rng.integers(0, 2**32, n, dtype=np.uint32)
                b = rng.integers(0, 2**32, n, dtype=np.uint32)
                scr = np.arange(n) % 3 != 0            # two thirds with screen-like magnitudes, as in the CPU test
                for v in (a, b):
                    e = (115 + (v >> 23) % 14).astype(np.uint32)
                    v[scr] = (v[scr] & np.uint32(0x807FFFFF)) | (e[scr] << np.uint32(23))
                a, b = a.view(np.float32), b.view(np.float32)
            a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
            want, got = np.empty_like(a), np.empty_like(a)
            lib.emu_libm_host(fn, a.ctypes.data, b.ctypes.data, want.ctypes.data, len(a))
            assert L.y4gpu_libm_eval(ctx, fn, a.ctypes.data, b.ctypes.data, got.ctypes.data, len(a)) == 0
            same = (want.view(np.uint32) == got.view(np.uint32)) | (np.isnan(want) & np.isnan(got))
            assert same.all(), (fn, int((~same).sum()), a[~same][:4], b[~same][:4])
    finally:
        L.y4gpu_destroy(ctx)
