"""The device restatement of the reference's samplers (csrc/y4_pt_sampling.cuh, host build: tests/emu) against the UNMODIFIED
reference in bulk: oracle/_ref/ref_kat_bulk evaluates Halton::lowDiscrepancySampling, Halton::getNext and riVdC / riS / riLp on
hundreds of thousands of inputs here (3 million each in the run recorded in profiles/r2_v16_fuzz_vs_reference.log), bit for bit.
The committed KATs (tests/golden/kat_sampling.json) cover a few hundred; this needs the reference build, i.e. this container."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import helpers

REF = os.path.join(helpers.ROOT, "oracle", "_ref", "ref_kat_bulk")
pytestmark = pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/ref_kat_bulk not built")
N = 400_000


def _ref(mode, arr, dtype, per):
    out = subprocess.run([REF, mode], input=arr.tobytes(), capture_output=True, check=True).stdout
    a = np.frombuffer(out, dtype)
    return a.reshape(-1, per) if per > 1 else a


def _indices(rng, n):
    """32-bit indices of every magnitude, plus the small ones and the ones just below 2^32 (wrap-around of the offsets)."""
    v = rng.integers(0, 2**32, n, dtype=np.uint64) >> rng.integers(0, 32, n).astype(np.uint64)
    v[: n // 10] = rng.integers(0, 4096, n // 10)
    v[n // 10: n // 5] = 2**32 - 1 - rng.integers(0, 4096, n // 10)
    return v.astype(np.uint32)


def test_low_discrepancy_sampling_bit_exact_in_bulk():
    rng, E = np.random.default_rng(3), helpers.emu()
    inp = np.ascontiguousarray(np.stack([rng.integers(1, 50, N, dtype=np.uint32), _indices(rng, N)], 1), np.uint32)
    want = _ref("lds", inp, np.float64, 1)
    for fn in (E.emu_lds_bulk, E.emu_lds_loop_bulk):          # the device's table form and the plain digit loop
        got = np.empty(N, np.float64)
        fn(inp.ctypes.data_as(C.c_void_p), C.c_uint64(N), got.ctypes.data_as(C.c_void_p))
        assert np.array_equal(want.view(np.uint64), got.view(np.uint64))


def test_halton_sequences_bit_exact_in_bulk():
    rng, E = np.random.default_rng(4), helpers.emu()
    base = rng.choice(np.array([2, 3, 5], np.uint32), N)
    inp = np.ascontiguousarray(np.stack([base, _indices(rng, N)], 1), np.uint32)
    want = _ref("halton", inp, np.float32, 4)
    got = np.empty((N, 4), np.float32)
    E.emu_halton_bulk(inp.ctypes.data_as(C.c_void_p), C.c_uint64(N), got.ctypes.data_as(C.c_void_p))
    assert np.array_equal(want.view(np.uint32), got.view(np.uint32))
    E.emu_halton2_at_bulk(inp.ctypes.data_as(C.c_void_p), C.c_uint64(N), got.ctypes.data_as(C.c_void_p))       # the closed form used for base 2
    m = base == 2
    assert np.array_equal(want.view(np.uint32)[m], got.view(np.uint32)[m])


def test_scrambled_radical_inverses_bit_exact_in_bulk():
    rng, E = np.random.default_rng(5), helpers.emu()
    r = rng.integers(0, 2**32, N, dtype=np.uint64).astype(np.uint32)
    r[::7] = 0
    inp = np.ascontiguousarray(np.stack([_indices(rng, N), r], 1), np.uint32)
    want = _ref("ri", inp, np.float32, 3)
    got = np.empty((N, 3), np.float32)
    E.emu_ri_bulk(inp.ctypes.data_as(C.c_void_p), C.c_uint64(N), got.ctypes.data_as(C.c_void_p))
    assert np.array_equal(want.view(np.uint32), got.view(np.uint32))
