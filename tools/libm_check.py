"""Compare the fp32 atanf / asinf / acosf / atan2f restatements of csrc/y4_libm.cuh (host build: tests/emu) with the C library
of this machine, bit for bit.  usage: python tools/libm_check.py [stride]   (stride 1 = every float; default 97)"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = C.CDLL(os.path.join(ROOT, "tests", "emu", "libemu.so"))
lib.emu_libm_mismatch.restype = C.c_uint64
lib.emu_libm_mismatch.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
lib.emu_libm_atan2_mismatch.restype = C.c_uint64
lib.emu_libm_atan2_mismatch.argtypes = [C.c_uint64, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]


def main():
    stride = int(sys.argv[1]) if len(sys.argv) > 1 else 97
    total = 0
    for fn, name, hi in ((0, "atanf", 0x7f800001), (1, "asinf", 0x3f800010), (2, "acosf", 0x3f800010)):
        bad = C.c_uint32(0)
        t = time.time()
        n = lib.emu_libm_mismatch(fn, 0, hi, stride, C.byref(bad))
        total += n
        print(f"{name}: {2 * ((hi + stride - 1) // stride)} arguments, {n} mismatches" + (f" (first {bad.value:#010x})" if n else "") + f", {time.time() - t:.1f} s")
    by, bx = C.c_uint32(0), C.c_uint32(0)
    t = time.time()
    pairs = 2_000_000_000 // stride + 1000
    n = lib.emu_libm_atan2_mismatch(pairs, 7, C.byref(by), C.byref(bx))
    total += n
    print(f"atan2f: {pairs} pairs, {n} mismatches" + (f" (first y {by.value:#010x} x {bx.value:#010x})" if n else "") + f", {time.time() - t:.1f} s")
    return 1 if total else 0


if __name__ == "__main__":
    sys.exit(main())
