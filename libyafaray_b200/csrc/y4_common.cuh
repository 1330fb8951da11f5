// Shared types and exact-arithmetic helpers for the B200 ray/path back end.
//
// All per-element logic in csrc/*.cuh is written as Y4_HD functions so that the
// same source compiles (a) into sm_100a kernels (the product, y4_*.cu) and (b)
// into a host-side *unit-test* emulation (tests/emu, g++ -DY4_HOST_EMU) used only
// to debug the builder/traversal logic where no GPU exists.  The emulation is
// never linked into liby4gpu.so.
#pragma once
#include <stdint.h>
#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define Y4_HD __host__ __device__ __forceinline__
#define Y4_D __device__ __forceinline__
#else
#define Y4_HD inline
#define Y4_D inline
#endif

// ---- exact (never-contracted) fp32 arithmetic -------------------------------------------
// The reference is x86-64 SSE2 code compiled without FMA contraction.
// Everything whose bits must match the reference (triangle test, shadow-ray origin advance,
// camera rays, sampling maths) goes through these so nvcc can never fuse a*b+c.
#if defined(__CUDA_ARCH__)
Y4_HD float xmul(float a, float b) { return __fmul_rn(a, b); }
Y4_HD float xadd(float a, float b) { return __fadd_rn(a, b); }
Y4_HD float xsub(float a, float b) { return __fsub_rn(a, b); }
Y4_HD float xdiv(float a, float b) { return __fdiv_rn(a, b); }
Y4_HD float xsqrt(float a) { return __fsqrt_rn(a); }
Y4_HD double xdmul(double a, double b) { return __dmul_rn(a, b); }
Y4_HD double xdadd(double a, double b) { return __dadd_rn(a, b); }
Y4_HD double xdsub(double a, double b) { return __dsub_rn(a, b); }
Y4_HD double xddiv(double a, double b) { return __ddiv_rn(a, b); }
#else
// host build: compiled with -ffp-contract=off (see tests/emu/Makefile)
Y4_HD float xmul(float a, float b) { return a * b; }
Y4_HD float xadd(float a, float b) { return a + b; }
Y4_HD float xsub(float a, float b) { return a - b; }
Y4_HD float xdiv(float a, float b) { return a / b; }
Y4_HD float xsqrt(float a) { return sqrtf(a); }
Y4_HD double xdmul(double a, double b) { return a * b; }
Y4_HD double xdadd(double a, double b) { return a + b; }
Y4_HD double xdsub(double a, double b) { return a - b; }
Y4_HD double xddiv(double a, double b) { return a / b; }
#endif

#if !defined(__CUDACC__)
struct uint2 { uint32_t x, y; };
struct alignas(16) float4 { float x, y, z, w; };
#endif

struct V3 { float x, y, z; };

Y4_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
// reference include/geometry/vector.h:108-111 (dot), :148-151 (cross), :61-62 (length)
Y4_HD float xdot(V3 a, V3 b) { return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z)); }
Y4_HD V3 xcross(V3 a, V3 b)
{
	return v3(xsub(xmul(a.y, b.z), xmul(a.z, b.y)), xsub(xmul(a.z, b.x), xmul(a.x, b.z)), xsub(xmul(a.x, b.y), xmul(a.y, b.x)));
}
Y4_HD V3 xsub3(V3 a, V3 b) { return v3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
Y4_HD V3 xadd3(V3 a, V3 b) { return v3(xadd(a.x, b.x), xadd(a.y, b.y), xadd(a.z, b.z)); }
Y4_HD V3 xscale(V3 a, float s) { return v3(xmul(s, a.x), xmul(s, a.y), xmul(s, a.z)); }
Y4_HD float xlen(V3 a) { return xsqrt(xdot(a, a)); }

// ---- C-ABI mirrored PODs (include/y4gpu.h) ---------------------------------------------
struct alignas(16) Y4Ray { float ox, oy, oz, tmin, dx, dy, dz, tmax; };   // 32 B, two float4 loads
struct alignas(16) Y4Hit { float t, bu, bv; int32_t prim; };               // 16 B, one float4 store

// Per-triangle visibility/material word.
enum : uint32_t
{
	Y4_TRI_VIS_CLOSEST = 1u,       // prim & material visibility in {Normal, VisibleNoShadows}
	Y4_TRI_VIS_SHADOW = 2u,        // prim & material visibility in {Normal, InvisibleShadowsOnly}
	Y4_TRI_VIS_TSHADOW = 4u,       // material visibility only (intersectTs, accelerator_kdtree.cc:1022)
	Y4_TRI_TRANSPARENT = 8u,       // material isTransparent()
	Y4_TRI_MAT_SHIFT = 8u          // material index in bits 8..31
};
#define Y4_TRI_DEFAULT_FLAGS (Y4_TRI_VIS_CLOSEST | Y4_TRI_VIS_SHADOW | Y4_TRI_VIS_TSHADOW)

enum { Y4_TRACE_CLOSEST = 0, Y4_TRACE_SHADOW = 1, Y4_TRACE_TSHADOW = 2 };

// Device triangle record, 48 B = three 128-bit loads, stored in BVH leaf order.
// e1/e2/eps are the reference's per-test temporaries (primitive_triangle.cc:47-49)
// hoisted to build time; they are fp32-deterministic so the bits are unchanged.
struct alignas(16) Y4Tri
{
	float v0x, v0y, v0z, eps;
	float e1x, e1y, e1z; uint32_t prim;
	float e2x, e2y, e2z; uint32_t flags;
};

Y4_HD uint32_t y4_f2u(float f) { union { float f; uint32_t u; } c; c.f = f; return c.u; }
Y4_HD float y4_u2f(uint32_t u) { union { float f; uint32_t u; } c; c.u = u; return c.f; }
