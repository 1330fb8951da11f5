// Sample sequences and approximated maths of the reference, restated for the device.
//
// Everything here must reproduce the reference's numbers so that renders are comparable
// sample for sample (BASELINE.json north_star).  References (libYafaRay tree):
//   riVdC / riS / riLp / fnv32ABuf      include/sampler/sample.h:103-151
//   Halton::setStart / getNext          include/sampler/halton.h:54-81
//   Halton::lowDiscrepancySampling      src/sampler/halton.cc:421-441 (+ Faure tables :26-405)
//   math::sin / cos (FAST_TRIG)         include/math/math.h:218-250
//   math::exp2 / log2 / exp (FAST_MATH) include/math/math.h:102-142,190-197
//   sample::cosHemisphere               include/sampler/sample.h:45-54
//   Vec3::createCoordsSystem            include/geometry/vector.h:262-276
//   math::addMod1                       include/math/math.h:331-336
//
// The reference mixes `long double` constants into float expressions (x87 80-bit
// intermediates).  The device has no 80-bit type; those intermediates are evaluated in
// fp64 (Y4_WIDE).  The only observable difference is a last-bit change of the float result
// when the exact product lies within ~2^-53 of a rounding boundary (probability ~2^-29 per
// operation); tests/test_pt_cpu.py quantifies it against KATs taken from the reference.
#pragma once
#include "y4_common.cuh"

#ifndef Y4_WIDE
#define Y4_WIDE double
#endif

#define Y4_PI ((Y4_WIDE) 3.14159265358979323846264338327950288L)
#define Y4_2PI ((Y4_WIDE) 6.28318530717958647692528676655900577L)
#define Y4_PI_2 ((Y4_WIDE) 1.57079632679489661923132169163975144L)
#define Y4_PI_4 ((Y4_WIDE) 0.785398163397448309615660845819875721L)   /* math::div_pi_by_4 */
#define Y4_1_PI ((Y4_WIDE) 0.31830988618379067153776752674502872L)
#define Y4_1_2PI ((Y4_WIDE) 0.15915494309189533576888376337251L)
#define Y4_4_PI ((Y4_WIDE) 1.2732395447351626861510701069801L)
#define Y4_4_PI2 ((Y4_WIDE) 0.40528473456935108577551785283891L)
#define Y4_LOG2E ((Y4_WIDE) 1.4426950408889634073599246810019L)
#define Y4_2M32 ((Y4_WIDE) 0.00000000023283064365386962890625L)   /* math::sample_mult_ratio */

Y4_HD float y4_clamp01(float v) { return fmaxf(0.f, fminf(1.f, v)); }   // std::max(0.f, std::min(1.f, v))

Y4_HD float y4_ri_vdc(uint32_t bits, uint32_t r)
{
	bits = (bits << 16) | (bits >> 16);
	bits = ((bits & 0x00ff00ffu) << 8) | ((bits & 0xff00ff00u) >> 8);
	bits = ((bits & 0x0f0f0f0fu) << 4) | ((bits & 0xf0f0f0f0u) >> 4);
	bits = ((bits & 0x33333333u) << 2) | ((bits & 0xccccccccu) >> 2);
	bits = ((bits & 0x55555555u) << 1) | ((bits & 0xaaaaaaaau) >> 1);
	return y4_clamp01((float) ((double) (bits ^ r) * (double) Y4_2M32));   // exact: power-of-two scale, one rounding
}
Y4_HD float y4_ri_s(uint32_t i, uint32_t r)
{
	for(uint32_t v = 1u << 31; i; i >>= 1, v ^= v >> 1)
		if(i & 1u) r ^= v;
	return y4_clamp01((float) ((double) r * (double) Y4_2M32));
}
// Larcher-Pillichshammer radical inverse, sample::riLp (sample.h): "for(v = 1 << 31; i; i >>= 1, v |= v >> 1) if(i & 1) r ^= v".
// Bit k of i flips the top k + 1 bits of r, so bit 31 - j of the result flips parity(i >> j) times: a suffix-XOR scan
// (five shift-xor steps) and one bit reversal instead of up to 32 loop trips (pinned against the loop in tests/test_pt_cpu.py).
Y4_HD uint32_t y4_brev32(uint32_t x)
{
#if defined(__CUDA_ARCH__)
	return __brev(x);
#else
	x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1); x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
	x = ((x >> 4) & 0x0f0f0f0fu) | ((x & 0x0f0f0f0fu) << 4); x = ((x >> 8) & 0x00ff00ffu) | ((x & 0x00ff00ffu) << 8);
	return (x >> 16) | (x << 16);
#endif
}
Y4_HD float y4_ri_lp(uint32_t i, uint32_t r)
{
	i ^= i >> 1; i ^= i >> 2; i ^= i >> 4; i ^= i >> 8; i ^= i >> 16;
	r ^= y4_brev32(i);
	return y4_clamp01((float) ((double) r * (double) Y4_2M32));
}
Y4_HD uint32_t y4_fnv32a(uint32_t value)
{
	uint32_t hash = 0x811c9dc5u;
	for(int k = 0; k < 4; ++k) { hash ^= (value >> (8 * k)) & 0xffu; hash *= 0x01000193u; }   // little-endian byte order
	return hash;
}

// ---- incremental Halton (fp64 recurrence) ------------------------------------------------------
struct Y4Halton { uint32_t base; double inv_base, value; };

Y4_HD Y4Halton y4_halton_start(uint32_t base, uint32_t start)
{
	Y4Halton h; h.base = base; h.inv_base = 1.0 / (double) base; h.value = 0.0;
	double factor = h.inv_base;
	while(start > 0)
	{
		h.value = xdadd(h.value, xdmul((double) (start % base), factor));
		start /= base;
		factor = xdmul(factor, h.inv_base);
	}
	return h;
}
Y4_HD float y4_halton_next(Y4Halton &h)
{
	const double r = xdsub(0.9999999999, h.value);
	if(h.inv_base < r) h.value = xdadd(h.value, h.inv_base);
	else
	{
		double hh = 0.0, hcur = h.inv_base;
		while(hcur >= r) { hh = hcur; hcur = xdmul(hcur, h.inv_base); }
		h.value = xdadd(h.value, xdsub(xdadd(hh, hcur), 1.0));
	}
	return y4_clamp01((float) h.value);
}

// Base-2 member of the sequence in closed form: the q-th y4_halton_next after y4_halton_start(2, start)
// is the radical inverse of start + q.  Every intermediate of the recurrence is a dyadic rational with at
// most 33 fractional bits (exact in fp64), and the 1e-10 guard in `r` never flips a comparison because the
// compared quantities differ by 0 or by at least 2^-33 (1.16e-10).  tests/test_pt_cpu.py checks it against
// the recurrence above (which is pinned to the reference by KATs).
Y4_HD float y4_halton2_at(uint32_t start, uint32_t q)
{
	uint64_t n = (uint64_t) start + (uint64_t) q, r = 0;
#if defined(__CUDA_ARCH__)
	r = __brevll(n);
#else
	for(int k = 0; k < 64; ++k) { r = (r << 1) | (n & 1ull); n >>= 1; }
#endif
	return y4_clamp01((float) ((double) r * 5.421010862427522170037264e-20));     // 2^-64, exact scale
}

// ---- Faure-permuted Halton, 50 dimensions -------------------------------------------------------
// Tables are generated (Faure's recursive construction, y4_faure.h) rather than transcribed;
// the generated permutations and the 9-digit reciprocals are pinned by KATs from the reference.
struct Y4LdsTables
{
	const uint8_t *sigma;       // concatenated permutations
	const uint16_t *offset;     // [50] start of dimension d in sigma
	const uint16_t *prime;      // [50] base of dimension d   (1, 2, 3, 5, 7, ...)
	const double *inv_prime;    // [50] reciprocal truncated to 9 decimals, as double
	const double *sigma_d;      // sigma as doubles
	const unsigned long long *magic;   // [50] floor(2^64 / prime) + 1
};

// the reference's loop as written (Halton::lowDiscrepancySampling, halton.cc:403-441); kept as the
// cross-check of the fast form below
Y4_HD double y4_lds_loop(const Y4LdsTables &tb, int dim, uint32_t n)
{
	double value = 0.0;
	const uint8_t *sigma = tb.sigma + tb.offset[dim];
	const uint32_t base = tb.prime[dim];
	const double f = tb.inv_prime[dim];
	double dn = (double) n;
	double factor = f;
	while(n > 0)
	{
		value = xdadd(value, xdmul((double) sigma[n % base], factor));
		dn = xdmul(dn, f);
		n = (uint32_t) dn;
		factor = xdmul(factor, f);
	}
	return value;
}

// Same arithmetic, cheaper digit extraction: n % base through an exact multiply-high (no integer
// division), permutation values read as doubles.
Y4_HD double y4_lds(const Y4LdsTables &tb, int dim, uint32_t n)
{
	double value = 0.0;
	const double *sigma = tb.sigma_d + tb.offset[dim];
	const uint32_t base = tb.prime[dim];
	const unsigned long long magic = tb.magic[dim];
	const uint32_t m_lo = (uint32_t) magic, m_hi = (uint32_t) (magic >> 32);
	const double f = tb.inv_prime[dim];
	double dn = (double) n;
	double factor = f;
	while(n > 0)
	{
		const uint32_t q = (uint32_t) (((uint64_t) n * m_hi + (((uint64_t) n * m_lo) >> 32)) >> 32);     // n / base
		value = xdadd(value, xdmul(sigma[n - q * base], factor));
		dn = xdmul(dn, f);
		n = (uint32_t) dn;
		factor = xdmul(factor, f);
	}
	return value;
}

// ---- FAST_TRIG / FAST_MATH -----------------------------------------------------------------------
Y4_HD float y4_fast_sin(float x)
{
	const float two_pi_f = (float) Y4_2PI;
	if((Y4_WIDE) x > Y4_2PI || (Y4_WIDE) x < -Y4_2PI) x = xsub(x, xmul((float) ((int) xmul(x, (float) Y4_1_2PI)), two_pi_f));
	if((Y4_WIDE) x < -Y4_PI) x = xadd(x, two_pi_f);
	else if((Y4_WIDE) x > Y4_PI) x = xsub(x, two_pi_f);
	const float a = (float) (Y4_4_PI * (Y4_WIDE) x);
	const float b = (float) ((Y4_4_PI2 * (Y4_WIDE) x) * (Y4_WIDE) fabsf(x));
	x = xsub(a, b);
	const float result = xadd(xmul(0.225f, xsub(xmul(x, fabsf(x)), x)), x);
	if(result <= -1.f) return -1.f;
	else if(result >= 1.f) return 1.f;
	return result;
}
Y4_HD float y4_fast_cos(float x) { return y4_fast_sin(xadd(x, (float) Y4_PI_2)); }

Y4_HD float y4_polyexp(float x)
{
	return xadd(xmul(x, xadd(xmul(x, xadd(xmul(x, xadd(xmul(x, xadd(xmul(x, 1.8775767e-3f), 8.9893397e-3f)), 5.5826318e-2f)), 2.4015361e-1f)), 6.9315308e-1f)), 9.9999994e-1f);
}
Y4_HD float y4_fast_exp2(float x)
{
	x = fminf(x, 129.00000f);
	x = fmaxf(x, -126.99999f);
	const int ipart = (int) xsub(x, 0.5f);
	const float fpart = xsub(x, (float) ipart);
	const float expipart = y4_u2f((uint32_t) ((ipart + 127) << 23));
	return xmul(expipart, y4_polyexp(fpart));
}
Y4_HD float y4_fast_exp(float a) { return y4_fast_exp2((float) (Y4_LOG2E * (Y4_WIDE) a)); }
// math::log2 / math::pow with FAST_MATH (math.h:125-141,172-179): exponent bits + a degree-5 polynomial on the mantissa
Y4_HD float y4_polylog(float x)
{
	return xadd(xmul(x, xadd(xmul(x, xadd(xmul(x, xadd(xmul(x, xadd(xmul(x, -3.4436006e-2f), 3.1821337e-1f)), -1.2315303f)), 2.5988452f)), -3.3241990f)), 3.1157899f);
}
Y4_HD float y4_fast_log2(float x)
{
	const uint32_t i = y4_f2u(x);
	const float e = (float) ((int) ((i & 0x7F800000u) >> 23) - 127);
	const float m = y4_u2f((i & 0x7FFFFFu) | 0x3F800000u);
	return xadd(xmul(y4_polylog(m), xsub(m, 1.0f)), e);
}
Y4_HD float y4_fast_pow(float a, float b) { return y4_fast_exp2(xmul(y4_fast_log2(a), b)); }

// ---- geometry helpers ----------------------------------------------------------------------------
Y4_HD V3 xneg(V3 a) { return v3(-a.x, -a.y, -a.z); }
// Vec3::normalize(), vector.h:201-210
Y4_HD V3 xnormalize(V3 a)
{
	float len = xdot(a, a);
	if(len != 0.f)
	{
		len = xdiv(1.f, xsqrt(len));
		a = v3(xmul(a.x, len), xmul(a.y, len), xmul(a.z, len));
	}
	return a;
}
Y4_HD void y4_coords_system(const V3 n, V3 &u, V3 &v)
{
	if(n.x == 0.f && n.y == 0.f)
	{
		u = (n.z < 0.f) ? v3(-1.f, 0.f, 0.f) : v3(1.f, 0.f, 0.f);
		v = v3(0.f, 1.f, 0.f);
	}
	else
	{
		const float d = xdiv(1.f, xsqrt(xadd(xmul(n.y, n.y), xmul(n.x, n.x))));
		u = v3(xmul(n.y, d), xmul(-n.x, d), 0.f);
		v = xcross(n, u);
	}
}
// (ru*cos + rv*sin)*sqrt(1-s1) + n*sqrt(s1); Vec3*float is {f*v[i]} (vector.h:118-121)
Y4_HD V3 y4_cos_hemisphere(const V3 n, const V3 ru, const V3 rv, float s1, float s2)
{
	if(s1 >= 1.0f) return n;
	const float z1 = s1;
	const float z2 = (float) ((Y4_WIDE) s2 * Y4_2PI);
	const V3 a = xadd3(xscale(ru, y4_fast_cos(z2)), xscale(rv, y4_fast_sin(z2)));
	return xadd3(xscale(a, xsqrt(xsub(1.f, z1))), xscale(n, xsqrt(z1)));
}
Y4_HD float y4_add_mod1(float a, float b) { const float s = xadd(a, b); return s > 1.f ? xsub(s, 1.f) : s; }

// Per-path hash RNG used ONLY where the reference draws from its rand()-seeded per-tile
// generator (Russian roulette, integrator_path_tracer.cc:249-255), which is not reproducible
// across runs/threads in the reference either (documented deviation: statistical parity).
Y4_HD float y4_hash_uniform(uint32_t a, uint32_t b, uint32_t c)
{
	uint32_t h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u ^ (c + 0x165667B1u) * 0xC2B2AE3Du;
	h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
	return (float) (h >> 8) * (1.f / 16777216.f);
}
