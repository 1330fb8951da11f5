// Inverse trigonometric functions with the bits of the host C library the reference links against.
//
// The reference calls std::atan2 / std::atan / std::asin / std::acos on floats in a few per-ray places
// (AngularCamera::shootRay, camera_angular.cc:59-78; Texture::sphereMap through math::acos, texture.cc:117-132).
// Those resolve to glibc's single-precision routines, which CUDA's atan2f / asinf / acosf do not reproduce bit for bit.
// This image ships glibc 2.39, whose flt-32 versions (sysdeps/ieee754/flt-32/s_atanf.c, e_atan2f.c, e_asinf.c,
// e_acosf.c) are the classic fdlibm argument reductions with polynomial / rational kernels, plain fp32 arithmetic and
// no fused multiply-adds on x86-64.  They are restated here from the published algorithm; tests/test_libm.py compares
// every function with the box's own libm over the whole argument range (strided in the default run, exhaustive on request).
// A libm of another version (glibc >= 2.41 switched these four to correctly rounded CORE-MATH code) would show up there.
#pragma once
#include "y4_common.cuh"

// atanf: reduce to |x| < 7/16 around the break points 0.5, 1, 1.5, inf; odd polynomial of degree 23 split in two chains
Y4_HD float y4_libm_atanf(float x)
{
	const float hi0 = 4.6364760399e-01f, hi1 = 7.8539812565e-01f, hi2 = 9.8279368877e-01f, hi3 = 1.5707962513e+00f;     // atan(0.5), atan(1), atan(1.5), atan(inf)
	const float lo0 = 5.0121582440e-09f, lo1 = 3.7748947079e-08f, lo2 = 3.4473217170e-08f, lo3 = 7.5497894159e-08f;     // their low parts
	const float a0 = 3.3333334327e-01f, a1 = -2.0000000298e-01f, a2 = 1.4285714924e-01f, a3 = -1.1111110449e-01f, a4 = 9.0908870101e-02f,
	            a5 = -7.6918758452e-02f, a6 = 6.6610731184e-02f, a7 = -5.8335702866e-02f, a8 = 4.9768779427e-02f, a9 = -3.6531571299e-02f,
	            a10 = 1.6285819933e-02f;
	const uint32_t hx = y4_f2u(x), ix = hx & 0x7fffffffu;
	const bool neg = (hx >> 31) != 0;
	int id;
	if(ix >= 0x4c000000u)                                  // |x| >= 2^25
	{
		if(ix > 0x7f800000u) return xadd(x, x);
		const float r = xadd(hi3, lo3);
		return neg ? -r : r;
	}
	if(ix < 0x3ee00000u)                                   // |x| < 7/16
	{
		if(ix < 0x31000000u) return x;                     // |x| < 2^-29
		id = -1;
	}
	else
	{
		x = fabsf(x);
		if(ix < 0x3f980000u)                               // |x| < 19/16
		{
			if(ix < 0x3f300000u) { id = 0; x = xdiv(xsub(xmul(2.f, x), 1.f), xadd(2.f, x)); }
			else { id = 1; x = xdiv(xsub(x, 1.f), xadd(x, 1.f)); }
		}
		else
		{
			if(ix < 0x401c0000u) { id = 2; x = xdiv(xsub(x, 1.5f), xadd(1.f, xmul(1.5f, x))); }
			else { id = 3; x = xdiv(-1.f, x); }
		}
	}
	const float z = xmul(x, x), w = xmul(z, z);
	const float s1 = xmul(z, xadd(a0, xmul(w, xadd(a2, xmul(w, xadd(a4, xmul(w, xadd(a6, xmul(w, xadd(a8, xmul(w, a10)))))))))));
	const float s2 = xmul(w, xadd(a1, xmul(w, xadd(a3, xmul(w, xadd(a5, xmul(w, xadd(a7, xmul(w, a9)))))))));
	if(id < 0) return xsub(x, xmul(x, xadd(s1, s2)));
	const float hi = id == 0 ? hi0 : (id == 1 ? hi1 : (id == 2 ? hi2 : hi3)), lo = id == 0 ? lo0 : (id == 1 ? lo1 : (id == 2 ? lo2 : lo3));
	const float r = xsub(hi, xsub(xsub(xmul(x, xadd(s1, s2)), lo), x));
	return neg ? -r : r;
}

// atan2f: quadrant from the signs, atanf(|y / x|) in between
Y4_HD float y4_libm_atan2f(float y, float x)
{
	const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f, pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f;
	const uint32_t hx = y4_f2u(x), hy = y4_f2u(y), ix = hx & 0x7fffffffu, iy = hy & 0x7fffffffu;
	if(ix > 0x7f800000u || iy > 0x7f800000u) return xadd(x, y);
	if(hx == 0x3f800000u) return y4_libm_atanf(y);
	const int m = (int) ((hy >> 31) & 1u) | (int) ((hx >> 30) & 2u);
	if(iy == 0)
	{
		if(m < 2) return y;
		return m == 2 ? xadd(pi, tiny) : xsub(-pi, tiny);
	}
	if(ix == 0) return (hy >> 31) ? xsub(-pi_o_2, tiny) : xadd(pi_o_2, tiny);
	if(ix == 0x7f800000u)
	{
		if(iy == 0x7f800000u)
		{
			if(m == 0) return xadd(pi_o_4, tiny);
			if(m == 1) return xsub(-pi_o_4, tiny);
			if(m == 2) return xadd(xmul(3.f, pi_o_4), tiny);
			return xsub(xmul(-3.f, pi_o_4), tiny);
		}
		if(m == 0) return 0.f;
		if(m == 1) return -0.f;
		return m == 2 ? xadd(pi, tiny) : xsub(-pi, tiny);
	}
	if(iy == 0x7f800000u) return (hy >> 31) ? xsub(-pi_o_2, tiny) : xadd(pi_o_2, tiny);
	const int k = ((int) iy - (int) ix) >> 23;
	float z;
	if(k > 60) z = xadd(pi_o_2, xmul(0.5f, pi_lo));
	else if((hx >> 31) && k < -60) z = 0.f;
	else z = y4_libm_atanf(fabsf(xdiv(y, x)));
	if(m == 0) return z;
	if(m == 1) return y4_u2f(y4_f2u(z) ^ 0x80000000u);
	if(m == 2) return xsub(pi, xsub(z, pi_lo));
	return xsub(xsub(z, pi_lo), pi);
}

// asinf: x + x^3 p(x^2) below 0.5, pi/2 - 2 asin(sqrt((1 - |x|) / 2)) above, with a split square root
Y4_HD float y4_libm_asinf(float x)
{
	const float pio2_hi = 1.57079637050628662109375f, pio2_lo = -4.37113900018624283e-8f, pio4_hi = 0.785398185253143310546875f;
	const float p0 = 1.666675248e-1f, p1 = 7.495297643e-2f, p2 = 4.547037598e-2f, p3 = 2.417951451e-2f, p4 = 4.216630880e-2f;
	const uint32_t hx = y4_f2u(x), ix = hx & 0x7fffffffu;
	if(ix == 0x3f800000u) return xadd(xmul(x, pio2_hi), xmul(x, pio2_lo));
	if(ix > 0x3f800000u) return xdiv(xsub(x, x), xsub(x, x));
	if(ix < 0x3f000000u)
	{
		if(ix < 0x32000000u) return x;
		const float t = xmul(x, x);
		const float w = xmul(t, xadd(p0, xmul(t, xadd(p1, xmul(t, xadd(p2, xmul(t, xadd(p3, xmul(t, p4)))))))));
		return xadd(x, xmul(x, w));
	}
	float w = xsub(1.f, fabsf(x));
	float t = xmul(w, 0.5f);
	float p = xmul(t, xadd(p0, xmul(t, xadd(p1, xmul(t, xadd(p2, xmul(t, xadd(p3, xmul(t, p4)))))))));
	const float s = xsqrt(t);
	if(ix >= 0x3f79999au) t = xsub(pio2_hi, xsub(xmul(2.f, xadd(s, xmul(s, p))), pio2_lo));
	else
	{
		w = y4_u2f(y4_f2u(s) & 0xfffff000u);
		const float c = xdiv(xsub(t, xmul(w, w)), xadd(s, w));
		const float r = p;
		p = xsub(xmul(xmul(2.f, s), r), xsub(pio2_lo, xmul(2.f, c)));
		const float q = xsub(pio4_hi, xmul(2.f, w));
		t = xsub(pio4_hi, xsub(p, q));
	}
	return (hx >> 31) ? -t : t;
}

// acosf: pi/2 - asin(x) below 0.5 with a rational kernel, 2 asin(sqrt((1 - |x|) / 2)) forms above
Y4_HD float y4_libm_acos_r(float z)
{
	const float s0 = 1.6666667163e-01f, s1 = -3.2556581497e-01f, s2 = 2.0121252537e-01f, s3 = -4.0055535734e-02f, s4 = 7.9153501429e-04f,
	            s5 = 3.4793309169e-05f, q1 = -2.4033949375e+00f, q2 = 2.0209457874e+00f, q3 = -6.8828397989e-01f, q4 = 7.7038154006e-02f;
	const float p = xmul(z, xadd(s0, xmul(z, xadd(s1, xmul(z, xadd(s2, xmul(z, xadd(s3, xmul(z, xadd(s4, xmul(z, s5)))))))))));
	const float q = xadd(1.f, xmul(z, xadd(q1, xmul(z, xadd(q2, xmul(z, xadd(q3, xmul(z, q4))))))));
	return xdiv(p, q);
}
Y4_HD float y4_libm_acosf(float x)
{
	const float pi = 3.1415925026e+00f, pio2_hi = 1.5707962513e+00f, pio2_lo = 7.5497894159e-08f;
	const uint32_t hx = y4_f2u(x), ix = hx & 0x7fffffffu;
	if(ix == 0x3f800000u) return (hx >> 31) ? xadd(pi, xmul(2.f, pio2_lo)) : 0.f;
	if(ix > 0x3f800000u) return xdiv(xsub(x, x), xsub(x, x));
	if(ix < 0x3f000000u)
	{
		if(ix <= 0x32800000u) return xadd(pio2_hi, pio2_lo);
		const float r = y4_libm_acos_r(xmul(x, x));
		return xsub(pio2_hi, xsub(x, xsub(pio2_lo, xmul(x, r))));
	}
	if(hx >> 31)
	{
		const float z = xmul(xadd(1.f, x), 0.5f);
		const float r = y4_libm_acos_r(z), s = xsqrt(z);
		const float w = xsub(xmul(r, s), pio2_lo);
		return xsub(pi, xmul(2.f, xadd(s, w)));
	}
	const float z = xmul(xsub(1.f, x), 0.5f);
	const float s = xsqrt(z);
	const float df = y4_u2f(y4_f2u(s) & 0xfffff000u);
	const float c = xdiv(xsub(z, xmul(df, df)), xadd(s, df));
	const float r = y4_libm_acos_r(z);
	const float w = xadd(xmul(r, s), c);
	return xmul(2.f, xadd(df, w));
}
