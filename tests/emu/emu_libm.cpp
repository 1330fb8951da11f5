// UNIT-TEST EMULATION ONLY: the fp32 inverse trigonometric restatements of libyafaray_b200/csrc/y4_libm.cuh against the
// C library this process is linked with (the one the reference calls through std::atan2 / std::asin / std::acos).
#include "../../libyafaray_b200/csrc/y4_libm.cuh"

static inline bool same_bits(float a, float b)
{
	if(a != a && b != b) return true;
	return y4_f2u(a) == y4_f2u(b);
}
static inline uint32_t mix(uint32_t x)
{
	x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
	return x;
}

extern "C" {
// fn 0 atanf, 1 asinf, 2 acosf: every `stride`-th bit pattern of [lo, hi) and its negation.  Returns the number of
// mismatches; the first offending pattern goes to *bad.
uint64_t emu_libm_mismatch(int fn, uint32_t lo, uint32_t hi, uint32_t stride, uint32_t *bad)
{
	uint64_t n_bad = 0;
	for(uint64_t b = lo; b < hi; b += stride)
		for(uint32_t sign = 0; sign < 2; ++sign)
		{
			const uint32_t u = (uint32_t) b | (sign << 31);
			const float x = y4_u2f(u);
			volatile float vx = x;
			float want, got;
			if(fn == 0) { want = atanf(vx); got = y4_libm_atanf(x); }
			else if(fn == 1) { want = asinf(vx); got = y4_libm_asinf(x); }
			else { want = acosf(vx); got = y4_libm_acosf(x); }
			if(!same_bits(want, got)) { if(!n_bad && bad) *bad = u; ++n_bad; }
		}
	return n_bad;
}
// atan2f over n pseudo-random (y, x) pairs: full-range bit patterns for a third of them, screen-like magnitudes
// (|v| in [2^-12, 4)) for the rest, plus the axis / zero / infinity cases.
uint64_t emu_libm_atan2_mismatch(uint64_t n, uint32_t seed, uint32_t *bad_y, uint32_t *bad_x)
{
	uint64_t n_bad = 0;
	const uint32_t special[] = {0x00000000u, 0x80000000u, 0x3f800000u, 0xbf800000u, 0x7f800000u, 0xff800000u, 0x00000001u, 0x7f7fffffu, 0x7fc00000u, 0x33800000u};
	const int ns = (int) (sizeof(special) / sizeof(special[0]));
	for(uint64_t i = 0; i < n + (uint64_t) ns * ns; ++i)
	{
		uint32_t uy, ux;
		if(i >= n) { uy = special[(i - n) / ns]; ux = special[(i - n) % ns]; }
		else
		{
			uy = mix((uint32_t) i * 2u + seed); ux = mix((uint32_t) i * 2u + 1u + seed * 0x9e3779b9u);
			if(i % 3)
			{
				uy = (uy & 0x807fffffu) | ((115u + (uy >> 23) % 14u) << 23);
				ux = (ux & 0x807fffffu) | ((115u + (ux >> 23) % 14u) << 23);
			}
		}
		const float y = y4_u2f(uy), x = y4_u2f(ux);
		volatile float vy = y, vx = x;
		if(!same_bits(atan2f(vy, vx), y4_libm_atan2f(y, x))) { if(!n_bad) { if(bad_y) *bad_y = uy; if(bad_x) *bad_x = ux; } ++n_bad; }
	}
	return n_bad;
}
// the C library itself, in bulk (what the device results of y4gpu_libm_eval are compared with)
void emu_libm_host(int fn, const float *a, const float *b, float *out, uint64_t n)
{
	for(uint64_t i = 0; i < n; ++i)
	{
		volatile float x = a[i];
		out[i] = fn == 0 ? atanf(x) : (fn == 1 ? asinf(x) : (fn == 2 ? acosf(x) : atan2f(x, b[i])));
	}
}
float emu_libm_eval(int fn, float a, float b)
{
	if(fn == 0) return y4_libm_atanf(a);
	if(fn == 1) return y4_libm_asinf(a);
	if(fn == 2) return y4_libm_acosf(a);
	return y4_libm_atan2f(a, b);
}
}
