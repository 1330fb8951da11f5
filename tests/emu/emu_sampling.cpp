// UNIT-TEST EMULATION ONLY: host build of the Y4_HD sampling/maths restatement
// (libyafaray_b200/csrc/y4_pt_sampling.cuh) so it can be checked against reference KATs.
#include "../../libyafaray_b200/csrc/y4_pt_sampling.cuh"
#include "../../libyafaray_b200/csrc/y4_faure.h"

static Y4FaureHost g_faure = y4_faure_build();
static Y4LdsTables tables()
{
	Y4LdsTables t; t.sigma = g_faure.sigma.data(); t.offset = g_faure.offset; t.prime = g_faure.prime; t.inv_prime = g_faure.inv_prime;
	t.sigma_d = g_faure.sigma_d.data(); t.magic = g_faure.magic;
	return t;
}

extern "C" {
double emu_lds(int dim, uint32_t n) { return y4_lds(tables(), dim, n); }
double emu_lds_loop(int dim, uint32_t n) { return y4_lds_loop(tables(), dim, n); }
float emu_halton2_at(uint32_t start, uint32_t q) { return y4_halton2_at(start, q); }
void emu_ri(uint32_t i, uint32_t r, float *out3, uint32_t *fnv)
{
	out3[0] = y4_ri_vdc(i, r); out3[1] = y4_ri_s(i, r); out3[2] = y4_ri_lp(i, r); *fnv = y4_fnv32a(i);
}
void emu_halton(uint32_t base, uint32_t start, int count, float *out)
{
	Y4Halton h = y4_halton_start(base, start);
	for(int k = 0; k < count; ++k) out[k] = y4_halton_next(h);
}
// bulk versions (tests/test_sampling_bulk.py: millions of inputs against oracle/_ref/ref_kat_bulk)
void emu_lds_bulk(const uint32_t *in, uint64_t n, double *out) { const Y4LdsTables t = tables(); for(uint64_t k = 0; k < n; ++k) out[k] = y4_lds(t, (int) in[2 * k], in[2 * k + 1]); }
void emu_lds_loop_bulk(const uint32_t *in, uint64_t n, double *out) { const Y4LdsTables t = tables(); for(uint64_t k = 0; k < n; ++k) out[k] = y4_lds_loop(t, (int) in[2 * k], in[2 * k + 1]); }
void emu_halton_bulk(const uint32_t *in, uint64_t n, float *out)
{
	for(uint64_t k = 0; k < n; ++k)
	{
		Y4Halton h = y4_halton_start(in[2 * k], in[2 * k + 1]);
		for(int j = 0; j < 4; ++j) out[4 * k + j] = y4_halton_next(h);
	}
}
void emu_halton2_at_bulk(const uint32_t *in, uint64_t n, float *out) { for(uint64_t k = 0; k < n; ++k) for(uint32_t j = 0; j < 4; ++j) out[4 * k + j] = y4_halton2_at(in[2 * k + 1], j + 1u); }
void emu_ri_bulk(const uint32_t *in, uint64_t n, float *out)
{
	for(uint64_t k = 0; k < n; ++k) { out[3 * k] = y4_ri_vdc(in[2 * k], in[2 * k + 1]); out[3 * k + 1] = y4_ri_s(in[2 * k], in[2 * k + 1]); out[3 * k + 2] = y4_ri_lp(in[2 * k], in[2 * k + 1]); }
}
void emu_trig(float x, float *out3) { out3[0] = y4_fast_sin(x); out3[1] = y4_fast_cos(x); out3[2] = y4_fast_exp(x * 0.5f); }
void emu_coshemi(const float *n, float s1, float s2, float *out9)
{
	V3 u, v;
	const V3 nn = v3(n[0], n[1], n[2]);
	y4_coords_system(nn, u, v);
	const V3 d = y4_cos_hemisphere(nn, u, v, s1, s2);
	out9[0] = u.x; out9[1] = u.y; out9[2] = u.z; out9[3] = v.x; out9[4] = v.y; out9[5] = v.z; out9[6] = d.x; out9[7] = d.y; out9[8] = d.z;
}
}
