// TEST INFRASTRUCTURE ONLY (oracle): dumps known-answer vectors from the UNMODIFIED reference
// (libYafaRay headers + oracle/_ref/libyafaray4_ref.a) for the sampler / approximated-maths
// functions the device code restates.  Output: JSON on stdout; floats are given as their
// uint32 bit patterns and doubles as uint64 so comparisons are bit-exact.
//   Halton::lowDiscrepancySampling      src/sampler/halton.cc:421-441
//   Halton::setStart/getNext            include/sampler/halton.h:54-81
//   sample::riVdC/riS/riLp/fnv32ABuf    include/sampler/sample.h:103-151
//   math::sin/cos/exp                   include/math/math.h:190-250
//   sample::cosHemisphere               include/sampler/sample.h:45-54
//   Vec3::createCoordsSystem            include/geometry/vector.h:262-276
// The committed copy lives in tests/golden/kat_sampling.json (made by tests/golden/make_golden.py).
#include "sampler/halton.h"
#include "sampler/sample.h"
#include "math/math.h"
#include "geometry/vector.h"
#include <cstdio>
#include <cstring>
#include <cstdint>

using namespace yafaray;

static uint32_t fb(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static uint64_t db(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
static uint64_t g_state = 0x1234567ull;
static uint32_t rnd() { g_state = g_state * 6364136223846793005ull + 1442695040888963407ull; return (uint32_t) (g_state >> 32); }
static float rndf() { return (float) (rnd() >> 8) * (1.f / 16777216.f); }

int main()
{
	printf("{\n");
	// lowDiscrepancySampling: all dims 1..49, assorted n
	printf("\"lds\": [");
	bool first = true;
	for(int dim = 1; dim < 50; ++dim)
		for(int k = 0; k < 24; ++k)
		{
			const uint32_t n = k < 8 ? (uint32_t) k : (k < 16 ? rnd() >> (rnd() % 24) : rnd());
			printf("%s[%d,%u,%llu]", first ? "" : ",", dim, n, (unsigned long long) db(Halton::lowDiscrepancySampling(dim, n)));
			first = false;
		}
	printf("],\n\"ri\": [");
	for(int k = 0; k < 256; ++k)
	{
		const uint32_t i = k < 16 ? (uint32_t) k : rnd(), r = (k & 1) ? rnd() : 0u;
		printf("%s[%u,%u,%u,%u,%u,%u]", k ? "," : "", i, r, fb(sample::riVdC(i, r)), fb(sample::riS(i, r)), fb(sample::riLp(i, r)), sample::fnv32ABuf(i));
	}
	printf("],\n\"halton\": [");
	for(int k = 0; k < 64; ++k)
	{
		const int base = (k & 1) ? 3 : 2;
		const uint32_t start = k < 8 ? (uint32_t) k : rnd();
		Halton h(base, start);
		printf("%s[%d,%u", k ? "," : "", base, start);
		for(int j = 0; j < 6; ++j) printf(",%u", fb(h.getNext()));
		printf("]");
	}
	printf("],\n\"trig\": [");
	for(int k = 0; k < 512; ++k)
	{
		float x = (rndf() - 0.5f) * (k < 384 ? 14.f : 200.f);
		if(k == 0) x = 0.f;
		printf("%s[%u,%u,%u,%u]", k ? "," : "", fb(x), fb(math::sin(x)), fb(math::cos(x)), fb(math::exp(x * 0.5f)));
	}
	printf("],\n\"coshemi\": [");
	for(int k = 0; k < 256; ++k)
	{
		Vec3 n{rndf() - 0.5f, rndf() - 0.5f, rndf() - 0.5f};
		if(k % 16 == 0) n = Vec3{0.f, 0.f, (k & 16) ? -1.f : 1.f};
		if(k % 16 == 1) n = Vec3{0.f, 1.f, 0.f};
		n.normalize();
		Vec3 u, v;
		std::tie(u, v) = Vec3::createCoordsSystem(n);
		const float s_1 = (k % 32 == 5) ? 1.f : rndf(), s_2 = rndf();
		const Vec3 d = sample::cosHemisphere(n, u, v, s_1, s_2);
		printf("%s[%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u,%u]", k ? "," : "", fb(n.x()), fb(n.y()), fb(n.z()), fb(s_1), fb(s_2),
			   fb(u.x()), fb(u.y()), fb(u.z()), fb(v.x()), fb(v.y()), fb(v.z()), fb(d.x()), fb(d.y()), fb(d.z()));
	}
	printf("]\n}\n");
	return 0;
}
