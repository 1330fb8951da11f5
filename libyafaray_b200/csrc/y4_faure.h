// Host-side generation of the tables behind Halton::lowDiscrepancySampling
// (/root/reference/src/sampler/halton.cc:26-441).  The reference stores them as literals;
// here they are *constructed*: Faure's permutations by his recursive rule
//   b even: sigma_b = (2*sigma_{b/2}, 2*sigma_{b/2} + 1)
//   b odd : sigma_b = sigma_{b-1} with values >= c=(b-1)/2 incremented and c inserted in the middle
// and the reciprocals as 1/p rounded to 9 decimals (the reference's `inv_prims` literals).
// tests/test_pt_cpu.py checks the generated tables against reference output for all 49 dims.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

struct Y4FaureHost
{
	std::vector<uint8_t> sigma;
	std::vector<double> sigma_d;            // the same permutation values as doubles (saves the int->fp64 conversion per digit)
	unsigned long long magic[50];           // floor(2^64 / prime) + 1: n / prime == mulhi64(n, magic) for every 32-bit n
	uint16_t offset[50];
	uint16_t prime[50];
	double inv_prime[50];
};

inline std::vector<int> y4_faure_perm(int b)
{
	if(b == 1) return {0};
	if(b == 2) return {0, 1};
	std::vector<int> r;
	if(b % 2 == 0)
	{
		const std::vector<int> h = y4_faure_perm(b / 2);
		for(int v : h) r.push_back(2 * v);
		for(int v : h) r.push_back(2 * v + 1);
	}
	else
	{
		const int c = (b - 1) / 2;
		std::vector<int> e = y4_faure_perm(b - 1);
		for(int &v : e) if(v >= c) ++v;
		r.assign(e.begin(), e.begin() + c);
		r.push_back(c);
		r.insert(r.end(), e.begin() + c, e.end());
	}
	return r;
}

inline Y4FaureHost y4_faure_build()
{
	Y4FaureHost t;
	int p = 1, dim = 0;
	while(dim < 50)
	{
		bool is_prime = true;
		for(int d = 2; d * d <= p; ++d) if(p % d == 0) { is_prime = false; break; }
		if(is_prime)
		{
			t.prime[dim] = (uint16_t) p;
			char buf[32];
			snprintf(buf, sizeof(buf), "%.9f", 1.0 / (double) p);
			t.inv_prime[dim] = strtod(buf, nullptr);
			t.offset[dim] = (uint16_t) t.sigma.size();
			// the reference's table list reuses the identity base-3 table for dimensions 0, 1, 2 (halton.cc:398-399)
			const std::vector<int> perm = dim < 3 ? std::vector<int>{0, 1, 2} : y4_faure_perm(p);
			for(int v : perm) { t.sigma.push_back((uint8_t) v); t.sigma_d.push_back((double) v); }
			t.magic[dim] = p > 1 ? (unsigned long long) ((((unsigned __int128) 1) << 64) / (unsigned) p) + 1ull : 0ull;
			++dim;
		}
		++p;
	}
	return t;
}
