// TEST INFRASTRUCTURE ONLY (oracle): bulk known-answer generator for the samplers of the UNMODIFIED reference.
// usage: ref_kat_bulk lds|halton|ri < binary input > binary output
//   lds:    records {int32 dim, uint32 n}            -> double  Halton::lowDiscrepancySampling(dim, n)      (src/sampler/halton.cc)
//   halton: records {uint32 base, uint32 start}      -> float[4] the next four values of Halton(base, start) (include/sampler/halton.h)
//   ri:     records {uint32 i, uint32 r}             -> float[3] riVdC, riS, riLp                           (include/sampler/sample.h)
#include "sampler/halton.h"
#include "sampler/sample.h"
#include <cstdio>
#include <cstring>
#include <cstdint>
#include <vector>

using namespace yafaray;

int main(int argc, char **argv)
{
	if(argc < 2) return 1;
	const std::string mode = argv[1];
	std::vector<uint32_t> in;
	uint32_t buf[4096];
	size_t got;
	while((got = fread(buf, 4, 4096, stdin)) > 0) in.insert(in.end(), buf, buf + got);
	const size_t n = in.size() / 2;
	if(mode == "lds")
	{
		std::vector<double> out(n);
		for(size_t k = 0; k < n; ++k) out[k] = Halton::lowDiscrepancySampling((int) in[2 * k], in[2 * k + 1]);
		fwrite(out.data(), 8, n, stdout);
	}
	else if(mode == "halton")
	{
		std::vector<float> out(4 * n);
		for(size_t k = 0; k < n; ++k) { Halton h((int) in[2 * k], in[2 * k + 1]); for(int j = 0; j < 4; ++j) out[4 * k + j] = h.getNext(); }
		fwrite(out.data(), 4, 4 * n, stdout);
	}
	else
	{
		std::vector<float> out(3 * n);
		for(size_t k = 0; k < n; ++k) { out[3 * k] = sample::riVdC(in[2 * k], in[2 * k + 1]); out[3 * k + 1] = sample::riS(in[2 * k], in[2 * k + 1]); out[3 * k + 2] = sample::riLp(in[2 * k], in[2 * k + 1]); }
		fwrite(out.data(), 4, 3 * n, stdout);
	}
	return 0;
}
