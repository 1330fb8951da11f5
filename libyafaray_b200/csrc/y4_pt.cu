// Wavefront path tracer back end (PathIntegrator on the GPU).  See y4_pt_*.cuh.
#include "y4_ctx.h"

extern "C" void y4gpu_pt_release(y4gpu_ctx *ctx)
{
	(void) ctx;
}
