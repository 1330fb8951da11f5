// C-ABI entry points (include/y4gpu.h): context lifetime, geometry upload, build, ray batches.
// No CPU fallback: every compute entry point returns Y4GPU_ERR_CUDA when CUDA is unusable.
#include "y4_ctx.h"
#include <cstring>
#include <cstdlib>

static std::string g_create_err;

static bool is_device_ptr(const void *p)
{
	cudaPointerAttributes a;
	if(cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
	return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

extern "C" {

int y4gpu_device_count(void)
{
	int n = 0;
	if(cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n;
}

int y4gpu_create(int device, y4gpu_ctx **out)
{
	if(!out) return Y4GPU_ERR_ARG;
	*out = nullptr;
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if(e != cudaSuccess || n == 0)
	{
		g_create_err = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") + " (this back end has no CPU fallback)";
		cudaGetLastError();
		return Y4GPU_ERR_CUDA;
	}
	if(device < 0 || device >= n) { g_create_err = "device index out of range"; return Y4GPU_ERR_ARG; }
	if((e = cudaSetDevice(device)) != cudaSuccess) { g_create_err = cudaGetErrorString(e); return Y4GPU_ERR_CUDA; }
	y4gpu_ctx *ctx = new y4gpu_ctx;
	ctx->device = device;
	cudaDeviceProp prop;
	if(cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
	bool ok = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) == cudaSuccess
		&& cudaEventCreate(&ctx->ev0) == cudaSuccess && cudaEventCreate(&ctx->ev1) == cudaSuccess
		&& cudaMalloc(&ctx->d_work_counter, 64) == cudaSuccess
		&& cudaMalloc(&ctx->d_visit_counters, 64) == cudaSuccess
		&& cudaMalloc(&ctx->d_consts, 64) == cudaSuccess;
	const uint32_t consts[4] = {0x4B000000u, 0u, 0u, 0u};
	ok = ok && cudaMemcpy(ctx->d_consts, consts, sizeof(consts), cudaMemcpyHostToDevice) == cudaSuccess;
	ctx->stream = ctx->own_stream;
	if(!ok)
	{
		g_create_err = std::string("context setup failed: ") + cudaGetErrorString(cudaGetLastError());
		delete ctx;
		return Y4GPU_ERR_CUDA;
	}
	if(const char *e = getenv("Y4GPU_FETCH_THRESHOLD")) ctx->fetch_threshold = atoi(e);
	if(const char *e = getenv("Y4GPU_POSTPONE_DIV")) ctx->postpone_div = atoi(e);
	if(const char *e = getenv("Y4GPU_MORTON_SIZED")) ctx->morton_sized = atoi(e);
	if(const char *e = getenv("Y4GPU_TREELET_DEFER")) { const int k = atoi(e); ctx->treelet_defer = k < 0 ? 0 : (k > 1024 ? 1024 : k); }
	if(const char *e = getenv("Y4GPU_TREELET")) { const int k = atoi(e); ctx->treelet_leaves = k < 0 ? 0 : (k > 4096 ? 4096 : k); }
	if(const char *e = getenv("Y4GPU_MIN_BLOCKS")) ctx->min_blocks = atoi(e);
	*out = ctx;
	return Y4GPU_OK;
}

void y4gpu_pt_release(y4gpu_ctx *ctx);   // y4_pt.cu
} void y4_pt_invalidate(y4gpu_ctx *ctx); extern "C" {

void y4gpu_destroy(y4gpu_ctx *ctx)
{
	if(!ctx) return;
	cudaSetDevice(ctx->device);
	if(ctx->stream) cudaStreamSynchronize(ctx->stream);
	y4gpu_pt_release(ctx);
	cudaFree(ctx->d_verts); cudaFree(ctx->d_flags); cudaFree(ctx->d_nodes); cudaFree(ctx->d_tris);
	cudaFree(ctx->d_rays); cudaFree(ctx->d_hits); cudaFree(ctx->d_work_counter); cudaFree(ctx->d_visit_counters); cudaFree(ctx->d_consts);
	cudaFree(ctx->d_mat_transp);
	if(ctx->ev0) cudaEventDestroy(ctx->ev0);
	if(ctx->ev1) cudaEventDestroy(ctx->ev1);
	if(ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
	delete ctx;
}

const char *y4gpu_last_error(const y4gpu_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

void *y4gpu_host_alloc(uint64_t bytes)
{
	void *p = nullptr;
	if(cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
	return p;
}
void y4gpu_host_free(void *p) { if(p) cudaFreeHost(p); }

int y4gpu_upload_triangles(y4gpu_ctx *ctx, const float *v_xyz, const uint32_t *flags, uint64_t n_tri)
{
	if(!ctx || (n_tri && !v_xyz)) return Y4GPU_ERR_ARG;
	if(n_tri > 0x3fffffffull) { ctx->err = "too many triangles (max 2^30-1)"; return Y4GPU_ERR_ARG; }
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	// a re-upload of the same size (animation frame, repeated render) reuses the device arrays
	const bool reuse = n_tri && n_tri == ctx->n_tri && ctx->d_verts && (!flags == !ctx->d_flags);
	ctx->built = false;
	y4_pt_invalidate(ctx);
	if(!reuse)
	{
		cudaFree(ctx->d_verts); cudaFree(ctx->d_flags);
		ctx->d_verts = nullptr; ctx->d_flags = nullptr; ctx->n_tri = 0;
	}
	if(n_tri == 0) return Y4GPU_OK;
	if(!reuse) Y4_CUDA(ctx, cudaMalloc(&ctx->d_verts, n_tri * 9 * sizeof(float)));
	Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_verts, v_xyz, n_tri * 9 * sizeof(float), cudaMemcpyDefault, ctx->stream));
	if(flags)
	{
		if(!reuse) Y4_CUDA(ctx, cudaMalloc(&ctx->d_flags, n_tri * sizeof(uint32_t)));
		Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_flags, flags, n_tri * sizeof(uint32_t), cudaMemcpyDefault, ctx->stream));
	}
	Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	ctx->n_tri = n_tri;
	return Y4GPU_OK;
}

int y4gpu_build(y4gpu_ctx *ctx, const y4gpu_build_params *params)
{
	if(!ctx) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	int leaf_max = params && params->leaf_max ? params->leaf_max : 1;
	y4_pt_invalidate(ctx);
	return y4_build_bvh(ctx, leaf_max);
}

int y4gpu_get_bound(const y4gpu_ctx *ctx, float out_bound[6])
{
	if(!ctx || !out_bound || !ctx->built) return Y4GPU_ERR_ARG;
	memcpy(out_bound, ctx->bound, sizeof(float) * 6);
	return Y4GPU_OK;
}

static int ensure_stage(y4gpu_ctx *ctx, uint64_t n, uint64_t hit_bytes)
{
	if(n > ctx->stage_cap)
	{
		cudaFree(ctx->d_rays); ctx->d_rays = nullptr; ctx->stage_cap = 0;
		Y4_CUDA(ctx, cudaMalloc(&ctx->d_rays, n * sizeof(Y4Ray)));
		ctx->stage_cap = n;
	}
	if(n * hit_bytes > ctx->stage_hit_bytes)
	{
		cudaFree(ctx->d_hits); ctx->d_hits = nullptr; ctx->stage_hit_bytes = 0;
		Y4_CUDA(ctx, cudaMalloc(&ctx->d_hits, n * hit_bytes));
		ctx->stage_hit_bytes = n * hit_bytes;
	}
	return Y4GPU_OK;
}

static int trace_impl(y4gpu_ctx *ctx, int kind, bool counted, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out)
{
	if(!ctx || (n && (!rays || !out))) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	if(!ctx->built) { ctx->err = "y4gpu_trace before y4gpu_build"; return Y4GPU_ERR_ARG; }
	if(n == 0) return Y4GPU_OK;
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	const bool rays_dev = is_device_ptr(rays), out_dev = is_device_ptr(out);
	if(!rays_dev || !out_dev) { const int rc = ensure_stage(ctx, n, sizeof(Y4Hit)); if(rc) return rc; }
	const Y4Ray *d_rays = reinterpret_cast<const Y4Ray *>(rays);
	Y4Hit *d_out = reinterpret_cast<Y4Hit *>(out);
	if(!rays_dev)
	{
		Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_rays, rays, n * sizeof(Y4Ray), cudaMemcpyHostToDevice, ctx->stream));
		d_rays = ctx->d_rays;
	}
	if(!out_dev) d_out = reinterpret_cast<Y4Hit *>(ctx->d_hits);
	const int rc = y4_launch_trace(ctx, kind, counted, d_rays, n, d_out);
	if(rc) return rc;
	if(!out_dev) Y4_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * sizeof(Y4Hit), cudaMemcpyDeviceToHost, ctx->stream));
	if(!out_dev || !rays_dev || counted) Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	if(counted)
	{
		unsigned long long v[2];
		Y4_CUDA(ctx, cudaMemcpy(v, ctx->d_visit_counters, sizeof(v), cudaMemcpyDeviceToHost));
		ctx->stats.nodes_visited = v[0]; ctx->stats.tris_tested = v[1];
	}
	return Y4GPU_OK;
}

int y4gpu_trace(y4gpu_ctx *ctx, int kind, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out) { return trace_impl(ctx, kind, false, rays, n, out); }
int y4gpu_trace_counted(y4gpu_ctx *ctx, int kind, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out) { return trace_impl(ctx, kind, true, rays, n, out); }

int y4gpu_trace_tshadow(y4gpu_ctx *ctx, const y4gpu_ray *rays, uint64_t n, int max_depth, const float *mat_transparency, uint32_t n_mat, y4gpu_tshit *out)
{
	if(!ctx || (n && (!rays || !out)) || (n_mat && !mat_transparency) || max_depth < 0) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	if(max_depth > 16) { ctx->err = "y4gpu_trace_tshadow: max_depth > 16 is not supported"; return Y4GPU_ERR_UNSUPPORTED; }
	if(!ctx->built) { ctx->err = "y4gpu_trace_tshadow before y4gpu_build"; return Y4GPU_ERR_ARG; }
	if(n == 0) return Y4GPU_OK;
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	const bool rays_dev = is_device_ptr(rays), out_dev = is_device_ptr(out);
	if(!rays_dev || !out_dev) { const int rc = ensure_stage(ctx, n, sizeof(y4gpu_tshit)); if(rc) return rc; }
	if(n_mat > ctx->mat_transp_cap)
	{
		cudaFree(ctx->d_mat_transp); ctx->d_mat_transp = nullptr; ctx->mat_transp_cap = 0;
		Y4_CUDA(ctx, cudaMalloc(&ctx->d_mat_transp, (size_t) n_mat * 3 * sizeof(float)));
		ctx->mat_transp_cap = n_mat;
	}
	if(n_mat) Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_mat_transp, mat_transparency, (size_t) n_mat * 3 * sizeof(float), cudaMemcpyDefault, ctx->stream));
	const Y4Ray *d_rays = reinterpret_cast<const Y4Ray *>(rays);
	y4gpu_tshit *d_out = out;
	if(!rays_dev)
	{
		Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_rays, rays, n * sizeof(Y4Ray), cudaMemcpyHostToDevice, ctx->stream));
		d_rays = ctx->d_rays;
	}
	if(!out_dev) d_out = reinterpret_cast<y4gpu_tshit *>(ctx->d_hits);
	const int rc = y4_launch_tshadow(ctx, d_rays, n, max_depth, ctx->d_mat_transp, n_mat, d_out);
	if(rc) return rc;
	if(!out_dev) Y4_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * sizeof(y4gpu_tshit), cudaMemcpyDeviceToHost, ctx->stream));
	Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return Y4GPU_OK;
}

int y4gpu_set_stream(y4gpu_ctx *ctx, void *cuda_stream)
{
	if(!ctx) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	ctx->stream = cuda_stream ? (cudaStream_t) cuda_stream : ctx->own_stream;
	return Y4GPU_OK;
}

int y4gpu_synchronize(y4gpu_ctx *ctx)
{
	if(!ctx) return Y4GPU_ERR_ARG;
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return Y4GPU_OK;
}

int y4gpu_get_stats(const y4gpu_ctx *ctx, y4gpu_stats *out)
{
	if(!ctx || !out) return Y4GPU_ERR_ARG;
	*out = ctx->stats;
	float ms = 0.f;
	if(ctx->stats.last_trace_rays && cudaEventQuery(ctx->ev1) == cudaSuccess && cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) == cudaSuccess) out->last_trace_ms = ms;
	else cudaGetLastError();
	return Y4GPU_OK;
}

} // extern "C"
