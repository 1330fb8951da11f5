// C-ABI entry points (include/y4gpu.h): context lifetime, geometry upload, build, ray batches.
// No CPU fallback: every compute entry point returns Y4GPU_ERR_CUDA when CUDA is unusable.
#include "y4_ctx.h"
#include <cstring>
#include <cstdlib>

static std::string g_create_err;
extern "C" { static void ray_sort_env(y4gpu_ctx *ctx); static void free_meshes(y4gpu_ctx *ctx); }

extern "C" { static void slot_free_all(y4gpu_ctx *ctx); }

static bool is_device_ptr(const void *p)
{
	cudaPointerAttributes a;
	if(cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
	return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

extern "C" {

uint64_t y4gpu_struct_size(int which)
{
	static const uint64_t sizes[] = {sizeof(y4gpu_build_params), sizeof(y4gpu_stats), sizeof(y4gpu_mesh), sizeof(y4gpu_instance), sizeof(y4gpu_pt_material),
	                                 sizeof(y4gpu_pt_texture), sizeof(y4gpu_pt_layer), sizeof(y4gpu_pt_light), sizeof(y4gpu_pt_camera), sizeof(y4gpu_pt_params),
	                                 sizeof(y4gpu_pt_stats), sizeof(y4gpu_pt_aa_noise), sizeof(y4gpu_ray), sizeof(y4gpu_hit), sizeof(y4gpu_tshit)};
	return which >= 0 && which < (int) (sizeof(sizes) / sizeof(sizes[0])) ? sizes[which] : 0;
}

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
	if(const char *e = getenv("Y4GPU_PT_FETCH_THRESHOLD")) ctx->pt_fetch_threshold = atoi(e);
	if(const char *e = getenv("Y4GPU_MORTON_SIZED")) ctx->morton_sized = atoi(e);
	if(const char *e = getenv("Y4GPU_TREELET_DEFER")) { const int k = atoi(e); ctx->treelet_defer = k < 0 ? 0 : (k > 1024 ? 1024 : k); }
	if(const char *e = getenv("Y4GPU_TREELET")) { const int k = atoi(e); ctx->treelet_leaves = k < 0 ? 0 : (k > 4096 ? 4096 : k); }
	if(const char *e = getenv("Y4GPU_MIN_BLOCKS")) ctx->min_blocks = atoi(e);
	ray_sort_env(ctx);
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
	slot_free_all(ctx);
	cudaFree(ctx->d_verts); cudaFree(ctx->d_flags); cudaFree(ctx->d_nodes); cudaFree(ctx->d_tris);
	cudaFree(ctx->d_rays); cudaFree(ctx->d_hits); cudaFree(ctx->d_work_counter); cudaFree(ctx->d_visit_counters); cudaFree(ctx->d_consts);
	cudaFree(ctx->d_mat_transp); cudaFree(ctx->sort_scratch.buf);
	cudaFree(ctx->d_mesh_verts); cudaFree(ctx->d_mesh_flags); y4_free_two_level(ctx);
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
	if(ctx->two_level) { free_meshes(ctx); ctx->n_tri = 0; }
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

static void free_meshes(y4gpu_ctx *ctx)
{
	cudaFree(ctx->d_mesh_verts); cudaFree(ctx->d_mesh_flags);
	ctx->d_mesh_verts = nullptr; ctx->d_mesh_flags = nullptr;
	ctx->mesh_first.clear(); ctx->h_inst.clear();
	y4_free_two_level(ctx);
	ctx->two_level = false;
}

int y4gpu_upload_instanced(y4gpu_ctx *ctx, const y4gpu_mesh *meshes, uint32_t n_mesh, const y4gpu_instance *instances, uint32_t n_inst)
{
	if(!ctx || (n_mesh && !meshes) || (n_inst && !instances)) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	ctx->built = false;
	y4_pt_invalidate(ctx);
	free_meshes(ctx);
	cudaFree(ctx->d_verts); cudaFree(ctx->d_flags); cudaFree(ctx->d_nodes); cudaFree(ctx->d_tris);
	ctx->d_verts = nullptr; ctx->d_flags = nullptr; ctx->d_nodes = nullptr; ctx->d_tris = nullptr; ctx->n_tri = 0; ctx->n_nodes = 0;
	std::vector<uint64_t> first(n_mesh + 1, 0);
	bool any_flags = false;
	for(uint32_t i = 0; i < n_mesh; ++i)
	{
		if(meshes[i].n_tri && !meshes[i].v_xyz) { ctx->err = "mesh without vertices"; return Y4GPU_ERR_ARG; }
		first[i + 1] = first[i] + meshes[i].n_tri;
		any_flags = any_flags || meshes[i].flags;
	}
	uint64_t total = 0;
	for(uint32_t i = 0; i < n_inst; ++i)
	{
		if(instances[i].mesh >= n_mesh) { ctx->err = "instance refers to a mesh that was not given"; return Y4GPU_ERR_ARG; }
		total += meshes[instances[i].mesh].n_tri;
		if(!instances[i].identity)
		{
			// the walk maps rays into object space: the matrix must be invertible (a collapsed instance can still be uploaded flattened)
			const float *m = instances[i].obj_to_world;
			const double det = (double) m[0] * ((double) m[5] * m[10] - (double) m[6] * m[9]) - (double) m[1] * ((double) m[4] * m[10] - (double) m[6] * m[8])
			                   + (double) m[2] * ((double) m[4] * m[9] - (double) m[5] * m[8]);
			double mag = 0.0;
			for(int k = 0; k < 12; ++k) { if(!(m[k] == m[k]) || fabs((double) m[k]) > 1e30) mag = -1.0; else if((k & 3) != 3 && mag >= 0.0) mag = fmax(mag, fabs((double) m[k])); }
			if(mag < 0.0 || !(fabs(det) > 1e-12 * mag * mag * mag)) { ctx->err = "singular or non-finite instance matrix (upload such a scene as flattened triangles)"; return Y4GPU_ERR_UNSUPPORTED; }
		}
	}
	if(total > 0x3fffffffull || first[n_mesh] > 0x3fffffffull) { ctx->err = "too many triangles (max 2^30-1)"; return Y4GPU_ERR_ARG; }
	ctx->two_level = true;
	if(first[n_mesh])
	{
		Y4_CUDA(ctx, cudaMalloc(&ctx->d_mesh_verts, first[n_mesh] * 9 * sizeof(float)));
		if(any_flags) Y4_CUDA(ctx, cudaMalloc(&ctx->d_mesh_flags, first[n_mesh] * sizeof(uint32_t)));
		for(uint32_t i = 0; i < n_mesh; ++i)
		{
			const uint64_t n = meshes[i].n_tri;
			if(!n) continue;
			Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_mesh_verts + 9ull * first[i], meshes[i].v_xyz, n * 9 * sizeof(float), cudaMemcpyDefault, ctx->stream));
			if(any_flags)
			{
				if(meshes[i].flags) Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_mesh_flags + first[i], meshes[i].flags, n * sizeof(uint32_t), cudaMemcpyDefault, ctx->stream));
				else
				{
					std::vector<uint32_t> def(n, (uint32_t) Y4_TRI_DEFAULT_FLAGS);
					Y4_CUDA(ctx, cudaMemcpy(ctx->d_mesh_flags + first[i], def.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice));
				}
			}
		}
		Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	}
	ctx->mesh_first = first;
	uint32_t first_prim = 0;
	ctx->h_inst.clear();
	for(uint32_t i = 0; i < n_inst; ++i)
	{
		if(meshes[instances[i].mesh].n_tri == 0) continue;          // an instance of an empty mesh owns no primitives
		ctx->h_inst.emplace_back();
		Y4Inst &in = ctx->h_inst.back();
		memset(&in, 0, sizeof(in));
		memcpy(in.m, instances[i].obj_to_world, sizeof(in.m));
		in.mesh = instances[i].mesh; in.identity = instances[i].identity ? 1u : 0u;
		in.n_tri = (uint32_t) meshes[in.mesh].n_tri; in.first_prim = first_prim; in.vert_first = (uint32_t) first[in.mesh];
		first_prim += in.n_tri;
	}
	ctx->n_tri = total;
	return Y4GPU_OK;
}

int y4gpu_build(y4gpu_ctx *ctx, const y4gpu_build_params *params)
{
	if(!ctx) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	Y4_CUDA(ctx, cudaSetDevice(ctx->device));
	int leaf_max = params && params->leaf_max ? params->leaf_max : 1;
	y4_pt_invalidate(ctx);
	return ctx->two_level ? y4_build_two_level(ctx, leaf_max) : y4_build_bvh(ctx, leaf_max);
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

// ---- slot pool: the re-entrant host-buffer path ---------------------------------------------------------
// The Accelerator virtuals are called concurrently by every render thread (integrator_tiled.cc:248-251).  A call with host
// buffers takes a private slot (streams, device staging, work counters, pinned staging), so concurrent calls share nothing
// but the read-only BVH; the context mutex is not held while the GPU works.  Errors are reported per call through the
// return code; the message goes to ctx->err under the statistics lock (last writer wins).
static void slot_error(y4gpu_ctx *ctx, const std::string &msg) { std::lock_guard<std::mutex> lk(ctx->stat_mtx); ctx->err = msg; }
#define Y4_SLOT_CUDA(ctx, call)                                                                  \
	do {                                                                                         \
		const cudaError_t e_ = (call);                                                           \
		if(e_ != cudaSuccess) { slot_error(ctx, std::string(#call) + ": " + cudaGetErrorString(e_)); return Y4GPU_ERR_CUDA; } \
	} while(0)

static int slot_acquire(y4gpu_ctx *ctx, Y4TraceSlot **out)
{
	{
		std::lock_guard<std::mutex> lk(ctx->slot_mtx);
		if(!ctx->free_slots.empty()) { *out = ctx->free_slots.back(); ctx->free_slots.pop_back(); return Y4GPU_OK; }
	}
	Y4TraceSlot *s = new Y4TraceSlot;
	bool ok = cudaEventCreate(&s->ev0) == cudaSuccess && cudaEventCreate(&s->ev1) == cudaSuccess && cudaEventCreateWithFlags(&s->ev_dep, cudaEventDisableTiming) == cudaSuccess
		&& cudaMalloc(&s->d_counters, Y4TraceSlot::kMaxChunks * sizeof(unsigned long long)) == cudaSuccess;
	for(int i = 0; ok && i < Y4TraceSlot::kStreams; ++i)
		ok = cudaStreamCreateWithFlags(&s->streams[i], cudaStreamNonBlocking) == cudaSuccess && cudaEventCreateWithFlags(&s->ev_join[i], cudaEventDisableTiming) == cudaSuccess;
	{
		std::lock_guard<std::mutex> lk(ctx->slot_mtx);
		ctx->all_slots.push_back(s);              // owned by the context from here on (freed in y4gpu_destroy)
	}
	if(!ok) { slot_error(ctx, std::string("trace slot setup failed: ") + cudaGetErrorString(cudaGetLastError())); return Y4GPU_ERR_CUDA; }
	*out = s;
	return Y4GPU_OK;
}
static void slot_release(y4gpu_ctx *ctx, Y4TraceSlot *s)
{
	std::lock_guard<std::mutex> lk(ctx->slot_mtx);
	ctx->free_slots.push_back(s);
}
static void slot_free_all(y4gpu_ctx *ctx)
{
	for(Y4TraceSlot *s : ctx->all_slots)
	{
		for(int i = 0; i < Y4TraceSlot::kStreams; ++i)
		{
			if(s->streams[i]) { cudaStreamSynchronize(s->streams[i]); cudaStreamDestroy(s->streams[i]); }
			if(s->ev_join[i]) cudaEventDestroy(s->ev_join[i]);
		}
		if(s->ev0) cudaEventDestroy(s->ev0);
		if(s->ev1) cudaEventDestroy(s->ev1);
		if(s->ev_dep) cudaEventDestroy(s->ev_dep);
		cudaFree(s->d_rays); cudaFree(s->d_hits); cudaFree(s->d_counters); cudaFree(s->d_mat_transp);
		for(int i = 0; i < Y4TraceSlot::kStreams; ++i) cudaFree(s->sort[i].buf);
		if(s->h_pinned) cudaFreeHost(s->h_pinned);
		delete s;
	}
	ctx->all_slots.clear(); ctx->free_slots.clear();
}
static int slot_ensure(y4gpu_ctx *ctx, Y4TraceSlot *s, uint64_t n, uint64_t hit_bytes)
{
	if(n > s->cap_rays)
	{
		cudaFree(s->d_rays); s->d_rays = nullptr; s->cap_rays = 0;
		const uint64_t cap = n < 4096 ? 4096 : n;
		Y4_SLOT_CUDA(ctx, cudaMalloc(&s->d_rays, cap * sizeof(Y4Ray)));
		s->cap_rays = cap;
	}
	if(n * hit_bytes > s->cap_hit_bytes)
	{
		cudaFree(s->d_hits); s->d_hits = nullptr; s->cap_hit_bytes = 0;
		const uint64_t cap = n * hit_bytes < 4096 * 32 ? 4096 * 32 : n * hit_bytes;
		Y4_SLOT_CUDA(ctx, cudaMalloc(&s->d_hits, cap));
		s->cap_hit_bytes = cap;
	}
	return Y4GPU_OK;
}

constexpr uint64_t kSmallBatch = 2048;          // up to here rays / hits are staged through the slot's pinned buffer
constexpr uint64_t kMinChunk = 1ull << 18;      // rays per pipelined chunk, at least

// Host rays in, host hits out.  Small batches: one stream, pinned staging.  Large batches: the batch is cut into chunks that
// go round-robin over the slot's streams (H2D copy -> kernel -> D2H copy per chunk), so the copies of one chunk overlap the
// traversal of another and both copy engines stay busy; the caller's buffers should be page-locked (y4gpu_host_alloc) for the
// copies to be asynchronous.  tshadow: mat != null.
// tuning knobs of the ray sort, read when a context is created and again by y4gpu_set_ray_sort
static void ray_sort_env(y4gpu_ctx *ctx)
{
	if(const char *e = getenv("Y4GPU_RAY_SORT")) { const int k = atoi(e); ctx->ray_sort = k < 0 ? 0 : (k > 2 ? 2 : k); }
	if(const char *e = getenv("Y4GPU_RAY_SORT_OBITS")) { const int k = atoi(e); ctx->ray_sort_obits = k < 1 ? 1 : (k > 8 ? 8 : k); }
	if(const char *e = getenv("Y4GPU_RAY_SORT_DIR_MAJOR")) ctx->ray_sort_dir_major = atoi(e) != 0;
	if(const char *e = getenv("Y4GPU_RAY_SORT_MIN_DISTINCT")) ctx->ray_sort_min_distinct = (float) atof(e);
	if(const char *e = getenv("Y4GPU_RAY_SORT_MIN_BVH_MB")) ctx->ray_sort_min_bvh_bytes = (uint64_t) atoll(e) << 20;
}

static int trace_host(y4gpu_ctx *ctx, int kind, const y4gpu_ray *rays, uint64_t n, void *out, uint64_t hit_bytes,
                      int max_depth, const float *mat_transparency, uint32_t n_mat, bool inner)
{
	Y4TraceSlot *s = nullptr;
	{ const int rc = slot_acquire(ctx, &s); if(rc) return rc; }
	struct Release { y4gpu_ctx *c; Y4TraceSlot *s; ~Release() { for(int i = 0; i < Y4TraceSlot::kStreams; ++i) cudaStreamSynchronize(s->streams[i]); slot_release(c, s); } } release{ctx, s};
	{ const int rc = slot_ensure(ctx, s, n, hit_bytes); if(rc) return rc; }
	const bool tshadow = kind == Y4_TRACE_TSHADOW;
	if(tshadow && n_mat)
	{
		if(n_mat > s->mat_transp_cap)
		{
			cudaFree(s->d_mat_transp); s->d_mat_transp = nullptr; s->mat_transp_cap = 0;
			Y4_SLOT_CUDA(ctx, cudaMalloc(&s->d_mat_transp, (size_t) n_mat * 3 * sizeof(float)));
			s->mat_transp_cap = n_mat;
		}
		Y4_SLOT_CUDA(ctx, cudaMemcpyAsync(s->d_mat_transp, mat_transparency, (size_t) n_mat * 3 * sizeof(float), cudaMemcpyDefault, s->streams[0]));
	}
	uint64_t n_chunks = 1, chunk = n;
	if(n >= 2 * kMinChunk)
	{
		n_chunks = n / kMinChunk;
		if(n_chunks > 16) n_chunks = 16;
		chunk = ((n + n_chunks - 1) / n_chunks + 31) & ~31ull;
		n_chunks = (n + chunk - 1) / chunk;
	}
	cudaStream_t s0 = s->streams[0];
	if(ctx->stream != ctx->own_stream)
	{
		// a caller-installed stream: whatever it queued before this call (e.g. an asynchronous copy into the ray buffer) comes first
		Y4_SLOT_CUDA(ctx, cudaEventRecord(s->ev_dep, ctx->stream));
		Y4_SLOT_CUDA(ctx, cudaStreamWaitEvent(s0, s->ev_dep, 0));
	}
	Y4_SLOT_CUDA(ctx, cudaMemsetAsync(s->d_counters, 0, n_chunks * sizeof(unsigned long long), s0));
	Y4_SLOT_CUDA(ctx, cudaEventRecord(s->ev0, s0));
	const int n_streams = n_chunks > 1 ? Y4TraceSlot::kStreams : 1;
	for(int i = 1; i < n_streams; ++i) Y4_SLOT_CUDA(ctx, cudaStreamWaitEvent(s->streams[i], s->ev0, 0));
	const bool small = n <= kSmallBatch;
	char *h_stage = nullptr;
	if(small)
	{
		const size_t need = n * (sizeof(Y4Ray) + hit_bytes);
		if(need > s->pinned_bytes)
		{
			if(s->h_pinned) cudaFreeHost(s->h_pinned);
			s->h_pinned = nullptr; s->pinned_bytes = 0;
			const size_t cap = kSmallBatch * (sizeof(Y4Ray) + 32);
			Y4_SLOT_CUDA(ctx, cudaMallocHost(&s->h_pinned, cap));
			s->pinned_bytes = cap;
		}
		h_stage = static_cast<char *>(s->h_pinned);
		memcpy(h_stage, rays, n * sizeof(Y4Ray));
	}
	// sorted chunks (y4gpu_set_ray_sort): decided once per batch on its first chunk; every stream sorts in its own scratch
	bool sorted = false;
	const bool may_sort = !tshadow && ctx->ray_sort != 0 && chunk >= (1u << 16);
	if(may_sort)
		for(int i = 0; i < n_streams; ++i) { const int rc = y4_sort_scratch_reserve(ctx, s->sort[i], chunk); if(rc) return rc; }
	for(uint64_t c = 0; c < n_chunks; ++c)
	{
		cudaStream_t st = s->streams[c % (uint64_t) n_streams];
		const uint64_t first = c * chunk, cnt = (n - first) < chunk ? (n - first) : chunk;
		Y4_SLOT_CUDA(ctx, cudaMemcpyAsync(s->d_rays + first, small ? (const void *) h_stage : (const void *) (rays + first), cnt * sizeof(Y4Ray), cudaMemcpyHostToDevice, st));
		char *d_hits = static_cast<char *>(s->d_hits) + first * hit_bytes;
		int rc;
		if(may_sort && c == 0) { rc = y4_ray_sort_wanted(ctx, s->d_rays, cnt, s->sort[0], st, sorted); if(rc) return rc; }
		if(sorted && cnt >= (1u << 16)) rc = y4_launch_trace_sorted(ctx, kind, s->d_rays + first, cnt, reinterpret_cast<Y4Hit *>(d_hits), s->d_counters + c, s->sort[c % (uint64_t) n_streams], st);
		else if(tshadow) rc = y4_launch_tshadow_on(ctx, s->d_rays + first, cnt, max_depth, s->d_mat_transp, n_mat, reinterpret_cast<y4gpu_tshit *>(d_hits), inner, s->d_counters + c, st);
		else rc = y4_launch_trace_on(ctx, kind, s->d_rays + first, cnt, reinterpret_cast<Y4Hit *>(d_hits), s->d_counters + c, st);
		if(rc) return rc;
		char *h_dst = small ? h_stage + n * sizeof(Y4Ray) : static_cast<char *>(out) + first * hit_bytes;
		Y4_SLOT_CUDA(ctx, cudaMemcpyAsync(h_dst, d_hits, cnt * hit_bytes, cudaMemcpyDeviceToHost, st));
	}
	for(int i = 1; i < n_streams; ++i)
	{
		Y4_SLOT_CUDA(ctx, cudaEventRecord(s->ev_join[i], s->streams[i]));
		Y4_SLOT_CUDA(ctx, cudaStreamWaitEvent(s0, s->ev_join[i], 0));
	}
	Y4_SLOT_CUDA(ctx, cudaEventRecord(s->ev1, s0));
	Y4_SLOT_CUDA(ctx, cudaStreamSynchronize(s0));
	if(small) memcpy(out, h_stage + n * sizeof(Y4Ray), n * hit_bytes);
	float ms = 0.f;
	cudaEventElapsedTime(&ms, s->ev0, s->ev1);
	{
		std::lock_guard<std::mutex> lk(ctx->stat_mtx);
		ctx->stats.kernel_launches += sorted ? 2 * n_chunks : n_chunks;
		ctx->stats.last_trace_sorted = sorted ? 1 : 0;
		ctx->stats.last_trace_rays = n;
		ctx->stats.last_trace_ms = ms;          // copies included (the whole pipelined batch)
		ctx->last_trace_async = false;
	}
	return Y4GPU_OK;
}

static int trace_impl(y4gpu_ctx *ctx, int kind, bool counted, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out)
{
	if(!ctx || (n && (!rays || !out))) return Y4GPU_ERR_ARG;
	if(kind != Y4_TRACE_CLOSEST && kind != Y4_TRACE_SHADOW) { slot_error(ctx, "bad trace kind"); return Y4GPU_ERR_ARG; }
	if(!ctx->built) { slot_error(ctx, "y4gpu_trace before y4gpu_build"); return Y4GPU_ERR_ARG; }
	if(n == 0) return Y4GPU_OK;
	if(cudaSetDevice(ctx->device) != cudaSuccess) { slot_error(ctx, "cudaSetDevice failed"); return Y4GPU_ERR_CUDA; }
	const bool rays_dev = is_device_ptr(rays), out_dev = is_device_ptr(out);
	// host buffers on both sides: the re-entrant path (synchronous; ordered after the work already queued on a caller-installed stream)
	if(!rays_dev && !out_dev && !counted)
		return trace_host(ctx, kind, rays, n, out, sizeof(Y4Hit), 0, nullptr, 0, false);
	std::lock_guard<std::mutex> lock(ctx->mtx);
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
	ctx->last_trace_async = true;
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

int y4gpu_set_ray_sort(y4gpu_ctx *ctx, int mode)
{
	if(!ctx || mode < 0 || mode > 2) return Y4GPU_ERR_ARG;
	std::lock_guard<std::mutex> lock(ctx->mtx);
	ctx->ray_sort = mode;
	const int keep = ctx->ray_sort;
	ray_sort_env(ctx);                 // the finer knobs (key bits, thresholds) follow the environment
	ctx->ray_sort = keep;
	return Y4GPU_OK;
}
int y4gpu_trace_counted(y4gpu_ctx *ctx, int kind, const y4gpu_ray *rays, uint64_t n, y4gpu_hit *out) { return trace_impl(ctx, kind, true, rays, n, out); }

static int tshadow_impl(y4gpu_ctx *ctx, const y4gpu_ray *rays, uint64_t n, int max_depth, const float *mat_transparency, uint32_t n_mat, y4gpu_tshit *out, bool inner)
{
	if(!ctx || (n && (!rays || !out)) || (n_mat && !mat_transparency) || max_depth < 0) return Y4GPU_ERR_ARG;
	if(max_depth > 16) { slot_error(ctx, "y4gpu_trace_tshadow: max_depth > 16 is not supported"); return Y4GPU_ERR_UNSUPPORTED; }
	if(!ctx->built) { slot_error(ctx, "y4gpu_trace_tshadow before y4gpu_build"); return Y4GPU_ERR_ARG; }
	if(n == 0) return Y4GPU_OK;
	if(cudaSetDevice(ctx->device) != cudaSuccess) { slot_error(ctx, "cudaSetDevice failed"); return Y4GPU_ERR_CUDA; }
	const bool rays_dev = is_device_ptr(rays), out_dev = is_device_ptr(out);
	if(!rays_dev && !out_dev)
		return trace_host(ctx, Y4_TRACE_TSHADOW, rays, n, out, sizeof(y4gpu_tshit), max_depth, mat_transparency, n_mat, inner);
	std::lock_guard<std::mutex> lock(ctx->mtx);
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
	Y4_CUDA(ctx, cudaMemsetAsync(ctx->d_work_counter, 0, sizeof(unsigned long long), ctx->stream));
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
	const int rc = y4_launch_tshadow_on(ctx, d_rays, n, max_depth, ctx->d_mat_transp, n_mat, d_out, inner, reinterpret_cast<unsigned long long *>(ctx->d_work_counter), ctx->stream);
	if(rc) return rc;
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
	ctx->stats.kernel_launches++; ctx->stats.last_trace_rays = n; ctx->last_trace_async = true;
	if(!out_dev) Y4_CUDA(ctx, cudaMemcpyAsync(out, d_out, n * sizeof(y4gpu_tshit), cudaMemcpyDeviceToHost, ctx->stream));
	Y4_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return Y4GPU_OK;
}

int y4gpu_trace_tshadow(y4gpu_ctx *ctx, const y4gpu_ray *rays, uint64_t n, int max_depth, const float *mat_transparency, uint32_t n_mat, y4gpu_tshit *out)
{
	return tshadow_impl(ctx, rays, n, max_depth, mat_transparency, n_mat, out, false);
}
int y4gpu_trace_tshadow_inner(y4gpu_ctx *ctx, const y4gpu_ray *rays, uint64_t n, int max_depth, const float *mat_transparency, uint32_t n_mat, y4gpu_tshit *out)
{
	return tshadow_impl(ctx, rays, n, max_depth, mat_transparency, n_mat, out, true);
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
	out->last_probe_distinct = ctx->last_probe_distinct;
	float ms = 0.f;
	// a batch queued on the context's stream is timed by its events once they have completed; the slot path stores its time directly
	if(ctx->last_trace_async)
	{
		if(ctx->stats.last_trace_rays && cudaEventQuery(ctx->ev1) == cudaSuccess && cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) == cudaSuccess) out->last_trace_ms = ms;
		else cudaGetLastError();
	}
	return Y4GPU_OK;
}

} // extern "C"
