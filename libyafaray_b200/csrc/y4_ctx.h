// Context object behind the C-ABI (include/y4gpu.h).  Host-side only.
#pragma once
#include <cuda_runtime.h>
#include <mutex>
#include <string>
#include <vector>
#include "y4_trace.cuh"
#include "../../include/y4gpu.h"

struct Y4Pt;   // wavefront path tracer state (y4_pt.cu)

// Everything one host-buffer ray batch needs, so that concurrent callers (the Accelerator virtuals are called from every
// render thread, integrator_tiled.cc:248-251) never share a stream, a staging buffer or a work counter.  Slots are pooled
// per context; a call takes one off the free list (or creates it) and puts it back when its streams have drained.
// device scratch of a sorted ray batch (y4_trace.cu: keys, indices, CUB temporary storage); grown on demand, never shrunk
struct Y4SortScratch { void *buf = nullptr; size_t cap = 0; };

struct Y4TraceSlot
{
	static constexpr int kStreams = 4, kMaxChunks = 64;
	cudaStream_t streams[kStreams] = {nullptr, nullptr, nullptr, nullptr};
	cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_dep = nullptr, ev_join[kStreams] = {nullptr, nullptr, nullptr, nullptr};
	Y4Ray *d_rays = nullptr; void *d_hits = nullptr;
	uint64_t cap_rays = 0, cap_hit_bytes = 0;
	unsigned long long *d_counters = nullptr;        // kMaxChunks work counters
	float *d_mat_transp = nullptr; uint32_t mat_transp_cap = 0;
	void *h_pinned = nullptr; size_t pinned_bytes = 0;   // staging of small batches (1-ray calls): truly asynchronous copies
	Y4SortScratch sort[kStreams];                        // sorted chunks: one scratch per stream
};

struct y4gpu_ctx
{
	int device = 0;
	int sm_count = 148;
	cudaStream_t stream = nullptr;      // stream in use (own_stream or the caller's)
	cudaStream_t own_stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	mutable std::mutex mtx;
	std::string err;

	// uploaded geometry (primitive order)
	float *d_verts = nullptr;
	uint32_t *d_flags = nullptr;
	uint64_t n_tri = 0;

	// accelerator
	Y4U4 *d_nodes = nullptr;
	Y4Tri *d_tris = nullptr;
	uint64_t n_nodes = 0;
	int bvh_levels = 0;                 // levels of the BVH8 (collapse rounds of the last build)
	bool built = false;
	float raw_bound[6] = {0, 0, 0, 0, 0, 0};
	float bound[6] = {0, 0, 0, 0, 0, 0};

	// two-level scene (y4gpu_upload_instanced): base meshes in object space + instances; n_tri is then the number of instanced
	// triangles (the primitive index space), d_verts / d_flags / d_nodes / d_tris stay empty
	bool two_level = false;
	float *d_mesh_verts = nullptr; uint32_t *d_mesh_flags = nullptr;      // all meshes, concatenated
	std::vector<uint64_t> mesh_first;                                      // [n_mesh + 1] first triangle of mesh i in the concatenation
	std::vector<Y4Inst> h_inst;                                            // as uploaded (node_off / tri_off / bmax filled by the build)
	Y4Inst *d_inst = nullptr;
	Y4U4 *d_tlas_nodes = nullptr; Y4Tri *d_tlas_leaves = nullptr;
	Y4U4 *d_blas_nodes = nullptr; Y4TriObj *d_blas_tris = nullptr;
	uint64_t n_tlas_nodes = 0, n_blas_nodes = 0, n_base_tris = 0;

	// staging for host-pointer batches and work counters
	Y4Ray *d_rays = nullptr;
	void *d_hits = nullptr;
	uint64_t stage_cap = 0, stage_hit_bytes = 0;
	uint32_t *d_work_counter = nullptr;
	uint32_t *d_consts = nullptr;                     // [0] = 0x4B000000, see y4_byte_m
	unsigned long long *d_visit_counters = nullptr;   // [2] nodes, tris
	float *d_mat_transp = nullptr;
	uint32_t mat_transp_cap = 0;

	// warp scheduling knobs of k_trace (env Y4GPU_FETCH_THRESHOLD / Y4GPU_POSTPONE_DIV override)
	int fetch_threshold = 20, postpone_div = 5, min_blocks = 3;
	int pt_fetch_threshold = 16;               // refill threshold of the path tracer's wavefront launches (env Y4GPU_PT_FETCH_THRESHOLD; C2 sweep: profiles/r2_v8_sweep_trace_knobs.log)
	int morton_sized = -1;     // key order: -1 build both binary trees, keep the smaller surface-area sum; 0 plain 21-bit-per-axis Morton keys; 1 extended keys with size bits (env Y4GPU_MORTON_SIZED)
	int morton_used = 0;       // what the last build chose
	int treelet_defer = 64;    // sub-ranges of a treelet with at most this many leaves are split by a second, wider launch (env Y4GPU_TREELET_DEFER; 0 = one pass)
	int treelet_leaves = 2048;  // SAH rebuild of the binary subtrees with at most this many leaves (0 = plain LBVH; env Y4GPU_TREELET, <= 4096)

	// ray sorting (y4gpu_set_ray_sort; env Y4GPU_RAY_SORT / _OBITS / _DIR_MAJOR / _MIN_DISTINCT / _MIN_BVH_MB at creation)
	int ray_sort = 2;                          // 0 never, 1 every batch of >= 65536 rays, 2 automatic (probe)
	int ray_sort_obits = 6, ray_sort_dir_major = 0;
	float ray_sort_min_distinct = 12.f;        // automatic mode: mean number of distinct coarse keys per 32 consecutive rays
	uint64_t ray_sort_min_bvh_bytes = 256ull << 20;
	float last_probe_distinct = 0.f;
	Y4SortScratch sort_scratch;

	y4gpu_stats stats = {};
	bool last_trace_async = false;             // the last batch ran on the context's stream and is timed by ev0 / ev1
	std::mutex stat_mtx;                       // guards `stats` updates of the slot path
	std::mutex slot_mtx;                       // guards the slot pool
	std::vector<Y4TraceSlot *> free_slots, all_slots;
	Y4Pt *pt = nullptr;

	Y4Scene2 view2() const
	{
		Y4Scene2 v; v.tlas_nodes = d_tlas_nodes; v.tlas_leaves = d_tlas_leaves; v.inst = d_inst; v.n_inst = (uint32_t) h_inst.size();
		v.blas_nodes = d_blas_nodes; v.blas_tris = d_blas_tris; v.consts = d_consts; return v;
	}
	Y4Bvh8 view() const { Y4Bvh8 v; v.nodes = d_nodes; v.tris = d_tris; v.n_tris = (uint32_t) n_tri; v.consts = d_consts; v.postpone_sp_limit = Y4_STACK_SIZE - bvh_levels; return v; }
};

#define Y4_CUDA(ctx, call)                                                                       \
	do {                                                                                         \
		const cudaError_t e_ = (call);                                                           \
		if(e_ != cudaSuccess)                                                                    \
		{                                                                                        \
			(ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                     \
			return Y4GPU_ERR_CUDA;                                                               \
		}                                                                                        \
	} while(0)

// y4_build.cu
int y4_build_bvh(y4gpu_ctx *ctx, int leaf_max);
int y4_build_two_level(y4gpu_ctx *ctx, int leaf_max);       // after y4gpu_upload_instanced
void y4_free_two_level(y4gpu_ctx *ctx);                     // the built two-level structures (not the uploaded meshes)
// y4_trace.cu
int y4_launch_trace(y4gpu_ctx *ctx, int kind, bool counted, const Y4Ray *d_rays, uint64_t n, Y4Hit *d_out);
int y4_launch_trace_wave(y4gpu_ctx *ctx, int kind, bool counted, const Y4Ray *d_rays, Y4Hit *d_out, uint64_t n_max,
                         const uint32_t *d_queue, const uint32_t *d_queue_count, unsigned long long *d_work_counter, unsigned long long *d_visit,
                         cudaStream_t stream, int max_blocks_per_sm);
int y4_launch_tshadow_wave(y4gpu_ctx *ctx, const Y4Ray *d_rays, Y4Hit *d_out, const uint32_t *d_queue, const uint32_t *d_queue_count, int max_depth,
                           const float *d_mat_transp, uint32_t n_mat, unsigned long long *d_work_counter, cudaStream_t stream, int max_blocks_per_sm);
int y4_launch_tshadow(y4gpu_ctx *ctx, const Y4Ray *d_rays, uint64_t n, int max_depth, const float *d_mat_transp, uint32_t n_mat, y4gpu_tshit *d_out);
// ray sorting: decision (automatic mode probes the batch: one small kernel + a synchronisation of `stream`) and the sorted launch
int y4_sort_scratch_reserve(y4gpu_ctx *ctx, Y4SortScratch &sc, uint64_t n);
int y4_ray_sort_wanted(y4gpu_ctx *ctx, const Y4Ray *d_rays, uint64_t n, Y4SortScratch &sc, cudaStream_t stream, bool &want);
int y4_launch_trace_sorted(y4gpu_ctx *ctx, int kind, const Y4Ray *d_rays, uint64_t n, Y4Hit *d_out, unsigned long long *d_work_counter,
                           Y4SortScratch &sc, cudaStream_t stream);
// launches on an explicit stream with an explicit (pre-zeroed) work counter: the re-entrant host-buffer path of y4_api.cu
int y4_launch_trace_on(y4gpu_ctx *ctx, int kind, const Y4Ray *d_rays, uint64_t n, Y4Hit *d_out, unsigned long long *d_work_counter, cudaStream_t stream);
int y4_launch_tshadow_on(y4gpu_ctx *ctx, const Y4Ray *d_rays, uint64_t n, int max_depth, const float *d_mat_transp, uint32_t n_mat, y4gpu_tshit *d_out,
                         bool inner, unsigned long long *d_work_counter, cudaStream_t stream);
