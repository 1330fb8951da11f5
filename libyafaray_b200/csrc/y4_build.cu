// GPU construction of the BVH8: triangle boxes -> Morton keys -> radix sort -> Karras
// hierarchy -> refit -> level-synchronous collapse into compressed 8-wide nodes.
// Replaces the reference's serial kd-tree build
// (/root/reference/src/accelerator/accelerator_kdtree.cc:47-123,420-628).
// Per-element logic lives in y4_bvh_build.cuh / y4_bvh8.cuh (shared with the unit-test
// emulation); this file holds the kernels and the host orchestration.
#include "y4_ctx.h"
#include <cub/device/device_radix_sort.cuh>
#include <vector>

namespace {

__device__ __forceinline__ uint32_t enc_f(float f)
{
	const uint32_t u = __float_as_uint(f);
	return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
inline float dec_f(uint32_t e)
{
	const uint32_t u = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
	float f; memcpy(&f, &u, 4); return f;
}

// 1. per-triangle boxes + scene bound (warp shuffle reduce, one atomic set per warp)
__global__ void __launch_bounds__(256) k_tri_boxes(const float *__restrict__ verts, int n, Y4Box *__restrict__ tb, uint32_t *__restrict__ scene_enc)
{
	Y4Box acc = y4_box_empty();
	for(int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
	{
		float v[9];
#pragma unroll
		for(int k = 0; k < 9; ++k) v[k] = __ldg(verts + 9ull * i + k);
		const Y4Box b = y4_tri_box(v);
		tb[i] = b;
		acc = y4_box_union(acc, b);
	}
#pragma unroll
	for(int o = 16; o > 0; o >>= 1)
	{
		acc.lox = fminf(acc.lox, __shfl_xor_sync(0xffffffffu, acc.lox, o));
		acc.loy = fminf(acc.loy, __shfl_xor_sync(0xffffffffu, acc.loy, o));
		acc.loz = fminf(acc.loz, __shfl_xor_sync(0xffffffffu, acc.loz, o));
		acc.hix = fmaxf(acc.hix, __shfl_xor_sync(0xffffffffu, acc.hix, o));
		acc.hiy = fmaxf(acc.hiy, __shfl_xor_sync(0xffffffffu, acc.hiy, o));
		acc.hiz = fmaxf(acc.hiz, __shfl_xor_sync(0xffffffffu, acc.hiz, o));
	}
	if((threadIdx.x & 31) == 0)
	{
		atomicMin(scene_enc + 0, enc_f(acc.lox)); atomicMin(scene_enc + 1, enc_f(acc.loy)); atomicMin(scene_enc + 2, enc_f(acc.loz));
		atomicMax(scene_enc + 3, enc_f(acc.hix)); atomicMax(scene_enc + 4, enc_f(acc.hiy)); atomicMax(scene_enc + 5, enc_f(acc.hiz));
	}
}

// 2. Morton keys
__global__ void __launch_bounds__(256) k_morton(const Y4Box *__restrict__ tb, int n, Y4Box scene, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals, int sized)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	keys[i] = sized ? y4_morton63_sized(tb[i], scene) : y4_morton63(tb[i], scene);
	vals[i] = (uint32_t) i;
}

// 4. Karras hierarchy, one thread per internal node
__global__ void __launch_bounds__(256) k_karras(const uint64_t *__restrict__ keys, int n, int *left, int *right, int *parent, int *first, int *last)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n - 1) return;
	y4_karras_node(keys, n, i, left, right, parent, first, last);
}

__device__ __forceinline__ Y4Box ld_box_cg(const Y4Box *p)
{
	const float *f = reinterpret_cast<const float *>(p);
	Y4Box b;
	b.lox = __ldcg(f + 0); b.loy = __ldcg(f + 1); b.loz = __ldcg(f + 2);
	b.hix = __ldcg(f + 3); b.hiy = __ldcg(f + 4); b.hiz = __ldcg(f + 5);
	return b;
}

// 5. leaf boxes + bottom-up refit; the second thread to reach a node owns it
__global__ void __launch_bounds__(256) k_refit(int n, const uint32_t *__restrict__ sorted_prim, const Y4Box *__restrict__ tb, Y4Box *box,
                                               const int *__restrict__ parent, const int *__restrict__ left, const int *__restrict__ right, int *flags)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if(k >= n) return;
	box[n - 1 + k] = tb[sorted_prim[k]];
	if(n == 1) return;
	int cur = parent[n - 1 + k];
	while(cur >= 0)
	{
		__threadfence();
		if(atomicAdd(flags + cur, 1) == 0) return;
		const Y4Box a = ld_box_cg(box + left[cur]), b = ld_box_cg(box + right[cur]);
		box[cur] = y4_box_union(a, b);
		cur = parent[cur];
	}
}

// 5a. SAH refinement: collect the maximal subtrees with <= max_leaves leaves, rebuild each with binned SAH (one thread per treelet)
__global__ void __launch_bounds__(256) k_treelet_roots(int n, int max_leaves, const int *__restrict__ parent, const int *__restrict__ first,
                                                       const int *__restrict__ last, int *list, uint32_t *count)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n - 1) return;
	if(y4_is_treelet_root(n, max_leaves, parent, first, last, i)) list[atomicAdd(count, 1u)] = i;
}
// One warp per treelet (CTA = 1 warp, persistent over the root list).  The leaf boxes of the treelet sit in shared memory
// (SoA) and stay where they are; the splits permute a 16-bit index array.  Per split: warp-reduced range box and centroid
// bounds, 3 x 16 bins filled with shared-memory atomics on order-preserving float encodings, the 45 candidate planes
// costed by the lanes, warp arg-min (lowest candidate on ties, like the sequential y4_treelet_sah), stable ballot
// partition.  Same decisions as y4_treelet_sah (y4_bvh_build.cuh), which the unit-test emulation runs.
__device__ __forceinline__ float dec_f_dev(uint32_t e) { return __uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e); }
__device__ __forceinline__ float warp_min(float v) { for(int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o)); return v; }
__device__ __forceinline__ float warp_max(float v) { for(int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o)); return v; }
constexpr int kBinWords = 7;     // lo xyz, hi xyz (encoded), count

// Two passes: the first one splits the big ranges of every treelet and hands the sub-ranges with <= defer_limit leaves
// (most of the splits, each with a fixed per-split cost) to `defer_list`; the second pass runs the same kernel over that
// list, one warp per small range, with far more warps in flight.
__global__ void __launch_bounds__(32) k_treelet_sah(Y4TreeletArgs t, const int *__restrict__ list, const uint32_t *__restrict__ count, int K,
                                                    int defer_limit, int *defer_list, uint32_t *defer_count)
{
	extern __shared__ __align__(16) unsigned char smem[];
	float *bx = reinterpret_cast<float *>(smem);                         // [6][K]
	uint32_t *prim = reinterpret_cast<uint32_t *>(bx + 6 * K);           // [K]
	uint32_t *bins = prim + K;                                           // [3][16][7]
	int *stack = reinterpret_cast<int *>(bins + 3 * Y4_SAH_BINS * kBinWords);   // [3 * 16]
	uint16_t *idx = reinterpret_cast<uint16_t *>(stack + 48);            // [K]
	uint16_t *idx2 = idx + K;                                            // [K]
	const int lane = threadIdx.x;
	const int n = t.n;
	Y4Box *leaf = t.box + (n - 1);
	const uint32_t n_roots = *count;
	for(uint32_t ti = blockIdx.x; ti < n_roots; ti += gridDim.x)
	{
		const int root = list[ti];
		const int lo = t.first[root], M = t.last[root] - lo + 1;
		__syncwarp();
		for(int k = lane; k < M; k += 32)
		{
			const Y4Box b = leaf[lo + k];
			bx[k] = b.lox; bx[K + k] = b.loy; bx[2 * K + k] = b.loz; bx[3 * K + k] = b.hix; bx[4 * K + k] = b.hiy; bx[5 * K + k] = b.hiz;
			prim[k] = t.sorted_prim[lo + k];
			idx[k] = (uint16_t) k;
		}
		__syncwarp();
		int sp = 0, a = 0, b = M - 1, self = root;
		for(;;)
		{
			const int m = b - a + 1;
			float rl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, rh[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
			float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
			for(int k = a + lane; k <= b; k += 32)
			{
				const int j = idx[k];
#pragma unroll
				for(int ax = 0; ax < 3; ++ax)
				{
					const float l = bx[ax * K + j], h = bx[(3 + ax) * K + j], c = 0.5f * (l + h);
					rl[ax] = fminf(rl[ax], l); rh[ax] = fmaxf(rh[ax], h); clo[ax] = fminf(clo[ax], c); chi[ax] = fmaxf(chi[ax], c);
				}
			}
#pragma unroll
			for(int ax = 0; ax < 3; ++ax) { rl[ax] = warp_min(rl[ax]); rh[ax] = warp_max(rh[ax]); clo[ax] = warp_min(clo[ax]); chi[ax] = warp_max(chi[ax]); }
			int g = a + (m >> 1) - 1;
			if(m > 2)
			{
				float scale[3];
#pragma unroll
				for(int ax = 0; ax < 3; ++ax) { const float e = chi[ax] - clo[ax]; scale[ax] = e > 0.f ? (float) Y4_SAH_BINS / e : 0.f; }
				for(int w = lane; w < 3 * Y4_SAH_BINS * kBinWords; w += 32) { const int f = w % kBinWords; bins[w] = f < 3 ? 0xffffffffu : 0u; }
				__syncwarp();
				for(int k = a + lane; k <= b; k += 32)
				{
					const int j = idx[k];
#pragma unroll
					for(int ax = 0; ax < 3; ++ax)
					{
						if(scale[ax] == 0.f) continue;
						const int q = y4_sah_bin(0.5f * (bx[ax * K + j] + bx[(3 + ax) * K + j]), clo[ax], scale[ax]);
						uint32_t *B = bins + (ax * Y4_SAH_BINS + q) * kBinWords;
						atomicMin(B + 0, enc_f(bx[j])); atomicMin(B + 1, enc_f(bx[K + j])); atomicMin(B + 2, enc_f(bx[2 * K + j]));
						atomicMax(B + 3, enc_f(bx[3 * K + j])); atomicMax(B + 4, enc_f(bx[4 * K + j])); atomicMax(B + 5, enc_f(bx[5 * K + j]));
						atomicAdd(B + 6, 1u);
					}
				}
				__syncwarp();
				float best = FLT_MAX; int best_c = 0x7fffffff, best_nl = 0;
				for(int c = lane; c < 3 * (Y4_SAH_BINS - 1); c += 32)
				{
					const int ax = c / (Y4_SAH_BINS - 1), q = c % (Y4_SAH_BINS - 1);
					if(scale[ax] == 0.f) continue;
					Y4Box L = y4_box_empty(), R = y4_box_empty(); int nl = 0, nr = 0;
					for(int w = 0; w < Y4_SAH_BINS; ++w)
					{
						const uint32_t *B = bins + (ax * Y4_SAH_BINS + w) * kBinWords;
						const int cw = (int) B[6];
						if(!cw) continue;
						Y4Box bb; bb.lox = dec_f_dev(B[0]); bb.loy = dec_f_dev(B[1]); bb.loz = dec_f_dev(B[2]); bb.hix = dec_f_dev(B[3]); bb.hiy = dec_f_dev(B[4]); bb.hiz = dec_f_dev(B[5]);
						if(w <= q) { L = y4_box_union(L, bb); nl += cw; } else { R = y4_box_union(R, bb); nr += cw; }
					}
					if(!nl || !nr) continue;
					const float cost = y4_box_half_area(L) * (float) nl + y4_box_half_area(R) * (float) nr;
					if(cost < best) { best = cost; best_c = c; best_nl = nl; }      // c ascends within a lane
				}
				for(int o = 16; o > 0; o >>= 1)
				{
					const float oc = __shfl_xor_sync(0xffffffffu, best, o);
					const int ocand = __shfl_xor_sync(0xffffffffu, best_c, o), onl = __shfl_xor_sync(0xffffffffu, best_nl, o);
					if(oc < best || (oc == best && ocand < best_c)) { best = oc; best_c = ocand; best_nl = onl; }
				}
				if(best_c != 0x7fffffff)
				{
					const int ax = best_c / (Y4_SAH_BINS - 1), q = best_c % (Y4_SAH_BINS - 1);
					int base_l = a, base_r = a + best_nl;
					for(int k0 = a; k0 <= b; k0 += 32)
					{
						const int k = k0 + lane;
						const bool valid = k <= b;
						const int j = valid ? idx[k] : 0;
						const bool fl = valid && y4_sah_bin(0.5f * (bx[ax * K + j] + bx[(3 + ax) * K + j]), clo[ax], scale[ax]) <= q;
						const uint32_t ml = __ballot_sync(0xffffffffu, fl), mr = __ballot_sync(0xffffffffu, valid && !fl);
						const uint32_t lt = (1u << lane) - 1u;
						if(fl) idx2[base_l + __popc(ml & lt)] = (uint16_t) j;
						else if(valid) idx2[base_r + __popc(mr & lt)] = (uint16_t) j;
						base_l += __popc(ml); base_r += __popc(mr);
					}
					__syncwarp();
					for(int k = a + lane; k <= b; k += 32) idx[k] = idx2[k];
					__syncwarp();
					g = a + best_nl - 1;
				}
			}
			if(lane == 0)
			{
				t.left[self] = (g == a) ? (n - 1 + lo + a) : (lo + g);
				t.right[self] = (g + 1 == b) ? (n - 1 + lo + b) : (lo + g + 1);
				t.first[self] = lo + a; t.last[self] = lo + b;
				Y4Box rb; rb.lox = rl[0]; rb.loy = rl[1]; rb.loz = rl[2]; rb.hix = rh[0]; rb.hiy = rh[1]; rb.hiz = rh[2];
				t.box[self] = rb;
			}
			// smaller internal child first, the other one on the stack (depth <= log2 K); uniform across the warp
			bool li = g > a, ri = g + 1 < b;
			if(defer_list)
			{
				if(li && g - a + 1 >= 3 && g - a + 1 <= defer_limit)
				{
					if(lane == 0) { t.first[lo + g] = lo + a; t.last[lo + g] = lo + g; defer_list[atomicAdd(defer_count, 1u)] = lo + g; }
					li = false;
				}
				if(ri && b - g >= 3 && b - g <= defer_limit)
				{
					if(lane == 0) { t.first[lo + g + 1] = lo + g + 1; t.last[lo + g + 1] = lo + b; defer_list[atomicAdd(defer_count, 1u)] = lo + g + 1; }
					ri = false;
				}
			}
			if(li && ri)
			{
				if((g - a) <= (b - g - 1)) { if(lane == 0) { stack[sp] = g + 1; stack[sp + 1] = b; stack[sp + 2] = lo + g + 1; } sp += 3; b = g; self = lo + g; }
				else { if(lane == 0) { stack[sp] = a; stack[sp + 1] = g; stack[sp + 2] = lo + g; } sp += 3; a = g + 1; self = lo + g + 1; }
			}
			else if(li) { b = g; self = lo + g; }
			else if(ri) { a = g + 1; self = lo + g + 1; }
			else
			{
				if(sp == 0) break;
				__syncwarp();
				sp -= 3;
				a = stack[sp]; b = stack[sp + 1]; self = stack[sp + 2];
				__syncwarp();
			}
		}
		__syncwarp();
		for(int k = lane; k < M; k += 32)
		{
			const int j = idx[k];
			Y4Box o; o.lox = bx[j]; o.loy = bx[K + j]; o.loz = bx[2 * K + j]; o.hix = bx[3 * K + j]; o.hiy = bx[4 * K + j]; o.hiz = bx[5 * K + j];
			leaf[lo + k] = o;
			t.sorted_prim[lo + k] = prim[j];
		}
	}
}
static size_t treelet_smem_bytes(int K) { return (size_t) K * (6 * 4 + 4 + 2 + 2) + (3 * Y4_SAH_BINS * kBinWords + 48) * 4; }

// 5b. surface-area sum of the binary tree's internal nodes: the part of the SAH cost that depends on the key order
// (the leaves are the same single triangles either way)
__global__ void __launch_bounds__(256) k_area_sum(const Y4Box *__restrict__ box, int n_internal, double *sum)
{
	double a = 0.0;
	for(int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_internal; i += gridDim.x * blockDim.x)
	{
		const Y4Box b = box[i];
		const double dx = (double) b.hix - b.lox, dy = (double) b.hiy - b.loy, dz = (double) b.hiz - b.loz;
		a += dx * dy + dy * dz + dz * dx;
	}
	for(int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
	__shared__ double part[8];
	if((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = a;
	__syncthreads();
	if(threadIdx.x == 0)
	{
		double t = 0.0;
		for(int w = 0; w < 8; ++w) t += part[w];
		atomicAdd(sum, t);
	}
}

// 6. one collapse level
__global__ void __launch_bounds__(128) k_collapse(Y4BinTree bt, int leaf_max, const uint32_t *__restrict__ items, uint32_t count,
                                                  const float *__restrict__ verts, const uint32_t *__restrict__ tri_flags, Y4CollapseOut out)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= count) return;
	y4_collapse_item(bt, leaf_max, items[2 * i], items[2 * i + 1], verts, tri_flags, out);
}

// Scratch arena of a series of small builds (the bottom-level trees of a two-level scene): one allocation serves every
// temporary of every build instead of ~20 cudaMalloc / cudaFree pairs per mesh.  Set per host thread around the series.
struct Arena
{
	char *base = nullptr; size_t cap = 0, used = 0;
	void *take(size_t bytes)
	{
		const size_t at = (used + 255) & ~(size_t) 255;
		if(!base || at + bytes > cap) return nullptr;
		used = at + bytes;
		return base + at;
	}
};
thread_local Arena *t_arena = nullptr;

template<typename T> struct DevBuf
{
	T *p = nullptr;
	bool owned = true;
	~DevBuf() { if(p && owned) cudaFree(p); }
	cudaError_t alloc(size_t n)
	{
		const size_t bytes = (n ? n : 1) * sizeof(T);
		if(t_arena) { p = static_cast<T *>(t_arena->take(bytes)); if(p) { owned = false; return cudaSuccess; } }
		owned = true;
		return cudaMalloc(&p, bytes);
	}
};

} // namespace

// One BVH8 over `n` triangles (world space for a flat scene and for the proxies of a top-level tree, object space for a
// base mesh).  The caller owns out.nodes / out.tris.
struct Y4Built
{
	Y4U4 *nodes = nullptr; Y4Tri *tris = nullptr;
	uint64_t n_nodes = 0;
	bool placed = false;               // nodes / tris were given by the caller (room for n nodes and n leaf records): no allocation, no shrink
	int levels = 0, morton_used = 0;
	float raw[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
};

static int build_core(y4gpu_ctx *ctx, const float *d_verts, const uint32_t *d_flags, int n, int leaf_max, Y4Built &res)
{
	cudaStream_t st = ctx->stream;
	DevBuf<Y4Box> tb, box;
	DevBuf<uint32_t> scene_enc, vals_a, vals_b, counters, items_a, items_b;
	DevBuf<uint64_t> keys_a, keys_b;
	DevBuf<int> left, right, parent, first, last, flags;
	DevBuf<char> sort_tmp;
	Y4_CUDA(ctx, tb.alloc(n)); Y4_CUDA(ctx, box.alloc(2ull * n));
	Y4_CUDA(ctx, scene_enc.alloc(6));
	Y4_CUDA(ctx, keys_a.alloc(n)); Y4_CUDA(ctx, keys_b.alloc(n)); Y4_CUDA(ctx, vals_a.alloc(n)); Y4_CUDA(ctx, vals_b.alloc(n));
	Y4_CUDA(ctx, left.alloc(n)); Y4_CUDA(ctx, right.alloc(n)); Y4_CUDA(ctx, first.alloc(n)); Y4_CUDA(ctx, last.alloc(n));
	Y4_CUDA(ctx, parent.alloc(2ull * n)); Y4_CUDA(ctx, flags.alloc(n));
	Y4_CUDA(ctx, counters.alloc(4));

	const uint32_t enc_init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
	Y4_CUDA(ctx, cudaMemcpyAsync(scene_enc.p, enc_init, sizeof(enc_init), cudaMemcpyHostToDevice, st));
	const int tpb = 256;
	const int grid_n = (n + tpb - 1) / tpb;
	k_tri_boxes<<<min(grid_n, ctx->sm_count * 8), tpb, 0, st>>>(d_verts, n, tb.p, scene_enc.p);
	uint32_t enc_out[6];
	Y4_CUDA(ctx, cudaMemcpyAsync(enc_out, scene_enc.p, sizeof(enc_out), cudaMemcpyDeviceToHost, st));
	Y4_CUDA(ctx, cudaStreamSynchronize(st));
	Y4Box scene;
	scene.lox = dec_f(enc_out[0]); scene.loy = dec_f(enc_out[1]); scene.loz = dec_f(enc_out[2]);
	scene.hix = dec_f(enc_out[3]); scene.hiy = dec_f(enc_out[4]); scene.hiz = dec_f(enc_out[5]);
	res.raw[0] = scene.lox; res.raw[1] = scene.loy; res.raw[2] = scene.loz;
	res.raw[3] = scene.hix; res.raw[4] = scene.hiy; res.raw[5] = scene.hiz;

	// Key order: plain 63-bit Morton codes of the box centres, or extended codes that interleave one size bit per octree
	// level (large triangles leave their neighbours' subtree early).  Which one gives the cheaper tree depends on the
	// scene (sized keys: city -8 % node visits, Cornell box +4.5 % pass time), so by default both binary trees are built
	// -- key generation, sort, hierarchy and refit are the cheap part of the build -- and the one with the smaller
	// internal-node surface area (the order-dependent part of the SAH cost) goes on to the collapse.
	DevBuf<double> area_sum;
	Y4_CUDA(ctx, area_sum.alloc(1));
	cub::DoubleBuffer<uint64_t> dk(keys_a.p, keys_b.p);
	cub::DoubleBuffer<uint32_t> dv(vals_a.p, vals_b.p);
	size_t tmp_bytes = 0;
	Y4_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, n, 0, 63, st));
	Y4_CUDA(ctx, sort_tmp.alloc(tmp_bytes));
	const uint64_t *keys = nullptr;
	const uint32_t *sorted_prim = nullptr;
	const bool choose = ctx->morton_sized < 0 || ctx->morton_sized > 1;
	int sized = choose ? 0 : ctx->morton_sized, built_sized = -1;
	double area[2] = {0.0, 0.0};
	for(int round = 0; round < 3; ++round)
	{
		if(choose && round == 1) sized = 1;
		if(choose && round == 2) sized = area[1] < area[0] ? 1 : 0;
		if(sized == built_sized) break;                                  // the wanted tree is the one in the buffers
		dk = cub::DoubleBuffer<uint64_t>(keys_a.p, keys_b.p);
		dv = cub::DoubleBuffer<uint32_t>(vals_a.p, vals_b.p);
		k_morton<<<grid_n, tpb, 0, st>>>(tb.p, n, scene, keys_a.p, vals_a.p, sized);
		Y4_CUDA(ctx, cub::DeviceRadixSort::SortPairs(sort_tmp.p, tmp_bytes, dk, dv, n, 0, 63, st));
		keys = dk.Current();
		sorted_prim = dv.Current();
		Y4_CUDA(ctx, cudaMemsetAsync(flags.p, 0, sizeof(int) * (size_t) n, st));
		Y4_CUDA(ctx, cudaMemsetAsync(parent.p, 0xff, sizeof(int) * 2ull * n, st));
		if(n > 1) k_karras<<<(n - 1 + tpb - 1) / tpb, tpb, 0, st>>>(keys, n, left.p, right.p, parent.p, first.p, last.p);
		k_refit<<<grid_n, tpb, 0, st>>>(n, sorted_prim, tb.p, box.p, parent.p, left.p, right.p, flags.p);
		ctx->stats.kernel_launches += 4;
		if(ctx->treelet_leaves >= 3 && n > 2)
		{
			// `flags` (the refit's arrival counters) is free again: it holds the treelet root list; counters[3] their number
			Y4_CUDA(ctx, cudaMemsetAsync(counters.p + 3, 0, sizeof(uint32_t), st));
			k_treelet_roots<<<(n - 1 + tpb - 1) / tpb, tpb, 0, st>>>(n, ctx->treelet_leaves, parent.p, first.p, last.p, flags.p, counters.p + 3);
			Y4TreeletArgs ta;
			ta.n = n; ta.left = left.p; ta.right = right.p; ta.first = first.p; ta.last = last.p; ta.box = box.p; ta.sorted_prim = const_cast<uint32_t *>(sorted_prim);
			// persistent grid of one-warp CTAs; the kernel reads the number of roots from the device
			const int K = ctx->treelet_leaves < n ? ctx->treelet_leaves : n;
			const size_t smem = treelet_smem_bytes(K);
			Y4_CUDA(ctx, cudaFuncSetAttribute(k_treelet_sah, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
			const int per_sm = (int) ((size_t) 200 * 1024 / smem) < 16 ? (int) ((size_t) 200 * 1024 / smem) : 16;
			const int want = (n + 2) / 3;                    // upper bound of the number of roots (each has >= 3 leaves)
			const int grid = want < ctx->sm_count * (per_sm > 0 ? per_sm : 1) ? want : ctx->sm_count * (per_sm > 0 ? per_sm : 1);
			const int defer_limit = ctx->treelet_defer;
			const bool two_pass = defer_limit >= 3 && K > defer_limit;
			int *defer_list = flags.p + n / 2;                // second half of the buffer: at most n / 3 deferred ranges
			Y4_CUDA(ctx, cudaMemsetAsync(counters.p + 2, 0, sizeof(uint32_t), st));
			k_treelet_sah<<<grid, 32, smem, st>>>(ta, flags.p, counters.p + 3, K, defer_limit, two_pass ? defer_list : nullptr, counters.p + 2);
			if(two_pass)
			{
				const size_t smem2 = treelet_smem_bytes(defer_limit);
				const int grid2 = want < ctx->sm_count * 16 ? want : ctx->sm_count * 16;
				k_treelet_sah<<<grid2, 32, smem2, st>>>(ta, defer_list, counters.p + 2, defer_limit, 0, nullptr, nullptr);
				ctx->stats.kernel_launches++;
			}
			ctx->stats.kernel_launches += 2;
		}
		built_sized = sized;
		if(!choose) break;
		if(round < 2)
		{
			Y4_CUDA(ctx, cudaMemsetAsync(area_sum.p, 0, sizeof(double), st));
			if(n > 1) k_area_sum<<<min((n - 1 + tpb - 1) / tpb, ctx->sm_count * 8), tpb, 0, st>>>(box.p, n - 1, area_sum.p);
			Y4_CUDA(ctx, cudaMemcpyAsync(&area[round], area_sum.p, sizeof(double), cudaMemcpyDeviceToHost, st));
			Y4_CUDA(ctx, cudaStreamSynchronize(st));
			ctx->stats.kernel_launches++;
		}
	}
	res.morton_used = built_sized;
	if(getenv("Y4GPU_DEBUG")) fprintf(stderr, "[y4gpu] build: n=%d key order %s (internal area plain %.6g, sized %.6g)\n", n, built_sized ? "sized" : "plain", area[0], area[1]);

	// collapse into BVH8
	const uint64_t max_nodes = (uint64_t) n;   // every wide node below the root owns >= 1 binary internal node
	Y4U4 *nodes = res.nodes; Y4Tri *tris = res.tris;
	if(!res.placed)
	{
		nodes = nullptr; tris = nullptr;
		Y4_CUDA(ctx, cudaMalloc(&nodes, max_nodes * Y4_NODE_WORDS * sizeof(Y4U4)));
		if(cudaMalloc(&tris, (size_t) n * sizeof(Y4Tri)) != cudaSuccess) { cudaFree(nodes); ctx->err = "cudaMalloc tris failed"; return Y4GPU_ERR_CUDA; }
		res.nodes = nodes; res.tris = tris;
	}
	Y4_CUDA(ctx, items_a.alloc(2ull * n + 2)); Y4_CUDA(ctx, items_b.alloc(2ull * n + 2));
	const uint32_t init_counters[4] = {1u, 0u, 0u, 0u};   // node_counter, tri_counter, next_count, pad
	Y4_CUDA(ctx, cudaMemcpyAsync(counters.p, init_counters, sizeof(init_counters), cudaMemcpyHostToDevice, st));
	const uint32_t root_item[2] = {0u, 0u};
	Y4_CUDA(ctx, cudaMemcpyAsync(items_a.p, root_item, sizeof(root_item), cudaMemcpyHostToDevice, st));
	Y4BinTree bt;
	bt.n = n; bt.left = left.p; bt.right = right.p; bt.first = first.p; bt.last = last.p; bt.box = box.p; bt.sorted_prim = sorted_prim;
	uint32_t *cur = items_a.p, *nxt = items_b.p;
	uint32_t count = 1;
	int levels = 0;
	while(count)
	{
		Y4CollapseOut out;
		out.nodes = nodes; out.tris = tris; out.node_counter = counters.p; out.tri_counter = counters.p + 1;
		out.next_items = nxt; out.next_count = counters.p + 2;
		Y4_CUDA(ctx, cudaMemsetAsync(counters.p + 2, 0, sizeof(uint32_t), st));
		k_collapse<<<(count + 127) / 128, 128, 0, st>>>(bt, leaf_max, cur, count, d_verts, d_flags, out);
		Y4_CUDA(ctx, cudaMemcpyAsync(&count, counters.p + 2, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		Y4_CUDA(ctx, cudaStreamSynchronize(st));
		uint32_t *t = cur; cur = nxt; nxt = t;
		ctx->stats.kernel_launches++;
		if(++levels > 256) { ctx->err = "collapse did not terminate"; return Y4GPU_ERR_CUDA; }
	}
	res.levels = levels;
	uint32_t fin[2];
	Y4_CUDA(ctx, cudaMemcpyAsync(fin, counters.p, sizeof(fin), cudaMemcpyDeviceToHost, st));
	Y4_CUDA(ctx, cudaStreamSynchronize(st));
	Y4_CUDA(ctx, cudaGetLastError());
	if(fin[1] != (uint32_t) n) { ctx->err = "collapse lost triangles"; return Y4GPU_ERR_CUDA; }
	res.n_nodes = fin[0];
	// shrink the node array to its final size so only the live BVH stays resident
	if(!res.placed && res.n_nodes < max_nodes)
	{
		Y4U4 *small = nullptr;
		if(cudaMalloc(&small, res.n_nodes * Y4_NODE_WORDS * sizeof(Y4U4)) == cudaSuccess)
		{
			Y4_CUDA(ctx, cudaMemcpyAsync(small, nodes, res.n_nodes * Y4_NODE_WORDS * sizeof(Y4U4), cudaMemcpyDeviceToDevice, st));
			Y4_CUDA(ctx, cudaStreamSynchronize(st));
			cudaFree(nodes);
			res.nodes = small;
		}
		else cudaGetLastError();
	}
	ctx->stats.kernel_launches += 1;
	return Y4GPU_OK;
}

// reference tree bound inflation, accelerator_kdtree.cc:81-85 (double arithmetic, stored as float)
static void set_bounds(y4gpu_ctx *ctx, const float raw[6])
{
	for(int a = 0; a < 6; ++a) ctx->raw_bound[a] = raw[a];
	for(int a = 0; a < 3; ++a)
	{
		const double foo = (double) (ctx->raw_bound[3 + a] - ctx->raw_bound[a]) * 0.001;
		ctx->bound[a] = (float) ((double) ctx->raw_bound[a] - foo);
		ctx->bound[3 + a] = (float) ((double) ctx->raw_bound[3 + a] + foo);
	}
}

int y4_build_bvh(y4gpu_ctx *ctx, int leaf_max)
{
	const int n = (int) ctx->n_tri;
	cudaStream_t st = ctx->stream;
	y4_free_two_level(ctx);
	if(ctx->d_nodes) { cudaFree(ctx->d_nodes); ctx->d_nodes = nullptr; }
	if(ctx->d_tris) { cudaFree(ctx->d_tris); ctx->d_tris = nullptr; }
	ctx->n_nodes = 0;
	for(int k = 0; k < 6; ++k) ctx->raw_bound[k] = ctx->bound[k] = 0.f;
	if(n == 0) { ctx->built = true; return Y4GPU_OK; }
	if(leaf_max < 1 || leaf_max > 3) { ctx->err = "leaf_max must be 1..3"; return Y4GPU_ERR_ARG; }
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
	Y4Built b;
	const int rc = build_core(ctx, ctx->d_verts, ctx->d_flags, n, leaf_max, b);
	ctx->d_nodes = b.nodes; ctx->d_tris = b.tris;
	if(rc) return rc;
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
	Y4_CUDA(ctx, cudaStreamSynchronize(st));
	set_bounds(ctx, b.raw);
	ctx->morton_used = b.morton_used;
	// the traversal kernels keep at most one pending sibling group per level plus postponed triangle groups, which are only
	// pushed while sp < Y4_STACK_SIZE - levels (Y4Bvh8::postpone_sp_limit), on a Y4_STACK_SIZE-entry stack: refuse a tree
	// they could overflow instead of corrupting memory
	ctx->bvh_levels = b.levels;
	if(b.levels > Y4_STACK_SIZE - 8)
	{
		ctx->err = "BVH8 is deeper than the traversal stack (degenerate input: very many coincident triangles)";
		return Y4GPU_ERR_UNSUPPORTED;
	}
	ctx->n_nodes = b.n_nodes;
	float ms = 0.f;
	cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
	ctx->stats.build_ms = ms;
	ctx->stats.n_tris = n; ctx->stats.n_nodes = ctx->n_nodes; ctx->stats.bvh_levels = (uint64_t) b.levels;
	ctx->stats.node_bytes = ctx->n_nodes * Y4_NODE_WORDS * sizeof(Y4U4);
	ctx->stats.tri_bytes = (uint64_t) n * sizeof(Y4Tri);
	ctx->built = true;
	return Y4GPU_OK;
}

// ---- two-level build: one bottom-level BVH8 per base mesh in object space, one top-level BVH8 over the instances ----
namespace {

// leaf records of a bottom-level tree: replace {v0, eps | e1, prim | e2, flags} by the three object-space vertices (Y4TriObj)
__global__ void __launch_bounds__(256) k_blas_leaves(Y4Tri *__restrict__ tris, int n, const float *__restrict__ verts)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	const uint32_t prim = tris[i].prim, flags = tris[i].flags;
	Y4TriObj o;
	const float *v = verts + 9ull * prim;
#pragma unroll
	for(int k = 0; k < 9; ++k) o.v[k] = v[k];
	o.prim_local = prim; o.flags = flags; o.pad0 = 0u;
	*reinterpret_cast<Y4TriObj *>(tris + i) = o;
}

// exact world box of every instance (all vertices through the reference's Matrix4 * Point3) as a proxy triangle with that box:
// v0 = lo, v1 = hi, v2 = (lo.x, hi.y, lo.z).  One CTA per instance.
__global__ void __launch_bounds__(256) k_inst_proxy(const Y4Inst *__restrict__ inst, const float *__restrict__ mesh_verts,
                                                    const unsigned long long *__restrict__ mesh_first, float *__restrict__ proxy)
{
	const Y4Inst in = inst[blockIdx.x];
	const float *v = mesh_verts + 9ull * mesh_first[in.mesh];
	float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
	for(uint32_t k = threadIdx.x; k < 3u * in.n_tri; k += blockDim.x)
	{
		V3 p = v3(v[3 * k], v[3 * k + 1], v[3 * k + 2]);
		if(!in.identity) p = y4_inst_point(in.m, p);
		lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
		hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
	}
	__shared__ float s_lo[3][8], s_hi[3][8];
#pragma unroll
	for(int a = 0; a < 3; ++a)
	{
#pragma unroll
		for(int o = 16; o > 0; o >>= 1) { lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o)); hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o)); }
		if((threadIdx.x & 31) == 0) { s_lo[a][threadIdx.x >> 5] = lo[a]; s_hi[a][threadIdx.x >> 5] = hi[a]; }
	}
	__syncthreads();
	if(threadIdx.x == 0)
	{
		for(int a = 0; a < 3; ++a) for(int w = 1; w < 8; ++w) { s_lo[a][0] = fminf(s_lo[a][0], s_lo[a][w]); s_hi[a][0] = fmaxf(s_hi[a][0], s_hi[a][w]); }
		float *q = proxy + 9ull * blockIdx.x;
		q[0] = s_lo[0][0]; q[1] = s_lo[1][0]; q[2] = s_lo[2][0]; q[3] = s_hi[0][0]; q[4] = s_hi[1][0]; q[5] = s_hi[2][0];
		q[6] = s_lo[0][0]; q[7] = s_hi[1][0]; q[8] = s_lo[2][0];
	}
}

// inverse of the affine map [A | t] in double, rounded to float
void inverse_affine(const float *m, float *mi)
{
	const double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
	const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
	const double id = 1.0 / det;
	const double r[9] = {(e * i - f * h) * id, (c * h - b * i) * id, (b * f - c * e) * id,
	                     (f * g - d * i) * id, (a * i - c * g) * id, (c * d - a * f) * id,
	                     (d * h - e * g) * id, (b * g - a * h) * id, (a * e - b * d) * id};
	for(int k = 0; k < 3; ++k)
	{
		mi[4 * k] = (float) r[3 * k]; mi[4 * k + 1] = (float) r[3 * k + 1]; mi[4 * k + 2] = (float) r[3 * k + 2];
		mi[4 * k + 3] = (float) -(r[3 * k] * (double) m[3] + r[3 * k + 1] * (double) m[7] + r[3 * k + 2] * (double) m[11]);
	}
}

} // namespace

void y4_free_two_level(y4gpu_ctx *ctx)
{
	cudaFree(ctx->d_inst); cudaFree(ctx->d_tlas_nodes); cudaFree(ctx->d_tlas_leaves); cudaFree(ctx->d_blas_nodes); cudaFree(ctx->d_blas_tris);
	ctx->d_inst = nullptr; ctx->d_tlas_nodes = nullptr; ctx->d_tlas_leaves = nullptr; ctx->d_blas_nodes = nullptr; ctx->d_blas_tris = nullptr;
	ctx->n_tlas_nodes = ctx->n_blas_nodes = 0;
}

int y4_build_two_level(y4gpu_ctx *ctx, int leaf_max)
{
	cudaStream_t st = ctx->stream;
	y4_free_two_level(ctx);
	if(ctx->d_nodes) { cudaFree(ctx->d_nodes); ctx->d_nodes = nullptr; }
	if(ctx->d_tris) { cudaFree(ctx->d_tris); ctx->d_tris = nullptr; }
	ctx->n_nodes = 0;
	for(int k = 0; k < 6; ++k) ctx->raw_bound[k] = ctx->bound[k] = 0.f;
	const size_t n_mesh = ctx->mesh_first.empty() ? 0 : ctx->mesh_first.size() - 1, n_inst = ctx->h_inst.size();
	if(n_inst == 0 || ctx->n_tri == 0) { ctx->built = true; return Y4GPU_OK; }
	if(leaf_max < 1 || leaf_max > 3) { ctx->err = "leaf_max must be 1..3"; return Y4GPU_ERR_ARG; }
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
	// bottom level: every mesh on its own, built straight into the pooled arrays (node / leaf indices stay mesh-relative: the
	// instance carries the offsets).  One scratch arena sized for the largest mesh serves the temporaries of all builds.
	std::vector<uint32_t> node_off(n_mesh, 0u), tri_off(n_mesh, 0u);
	std::vector<float> bmax(n_mesh, 0.f);
	uint64_t nodes_total = 0, tris_total = 0, tris_all = ctx->mesh_first[n_mesh], n_max = 0;
	for(size_t i = 0; i < n_mesh; ++i) n_max = std::max<uint64_t>(n_max, ctx->mesh_first[i + 1] - ctx->mesh_first[i]);
	int blas_levels = 0, rc = Y4GPU_OK;
	Y4U4 *pool_nodes = nullptr;
	if(cudaMalloc(&pool_nodes, (tris_all ? tris_all : 1) * Y4_NODE_WORDS * sizeof(Y4U4)) != cudaSuccess || cudaMalloc(&ctx->d_blas_tris, (tris_all ? tris_all : 1) * sizeof(Y4TriObj)) != cudaSuccess)
	{
		cudaFree(pool_nodes); ctx->err = "cudaMalloc of the pooled bottom-level trees failed"; return Y4GPU_ERR_CUDA;
	}
	Arena arena;
	size_t sort_tmp_bytes = 0;
	{
		cub::DoubleBuffer<uint64_t> qk; cub::DoubleBuffer<uint32_t> qv;
		cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp_bytes, qk, qv, (int) n_max, 0, 63, st);
	}
	arena.cap = (size_t) n_max * 200 + sort_tmp_bytes + (64u << 10);
	if(cudaMalloc(&arena.base, arena.cap) != cudaSuccess) { arena.base = nullptr; arena.cap = 0; cudaGetLastError(); }      // (falls back to per-buffer allocations)
	t_arena = &arena;
	for(size_t i = 0; i < n_mesh && !rc; ++i)
	{
		const uint64_t first = ctx->mesh_first[i], n = ctx->mesh_first[i + 1] - first;
		if(n == 0) continue;
		arena.used = 0;
		Y4Built b;
		b.placed = true; b.nodes = pool_nodes + nodes_total * Y4_NODE_WORDS; b.tris = reinterpret_cast<Y4Tri *>(ctx->d_blas_tris + tris_total);
		rc = build_core(ctx, ctx->d_mesh_verts + 9ull * first, ctx->d_mesh_flags ? ctx->d_mesh_flags + first : nullptr, (int) n, leaf_max, b);
		if(rc) break;
		k_blas_leaves<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(b.tris, (int) n, ctx->d_mesh_verts + 9ull * first);
		node_off[i] = (uint32_t) nodes_total; tri_off[i] = (uint32_t) tris_total;
		nodes_total += b.n_nodes; tris_total += n;
		if(b.levels > blas_levels) blas_levels = b.levels;
		for(int k = 0; k < 6; ++k) bmax[i] = fmaxf(bmax[i], fabsf(b.raw[k]));
	}
	t_arena = nullptr;
	cudaStreamSynchronize(st);
	cudaFree(arena.base);
	if(rc) { cudaFree(pool_nodes); return rc; }
	// the node pool was sized for the worst case (one node per triangle): keep the live part only
	if(cudaMalloc(&ctx->d_blas_nodes, (nodes_total ? nodes_total : 1) * Y4_NODE_WORDS * sizeof(Y4U4)) != cudaSuccess) { cudaGetLastError(); ctx->d_blas_nodes = pool_nodes; }
	else
	{
		Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_blas_nodes, pool_nodes, nodes_total * Y4_NODE_WORDS * sizeof(Y4U4), cudaMemcpyDeviceToDevice, st));
		Y4_CUDA(ctx, cudaStreamSynchronize(st));
		cudaFree(pool_nodes);
	}
	ctx->n_blas_nodes = nodes_total; ctx->n_base_tris = tris_total;
	// instances: offsets of their mesh, inverse transforms; device copy; exact world boxes as proxy triangles; top-level tree
	for(Y4Inst &in : ctx->h_inst)
	{
		in.node_off = node_off[in.mesh]; in.tri_off = tri_off[in.mesh]; in.bmax = bmax[in.mesh];
		if(in.identity) { const float id[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0}; memcpy(in.mi, id, sizeof(id)); }
		else inverse_affine(in.m, in.mi);
	}
	Y4_CUDA(ctx, cudaMalloc(&ctx->d_inst, n_inst * sizeof(Y4Inst)));
	Y4_CUDA(ctx, cudaMemcpyAsync(ctx->d_inst, ctx->h_inst.data(), n_inst * sizeof(Y4Inst), cudaMemcpyHostToDevice, st));
	DevBuf<float> proxy; DevBuf<unsigned long long> d_first;
	Y4_CUDA(ctx, proxy.alloc(9 * n_inst)); Y4_CUDA(ctx, d_first.alloc(n_mesh + 1));
	Y4_CUDA(ctx, cudaMemcpyAsync(d_first.p, ctx->mesh_first.data(), (n_mesh + 1) * sizeof(unsigned long long), cudaMemcpyHostToDevice, st));
	k_inst_proxy<<<(unsigned) n_inst, 256, 0, st>>>(ctx->d_inst, ctx->d_mesh_verts, d_first.p, proxy.p);
	Y4Built top;
	rc = build_core(ctx, proxy.p, nullptr, (int) n_inst, leaf_max, top);
	ctx->d_tlas_nodes = top.nodes; ctx->d_tlas_leaves = top.tris;
	if(rc) return rc;
	Y4_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
	Y4_CUDA(ctx, cudaStreamSynchronize(st));
	Y4_CUDA(ctx, cudaGetLastError());
	ctx->n_tlas_nodes = top.n_nodes;
	set_bounds(ctx, top.raw);            // the union of the instances' exact world boxes = the flattened scene's bound
	ctx->bvh_levels = top.levels + blas_levels;
	// stack: one pending group per level of either tree, plus per instance entry a pending node group, a pending leaf group and the marker
	if(top.levels + blas_levels + 3 > Y4_STACK_SIZE - 8)
	{
		ctx->err = "two-level BVH8 is deeper than the traversal stack";
		return Y4GPU_ERR_UNSUPPORTED;
	}
	float ms = 0.f;
	cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
	ctx->stats.build_ms = ms;
	ctx->stats.n_tris = ctx->n_tri; ctx->stats.n_nodes = nodes_total + top.n_nodes; ctx->stats.bvh_levels = (uint64_t) ctx->bvh_levels;
	ctx->stats.node_bytes = (nodes_total + top.n_nodes) * Y4_NODE_WORDS * sizeof(Y4U4);
	ctx->stats.tri_bytes = tris_total * sizeof(Y4TriObj) + n_inst * (sizeof(Y4Tri) + sizeof(Y4Inst));
	ctx->stats.kernel_launches += n_mesh + 1;
	ctx->built = true;
	return Y4GPU_OK;
}
