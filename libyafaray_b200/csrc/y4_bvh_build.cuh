// Binary LBVH construction, per-element logic (Y4_HD so the unit-test emulation can run it).
//
// Replaces the reference's serial SAH kd-tree build
// (/root/reference/src/accelerator/accelerator_kdtree.cc:47-123,420-628) with a
// data-parallel pipeline: triangle AABBs -> 63-bit Morton keys -> radix sort ->
// Karras 2012 hierarchy (one thread per internal node) -> bottom-up refit.
// The binary tree is only an intermediate: y4_bvh8.cuh collapses it to the
// compressed 8-wide BVH the traversal kernels walk.
//
// Node index space: [0, n-1) internal nodes, [n-1, 2n-1) leaves (leaf k == sorted
// triangle k).  Boxes are float4 pairs (w unused) so kernels move them as 128-bit words.
#pragma once
#include "y4_common.cuh"

struct Y4Box { float lox, loy, loz, hix, hiy, hiz; };

Y4_HD Y4Box y4_box_empty()
{
	Y4Box b; b.lox = b.loy = b.loz = FLT_MAX; b.hix = b.hiy = b.hiz = -FLT_MAX; return b;
}
Y4_HD Y4Box y4_box_union(const Y4Box &a, const Y4Box &b)
{
	Y4Box r;
	r.lox = fminf(a.lox, b.lox); r.loy = fminf(a.loy, b.loy); r.loz = fminf(a.loz, b.loz);
	r.hix = fmaxf(a.hix, b.hix); r.hiy = fmaxf(a.hiy, b.hiy); r.hiz = fmaxf(a.hiz, b.hiz);
	return r;
}
Y4_HD float y4_box_half_area(const Y4Box &b)
{
	const float dx = b.hix - b.lox, dy = b.hiy - b.loy, dz = b.hiz - b.loz;
	return dx * dy + dy * dz + dz * dx;
}
// Exact min/max of the three vertices (reference FacePrimitive::getBound, primitive_face.cc:113-128).
Y4_HD Y4Box y4_tri_box(const float *v)
{
	Y4Box b;
	b.lox = fminf(v[0], fminf(v[3], v[6])); b.hix = fmaxf(v[0], fmaxf(v[3], v[6]));
	b.loy = fminf(v[1], fminf(v[4], v[7])); b.hiy = fmaxf(v[1], fmaxf(v[4], v[7]));
	b.loz = fminf(v[2], fminf(v[5], v[8])); b.hiz = fmaxf(v[2], fmaxf(v[5], v[8]));
	return b;
}

// Spread the low 21 bits of x so that there are two zero bits between each.
Y4_HD uint64_t y4_expand21(uint64_t x)
{
	x &= 0x1fffffull;
	x = (x | x << 32) & 0x1f00000000ffffull;
	x = (x | x << 16) & 0x1f0000ff0000ffull;
	x = (x | x << 8) & 0x100f00f00f00f00full;
	x = (x | x << 4) & 0x10c30c30c30c30c3ull;
	x = (x | x << 2) & 0x1249249249249249ull;
	return x;
}
// 63-bit Morton key of a centroid normalised to the scene box.
Y4_HD uint64_t y4_morton63(const Y4Box &tri, const Y4Box &scene)
{
	const float ex = scene.hix - scene.lox, ey = scene.hiy - scene.loy, ez = scene.hiz - scene.loz;
	const float cx = 0.5f * (tri.lox + tri.hix), cy = 0.5f * (tri.loy + tri.hiy), cz = 0.5f * (tri.loz + tri.hiz);
	float nx = ex > 0.f ? (cx - scene.lox) / ex : 0.f;
	float ny = ey > 0.f ? (cy - scene.loy) / ey : 0.f;
	float nz = ez > 0.f ? (cz - scene.loz) / ez : 0.f;
	nx = fminf(fmaxf(nx, 0.f), 1.f); ny = fminf(fmaxf(ny, 0.f), 1.f); nz = fminf(fmaxf(nz, 0.f), 1.f);
	const uint64_t qx = (uint64_t) fminf(nx * 2097152.f, 2097151.f);
	const uint64_t qy = (uint64_t) fminf(ny * 2097152.f, 2097151.f);
	const uint64_t qz = (uint64_t) fminf(nz * 2097152.f, 2097151.f);
	return (y4_expand21(qx) << 2) | (y4_expand21(qy) << 1) | y4_expand21(qz);
}

// Extended Morton key (Vinkler et al. 2017): one bit of the primitive's SIZE after every xyz triple, so that a triangle
// much larger than its neighbours leaves their subtree at the octree level whose cells are about its own size instead
// of inflating every box down to a deep leaf.  15 levels of [x y z s] + one more xyz triple = 63 bits.  Scenes of
// uniformly small triangles have all-zero size bits and keep plain Morton order (at 16 instead of 21 bits per axis).
Y4_HD uint64_t y4_morton63_sized(const Y4Box &tri, const Y4Box &scene)
{
	const float ex = scene.hix - scene.lox, ey = scene.hiy - scene.loy, ez = scene.hiz - scene.loz;
	const float cx = 0.5f * (tri.lox + tri.hix), cy = 0.5f * (tri.loy + tri.hiy), cz = 0.5f * (tri.loz + tri.hiz);
	float nx = ex > 0.f ? (cx - scene.lox) / ex : 0.f;
	float ny = ey > 0.f ? (cy - scene.loy) / ey : 0.f;
	float nz = ez > 0.f ? (cz - scene.loz) / ez : 0.f;
	nx = fminf(fmaxf(nx, 0.f), 1.f); ny = fminf(fmaxf(ny, 0.f), 1.f); nz = fminf(fmaxf(nz, 0.f), 1.f);
	const uint32_t qx = (uint32_t) fminf(nx * 65536.f, 65535.f), qy = (uint32_t) fminf(ny * 65536.f, 65535.f), qz = (uint32_t) fminf(nz * 65536.f, 65535.f);
	// size relative to the cells of each level, per axis (the normalisation above is per axis too): the largest relative extent
	float sx = ex > 0.f ? (tri.hix - tri.lox) / ex : 0.f, sy = ey > 0.f ? (tri.hiy - tri.loy) / ey : 0.f, sz = ez > 0.f ? (tri.hiz - tri.loz) / ez : 0.f;
	const float sn = fminf(fmaxf(fmaxf(sx, sy), sz), 1.f);
	const uint32_t qs = (uint32_t) fminf(sn * 32768.f, 32767.f);
	uint64_t key = 0;
	for(int l = 0; l < 15; ++l)
	{
		key = (key << 4) | (uint64_t) ((((qx >> (15 - l)) & 1u) << 3) | (((qy >> (15 - l)) & 1u) << 2) | (((qz >> (15 - l)) & 1u) << 1) | ((qs >> (14 - l)) & 1u));
	}
	key = (key << 3) | (uint64_t) (((qx & 1u) << 2) | ((qy & 1u) << 1) | (qz & 1u));
	return key;
}

Y4_HD int y4_clz64(uint64_t x)
{
#if defined(__CUDA_ARCH__)
	return __clzll((long long) x);
#else
	return x ? __builtin_clzll(x) : 64;
#endif
}
Y4_HD int y4_clz32(uint32_t x)
{
#if defined(__CUDA_ARCH__)
	return __clz((int) x);
#else
	return x ? __builtin_clz(x) : 32;
#endif
}

// Karras' delta: common-prefix length of keys i and j, ties broken by index; -1 outside [0,n).
Y4_HD int y4_delta(const uint64_t *keys, int n, int i, int j)
{
	if(j < 0 || j >= n) return -1;
	const uint64_t x = keys[i] ^ keys[j];
	if(x == 0) return 64 + y4_clz32((uint32_t) i ^ (uint32_t) j);
	return y4_clz64(x);
}

// One internal node of the Karras hierarchy.  Writes its children, their parent links and
// the sorted-leaf range [first, last] it covers.
Y4_HD void y4_karras_node(const uint64_t *keys, int n, int i, int *left, int *right, int *parent, int *range_first, int *range_last)
{
	const int d = (y4_delta(keys, n, i, i + 1) - y4_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
	const int dmin = y4_delta(keys, n, i, i - d);
	int lmax = 2;
	while(y4_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
	int l = 0;
	for(int t = lmax >> 1; t >= 1; t >>= 1)
		if(y4_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
	const int j = i + l * d;
	const int dnode = y4_delta(keys, n, i, j);
	int s = 0;
	for(int t = (l + 1) >> 1; ; t = (t + 1) >> 1)
	{
		if(y4_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
		if(t <= 1) break;
	}
	const int gamma = i + s * d + (d < 0 ? -1 : 0);
	const int lo = i < j ? i : j, hi = i < j ? j : i;
	const int lc = (lo == gamma) ? (n - 1 + gamma) : gamma;
	const int rc = (hi == gamma + 1) ? (n - 1 + gamma + 1) : (gamma + 1);
	left[i] = lc; right[i] = rc;
	parent[lc] = i; parent[rc] = i;
	range_first[i] = lo; range_last[i] = hi;
	if(i == 0) parent[0] = -1;
}

// ---- SAH refinement of the LBVH: treelet rebuild ----------------------------------------------------------
// The Karras tree splits every range at a Morton-prefix boundary (a spatial median), whatever the triangle
// distribution inside.  Every maximal subtree with at most `max_leaves` leaves (a "treelet") is rebuilt top-down with
// the binned surface-area heuristic (16 bins per axis, cost = A_left * N_left + A_right * N_right): the leaves of its
// range are re-ordered in place and its internal nodes re-linked.  Any binary tree over a contiguous range of sorted
// leaves fits the Karras index space -- the left child of a split after leaf g is internal node g (or leaf g), the
// right child internal node g + 1 (or leaf g + 1) -- so the tree above the treelet root, the root's own index and its
// box stay as they are and the collapse to BVH8 runs unchanged.  (The kd-tree the reference builds is a full SAH tree,
// accelerator_kdtree.cc:136-410; this recovers the SAH decisions where most nodes are, in the lower levels.)
// One thread per treelet, sequential code (also run by the unit-test emulation); treelets are disjoint, so the
// result does not depend on scheduling.  `parent` links are not maintained: nothing after the refit reads them.
#define Y4_SAH_BINS 16
struct Y4TreeletArgs
{
	int n;
	int *left, *right, *first, *last;   // [n-1]
	Y4Box *box;                         // [2n-1]
	uint32_t *sorted_prim;              // [n]
};
Y4_HD bool y4_is_treelet_root(int n, int max_leaves, const int *parent, const int *first, const int *last, int i)
{
	const int m = last[i] - first[i] + 1;
	if(m < 3 || m > max_leaves) return false;
	if(i == 0) return true;
	const int p = parent[i];
	return last[p] - first[p] + 1 > max_leaves;
}
Y4_HD float y4_box_centre(const Y4Box &b, int axis)
{
	return axis == 0 ? 0.5f * (b.lox + b.hix) : (axis == 1 ? 0.5f * (b.loy + b.hiy) : 0.5f * (b.loz + b.hiz));
}
Y4_HD int y4_sah_bin(float c, float lo, float scale)
{
	const int b = (int) ((c - lo) * scale);
	return b < 0 ? 0 : (b > Y4_SAH_BINS - 1 ? Y4_SAH_BINS - 1 : b);
}
Y4_HD void y4_treelet_sah(const Y4TreeletArgs &t, int root)
{
	const int n = t.n;
	Y4Box *leaf = t.box + (n - 1);                   // leaf k of the sorted order
	int stack[3 * 24];
	int sp = 0;
	int a = t.first[root], b = t.last[root], self = root;
	for(;;)
	{
		const int m = b - a + 1;
		// range box and centroid bounds
		Y4Box rb = y4_box_empty();
		float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
		for(int k = a; k <= b; ++k)
		{
			const Y4Box lb = leaf[k];
			rb = y4_box_union(rb, lb);
			for(int ax = 0; ax < 3; ++ax) { const float c = y4_box_centre(lb, ax); clo[ax] = fminf(clo[ax], c); chi[ax] = fmaxf(chi[ax], c); }
		}
		int g = a + (m >> 1) - 1;                    // fall-back: middle of the current order
		if(m > 2)
		{
			float scale[3];
			for(int ax = 0; ax < 3; ++ax) { const float e = chi[ax] - clo[ax]; scale[ax] = e > 0.f ? (float) Y4_SAH_BINS / e : 0.f; }
			Y4Box bins[3][Y4_SAH_BINS];
			int cnt[3][Y4_SAH_BINS];
			for(int ax = 0; ax < 3; ++ax)
				for(int q = 0; q < Y4_SAH_BINS; ++q) { bins[ax][q] = y4_box_empty(); cnt[ax][q] = 0; }
			for(int k = a; k <= b; ++k)
			{
				const Y4Box lb = leaf[k];
				for(int ax = 0; ax < 3; ++ax)
				{
					if(scale[ax] == 0.f) continue;
					const int q = y4_sah_bin(y4_box_centre(lb, ax), clo[ax], scale[ax]);
					bins[ax][q] = y4_box_union(bins[ax][q], lb);
					++cnt[ax][q];
				}
			}
			float best = FLT_MAX; int best_ax = -1, best_q = -1, best_nl = 0;
			for(int ax = 0; ax < 3; ++ax)
			{
				if(scale[ax] == 0.f) continue;
				float right_cost[Y4_SAH_BINS];       // A_right * N_right of the split after bin q
				Y4Box acc = y4_box_empty(); int c = 0;
				for(int q = Y4_SAH_BINS - 1; q >= 1; --q)
				{
					if(cnt[ax][q]) { acc = y4_box_union(acc, bins[ax][q]); c += cnt[ax][q]; }
					right_cost[q - 1] = c ? y4_box_half_area(acc) * (float) c : FLT_MAX;
				}
				acc = y4_box_empty(); c = 0;
				for(int q = 0; q < Y4_SAH_BINS - 1; ++q)
				{
					if(cnt[ax][q]) { acc = y4_box_union(acc, bins[ax][q]); c += cnt[ax][q]; }
					if(c == 0 || right_cost[q] == FLT_MAX) continue;
					const float cost = y4_box_half_area(acc) * (float) c + right_cost[q];
					if(cost < best) { best = cost; best_ax = ax; best_q = q; best_nl = c; }
				}
			}
			if(best_ax >= 0)
			{
				// in-place partition: bins <= best_q to the front
				int i = a, j = b;
				while(i <= j)
				{
					if(y4_sah_bin(y4_box_centre(leaf[i], best_ax), clo[best_ax], scale[best_ax]) <= best_q) { ++i; continue; }
					if(y4_sah_bin(y4_box_centre(leaf[j], best_ax), clo[best_ax], scale[best_ax]) > best_q) { --j; continue; }
					const Y4Box tb = leaf[i]; leaf[i] = leaf[j]; leaf[j] = tb;
					const uint32_t tp = t.sorted_prim[i]; t.sorted_prim[i] = t.sorted_prim[j]; t.sorted_prim[j] = tp;
					++i; --j;
				}
				g = a + best_nl - 1;
			}
		}
		const int lc = (g == a) ? (n - 1 + a) : g;
		const int rc = (g + 1 == b) ? (n - 1 + b) : (g + 1);
		t.left[self] = lc; t.right[self] = rc; t.first[self] = a; t.last[self] = b; t.box[self] = rb;
		// continue with the smaller internal child, keep the other one for later
		const bool li = g > a, ri = g + 1 < b;
		if(li && ri)
		{
			const bool left_small = (g - a) <= (b - g - 1);
			if(left_small) { stack[sp++] = g + 1; stack[sp++] = b; stack[sp++] = g + 1; b = g; self = g; }
			else { stack[sp++] = a; stack[sp++] = g; stack[sp++] = g; a = g + 1; self = g + 1; }
		}
		else if(li) { b = g; self = g; }
		else if(ri) { a = g + 1; self = g + 1; }
		else
		{
			if(sp == 0) break;
			self = stack[--sp]; b = stack[--sp]; a = stack[--sp];
		}
	}
}
