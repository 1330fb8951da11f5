// Compressed 8-wide BVH ("BVH8"): node layout, collapse from the binary LBVH, encoding.
//
// Layout follows the published compressed-wide-BVH idea (Ylitie, Karras, Laine 2017):
// an 80-byte node = five 128-bit words holding a quantisation frame (origin + per-axis
// power-of-two scale), 8 child boxes as 8-bit fixed point, an internal-child mask and a
// per-slot meta byte.  Child slots are assigned so that `slot ^ octant` gives a
// front-to-back visiting order for any ray direction octant.
//
//   word0: p.x p.y p.z | ex | ey<<8 | ez<<16 | imask<<24
//   word1: child_base  tri_base  meta[0..3]  meta[4..7]
//   word2: qlo_x[0..3] qlo_x[4..7] qlo_y[0..3] qlo_y[4..7]
//   word3: qlo_z[0..3] qlo_z[4..7] qhi_x[0..3] qhi_x[4..7]
//   word4: qhi_y[0..3] qhi_y[4..7] qhi_z[0..3] qhi_z[4..7]
//
// meta byte: internal child -> 0b001_sssss with sssss = 24 + slot; leaf child ->
// unary triangle count in the top 3 bits (001 / 011 / 111) and the offset of its first
// triangle relative to tri_base in the low 5 bits (0..23); empty slot -> 0.
#pragma once
#include "y4_common.cuh"
#include "y4_tri.cuh"
#include "y4_bvh_build.cuh"

struct alignas(16) Y4U4 { uint32_t x, y, z, w; };   // one 128-bit node word (uint4 on device)
#define Y4_NODE_WORDS 5

#if defined(__CUDA_ARCH__)
Y4_HD uint32_t y4_atomic_add(uint32_t *p, uint32_t v) { return atomicAdd(p, v); }
#else
Y4_HD uint32_t y4_atomic_add(uint32_t *p, uint32_t v) { const uint32_t o = *p; *p = o + v; return o; }
#endif

// Read-only view of the binary tree used by the collapse pass.
struct Y4BinTree
{
	int n;                        // triangles == leaves
	const int *left, *right;      // [n-1] children in the unified index space
	const int *first, *last;      // [n-1] sorted-leaf range of each internal node
	const Y4Box *box;             // [2n-1]
	const uint32_t *sorted_prim;  // [n] original triangle index of sorted leaf k
};

Y4_HD int y4_bt_count(const Y4BinTree &bt, int idx) { return idx >= bt.n - 1 ? 1 : bt.last[idx] - bt.first[idx] + 1; }
Y4_HD int y4_bt_first(const Y4BinTree &bt, int idx) { return idx >= bt.n - 1 ? idx - (bt.n - 1) : bt.first[idx]; }

// Conservative padding applied to every child box before quantisation so that slab-test
// rounding can never cull a triangle whose exact Moller-Trumbore test accepts the ray.
Y4_HD Y4Box y4_box_pad(const Y4Box &b)
{
	const float k = 1.f / 2097152.f;   // 2^-21 of the coordinate magnitude
	Y4Box r;
	const float px = fmaxf(fabsf(b.lox), fabsf(b.hix)) * k + 1e-30f;
	const float py = fmaxf(fabsf(b.loy), fabsf(b.hiy)) * k + 1e-30f;
	const float pz = fmaxf(fabsf(b.loz), fabsf(b.hiz)) * k + 1e-30f;
	r.lox = b.lox - px; r.loy = b.loy - py; r.loz = b.loz - pz;
	r.hix = b.hix + px; r.hiy = b.hiy + py; r.hiz = b.hiz + pz;
	return r;
}

// Smallest biased exponent e (scale = 2^(e-127)) with lo + 255*scale >= hi.
Y4_HD uint32_t y4_quant_exponent(float lo, float hi)
{
	float ext = hi - lo;
	if(!(ext > 1e-30f)) ext = 1e-30f;
	int k;
	frexpf(ext / 255.f, &k);          // ext/255 = m * 2^k, m in [0.5, 1)  =>  2^k >= ext/255
	int e = k + 127;
	if(e < 1) e = 1;
	if(e > 254) e = 254;
	while(e < 254 && lo + 255.f * y4_u2f((uint32_t) e << 23) < hi) ++e;
	return (uint32_t) e;
}
Y4_HD uint32_t y4_quant_lo(float v, float p, float scale)
{
	float q = floorf((v - p) / scale);
	if(q < 0.f) q = 0.f;
	if(q > 255.f) q = 255.f;
	while(q > 0.f && p + q * scale > v) q -= 1.f;
	return (uint32_t) q;
}
Y4_HD uint32_t y4_quant_hi(float v, float p, float scale)
{
	float q = ceilf((v - p) / scale);
	if(q < 0.f) q = 0.f;
	if(q > 255.f) q = 255.f;
	while(q < 255.f && p + q * scale < v) q += 1.f;
	return (uint32_t) q;
}

// ---- collapse: one work item = (binary internal node, wide node index) --------------------
// Greedy surface-area opening until 8 slots, octant-order slot assignment, allocation of
// the contiguous child-node and triangle ranges with two atomics, node encoding and leaf
// triangle records.  Children that stay internal are appended to the next level's queue.
struct Y4CollapseOut
{
	Y4U4 *nodes;             // [max_nodes * 5]
	Y4Tri *tris;             // [n] leaf-ordered triangle records
	uint32_t *node_counter;  // next free wide node (starts at 1: root is node 0)
	uint32_t *tri_counter;   // next free triangle slot
	uint32_t *next_items;    // [2 * capacity] pairs (binary node, wide node) for the next level
	uint32_t *next_count;
};

Y4_HD void y4_collapse_item(const Y4BinTree &bt, int leaf_max, uint32_t bnode, uint32_t wnode,
                            const float *verts, const uint32_t *tri_flags, const Y4CollapseOut &out)
{
	int ent[8];
	int n_ent;
	if(bt.n == 1) { ent[0] = 0; n_ent = 1; }            // single-triangle scene: leaf index 0 == n-1+0
	else { ent[0] = bt.left[bnode]; ent[1] = bt.right[bnode]; n_ent = 2; }
	// open the largest-area openable entry until 8 slots are used
	while(n_ent < 8)
	{
		int best = -1; float best_area = -1.f;
		for(int k = 0; k < n_ent; ++k)
		{
			if(y4_bt_count(bt, ent[k]) <= leaf_max) continue;   // becomes a leaf slot
			const float a = y4_box_half_area(bt.box[ent[k]]);
			if(a > best_area) { best_area = a; best = k; }
		}
		if(best < 0) break;
		const int b = ent[best];
		ent[best] = bt.left[b];
		ent[n_ent++] = bt.right[b];
	}
	// padded child boxes and the node frame
	Y4Box cb[8];
	Y4Box nb = y4_box_empty();
	for(int k = 0; k < n_ent; ++k) { cb[k] = y4_box_pad(bt.box[ent[k]]); nb = y4_box_union(nb, cb[k]); }
	const float ncx = 0.5f * (nb.lox + nb.hix), ncy = 0.5f * (nb.loy + nb.hiy), ncz = 0.5f * (nb.loz + nb.hiz);
	// slot assignment: slot s "prefers" the child lying farthest along D_s, where bit2/1/0 of s
	// select +x/+y/+z (set) or -x/-y/-z (clear).  Greedy global-max auction over (child, slot).
	float cost[8][8];
	for(int k = 0; k < n_ent; ++k)
	{
		const float cx = 0.5f * (cb[k].lox + cb[k].hix) - ncx, cy = 0.5f * (cb[k].loy + cb[k].hiy) - ncy, cz = 0.5f * (cb[k].loz + cb[k].hiz) - ncz;
		for(int s = 0; s < 8; ++s)
			cost[k][s] = ((s & 4) ? cx : -cx) + ((s & 2) ? cy : -cy) + ((s & 1) ? cz : -cz);
	}
	int slot_of[8], child_in_slot[8];
	for(int s = 0; s < 8; ++s) child_in_slot[s] = -1;
	for(int k = 0; k < 8; ++k) slot_of[k] = -1;
	for(int round = 0; round < n_ent; ++round)
	{
		int bk = -1, bs = -1; float bc = -FLT_MAX;
		for(int k = 0; k < n_ent; ++k)
		{
			if(slot_of[k] >= 0) continue;
			for(int s = 0; s < 8; ++s)
			{
				if(child_in_slot[s] >= 0) continue;
				if(bk < 0 || cost[k][s] > bc) { bc = cost[k][s]; bk = k; bs = s; }
			}
		}
		slot_of[bk] = bs; child_in_slot[bs] = bk;
	}
	// count internal children / triangles, allocate contiguous ranges
	uint32_t n_inner = 0, n_tris = 0;
	for(int s = 0; s < 8; ++s)
	{
		const int k = child_in_slot[s];
		if(k < 0) continue;
		const int c = y4_bt_count(bt, ent[k]);
		if(c <= leaf_max) n_tris += (uint32_t) c; else ++n_inner;
	}
	const uint32_t child_base = n_inner ? y4_atomic_add(out.node_counter, n_inner) : 0u;
	const uint32_t tri_base = n_tris ? y4_atomic_add(out.tri_counter, n_tris) : 0u;
	const uint32_t q_base = n_inner ? y4_atomic_add(out.next_count, n_inner) : 0u;
	// quantisation frame
	const uint32_t ex = y4_quant_exponent(nb.lox, nb.hix), ey = y4_quant_exponent(nb.loy, nb.hiy), ez = y4_quant_exponent(nb.loz, nb.hiz);
	const float sx = y4_u2f(ex << 23), sy = y4_u2f(ey << 23), sz = y4_u2f(ez << 23);
	uint32_t imask = 0, meta[8], q[6][8];
	uint32_t inner_i = 0, tri_off = 0;
	for(int s = 0; s < 8; ++s)
	{
		const int k = child_in_slot[s];
		if(k < 0)
		{
			meta[s] = 0;
			q[0][s] = q[1][s] = q[2][s] = 255u; q[3][s] = q[4][s] = q[5][s] = 0u;   // inverted box: never hit
			continue;
		}
		q[0][s] = y4_quant_lo(cb[k].lox, nb.lox, sx); q[1][s] = y4_quant_lo(cb[k].loy, nb.loy, sy); q[2][s] = y4_quant_lo(cb[k].loz, nb.loz, sz);
		q[3][s] = y4_quant_hi(cb[k].hix, nb.lox, sx); q[4][s] = y4_quant_hi(cb[k].hiy, nb.loy, sy); q[5][s] = y4_quant_hi(cb[k].hiz, nb.loz, sz);
		const int c = y4_bt_count(bt, ent[k]);
		if(c <= leaf_max)
		{
			meta[s] = (((1u << c) - 1u) << 5) | tri_off;
			const int first = y4_bt_first(bt, ent[k]);
			for(int t = 0; t < c; ++t)
			{
				const uint32_t prim = bt.sorted_prim[first + t];
				out.tris[tri_base + tri_off + (uint32_t) t] = y4_tri_setup(verts + 9ull * prim, prim, tri_flags ? tri_flags[prim] : (uint32_t) Y4_TRI_DEFAULT_FLAGS);
			}
			tri_off += (uint32_t) c;
		}
		else
		{
			imask |= 1u << s;
			meta[s] = (1u << 5) | (24u + (uint32_t) s);
			out.next_items[2 * (q_base + inner_i)] = (uint32_t) ent[k];
			out.next_items[2 * (q_base + inner_i) + 1] = child_base + inner_i;
			++inner_i;
		}
	}
	Y4U4 w0, w1, w2, w3, w4;
	w0.x = y4_f2u(nb.lox); w0.y = y4_f2u(nb.loy); w0.z = y4_f2u(nb.loz); w0.w = ex | (ey << 8) | (ez << 16) | (imask << 24);
#define Y4_PACK4(a, o) ((a)[o] | ((a)[o + 1] << 8) | ((a)[o + 2] << 16) | ((a)[o + 3] << 24))
	w1.x = child_base; w1.y = tri_base; w1.z = Y4_PACK4(meta, 0); w1.w = Y4_PACK4(meta, 4);
	w2.x = Y4_PACK4(q[0], 0); w2.y = Y4_PACK4(q[0], 4); w2.z = Y4_PACK4(q[1], 0); w2.w = Y4_PACK4(q[1], 4);
	w3.x = Y4_PACK4(q[2], 0); w3.y = Y4_PACK4(q[2], 4); w3.z = Y4_PACK4(q[3], 0); w3.w = Y4_PACK4(q[3], 4);
	w4.x = Y4_PACK4(q[4], 0); w4.y = Y4_PACK4(q[4], 4); w4.z = Y4_PACK4(q[5], 0); w4.w = Y4_PACK4(q[5], 4);
#undef Y4_PACK4
	Y4U4 *dst = out.nodes + (size_t) wnode * Y4_NODE_WORDS;
	dst[0] = w0; dst[1] = w1; dst[2] = w2; dst[3] = w3; dst[4] = w4;
}
