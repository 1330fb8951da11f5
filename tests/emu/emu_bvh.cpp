// UNIT-TEST EMULATION ONLY.  Compiles the Y4_HD per-element logic of
// libyafaray_b200/csrc/*.cuh with g++ and runs the "kernels" as serial loops, so the BVH
// builder / BVH8 encoder / traversal logic can be debugged on a machine without a GPU.
// Never linked into liby4gpu.so; the product path has no CPU fallback.
#include "../../libyafaray_b200/csrc/y4_trace.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

struct EmuBvh
{
	std::vector<Y4U4> nodes;
	std::vector<Y4Tri> tris;
	uint32_t n_nodes = 0;
	int n = 0;
};

void emu_build(const float *verts, const uint32_t *flags, int n, int leaf_max, EmuBvh &out)
{
	out.n = n;
	if(n == 0) return;
	std::vector<Y4Box> tb(n);
	Y4Box scene = y4_box_empty();
	for(int i = 0; i < n; ++i) { tb[i] = y4_tri_box(verts + 9ull * i); scene = y4_box_union(scene, tb[i]); }
	std::vector<std::pair<uint64_t, uint32_t>> kv(n);
	for(int i = 0; i < n; ++i) kv[i] = {y4_morton63_sized(tb[i], scene), (uint32_t) i};
	std::stable_sort(kv.begin(), kv.end(), [](const std::pair<uint64_t, uint32_t> &a, const std::pair<uint64_t, uint32_t> &b) { return a.first < b.first; });
	std::vector<uint64_t> keys(n);
	std::vector<uint32_t> sorted_prim(n);
	for(int i = 0; i < n; ++i) { keys[i] = kv[i].first; sorted_prim[i] = kv[i].second; }
	const int n_int = n > 1 ? n - 1 : 0;
	std::vector<int> left(n_int + 1), right(n_int + 1), first(n_int + 1), last(n_int + 1), parent(2 * n), flag(n_int + 1, 0);
	std::vector<Y4Box> box(2 * n);
	for(int i = 0; i < n_int; ++i) y4_karras_node(keys.data(), n, i, left.data(), right.data(), parent.data(), first.data(), last.data());
	for(int k = 0; k < n; ++k) box[n - 1 + k] = tb[sorted_prim[k]];
	// bottom-up refit, second arrival proceeds
	for(int k = 0; k < n && n > 1; ++k)
	{
		int cur = parent[n - 1 + k];
		while(cur >= 0)
		{
			if(flag[cur]++ == 0) break;
			box[cur] = y4_box_union(box[left[cur]], box[right[cur]]);
			cur = parent[cur];
		}
	}
	// SAH treelet refinement (k_treelet_roots + k_treelet_sah of y4_build.cu); EMU_TREELET overrides the default size
	{
		const char *e = getenv("EMU_TREELET");
		const int max_leaves = e ? atoi(e) : 512;
		if(max_leaves >= 3 && n > 2)
		{
			std::vector<int> roots;
			for(int i = 0; i < n - 1; ++i) if(y4_is_treelet_root(n, max_leaves, parent.data(), first.data(), last.data(), i)) roots.push_back(i);
			Y4TreeletArgs ta;
			ta.n = n; ta.left = left.data(); ta.right = right.data(); ta.first = first.data(); ta.last = last.data(); ta.box = box.data(); ta.sorted_prim = sorted_prim.data();
			for(int r : roots) y4_treelet_sah(ta, r);
		}
	}
	Y4BinTree bt;
	bt.n = n; bt.left = left.data(); bt.right = right.data(); bt.first = first.data(); bt.last = last.data();
	bt.box = box.data(); bt.sorted_prim = sorted_prim.data();
	out.nodes.assign((size_t) std::max(1, n) * Y4_NODE_WORDS, Y4U4{0, 0, 0, 0});
	out.tris.resize(n);
	uint32_t node_counter = 1, tri_counter = 0;
	std::vector<uint32_t> cur_items{0u, 0u}, next_items(2 * (size_t) n + 2);
	uint32_t cur_count = 1;
	while(cur_count)
	{
		uint32_t next_count = 0;
		Y4CollapseOut co;
		co.nodes = out.nodes.data(); co.tris = out.tris.data(); co.node_counter = &node_counter; co.tri_counter = &tri_counter;
		co.next_items = next_items.data(); co.next_count = &next_count;
		for(uint32_t i = 0; i < cur_count; ++i) y4_collapse_item(bt, leaf_max, cur_items[2 * i], cur_items[2 * i + 1], verts, flags, co);
		cur_items.assign(next_items.begin(), next_items.begin() + 2 * (size_t) next_count);
		cur_count = next_count;
	}
	out.n_nodes = node_counter;
	if(tri_counter != (uint32_t) n) fprintf(stderr, "emu: tri_counter %u != n %d\n", tri_counter, n);
}

extern "C" {

// Build + trace; returns number of wide nodes.  stats_out (nullable) = [sum nodes visited, sum tris tested].
int emu_build_trace(const float *verts, const uint32_t *flags, int n, const float *rays, int m, int kind, int leaf_max,
                    Y4Hit *hits, uint64_t *stats_out)
{
	EmuBvh b;
	emu_build(verts, flags, n, leaf_max, b);
	Y4Bvh8 bvh; bvh.nodes = b.nodes.data(); bvh.tris = b.tris.data(); bvh.n_tris = (uint32_t) n; bvh.consts = nullptr; bvh.postpone_sp_limit = 0;
	uint64_t sn = 0, st = 0;
	for(int i = 0; i < m; ++i)
	{
		Y4Ray r; memcpy(&r, rays + 8ull * i, sizeof(r));
		const Y4PreparedRay pr = y4_prepare_ray(r, kind);
		uint2 stack[Y4_STACK_SIZE];
		uint32_t filt[16], filt_mat[16]; int n_filt = 0;
		Y4TraceStats s{0, 0};
		if(kind == Y4_TRACE_CLOSEST) y4_trace_ray<Y4_TRACE_CLOSEST>(bvh, pr, hits[i], stack, &s, filt, filt_mat, &n_filt, 0);
		else if(kind == Y4_TRACE_SHADOW) y4_trace_ray<Y4_TRACE_SHADOW>(bvh, pr, hits[i], stack, &s, filt, filt_mat, &n_filt, 0);
		else { y4_trace_ray<Y4_TRACE_TSHADOW>(bvh, pr, hits[i], stack, &s, filt, filt_mat, &n_filt, 8); hits[i].bu = (float) n_filt; }
		sn += s.nodes; st += s.tris;
	}
	if(stats_out) { stats_out[0] = sn; stats_out[1] = st; }
	return (int) b.n_nodes;
}

// Two-level scene: `n_mesh` base meshes in object space (vertices concatenated, mesh_tris[i] triangles each),
// `n_inst` instances {mesh, identity, 12 matrix floats (rows of Matrix4)} in primitive order.  Builds a bottom-level BVH8 per
// mesh and a top-level BVH8 over the instances' exact world boxes (the product does the same in y4_build.cu) and traces.
void emu_inverse_affine(const float *m, float *mi)
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

int emu_build_trace2(const float *mesh_verts, const uint32_t *mesh_flags, const int *mesh_tris, int n_mesh,
                     const int *inst_mesh, const int *inst_identity, const float *inst_matrix, int n_inst,
                     const float *rays, int m, int kind, Y4Hit *hits, uint64_t *stats_out)
{
	std::vector<size_t> mesh_first(n_mesh + 1, 0);
	for(int i = 0; i < n_mesh; ++i) mesh_first[i + 1] = mesh_first[i] + (size_t) mesh_tris[i];
	// bottom level
	std::vector<Y4U4> blas_nodes; std::vector<Y4TriObj> blas_tris;
	std::vector<uint32_t> node_off(n_mesh), tri_off(n_mesh);
	std::vector<float> bmax(n_mesh, 0.f);
	for(int i = 0; i < n_mesh; ++i)
	{
		const float *v = mesh_verts + 9 * mesh_first[i];
		const uint32_t *f = mesh_flags ? mesh_flags + mesh_first[i] : nullptr;
		EmuBvh b;
		emu_build(v, f, mesh_tris[i], 3, b);
		node_off[i] = (uint32_t) (blas_nodes.size() / Y4_NODE_WORDS); tri_off[i] = (uint32_t) blas_tris.size();
		blas_nodes.insert(blas_nodes.end(), b.nodes.begin(), b.nodes.begin() + (size_t) b.n_nodes * Y4_NODE_WORDS);
		for(const Y4Tri &t : b.tris)
		{
			Y4TriObj o; memcpy(o.v, v + 9ull * t.prim, 36); o.prim_local = t.prim; o.flags = t.flags; o.pad0 = 0u;
			blas_tris.push_back(o);
		}
		for(size_t k = 0; k < 9 * (size_t) mesh_tris[i]; ++k) bmax[i] = std::max(bmax[i], std::fabs(v[k]));
	}
	// instances + their exact world boxes as proxy triangles (v0 = lo, v1 = hi, v2 = (lo.x, hi.y, lo.z): the same box)
	std::vector<Y4Inst> inst(n_inst);
	std::vector<float> proxy(9 * (size_t) n_inst);
	uint32_t first_prim = 0;
	for(int i = 0; i < n_inst; ++i)
	{
		Y4Inst &in = inst[i];
		memset(&in, 0, sizeof(in));
		memcpy(in.m, inst_matrix + 12 * i, 48);
		emu_inverse_affine(in.m, in.mi);
		const int me = inst_mesh[i];
		in.node_off = node_off[me]; in.tri_off = tri_off[me]; in.first_prim = first_prim; in.identity = inst_identity[i] ? 1u : 0u;
		in.bmax = bmax[me]; in.n_tri = (uint32_t) mesh_tris[me]; in.mesh = (uint32_t) me; in.vert_first = (uint32_t) mesh_first[me];
		first_prim += in.n_tri;
		Y4Box wb = y4_box_empty();
		const float *v = mesh_verts + 9 * mesh_first[me];
		for(size_t k = 0; k < 3 * (size_t) mesh_tris[me]; ++k)
		{
			V3 p = v3(v[3 * k], v[3 * k + 1], v[3 * k + 2]);
			if(!in.identity) p = y4_inst_point(in.m, p);
			wb.lox = fminf(wb.lox, p.x); wb.loy = fminf(wb.loy, p.y); wb.loz = fminf(wb.loz, p.z);
			wb.hix = fmaxf(wb.hix, p.x); wb.hiy = fmaxf(wb.hiy, p.y); wb.hiz = fmaxf(wb.hiz, p.z);
		}
		float *q = proxy.data() + 9ull * i;
		q[0] = wb.lox; q[1] = wb.loy; q[2] = wb.loz; q[3] = wb.hix; q[4] = wb.hiy; q[5] = wb.hiz; q[6] = wb.lox; q[7] = wb.hiy; q[8] = wb.loz;
	}
	EmuBvh top;
	emu_build(proxy.data(), nullptr, n_inst, 3, top);
	Y4Scene2 sc;
	sc.tlas_nodes = top.nodes.data(); sc.tlas_leaves = top.tris.data(); sc.inst = inst.data(); sc.n_inst = (uint32_t) n_inst;
	sc.blas_nodes = blas_nodes.data(); sc.blas_tris = blas_tris.data(); sc.consts = nullptr;
	uint64_t sn = 0, st = 0;
	for(int i = 0; i < m; ++i)
	{
		Y4Ray r; memcpy(&r, rays + 8ull * i, sizeof(r));
		const Y4PreparedRay pr = y4_prepare_ray(r, kind);
		uint2 stack[Y4_STACK_SIZE];
		Y4TraceStats s{0, 0};
		if(kind == Y4_TRACE_CLOSEST) y4_trace_ray2<Y4_TRACE_CLOSEST>(sc, pr, hits[i], stack, &s);
		else if(kind == Y4_TRACE_SHADOW) y4_trace_ray2<Y4_TRACE_SHADOW>(sc, pr, hits[i], stack, &s);
		else { uint32_t filt[16], filt_mat[16]; int n_filt = 0; y4_trace_ray2<Y4_TRACE_TSHADOW>(sc, pr, hits[i], stack, &s, filt, filt_mat, &n_filt, 8); hits[i].bu = (float) n_filt; }
		sn += s.nodes; st += s.tris;
	}
	if(stats_out) { stats_out[0] = sn; stats_out[1] = st; }
	return (int) (top.n_nodes + blas_nodes.size() / Y4_NODE_WORDS);
}

// Brute force over all triangles in primitive order with the same exact triangle test.
void emu_brute(const float *verts, const uint32_t *flags, int n, const float *rays, int m, int kind, Y4Hit *hits)
{
	std::vector<Y4Tri> tr(n);
	for(int i = 0; i < n; ++i) tr[i] = y4_tri_setup(verts + 9ull * i, (uint32_t) i, flags ? flags[i] : (uint32_t) Y4_TRI_DEFAULT_FLAGS);
	const uint32_t vis = kind == Y4_TRACE_CLOSEST ? Y4_TRI_VIS_CLOSEST : (kind == Y4_TRACE_SHADOW ? Y4_TRI_VIS_SHADOW : Y4_TRI_VIS_TSHADOW);
	for(int i = 0; i < m; ++i)
	{
		Y4Ray r; memcpy(&r, rays + 8ull * i, sizeof(r));
		const Y4PreparedRay pr = y4_prepare_ray(r, kind);
		Y4Hit h{0.f, 0.f, 0.f, -1};
		float t_hi = pr.t_hi;
		for(int k = 0; k < n; ++k)
		{
			float t, u, v;
			if(!y4_tri_intersect(pr.o, pr.d, v3(tr[k].v0x, tr[k].v0y, tr[k].v0z), v3(tr[k].e1x, tr[k].e1y, tr[k].e1z), v3(tr[k].e2x, tr[k].e2y, tr[k].e2z), tr[k].eps, t, u, v)) continue;
			if(!(t < t_hi && t >= pr.t_lo) || !(tr[k].flags & vis)) continue;
			h.t = t; h.bu = u; h.bv = v; h.prim = k;
			if(kind != Y4_TRACE_CLOSEST) break;
			t_hi = t;
		}
		hits[i] = h;
	}
}
}
