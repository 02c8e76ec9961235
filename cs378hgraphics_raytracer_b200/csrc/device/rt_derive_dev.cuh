// rt_derive_dev.cuh - the traversal layouts (rt_scene.cuh: DNode2, DTree, DTri, slacks) built ON
// THE DEVICE from the raw arena sections, so that rt_upload_scene is one host->device copy of the
// arena plus a few streaming kernels (the host version, rt_derive.hpp, is kept for tests/hostsim and
// as the checker behind rt_selftest_derive).  The per-element range checks of rt_arena_validate
// that are O(scene size) run here as well: a bad index sets *err and the upload is rejected.
#pragma once
#include <cub/device/device_scan.cuh>

namespace rtd {

enum { DERR_NODE = 1, DERR_LEAF = 2, DERR_ITEM = 4, DERR_FACE = 8, DERR_VMAT = 16 };

// mesh whose [off, off+cnt) range of nodes (by_node) or leaf items contains idx; meshes are stored
// in ascending offset order (checked on the host)
__device__ __forceinline__ int mesh_of(const rt_mesh* ms, int n_mesh, uint64_t idx, bool by_node)
{
	int lo = 0, hi = n_mesh - 1;
	while (lo < hi) {
		const int mid = (lo + hi + 1) >> 1;
		const uint64_t off = by_node ? ms[mid].node_off : ms[mid].item_off;
		if (off <= idx) lo = mid;
		else hi = mid - 1;
	}
	return lo;
}

__global__ void __launch_bounds__(256) k_derive_flags(const rt_node* n, long long cnt, int* flag)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k < cnt) flag[k] = n[k].right >= 0 ? 1 : 0;
}

// pre[k] = number of internal nodes before node k (exclusive scan of the flags).  One thread per
// node of the concatenated trees (n_mesh > 0) or of the single scene tree (n_mesh == 0).
__global__ void __launch_bounds__(256) k_derive_nodes2(const rt_node* n, long long cnt, const int* pre, const rt_mesh* ms, int n_mesh, long long scene_items, DNode2* out, unsigned* err)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= cnt) return;
	long long base = 0, tree_cnt = cnt, items = scene_items;
	if (n_mesh > 0) {
		const rt_mesh& M = ms[mesh_of(ms, n_mesh, (uint64_t)k, true)];
		base = (long long)M.node_off;
		tree_cnt = (long long)M.node_cnt;
		items = (long long)M.item_cnt;
	}
	const rt_node& me = n[k];
	if (me.right < 0) {
		if (me.first < 0 || me.count < 0 || me.count > 3 || (long long)me.first + me.count > items || me.first >= (1 << 29)) atomicOr(err, DERR_LEAF);
		return;
	}
	const long long rel = k - base;
	if (rel + 1 >= tree_cnt || me.right <= rel + 1 || me.right >= tree_cnt) { atomicOr(err, DERR_NODE); return; }
	const int pb = pre[base];
	const rt_node& a = n[k + 1];
	const rt_node& b = n[base + me.right];
	DNode2 d;
#pragma unroll
	for (int c = 0; c < 3; c++) { d.lo0[c] = a.bmin[c]; d.hi0[c] = a.bmax[c]; d.lo1[c] = b.bmin[c]; d.hi1[c] = b.bmax[c]; }
	d.c0 = a.right >= 0 ? pre[k + 1] - pb : leaf_ref(a.first, a.count);
	d.c1 = b.right >= 0 ? pre[base + me.right] - pb : leaf_ref(b.first, b.count);
#pragma unroll
	for (int c = 0; c < 6; c++) d.pad[c] = 0;
	out[pre[k]] = d;
}

// float image of the wide nodes (round to nearest, like the host's (float) casts)
__global__ void __launch_bounds__(256) k_derive_nodes2f(const DNode2* in, long long cnt, DNode2f* out)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= cnt) return;
	const DNode2& a = in[k];
	DNode2f f;
#pragma unroll
	for (int c = 0; c < 3; c++) { f.lo0[c] = (float)a.lo0[c]; f.hi0[c] = (float)a.hi0[c]; f.lo1[c] = (float)a.lo1[c]; f.hi1[c] = (float)a.hi1[c]; }
	f.c0 = a.c0;
	f.c1 = a.c1;
	f.pad[0] = f.pad[1] = 0;
	out[k] = f;
}

// float image of the candidates' world boxes, in leaf item order (bad item indices are reported by
// k_derive_check; their slot is left zeroed)
__global__ void __launch_bounds__(256) k_derive_itembox(const int32_t* sitems, long long n_sitems, const rt_object* objs, long long n_obj, DItemBox* out, DPrimF* outp)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= n_sitems) return;
	DItemBox b;
	const int32_t obj = sitems[k];
	const bool ok = obj >= 0 && obj < n_obj;
#pragma unroll
	for (int c = 0; c < 3; c++) {
		b.lo[c] = ok ? (float)objs[obj].bbox[c] : 0.0f;
		b.hi[c] = ok ? (float)objs[obj].bbox[3 + c] : 0.0f;
	}
	b.obj = ok ? obj : 0;
	b.type = ok ? objs[obj].type : 0;
	out[k] = b;
	// the object-space float test's record of the same item (rt_prims.cuh: derive_primf, the host build runs the same function)
	DPrimF p;
	if (ok) derive_primf(objs[obj], b, p);
	else { for (int j = 0; j < 9; j++) p.m[j] = 0.0f; p.lc[0] = p.lc[1] = p.lc[2] = 0.0f; p.kind = 0; p.pad[0] = p.pad[1] = p.pad[2] = 0; }
	outp[k] = p;
}

// breadth-first top table of the scene tree (rt_scene.cuh: RT_REF_TOP); one thread, at most `cap` nodes
__global__ void k_derive_top(const DNode2f* nodesf, int root, int cap, DNode2f* out, int* n_out)
{
	if (threadIdx.x || blockIdx.x) return;
	int n = 0;
	if (root >= 0 && cap > 0) {
		out[0] = nodesf[root];
		out[0].pad[0] = root;
		n = 1;
		for (int j = 0; j < n; j++) {
			DNode2f nd = out[j];
			if (nd.c0 >= 0 && n < cap) { out[n] = nodesf[nd.c0]; out[n].pad[0] = nd.c0; nd.c0 = RT_REF_TOP | n; n++; }
			if (nd.c1 >= 0 && n < cap) { out[n] = nodesf[nd.c1]; out[n].pad[0] = nd.c1; nd.c1 = RT_REF_TOP | n; n++; }
			out[j].c0 = nd.c0;
			out[j].c1 = nd.c1;
		}
	}
	*n_out = n;
}

__device__ __forceinline__ double dev_box_extent(const rt_node& n)
{
	double m = 1.0;
	for (int a = 0; a < 3; a++) {
		if (fabs(n.bmin[a]) > m) m = fabs(n.bmin[a]);
		if (fabs(n.bmax[a]) > m) m = fabs(n.bmax[a]);
	}
	return m;
}

__global__ void __launch_bounds__(256) k_derive_mesh(const rt_mesh* ms, int n_mesh, const rt_node* mnodes, const int* pre, DTree* mtrees, double* mesh_slack)
{
	const int m = blockIdx.x * 256 + threadIdx.x;
	if (m >= n_mesh) return;
	const rt_mesh& M = ms[m];
	const rt_node& root = mnodes[M.node_off];
	DTree T;
	for (int a = 0; a < 3; a++) { T.lo[a] = root.bmin[a]; T.hi[a] = root.bmax[a]; }
	T.root = root.right >= 0 ? 0 : leaf_ref(root.first, root.count);
	T.pad = 0;
	T.node_off = pre[M.node_off];
	mtrees[m] = T;
	mesh_slack[m] = 1e-9 * dev_box_extent(root);
}

// one thread per mesh leaf item: the triangle record of the face it names
__global__ void __launch_bounds__(256) k_derive_tris(const rt_mesh* ms, int n_mesh, const int32_t* mitems, const rt_face* faces, const double* verts, long long n_items, DTri* tris, unsigned* err)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= n_items) return;
	const rt_mesh& M = ms[mesh_of(ms, n_mesh, (uint64_t)k, false)];
	const int32_t fi = mitems[k];
	DTri T;
	if (fi < 0 || (uint64_t)fi >= M.face_cnt) {
		atomicOr(err, DERR_ITEM);
		return;
	}
	const rt_face& f = faces[M.face_off + fi];
	if (f.ids[0] < 0 || f.ids[1] < 0 || f.ids[2] < 0 || (uint64_t)f.ids[0] >= M.vert_cnt || (uint64_t)f.ids[1] >= M.vert_cnt || (uint64_t)f.ids[2] >= M.vert_cnt) {
		atomicOr(err, DERR_FACE);
		return;
	}
#pragma unroll
	for (int c = 0; c < 3; c++) {
		T.v0[c] = verts[3 * (M.vert_off + f.ids[0]) + c];
		T.v1[c] = verts[3 * (M.vert_off + f.ids[1]) + c];
		T.v2[c] = verts[3 * (M.vert_off + f.ids[2]) + c];
		T.n[c] = f.normal[c];
	}
	tris[k] = T;
}

// ---- tree depth (the walk's deferred-subtree stack holds RT_STACK entries) ------------------------
// parent links, then pointer jumping: after r rounds anc[k] is the 2^r-th ancestor (or the root)
// and dist[k] the number of edges climbed.
__global__ void __launch_bounds__(256) k_depth_self(long long cnt, int* anc, int* dist)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k < cnt) { anc[k] = (int)k; dist[k] = 0; }
}
__global__ void __launch_bounds__(256) k_depth_init(const rt_node* n, long long cnt, const rt_mesh* ms, int n_mesh, int* anc, int* dist)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= cnt) return;
	long long base = 0;
	if (n_mesh > 0) base = (long long)ms[mesh_of(ms, n_mesh, (uint64_t)k, true)].node_off;
	const rt_node& me = n[k];
	if (me.right >= 0) { // links were range-checked by k_derive_nodes2 (err set otherwise; results unused then)
		const long long r = base + me.right;
		if (k + 1 < cnt) { anc[k + 1] = (int)k; dist[k + 1] = 1; }
		if (r < cnt && r > k + 1) { anc[r] = (int)k; dist[r] = 1; }
	}
}
__global__ void __launch_bounds__(256) k_depth_jump(long long cnt, const int* anc_in, const int* dist_in, int* anc_out, int* dist_out)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= cnt) return;
	const int p = anc_in[k];
	anc_out[k] = anc_in[p];
	dist_out[k] = dist_in[k] + dist_in[p];
}
// max_depth = deepest node (root = 1); bit 31 set if some node is still not at its root after the jumps
__global__ void __launch_bounds__(256) k_depth_max(long long cnt, const int* anc, const int* dist, unsigned* max_depth)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k >= cnt) return;
	const int p = anc[k];
	unsigned v = (unsigned)dist[k] + 1u;
	if (anc[p] != p) v |= 0x80000000u;
	atomicMax(max_depth, v);
}

// remaining O(n) range checks: scene leaf items name objects, per-vertex materials name materials
__global__ void __launch_bounds__(256) k_derive_check(const int32_t* sitems, long long n_sitems, long long n_obj, const int32_t* vmats, long long n_vmats, long long n_mat, unsigned* err)
{
	const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
	if (k < n_sitems && (sitems[k] < 0 || sitems[k] >= n_obj)) atomicOr(err, DERR_ITEM);
	if (k < n_vmats && (vmats[k] < 0 || vmats[k] >= n_mat)) atomicOr(err, DERR_VMAT);
}

} // namespace rtd
