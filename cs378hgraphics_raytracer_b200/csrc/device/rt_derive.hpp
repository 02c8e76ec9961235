// rt_derive.hpp - host-side preparation of the traversal layouts derived from an arena
// (rt_scene.cuh): the per-leaf-item triangle records and the pruning slacks.  Used by
// rt_api.cu before the upload (and by tests/hostsim to run the same per-ray code on the CPU).
#pragma once
#include <cmath>
#include <vector>
#include "rt_arena_writer.hpp"
#include "rt_scene.cuh"
#include "rt_prims.cuh" // derive_primf

namespace rtd {

struct Derived {
	std::vector<DNode2> snodes2, mnodes2;
	std::vector<DNode2f> snodes2f, mnodes2f;
	std::vector<DItemBox> sboxf;
	std::vector<DPrimF> sprimf;
	DTree stree;
	std::vector<DTree> mtrees;
	std::vector<DTri> tris;
	std::vector<double> mesh_slack;
	double scene_slack = 1e-9;
	int max_scene_depth = 0, max_mesh_depth = 0;
};

inline int tree_depth(const rt_node* n, uint64_t cnt)
{
	// pre-order array: depth of node k = number of open ancestors
	std::vector<int> depth(cnt, 0);
	int best = cnt ? 1 : 0;
	for (uint64_t k = 0; k < cnt; k++) {
		if (k == 0) depth[0] = 1;
		if (n[k].right >= 0) {
			depth[k + 1] = depth[k] + 1;
			depth[n[k].right] = depth[k] + 1;
			if (depth[k] + 1 > best) best = depth[k] + 1;
		}
	}
	return best;
}

// pre-order rt_node array (internal: child0 = k+1, child1 = right) -> DNode2 array of the internal
// nodes only, same pre-order; leaves become references
inline void build_nodes2(const rt_node* n, uint64_t cnt, std::vector<DNode2>& out, DTree& tree)
{
	tree.node_off = (int64_t)out.size();
	for (int a = 0; a < 3; a++) { tree.lo[a] = cnt ? n[0].bmin[a] : 0.0; tree.hi[a] = cnt ? n[0].bmax[a] : 0.0; }
	tree.pad = 0;
	std::vector<int32_t> ref(cnt);
	int32_t internal = 0;
	for (uint64_t k = 0; k < cnt; k++) ref[k] = n[k].right >= 0 ? internal++ : leaf_ref(n[k].first, n[k].count);
	tree.root = cnt ? ref[0] : leaf_ref(0, 0);
	out.resize(out.size() + internal);
	DNode2* dst = out.data() + tree.node_off;
	for (uint64_t k = 0; k < cnt; k++) {
		if (n[k].right < 0) continue;
		DNode2& d = dst[ref[k]];
		const rt_node& a = n[k + 1];
		const rt_node& b = n[n[k].right];
		for (int c = 0; c < 3; c++) { d.lo0[c] = a.bmin[c]; d.hi0[c] = a.bmax[c]; d.lo1[c] = b.bmin[c]; d.hi1[c] = b.bmax[c]; }
		d.c0 = ref[k + 1];
		d.c1 = ref[n[k].right];
		for (int c = 0; c < 6; c++) d.pad[c] = 0;
	}
}

inline void nodes_to_float(const std::vector<DNode2>& in, std::vector<DNode2f>& out)
{
	out.resize(in.size());
	for (size_t k = 0; k < in.size(); k++) {
		const DNode2& a = in[k];
		DNode2f& f = out[k];
		for (int c = 0; c < 3; c++) { f.lo0[c] = (float)a.lo0[c]; f.hi0[c] = (float)a.hi0[c]; f.lo1[c] = (float)a.lo1[c]; f.hi1[c] = (float)a.hi1[c]; }
		f.c0 = a.c0;
		f.c1 = a.c1;
		f.pad[0] = f.pad[1] = 0;
	}
}

// DScene::kt_unit: every constant kt lies in [0,1]^3 (texels always do), so a shadow attenuation - a
// product of kt ^ distance - is finite and 0 * attenuation is exactly 0
inline int kt_all_unit(const rt_material* m, uint64_t cnt)
{
	for (uint64_t i = 0; i < cnt; i++)
		for (int c = 0; c < 3; c++)
			if (!(m[i].k[RT_KT][c] >= 0.0 && m[i].k[RT_KT][c] <= 1.0)) return 0;
	return 1;
}

inline double box_extent(const rt_node& n)
{
	double m = 1.0;
	for (int a = 0; a < 3; a++) {
		if (std::fabs(n.bmin[a]) > m) m = std::fabs(n.bmin[a]);
		if (std::fabs(n.bmax[a]) > m) m = std::fabs(n.bmax[a]);
	}
	return m;
}

inline void derive(const rt::ArenaView& A, Derived& D)
{
	const rt_mesh* meshes = A.ptr<rt_mesh>(RT_SEC_MESHES);
	const rt_face* faces = A.ptr<rt_face>(RT_SEC_FACES);
	const double* verts = A.ptr<double>(RT_SEC_VERTS);
	const int32_t* mitems = A.ptr<int32_t>(RT_SEC_MESH_ITEMS);
	const rt_node* mnodes = A.ptr<rt_node>(RT_SEC_MESH_NODES);
	D.tris.resize(A.count(RT_SEC_MESH_ITEMS));
	D.mesh_slack.resize(A.count(RT_SEC_MESHES));
	D.mtrees.resize(A.count(RT_SEC_MESHES));
	for (uint64_t m = 0; m < A.count(RT_SEC_MESHES); m++) {
		const rt_mesh& M = meshes[m];
		for (uint64_t k = 0; k < M.item_cnt; k++) {
			const rt_face& f = faces[M.face_off + mitems[M.item_off + k]];
			DTri& T = D.tris[M.item_off + k];
			for (int c = 0; c < 3; c++) {
				T.v0[c] = verts[3 * (M.vert_off + f.ids[0]) + c];
				T.v1[c] = verts[3 * (M.vert_off + f.ids[1]) + c];
				T.v2[c] = verts[3 * (M.vert_off + f.ids[2]) + c];
				T.n[c] = f.normal[c];
			}
		}
		// slack: far above the rounding of a computed hit point (ulp(extent) ~ 2e-16*extent),
		// far below any geometric feature
		D.mesh_slack[m] = 1e-9 * box_extent(mnodes[M.node_off]);
		build_nodes2(mnodes + M.node_off, M.node_cnt, D.mnodes2, D.mtrees[m]);
		int d = tree_depth(mnodes + M.node_off, M.node_cnt);
		if (d > D.max_mesh_depth) D.max_mesh_depth = d;
	}
	const rt_node* snodes = A.ptr<rt_node>(RT_SEC_SCENE_NODES);
	D.scene_slack = 1e-9 * box_extent(snodes[0]);
	build_nodes2(snodes, A.count(RT_SEC_SCENE_NODES), D.snodes2, D.stree);
	D.max_scene_depth = tree_depth(snodes, A.count(RT_SEC_SCENE_NODES));
	nodes_to_float(D.snodes2, D.snodes2f);
	nodes_to_float(D.mnodes2, D.mnodes2f);
	const int32_t* sitems = A.ptr<int32_t>(RT_SEC_SCENE_ITEMS);
	const rt_object* objs = A.ptr<rt_object>(RT_SEC_OBJECTS);
	D.sboxf.resize(A.count(RT_SEC_SCENE_ITEMS));
	for (size_t k = 0; k < D.sboxf.size(); k++) {
		DItemBox& b = D.sboxf[k];
		std::memset(&b, 0, sizeof(b));
		const int32_t obj = sitems[k];
		if (obj < 0 || (uint64_t)obj >= A.count(RT_SEC_OBJECTS)) continue;
		for (int c = 0; c < 3; c++) { b.lo[c] = (float)objs[obj].bbox[c]; b.hi[c] = (float)objs[obj].bbox[3 + c]; }
		b.obj = obj;
		b.type = objs[obj].type;
	}
	D.sprimf.resize(D.sboxf.size());
	for (size_t k = 0; k < D.sboxf.size(); k++) {
		const int32_t obj = sitems[k];
		if (obj < 0 || (uint64_t)obj >= A.count(RT_SEC_OBJECTS)) { std::memset(&D.sprimf[k], 0, sizeof(DPrimF)); continue; }
		derive_primf(objs[obj], D.sboxf[k], D.sprimf[k]);
	}
}

} // namespace rtd
