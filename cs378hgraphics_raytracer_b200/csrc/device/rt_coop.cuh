// rt_coop.cuh - Scene::intersect for ONE ray by a whole warp.
//
// The closest-hit stage gives every ray a node budget (rt_traverse.cuh).  Rays that exhaust it are
// almost always rays with a zero direction component: the reference's slab test skips such an axis
// altogether (bbox.cc:43 `if (d[axis] == 0) continue`), so every box along that axis counts as hit
// and the walk fans out over thousands of nodes - e.g. the centre pixel column of a symmetric
// camera and its mirror images.  One thread would need milliseconds for such a ray and the rest of
// its warp (and, at the end of a launch, the whole GPU) would wait for it.
//
// Here the warp shares one stack in shared memory: each step the top (most recently pushed, i.e.
// nearest) up to 32 entries are popped, one per lane; node lanes test their two child boxes, leaf
// lanes their triangles; surviving children are pushed back with a ballot scan; the culling limit
// is the warp-wide minimum of the lanes' best hits.  The winner is argmin (t, leaf position) - the
// reference's first-strictly-smaller rule - which does not depend on the visiting order, so the
// result is the one the single-thread walk gives, bit for bit.  The scene level (few nodes) is
// walked by all lanes in lockstep; trimesh candidates are walked cooperatively.
#pragma once

namespace rtd {

#define RT_COOP_CAP 1536       // stack entries per warp (18 KB of shared memory)
#define RT_COOP_WARPS 1        // warps per block of the heavy-ray kernel (47 KB of shared memory per warp)

struct CoopStack {
	int ref[RT_COOP_CAP];
	double ent[RT_COOP_CAP];
};

__device__ __forceinline__ double warp_min(double v)
{
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		const double u = __shfl_xor_sync(0xffffffffu, v, o);
		v = u < v ? u : v;
	}
	return v;
}

// Trimesh::intersectLocal (trimesh.cpp:79-95) over the mesh's tree by the whole warp.
// o, d, I: the ray in the mesh's object space (the same in every lane); lim: object-space limit.
// On return have / best_t / best_pos are warp-uniform.
__device__ __noinline__ void coop_mesh(const DTree& T, const DNode2* nodes, const DNode2f* nodesf, const DTri* tris, const d3& o, const d3& d, const RayInv& I, const RayF& F, double lim,
                                       CoopStack& K, bool& have, double& best_t, int& best_pos)
{
	const unsigned lane = threadIdx.x & 31;
	const unsigned hi_mask = lane == 31 ? 0u : (0xffffffffu << (lane + 1));
	bool my_have = false;
	double my_t = RT_T_INF;
	int my_pos = 0x7fffffff;
	have = false;
	best_t = RT_T_INF;
	best_pos = 0x7fffffff;
	double er;
	if (!box_hit(T.lo, T.hi, o, d, I, er) || er > lim) return;
	int n = 1;
	if (lane == 0) { K.ref[0] = T.root; K.ent[0] = er; }
	__syncwarp();
	while (n > 0) {
		const int take = n > RT_COOP_CAP - 128 ? 1 : (n < 32 ? n : 32);
		const bool mine = (int)lane < take;
		int ref = RT_REF_NONE;
		double e = 0.0;
		if (mine) { ref = K.ref[n - 1 - lane]; e = K.ent[n - 1 - lane]; }
		n -= take;
		__syncwarp();
		const double m = warp_min(my_t);
		if (m < lim) lim = m;
		if (mine && e > lim) ref = RT_REF_NONE;
		int c0 = 0, c1 = 0, nc = 0;
		double ce0 = 0.0, ce1 = 0.0;
		if (ref != RT_REF_NONE) {
			if (ref >= 0) {
				const DNode2f& nd = nodesf[ref];
				double e0, e1;
				bool h0, h1;
				box2_hit(nd, nodes[ref], o, d, I, F, h0, h1, e0, e1);
				h0 = h0 && !(e0 > lim);
				h1 = h1 && !(e1 > lim);
				if (h0 && h1) {
					// far child first: the near one ends up on top
					const bool swap = e1 < e0;
					c0 = swap ? nd.c0 : nd.c1; ce0 = swap ? e0 : e1;
					c1 = swap ? nd.c1 : nd.c0; ce1 = swap ? e1 : e0;
					nc = 2;
				} else if (h0) { c0 = nd.c0; ce0 = e0; nc = 1; }
				else if (h1) { c0 = nd.c1; ce0 = e1; nc = 1; }
			} else {
				const int first = leaf_first(ref), count = leaf_count(ref);
#pragma unroll 1
				for (int i = 0; i < count; i++) {
					const int pos = first + i;
					double t;
					d3 bary;
					if (!tri_hit<false>(tris[pos], o, d, t, bary)) continue;
					if (!my_have || t < my_t || (t == my_t && pos < my_pos)) {
						my_have = true;
						my_t = t;
						my_pos = pos;
					}
				}
			}
		}
		// push: lanes in reverse order, so that lane 0's children (it held the top entry) end on top
		const unsigned b1 = __ballot_sync(0xffffffffu, nc >= 1), b2 = __ballot_sync(0xffffffffu, nc == 2);
		const int at = n + __popc(b1 & hi_mask) + __popc(b2 & hi_mask);
		if (nc >= 1) { K.ref[at] = c0; K.ent[at] = ce0; }
		if (nc == 2) { K.ref[at + 1] = c1; K.ent[at + 1] = ce1; }
		n += __popc(b1) + __popc(b2);
		__syncwarp();
	}
	// argmin (t, position) over the lanes
#pragma unroll
	for (int o2 = 16; o2 > 0; o2 >>= 1) {
		const double ot = __shfl_xor_sync(0xffffffffu, my_t, o2);
		const int op = __shfl_xor_sync(0xffffffffu, my_pos, o2);
		const bool oh = __shfl_xor_sync(0xffffffffu, my_have ? 1 : 0, o2) != 0;
		if (oh && (!my_have || ot < my_t || (ot == my_t && op < my_pos))) {
			my_have = true;
			my_t = ot;
			my_pos = op;
		}
	}
	have = my_have;
	best_t = my_t;
	best_pos = my_pos;
}

// Scene-level shared state of one warp: its own stack (the mesh walk uses the other one) and the
// list of trimesh candidates found so far, walked cooperatively one after the other (phase B).
#define RT_COOP_SCENE_CAP 1024
#define RT_COOP_MESH_LIST 128
struct CoopScene {
	int ref[RT_COOP_SCENE_CAP];
	double ent[RT_COOP_SCENE_CAP];
	int mesh_rank[RT_COOP_MESH_LIST];
	double mesh_entry[RT_COOP_MESH_LIST];
};

// Phase B: the listed trimesh candidates, each by the whole warp.  C is warp-uniform.
__device__ __noinline__ void coop_mesh_list(const DScene& S, const d3& ow, const d3& dw, const CoopScene& A, int n_list, TravClosest& C, CoopStack& K)
{
	for (int j = 0; j < n_list; j++) {
		const double lim = C.have ? C.best_t : RT_T_INF;
		if (A.mesh_entry[j] > lim) continue;
		const int rank = A.mesh_rank[j];
		const int obj = S.sitems[rank];
		const rt_object& ob = S.objs[obj];
		LocalRay L;
		to_local(L, ob, ow, dw);
		const DTree& T = S.mtrees[ob.mesh];
		const int item_off = (int)S.meshes[ob.mesh].item_off;
		RayInv Il;
		RayF Fl;
		make_inv(Il, Fl, L.o, L.d, S.mesh_slack[ob.mesh], S.mesh_slack[ob.mesh] * 1e9);
		bool have;
		double bt;
		int bp;
		coop_mesh(T, S.mnodes2 + T.node_off, S.mnodes2f + T.node_off, S.tris + item_off, L.o, L.d, Il, Fl, lim_to_local(lim, L.len), K, have, bt, bp);
		if (have) {
			const double t = bt / L.len;
			if (!C.have || t < C.best_t || (t == C.best_t && rank < C.best_rank)) {
				C.have = true;
				C.best_t = t;
				C.best_rank = rank;
				C.best_obj = obj;
				C.best_fpos = item_off + bp;
			}
		}
	}
}

// Scene::intersect (scene.cpp:157-180) for one ray by one warp; all lanes hold the same ray and
// return the same C.  Phase A walks the scene-level tree cooperatively (as coop_mesh does a mesh):
// analytic candidates are tested by the lane that popped their leaf and kept as that lane's best;
// trimesh candidates are only listed.  Phase B walks the listed meshes.  The final winner is
// argmin (t, rank) over both groups - the reference's rule, independent of the visiting order.
__device__ __forceinline__ void coop_closest(const DScene& S, const d3& ow, const d3& dw, TravClosest& C, CoopStack& K, CoopScene& A)
{
	const unsigned lane = threadIdx.x & 31;
	const unsigned lt_mask = (1u << lane) - 1u;
	const unsigned hi_mask = lane == 31 ? 0u : (0xffffffffu << (lane + 1));
	RayInv I;
	RayF F;
	make_inv(I, F, ow, dw, S.scene_slack, S.scene_slack * 1e9);
	double e;
	if (!box_hit(S.stree.lo, S.stree.hi, ow, dw, I, e)) return;
	TravClosest mine; // this lane's best analytic candidate
	mine.have = false;
	mine.best_t = RT_T_INF;
	mine.best_rank = 0x7fffffff;
	mine.best_obj = -1;
	mine.best_fpos = -1;
	int n = 1, n_list = 0;
	if (lane == 0) { A.ref[0] = S.stree.root; A.ent[0] = e; }
	__syncwarp();
	while (n > 0) {
		const int take = n > RT_COOP_SCENE_CAP - 128 ? 1 : (n < 32 ? n : 32);
		const bool has = (int)lane < take;
		int ref = RT_REF_NONE;
		double en = 0.0;
		if (has) { ref = A.ref[n - 1 - lane]; en = A.ent[n - 1 - lane]; }
		n -= take;
		__syncwarp();
		double lim = warp_min(mine.best_t);
		if (C.have && C.best_t < lim) lim = C.best_t;
		if (has && en > lim) ref = RT_REF_NONE;
		int c0 = 0, c1 = 0, nc = 0;
		double ce0 = 0.0, ce1 = 0.0;
		if (ref != RT_REF_NONE && ref >= 0) {
			const DNode2f& nd = S.snodes2f[ref];
			double e0, e1;
			bool h0, h1;
			box2_hit(nd, S.snodes2[ref], ow, dw, I, F, h0, h1, e0, e1);
			h0 = h0 && !(e0 > lim);
			h1 = h1 && !(e1 > lim);
			if (h0 && h1) {
				const bool swap = e1 < e0;
				c0 = swap ? nd.c0 : nd.c1; ce0 = swap ? e0 : e1;
				c1 = swap ? nd.c1 : nd.c0; ce1 = swap ? e1 : e0;
				nc = 2;
			} else if (h0) { c0 = nd.c0; ce0 = e0; nc = 1; }
			else if (h1) { c0 = nd.c1; ce0 = e1; nc = 1; }
		}
		// leaves: up to three candidates per lane; the loop is uniform so that meshes can be listed with a ballot
		const bool leaf = ref != RT_REF_NONE && ref < 0;
		const int first = leaf ? leaf_first(ref) : 0, count = leaf ? leaf_count(ref) : 0;
#pragma unroll 1
		for (int i = 0; i < 3; i++) {
			bool is_mesh = false;
			double eb = 0.0;
			const int rank = first + i;
			if (i < count && !box_miss_f(S.sboxf[rank], F)) {
				const int obj = S.sboxf[rank].obj;
				const rt_object& ob = S.objs[obj];
				if (box_hit(ob.bbox, ob.bbox + 3, ow, dw, I, eb) && !(eb > lim)) {
					if (ob.type == RT_OBJ_TRIMESH) is_mesh = true;
					else {
						LocalRay L;
						to_local(L, ob, ow, dw);
						prim_closest(ob, L, rank, obj, mine);
					}
				}
			}
			const unsigned mm = __ballot_sync(0xffffffffu, is_mesh);
			if (mm) {
				if (n_list + __popc(mm) > RT_COOP_MESH_LIST) { // list full: walk what is listed, then go on
					__syncwarp();
					coop_mesh_list(S, ow, dw, A, n_list, C, K);
					n_list = 0;
				}
				if (is_mesh) { const int at = n_list + __popc(mm & lt_mask); A.mesh_rank[at] = rank; A.mesh_entry[at] = eb; }
				n_list += __popc(mm);
				__syncwarp();
			}
		}
		const unsigned b1 = __ballot_sync(0xffffffffu, nc >= 1), b2 = __ballot_sync(0xffffffffu, nc == 2);
		const int at = n + __popc(b1 & hi_mask) + __popc(b2 & hi_mask);
		if (nc >= 1) { A.ref[at] = c0; A.ent[at] = ce0; }
		if (nc == 2) { A.ref[at + 1] = c1; A.ent[at + 1] = ce1; }
		n += __popc(b1) + __popc(b2);
		__syncwarp();
	}
	// the lanes' analytic winners: argmin (t, rank)
#pragma unroll
	for (int o2 = 16; o2 > 0; o2 >>= 1) {
		const double ot = __shfl_xor_sync(0xffffffffu, mine.best_t, o2);
		const int orank = __shfl_xor_sync(0xffffffffu, mine.best_rank, o2);
		const int oobj = __shfl_xor_sync(0xffffffffu, mine.best_obj, o2);
		const bool oh = __shfl_xor_sync(0xffffffffu, mine.have ? 1 : 0, o2) != 0;
		if (oh && (!mine.have || ot < mine.best_t || (ot == mine.best_t && orank < mine.best_rank))) {
			mine.have = true;
			mine.best_t = ot;
			mine.best_rank = orank;
			mine.best_obj = oobj;
		}
	}
	if (mine.have && (!C.have || mine.best_t < C.best_t || (mine.best_t == C.best_t && mine.best_rank < C.best_rank))) {
		C.have = true;
		C.best_t = mine.best_t;
		C.best_rank = mine.best_rank;
		C.best_obj = mine.best_obj;
		C.best_fpos = -1;
	}
	coop_mesh_list(S, ow, dw, A, n_list, C, K);
}

} // namespace rtd
