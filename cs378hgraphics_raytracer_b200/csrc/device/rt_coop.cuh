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
#define RT_COOP_WARPS 2        // warps per block of the heavy-ray kernel

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

// Scene::intersect (scene.cpp:157-180) for one ray by one warp; all lanes hold the same ray and
// return the same C.
__device__ __forceinline__ void coop_closest(const DScene& S, const d3& ow, const d3& dw, TravClosest& C, CoopStack& K)
{
	int stk[RT_STACK];
	double stke[RT_STACK];
	RayInv I;
	RayF F;
	make_inv(I, F, ow, dw, S.scene_slack, S.scene_slack * 1e9);
	double lim = RT_T_INF;
	double e;
	if (!box_hit(S.stree.lo, S.stree.hi, ow, dw, I, e)) return;
	int sp = 0;
	int cur = S.stree.root;
	for (;;) {
#pragma unroll 1
		while (cur >= 0) {
			const DNode2f& n = S.snodes2f[cur];
			double e0, e1;
			bool h0, h1;
			box2_hit(n, S.snodes2[cur], ow, dw, I, F, h0, h1, e0, e1);
			h0 = h0 && !(e0 > lim);
			h1 = h1 && !(e1 > lim);
			const int c0 = n.c0, c1 = n.c1;
			if (h0 && h1) {
				const bool swap = e1 < e0;
				stk[sp] = swap ? c0 : c1;
				stke[sp] = swap ? e0 : e1;
				sp++;
				cur = swap ? c1 : c0;
			} else if (h0) cur = c0;
			else if (h1) cur = c1;
			else cur = RT_REF_NONE;
		}
		if (cur != RT_REF_NONE) {
			const int first = leaf_first(cur), count = leaf_count(cur);
#pragma unroll 1
			for (int i = 0; i < count; i++) {
				const int rank = first + i;
				const int obj = S.sitems[rank];
				const rt_object& ob = S.objs[obj];
				double eb;
				if (!box_hit(ob.bbox, ob.bbox + 3, ow, dw, I, eb)) continue;
				if (eb > lim) continue;
				LocalRay L;
				to_local(L, ob, ow, dw);
				if (ob.type == RT_OBJ_TRIMESH) {
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
				} else {
					prim_closest(ob, L, rank, obj, C);
				}
				if (C.have) lim = C.best_t;
			}
		}
		for (;;) {
			if (sp == 0) return;
			sp--;
			if (stke[sp] > lim) continue;
			cur = stk[sp];
			break;
		}
	}
}

} // namespace rtd
