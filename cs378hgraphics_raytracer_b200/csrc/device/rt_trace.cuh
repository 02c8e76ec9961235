// rt_trace.cuh - closest-hit query: Scene::intersect re-designed for one GPU thread per ray.
//
// Reference semantics (ray/src/scene/scene.cpp:157-180, kdTree.h:100-117, scene.cpp:13-38,
// SceneObjects/trimesh.cpp:79-95,119-188): collect EVERY leaf item under boxes the ray touches in
// DFS order, test them all, keep the first candidate with a strictly smaller t.  That winner is
// argmin over candidates of (t, candidate rank).  Here the trees are walked front to back with a
// short per-thread stack and subtrees that provably cannot beat the current best are skipped; ties
// in t are broken by rank so the winner is the reference's.  A trimesh is one scene-level candidate
// whose own winner is decided in object space by (t_local, leaf item position), as in the
// reference's two-stage comparison.  With PRUNE=false the walk is the reference's exhaustive one;
// tests compare both.
#pragma once
#include "rt_prims.cuh"

namespace rtd {

#define RT_STACK 64
#define RT_T_INF 1.0e308

struct Hit {
	int32_t obj;  // index in Scene::objects, -1 = miss
	int32_t face; // index in Trimesh::faces, -1 for analytic primitives
	int32_t fpos; // global mesh item position of the face (tris[] index)
	int32_t boxface; // Box: face index 0..5 (for the bump-mapped normal)
	double t;
	d3 N;
	d2 uv;
	d3 bary;
};

struct LocalRay {
	d3 o, d;
	double len;
};

// Geometry::intersect's world -> object ray (reference scene.cpp:17-20)
RT_HD LocalRay to_local(const rt_object& ob, const d3& o, const d3& d)
{
	LocalRay L;
	L.o = xform_point(ob.inv, o);
	d3 dir = xform_point(ob.inv, o + d) - L.o;
	L.len = length(dir);
	L.d = normalize(dir);
	return L;
}

// TrimeshFace::intersectLocal (reference SceneObjects/trimesh.cpp:119-188): plane, three edge
// functions with the BTTC cut-off (edge-exact hits are misses), area check.
template <bool FULL> RT_HD bool tri_hit(const DTri& T, const d3& o, const d3& d, double& t_out, d3& bary)
{
	const d3 n = ld3(T.n), v0 = ld3(T.v0), v1 = ld3(T.v1), v2 = ld3(T.v2);
	double t = dot(n, d);
	if (t < RT_EPS32 && t > -RT_EPS32) return false;
	t = dot(v0 - o, n) / t;
	if (t < RT_EPS32) return false;
	const d3 P = o + t * d;
	if (dot(cross(v1 - v0, P - v0), n) < RT_EPS32) return false;
	if (dot(cross(v2 - v1, P - v1), n) < RT_EPS32) return false;
	if (dot(cross(v0 - v2, P - v2), n) < RT_EPS32) return false;
	const double area = dot(cross(v1 - v0, v2 - v0), n);
	if (area < RT_EPS32 && area > -RT_EPS32) return false;
	t_out = t;
	if (FULL) {
		const double bu = dot(cross(v1 - P, v2 - P), n);
		const double bv = dot(cross(v2 - P, v0 - P), n);
		bary.x = bu / area;
		bary.y = bv / area;
		bary.z = 1 - bary.x - bary.y;
	}
	return true;
}

RT_HD d3 slack_vec(const d3& d, double slack)
{
	return mk3(slack / fabs(d.x), slack / fabs(d.y), slack / fabs(d.z)); // inf for d==0: that axis is skipped
}

// Trimesh::intersectLocal over the per-mesh tree.  cut = largest t_local that could still matter
// to the caller (RT_T_INF when nothing is known).  Returns the winner (t_local, item position).
template <bool PRUNE>
RT_HD bool mesh_closest(const DScene& S, const rt_mesh& M, double slack, const d3& o, const d3& d, double cut,
                        int* stk, double* stke, int sp0, double& best_t, int& best_pos)
{
	const rt_node* nodes = S.mnodes + M.node_off;
	const DTri* tris = S.tris + M.item_off;
	const d3 sl = PRUNE ? slack_vec(d, slack) : mk3(0, 0, 0);
	bool have = false;
	best_t = RT_T_INF;
	best_pos = 0x7fffffff;
	double lim = cut; // prune bound: min(cut, best_t)
	double e;
	if (!box_hit(nodes[0].bmin, nodes[0].bmax, o, d, sl, e)) return false;
	if (PRUNE && e > lim) return false;
	int sp = sp0;
	int node = 0;
	for (;;) {
		const rt_node& n = nodes[node];
		if (n.right < 0) {
			for (int i = 0; i < n.count; i++) {
				const int pos = n.first + i;
				double t;
				d3 unused;
				if (tri_hit<false>(tris[pos], o, d, t, unused)) {
					if (!have || t < best_t || (t == best_t && pos < best_pos)) {
						have = true;
						best_t = t;
						best_pos = pos;
						if (t < lim) lim = t;
					}
				}
			}
		} else {
			const int c0 = node + 1, c1 = n.right;
			double e0, e1;
			bool h0 = box_hit(nodes[c0].bmin, nodes[c0].bmax, o, d, sl, e0);
			bool h1 = box_hit(nodes[c1].bmin, nodes[c1].bmax, o, d, sl, e1);
			if (PRUNE) { h0 = h0 && !(e0 > lim); h1 = h1 && !(e1 > lim); }
			if (h0 && h1) {
				const bool swap = PRUNE && e1 < e0;
				stk[sp] = swap ? c0 : c1;
				stke[sp] = swap ? e0 : e1;
				sp++;
				node = swap ? c1 : c0;
				continue;
			}
			if (h0) { node = c0; continue; }
			if (h1) { node = c1; continue; }
		}
		// pop
		for (;;) {
			if (sp == sp0) return have;
			sp--;
			node = stk[sp];
			if (!PRUNE || !(stke[sp] > lim)) break;
		}
	}
}

// One scene-level candidate: Geometry::intersect = world-bbox pretest (scene.cpp:15, a real
// filter: leaf boxes are unions) + object-space primitive test; returns t in world units.
template <bool PRUNE>
RT_HD bool object_closest(const DScene& S, int obj, const d3& o, const d3& d, const d3& sl, double lim,
                          int* stk, double* stke, int sp, double& t_world, int& fpos)
{
	const rt_object& ob = S.objs[obj];
	double e;
	if (!box_hit(ob.bbox, ob.bbox + 3, o, d, sl, e)) return false;
	if (PRUNE && e > lim) return false;
	const LocalRay L = to_local(ob, o, d);
	LocalHit h;
	int bf;
	bool hit = false;
	fpos = -1;
	switch (ob.type) {
	case RT_OBJ_SPHERE: hit = sphere_hit<false>(L.o, L.d, h); break;
	case RT_OBJ_BOX: hit = box_prim_hit<false>(L.o, L.d, h, bf); break;
	case RT_OBJ_SQUARE: hit = square_hit<false>(L.o, L.d, h); break;
	case RT_OBJ_CYLINDER: hit = cylinder_hit<false>(L.o, L.d, h); break;
	case RT_OBJ_CONE: hit = cone_hit<false>(L.o, L.d, ob.cone, ob.capped, h); break;
	case RT_OBJ_TRIMESH: {
		const rt_mesh& M = S.meshes[ob.mesh];
		// anything with t_world >= lim loses at the scene level: convert to object space, inflated
		const double cut = (PRUNE && lim < RT_T_INF) ? (lim * L.len) * (1.0 + 1e-9) + 1e-300 : RT_T_INF;
		int pos;
		hit = mesh_closest<PRUNE>(S, M, S.mesh_slack[ob.mesh], L.o, L.d, cut, stk, stke, sp, h.t, pos);
		fpos = (int)M.item_off + pos;
		break;
	}
	default: break;
	}
	if (!hit) return false;
	t_world = h.t / L.len;
	return true;
}

// Scene::intersect.  On return out.obj/out.fpos/out.t identify the winner; finalize_hit fills the rest.
template <bool PRUNE> RT_HD bool closest_hit(const DScene& S, const d3& o, const d3& d, Hit& out)
{
	int stk[RT_STACK];
	double stke[RT_STACK];
	const rt_node* nodes = S.snodes;
	const d3 sl = PRUNE ? slack_vec(d, S.scene_slack) : mk3(0, 0, 0);
	bool have = false;
	double best_t = RT_T_INF;
	int best_rank = 0x7fffffff, best_obj = -1, best_fpos = -1;
	out.obj = -1;
	out.face = -1;
	out.fpos = -1;
	out.t = 1000.0; // scene.cpp:174-175
	double e;
	if (!box_hit(nodes[0].bmin, nodes[0].bmax, o, d, sl, e)) return false;
	int sp = 0, node = 0;
	for (;;) {
		const rt_node& n = nodes[node];
		bool descend = false;
		if (n.right < 0) {
			for (int i = 0; i < n.count; i++) {
				const int rank = n.first + i;
				const int obj = S.sitems[rank];
				double t;
				int fpos;
				if (object_closest<PRUNE>(S, obj, o, d, sl, best_t, stk, stke, sp, t, fpos)) {
					if (!have || t < best_t || (t == best_t && rank < best_rank)) {
						have = true;
						best_t = t;
						best_rank = rank;
						best_obj = obj;
						best_fpos = fpos;
					}
				}
			}
		} else {
			const int c0 = node + 1, c1 = n.right;
			double e0, e1;
			bool h0 = box_hit(nodes[c0].bmin, nodes[c0].bmax, o, d, sl, e0);
			bool h1 = box_hit(nodes[c1].bmin, nodes[c1].bmax, o, d, sl, e1);
			if (PRUNE) { h0 = h0 && !(e0 > best_t); h1 = h1 && !(e1 > best_t); }
			if (h0 && h1) {
				const bool swap = PRUNE && e1 < e0;
				stk[sp] = swap ? c0 : c1;
				stke[sp] = swap ? e0 : e1;
				sp++;
				node = swap ? c1 : c0;
				descend = true;
			} else if (h0) { node = c0; descend = true; }
			else if (h1) { node = c1; descend = true; }
		}
		if (descend) continue;
		bool done = false;
		for (;;) {
			if (sp == 0) { done = true; break; }
			sp--;
			node = stk[sp];
			if (!PRUNE || !(stke[sp] > best_t)) break;
		}
		if (done) break;
	}
	if (!have) return false;
	out.obj = best_obj;
	out.fpos = best_fpos;
	out.t = best_t;
	return true;
}

} // namespace rtd
