// rt_trace.cuh - hit record, world->object ray transform and the triangle test.  The tree walk
// itself (closest-hit and all-hits in one engine) is rt_traverse.cuh.
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
RT_COLD void to_local(LocalRay& L, const rt_object& ob, const d3& o, const d3& d)
{
	L.o = xform_point(ob.inv, o);
	d3 dir = xform_point(ob.inv, o + d) - L.o;
	L.len = length(dir);
	L.d = normalize(dir);
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

} // namespace rtd
