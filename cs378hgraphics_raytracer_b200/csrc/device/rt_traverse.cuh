// rt_traverse.cuh - ONE tree-walk engine for both queries of the trace loop:
//   closest hit  = Scene::intersect       (reference ray/src/scene/scene.cpp:157-180)
//   all hits     = Scene::intersectList   (scene.cpp:182-197, feeding Light::srsAttenuation)
// over the scene-level tree and, nested, the per-mesh trees (kdTree.h:100-117, scene.cpp:13-64,
// SceneObjects/trimesh.cpp:79-105).
//
// The reference collects EVERY leaf item under boxes the ray touches, in DFS order, tests them all
// and (closest hit) keeps the first candidate with a strictly smaller t, i.e. argmin of (t, rank).
// Here each GPU thread walks front to back with a short stack, skips subtrees whose conservative
// entry distance exceeds `lim` (the best t so far, or the shadow cut), and breaks ties in t by
// candidate rank, so the winner is the reference's; box membership itself is decided by the
// reference's exact slab arithmetic.  A trimesh is one scene-level candidate whose own winner is
// chosen in object space by (t_local, leaf item position) - the reference's two-stage comparison.
//
// One loop serves both tree levels and both query kinds: a warp whose lanes hold a mix of camera,
// secondary and shadow rays still executes the same box-test code together.  The loop has three
// re-converging sections (while-while form): descend internal nodes / process one leaf / pop.
#pragma once
#include "rt_shadow.cuh"
#if defined(RT_COUNT_WORK) && !defined(__CUDACC__)
#include <cstdio>
#include <cstdlib>
#endif

namespace rtd {

enum { TRAV_CLOSEST = 0, TRAV_SHADOW = 1,
       TRAV_ANYHIT = 2 }; // shadow query answered by classification only (AnyHit below): no hit list, no kt, no sorting

// Work counters for the cost model (scenes/cost_model.py): host build of tests/hostsim only.
#if defined(RT_COUNT_WORK) && !defined(__CUDACC__)
struct TravWork {
	unsigned long long walks[2], nodes[2], tris[2], objs[2]; // per query kind, summed
	unsigned long long hist[2][32];                           // walks by floor(log2(nodes + 1))
	unsigned long long max_nodes[2];
};
inline thread_local TravWork g_work;
inline unsigned long long g_work_dump = ~0ull; // print rays whose walk visits more nodes than this
inline thread_local unsigned* g_walk_log = nullptr; // when set: (mode, nodes) of every walk, g_walk_log[0] = entries used, g_walk_log[1] = capacity
struct WorkScope {
	int mode;
	unsigned long long n0, t0;
	const d3 &o, &d;
	WorkScope(int m, const d3& o_, const d3& d_) : mode(m), n0(g_work.nodes[m]), t0(g_work.tris[m]), o(o_), d(d_) { g_work.walks[m]++; }
	~WorkScope()
	{
		const unsigned long long n = g_work.nodes[mode] - n0;
		if (g_walk_log && g_walk_log[0] < g_walk_log[1]) { g_walk_log[2 + 2 * g_walk_log[0]] = (unsigned)mode; g_walk_log[3 + 2 * g_walk_log[0]] = (unsigned)n; g_walk_log[0]++; }
		if (n > g_work_dump) fprintf(stderr, "walk mode %d nodes %llu tris %llu o %.17g %.17g %.17g d %.17g %.17g %.17g\n", mode, n, g_work.tris[mode] - t0, o.x, o.y, o.z, d.x, d.y, d.z);
		int b = 0;
		while ((n + 1) >> (b + 1)) b++;
		g_work.hist[mode][b]++;
		if (n > g_work.max_nodes[mode]) g_work.max_nodes[mode] = n;
	}
};
#define RT_CNT(f, m) (g_work.f[m]++)
#define RT_WORK_SCOPE(m) WorkScope work_scope_(m, ow, dw)
#else
#define RT_CNT(f, m) ((void)0)
#define RT_WORK_SCOPE(m) ((void)0)
#endif

// 1-based index of the lowest set bit
RT_HD int rt_ffs(unsigned v)
{
#ifdef __CUDA_ARCH__
	return __ffs((int)v);
#else
	return __builtin_ffs((int)v);
#endif
}

struct TravClosest {
	bool have;
	double best_t;
	int best_rank, best_obj, best_fpos;
};

// world-space limit -> object space, inflated so rounding in the conversion can never cull a winner
RT_HD double lim_to_local(double lim, double len) { return lim < RT_T_INF ? (lim * len) * (1.0 + 1e-9) + 1e-300 : RT_T_INF; }

// TRAV_ANYHIT: Light::srsAttenuation (light.cpp:21-53) decided without its hit list wherever that is possible.
//   SH_DEAD : some entered non-transmissive surface lies at t <= sure_t.  srsAttenuation is then exactly 0 whatever else the
//             ray meets (the any-hit argument of rt_shadow.cuh: sbuf_add) - the walk stops at once.
//   SH_OTHER: some other hit lies within the light's distance (a transmissive surface, a surface in the far half of the
//             segment, an exit from an object the origin is inside of): the ray needs the complete sorted walk (TRAV_SHADOW).
//   neither : nothing is hit before the light: the attenuation is exactly (1,1,1).
// A second exact shortcut, for rays that start just INSIDE an opaque object (the shadow ray's origin is backed up along the view
// ray, light.cpp:16-20, and may land under the surface it left): if the NEAREST hit of all is an exit (N.d > 0) from a medium
// whose kt is exactly (0,0,0), at t <= sure_t, the sorted walk multiplies the attenuation by pow(0, t - 0) = 0 at its first step
// (light.cpp:50; the limit check before it cannot fire at t <= 0.49 D) and the result is exactly 0.  "Nearest" must be strict -
// a second hit at the same t would be processed with delta 0 somewhere - so the walk tracks t_x, the nearest such exit, and t_o,
// the nearest hit of any other kind; beyond t_x nothing needs to be looked at any more (the pruning limit drops to t_x).
enum { SH_DEAD = 1, SH_OTHER = 2,
       SH_NEWLIM = 4 }; // t_x has dropped since the walk last looked: its pruning limit follows
struct AnyHit {
	double limit;  // hits beyond this cannot change the result (past a point light); the pruning limit of the walk
	double sure_t; // 0.49 D for a point light, infinity for a directional one
	double t_x;    // nearest exit from a kt = 0 medium at t <= sure_t
	double t_o;    // nearest hit of any other kind
	unsigned fl;
	RT_HD double prune() const { return t_x < limit ? t_x : limit; }
	RT_HD bool settled_by_exit() const { return t_x < t_o; } // (t_x finite: some such exit exists and is strictly nearest)
};
RT_HD void anyhit_classify(AnyHit& Y, double t, bool inside, bool trans, bool kt_zero)
{
	if (!(t <= Y.limit)) return;
	if (!inside && !trans && t <= Y.sure_t) { Y.fl |= SH_DEAD; return; }
	Y.fl |= SH_OTHER;
	if (inside && kt_zero && t <= Y.sure_t) { if (t < Y.t_x) { Y.t_x = t; Y.fl |= SH_NEWLIM; } }
	else if (t < Y.t_o) Y.t_o = t;
}
RT_HD bool kt_is_zero(const double* k) { return k[0] == 0.0 && k[1] == 0.0 && k[2] == 0.0; }
// the hits of an analytic primitive (object space -> world), classified
struct AnyEmit {
	const DScene& S;
	const rt_object& ob;
	const d3& dw;
	double len;
	AnyHit& Y;
	RT_HD void operator()(double tl, d3 Nl, d2 uv, int) const
	{
		if (Y.fl & SH_DEAD) return; // settled by an earlier hit of this primitive (its entry): the exit does not matter any more
		const double t = tl / len;
		if (!(t <= Y.limit)) return;
		const d3 N = world_normal(S, ob, Nl, uv);
		const rt_material& m = S.mats[ob.material];
		const bool inside = dot(N, dw) > 0;
		// kt lies on another cache line of the material record than the flags: read it for exits only
		anyhit_classify(Y, t, inside, m.trans != 0, inside && m.tex[RT_KT] < 0 && kt_is_zero(m.k[RT_KT]));
	}
};
RT_COLD void prim_anyhit(const DScene& S, const rt_object& ob, const LocalRay& L, const d3& dw, AnyHit& Y)
{
	const AnyEmit emit{S, ob, dw, L.len, Y};
	switch (ob.type) {
	case RT_OBJ_SPHERE: sphere_list(L.o, L.d, emit); break;
	case RT_OBJ_BOX: box_list(L.o, L.d, emit); break;
	case RT_OBJ_SQUARE: {
		LocalHit h;
		h.uv.x = h.uv.y = 0.0;
		if (square_hit<true>(L.o, L.d, h)) emit(h.t, h.N, h.uv, 0);
		break;
	}
	case RT_OBJ_CYLINDER: cylinder_list(L.o, L.d, emit); break;
	case RT_OBJ_CONE: cone_list(L.o, L.d, ob.cone, ob.capped, emit); break;
	default: break;
	}
}
// a mesh face that the short triangle test reported hit: TrimeshFace::intersectLocal's normal (trimesh.cpp:146-174), classified
RT_COLD void face_anyhit(const DScene& S, int obj, int pos, double len, const d3& o, const d3& d, const d3& dw, AnyHit& Y)
{
	const rt_object& ob = S.objs[obj];
	const rt_mesh& M = S.meshes[ob.mesh];
	const int fpos = (int)M.item_off + pos;
	d3 bary;
	double tl;
	tri_hit<true>(S.tris[fpos], o, d, tl, bary);
	const double t = tl / len;
	if (!(t <= Y.limit)) return;
	const rt_face& f = S.faces[M.face_off + S.mitems[fpos]];
	d3 Nl;
	if (M.vnorm_off >= 0) {
		const double* n0 = S.vnormals + 3 * (M.vnorm_off + f.ids[0]);
		const double* n1 = S.vnormals + 3 * (M.vnorm_off + f.ids[1]);
		const double* n2 = S.vnormals + 3 * (M.vnorm_off + f.ids[2]);
		Nl = normalize(mk3(n0[0] * bary.x + n1[0] * bary.y + n2[0] * bary.z,
		                   n0[1] * bary.x + n1[1] * bary.y + n2[1] * bary.z,
		                   n0[2] * bary.x + n1[2] * bary.y + n2[2] * bary.z));
	} else {
		Nl = ld3(f.normal);
	}
	const d3 N = normalize(mul33(ob.normi, Nl));
	// eval_kt's `trans`: false for a mesh with per-vertex materials (the blended Material's flags are never set), else the object's
	const bool inside = dot(N, dw) > 0;
	bool trans = false, kt_zero = false;
	if (!inside) {
		trans = M.vmat_off >= 0 ? false : S.mats[ob.material].trans != 0;
	} else if (M.vmat_off >= 0) { // kt = barycentric blend of the three vertex materials (eval_kt): zero iff the blend is
		const rt_material& a = S.mats[S.vmats[M.vmat_off + f.ids[0]]];
		const rt_material& b = S.mats[S.vmats[M.vmat_off + f.ids[1]]];
		const rt_material& c = S.mats[S.vmats[M.vmat_off + f.ids[2]]];
		const d3 kt = blend3(a.k[RT_KT], b.k[RT_KT], c.k[RT_KT], bary);
		kt_zero = kt.x == 0.0 && kt.y == 0.0 && kt.z == 0.0;
	} else {
		const rt_material& m = S.mats[ob.material];
		trans = m.trans != 0;
		kt_zero = m.tex[RT_KT] < 0 && kt_is_zero(m.k[RT_KT]);
	}
	anyhit_classify(Y, t, inside, trans, kt_zero);
}

// emit one all-hits record of an analytic primitive (object space -> world)
struct ListEmit {
	const DScene& S;
	const rt_object& ob;
	const d3& dw;
	double len;
	unsigned rank;
	SBuf& B;
	RT_HD void operator()(double tl, d3 Nl, d2 uv, int sub) const
	{
		const double t = tl / len;
		if (!(t <= B.cut)) return;
		const rt_material& m = S.mats[ob.material];
		const d3 N = world_normal(S, ob, Nl, uv);
		sbuf_add(B, t, ((unsigned long long)rank << 32) | (unsigned)sub, dot(N, dw) > 0, m.trans != 0, mat_param(S, m, RT_KT, uv));
	}
};

// Scene-level candidate that is an analytic primitive, closest-hit query: the object-space test
// and the (t, rank) comparison of Scene::intersect's loop (scene.cpp:163-172).
RT_COLD void prim_closest(const rt_object& ob, const LocalRay& L, int rank, int obj, TravClosest& C)
{
	LocalHit h;
	int bf;
	bool hit = false;
	switch (ob.type) {
	case RT_OBJ_SPHERE: hit = sphere_hit<false>(L.o, L.d, h); break;
	case RT_OBJ_BOX: hit = box_prim_hit<false>(L.o, L.d, h, bf); break;
	case RT_OBJ_SQUARE: hit = square_hit<false>(L.o, L.d, h); break;
	case RT_OBJ_CYLINDER: hit = cylinder_hit<false>(L.o, L.d, h); break;
	case RT_OBJ_CONE: hit = cone_hit<false>(L.o, L.d, ob.cone, ob.capped, h); break;
	default: break;
	}
	if (!hit) return;
	const double t = h.t / L.len;
	if (!C.have || t < C.best_t || (t == C.best_t && rank < C.best_rank)) {
		C.have = true;
		C.best_t = t;
		C.best_rank = rank;
		C.best_obj = obj;
		C.best_fpos = -1;
	}
}

// The same candidate for the all-hits query (Geometry::intersectList, scene.cpp:41-64)
RT_COLD void prim_list(const DScene& S, const rt_object& ob, const LocalRay& L, const d3& dw, int rank, SBuf& B)
{
	const ListEmit emit{S, ob, dw, L.len, (unsigned)rank, B};
	switch (ob.type) {
	case RT_OBJ_SPHERE: sphere_list(L.o, L.d, emit); break;
	case RT_OBJ_BOX: box_list(L.o, L.d, emit); break;
	case RT_OBJ_SQUARE: {
		LocalHit h;
		h.uv.x = h.uv.y = 0.0;
		if (square_hit<true>(L.o, L.d, h)) emit(h.t, h.N, h.uv, 0);
		break;
	}
	case RT_OBJ_CYLINDER: cylinder_list(L.o, L.d, emit); break;
	case RT_OBJ_CONE: cone_list(L.o, L.d, ob.cone, ob.capped, emit); break;
	default: break;
	}
}

// all-hits record of a mesh face (TrimeshFace::intersectLocal's isect, trimesh.cpp:146-174).
// The walk found the hit with the short form of the triangle test; the full form is re-run here
// for the barycentrics (same arithmetic, same t).
RT_COLD void emit_face(const DScene& S, int obj, unsigned rank, int pos, double len, const d3& o, const d3& d, const d3& dw, SBuf& B)
{
	const rt_object& ob = S.objs[obj];
	const rt_mesh& M = S.meshes[ob.mesh];
	Hit h;
	h.obj = obj;
	h.fpos = (int)M.item_off + pos;
	h.face = S.mitems[h.fpos];
	h.uv.x = h.uv.y = 0.0;
	double tl;
	tri_hit<true>(S.tris[h.fpos], o, d, tl, h.bary);
	const double t = tl / len;
	if (!(t <= B.cut)) return;
	const rt_face& f = S.faces[M.face_off + h.face];
	d3 Nl;
	if (M.vnorm_off >= 0) {
		const double* n0 = S.vnormals + 3 * (M.vnorm_off + f.ids[0]);
		const double* n1 = S.vnormals + 3 * (M.vnorm_off + f.ids[1]);
		const double* n2 = S.vnormals + 3 * (M.vnorm_off + f.ids[2]);
		Nl = normalize(mk3(n0[0] * h.bary.x + n1[0] * h.bary.y + n2[0] * h.bary.z,
		                   n0[1] * h.bary.x + n1[1] * h.bary.y + n2[1] * h.bary.z,
		                   n0[2] * h.bary.x + n1[2] * h.bary.y + n2[2] * h.bary.z));
	} else {
		Nl = ld3(f.normal);
	}
	const d3 N = normalize(mul33(ob.normi, Nl));
	d3 kt;
	bool trans;
	eval_kt(S, h, kt, trans);
	sbuf_add(B, t, ((unsigned long long)rank << 32) | (unsigned)pos, dot(N, dw) > 0, trans, kt);
}

// State of a walk that is inside a mesh (level 1).
struct MeshVisit {
	double len;      // |M^-1 d|: t_world = t_local / len
	int obj, rank;   // the scene-level candidate being visited
	int item_off;    // first leaf item of this mesh in tris[] / mitems[]
	bool have;       // Trimesh::intersectLocal's running winner (object space)
	double best_t;
	int best_pos;
};

// ---- the walk, split in two -------------------------------------------------------------------
// HOT: the node loop.  Everything it needs is single precision and lives in registers: the ray's float
// image (RayF), the float node array, the limit rounded UP to float, the stack of (reference, float
// entry distance) pairs - one 8-byte store per push.  No double-precision instruction and no local
// memory access other than the stack itself.
// COLD: everything that needs the reference's FP64 arithmetic - ambiguous boxes, leaf candidates,
// entering and leaving a mesh - is out of line and works on `Walk`, the double-precision side of the
// ray, which lives in local memory and is touched at leaves only (a walk visits ~10 nodes per leaf).
struct StackEnt {
	int ref;
	float e;
};
struct Walk {
	d3 ow, dw;           // the ray, world space
	d3 o, d;             // the ray in the current space (world at level 0, the mesh's object space at level 1)
	RayInv I;            // its double-precision inverse
	double lim;          // best t so far / shadow cut, current space
	MeshVisit V;
	const DNode2* nodes; // double images of the current tree's nodes
	const DTri* tris;    // the current mesh's triangles
	TravClosest* C;
	SBuf* B;
	AnyHit* Y;
	int mode;
	bool prune;
};

// set-up: inverse, float image, root box.  Returns the first reference to visit (RT_REF_NONE: the
// ray misses the scene's box)
RT_COLD int walk_begin(const DScene& S, Walk& W, RayF& F, float& limf)
{
	W.o = W.ow;
	W.d = W.dw;
	make_inv(W.I, F, W.o, W.d, W.prune ? S.scene_slack : 0.0, S.scene_slack * 1e9);
	W.nodes = S.snodes2;
	W.tris = S.tris;
	W.lim = W.mode == TRAV_CLOSEST ? RT_T_INF : (W.mode == TRAV_SHADOW ? W.B->cut : W.Y->limit);
	limf = f_up(W.lim);
	W.V.len = 1.0;
	W.V.obj = -1;
	W.V.rank = 0;
	W.V.item_off = 0;
	W.V.have = false;
	W.V.best_t = RT_T_INF;
	W.V.best_pos = 0x7fffffff;
	double e;
	if (!box_hit(S.stree.lo, S.stree.hi, W.o, W.d, W.I, e)) return RT_REF_NONE;
	return S.stree.root;
}

// a child of node `cur` whose float test was ambiguous: the reference's decision from the double box.
// Returned by value ({hit, entry distance}) so that the node loop's variables never have their address taken.
RT_COLD StackEnt walk_resolve(const Walk& W, int cur, int child)
{
	const DNode2& n = W.nodes[cur];
	StackEnt r;
	r.ref = boxf_resolve(child ? n.lo1 : n.lo0, child ? n.hi1 : n.hi0, W.o, W.d, W.I, r.e);
	return r;
}

// one 64-byte float node through the read-only path, four 128-bit loads
RT_HD DNode2f ld_node2f(const DNode2f* p)
{
#if defined(__CUDA_ARCH__) && !defined(RT_NODE_PLAIN_LOAD)
	const float4* q = reinterpret_cast<const float4*>(p);
	const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
	DNode2f n;
	n.lo0[0] = a.x; n.lo0[1] = a.y; n.lo0[2] = a.z; n.hi0[0] = a.w;
	n.hi0[1] = b.x; n.hi0[2] = b.y; n.lo1[0] = b.z; n.lo1[1] = b.w;
	n.lo1[2] = c.x; n.hi1[0] = c.y; n.hi1[1] = c.z; n.hi1[2] = c.w;
	n.c0 = __float_as_int(d.x); n.c1 = __float_as_int(d.y);
	n.pad[0] = n.pad[1] = 0;
	return n;
#else
	return *p;
#endif
}

// a mesh leaf: its (at most three) triangles
template <int FIXED_MODE> RT_COLD void walk_leaf_tris(const DScene& S, Walk& W, int first, int count, float& limf, int& sp)
{
	const int mode = FIXED_MODE >= 0 ? FIXED_MODE : W.mode;
#pragma unroll 1
	for (int i = 0; i < count; i++) {
		const int pos = first + i;
		double t;
		d3 bary;
		RT_CNT(tris, mode);
		if (!tri_hit<false>(W.tris[pos], W.o, W.d, t, bary)) continue;
		if (mode == TRAV_CLOSEST) {
			MeshVisit& V = W.V;
			if (!V.have || t < V.best_t || (t == V.best_t && pos < V.best_pos)) {
				V.have = true;
				V.best_t = t;
				V.best_pos = pos;
				if (t < W.lim) W.lim = t;
			}
		} else if (mode == TRAV_ANYHIT) {
			face_anyhit(S, W.V.obj, pos, W.V.len, W.o, W.d, W.dw, *W.Y);
			if (W.Y->fl & SH_DEAD) { sp = 0; break; } // settled: leave through the pop section's one way out
			if (W.Y->fl & SH_NEWLIM) { W.Y->fl &= ~SH_NEWLIM; W.lim = lim_to_local(W.Y->prune(), W.V.len); }
		} else {
			emit_face(S, W.V.obj, (unsigned)W.V.rank, pos, W.V.len, W.o, W.d, W.dw, *W.B);
			if (W.B->dead) { sp = 0; break; } // any-hit exit (rt_shadow.cuh): leave through the pop section's one way out
			W.lim = lim_to_local(W.B->cut, W.V.len);
		}
	}
	limf = f_up(W.lim);
}

// A trimesh candidate (item i of the scene leaf [first, first + count)) is ENTERED: the rest of the leaf and an
// exit marker go on the stack and the walk continues in the mesh's object space (L = the ray there).
// Returns the mesh's root reference, or RT_REF_NONE when the ray misses the root box.
RT_COLD int walk_enter_mesh(const DScene& S, Walk& W, const LocalRay& L, int first, int count, int i, RayF& F, const DNode2f*& nodesf, int& level, int& sp, StackEnt* stk)
{
	const int rank = first + i;
	const int obj = S.sboxf[rank].obj;
	const rt_object& ob = S.objs[obj];
	if (i + 1 < count) { stk[sp].ref = leaf_ref(rank + 1, count - i - 1); stk[sp].e = -INFINITY; sp++; }
	stk[sp].ref = RT_REF_MESH_EXIT;
	stk[sp].e = -INFINITY;
	sp++;
	const DTree& T = S.mtrees[ob.mesh];
	MeshVisit& V = W.V;
	V.obj = obj;
	V.rank = rank;
	V.len = L.len;
	V.item_off = (int)S.meshes[ob.mesh].item_off;
	V.have = false;
	V.best_t = RT_T_INF;
	V.best_pos = 0x7fffffff;
	W.tris = S.tris + V.item_off;
	W.nodes = S.mnodes2 + T.node_off;
	nodesf = S.mnodes2f + T.node_off;
	W.o = L.o;
	W.d = L.d;
	make_inv(W.I, F, W.o, W.d, W.prune ? S.mesh_slack[ob.mesh] : 0.0, S.mesh_slack[ob.mesh] * 1e9);
	level = 1;
	W.lim = lim_to_local(W.lim, V.len);
	double er;
	if (box_hit(T.lo, T.hi, W.o, W.d, W.I, er) && !(W.prune && er > W.lim)) return T.root;
	return RT_REF_NONE;
}

// a scene-level leaf: the candidates (bit i of `pend`) that the float test of their boxes could not
// rule out.  Two passes: first the reference's own box test of all of them (Geometry::intersect's
// world-bbox pretest, scene.cpp:15,43 - a real filter, leaf boxes are unions), then the object-space
// tests of the survivors; postponing a test can only weaken the distance cull, never the result.
// A trimesh candidate is ENTERED: the rest of the leaf and an exit marker go on the stack and the
// walk continues in the mesh's object space; the return value is the mesh's root reference then.
template <int FIXED_MODE> RT_COLD int walk_leaf_items(const DScene& S, Walk& W, int first, int count, unsigned pend, RayF& F, const DNode2f*& nodesf, float& limf, int& level, int& sp, StackEnt* stk)
{
	const int mode = FIXED_MODE >= 0 ? FIXED_MODE : W.mode;
	const unsigned sure = pend >> 8; // the float test already proved the double test's "hit" (boxf, result 1)
	pend &= 0xffu;
	unsigned keep = 0;
#pragma unroll 1
	while (pend) {
		const int i = rt_ffs(pend) - 1;
		pend &= pend - 1;
		if (!((sure >> i) & 1u)) {
			const rt_object& ob = S.objs[S.sboxf[first + i].obj];
			double eb;
			if (!box_hit(ob.bbox, ob.bbox + 3, W.o, W.d, W.I, eb)) continue;
			if (W.prune && eb > W.lim) continue;
		}
		keep |= 1u << i;
	}
	int next = RT_REF_NONE;
#pragma unroll 1
	while (keep) {
		const int i = rt_ffs(keep) - 1;
		keep &= keep - 1;
		const int rank = first + i;
		const DItemBox& ib = S.sboxf[rank];
		// object-space float test: a line that certainly misses the shape (rotated boxes and cylinders fill their world
		// boxes badly) never reaches the object record or the double-precision code
		if (ib.type != RT_OBJ_TRIMESH && prim_miss_f(S.sprimf[rank], ib, W.o, W.d)) continue;
		const int obj = ib.obj;
		const rt_object& ob = S.objs[obj];
		LocalRay L;
		to_local(L, ob, W.o, W.d);
		if (ob.type == RT_OBJ_TRIMESH) {
			next = walk_enter_mesh(S, W, L, first, count, i, F, nodesf, level, sp, stk);
			break;
		}
		if (mode == TRAV_CLOSEST) {
			prim_closest(ob, L, rank, obj, *W.C);
			if (W.C->have) W.lim = W.C->best_t;
		} else if (mode == TRAV_ANYHIT) {
			prim_anyhit(S, ob, L, W.dw, *W.Y);
			if (W.Y->fl & SH_DEAD) { sp = 0; break; }
			if (W.Y->fl & SH_NEWLIM) { W.Y->fl &= ~SH_NEWLIM; W.lim = W.Y->prune(); }
		} else {
			prim_list(S, ob, L, W.dw, rank, *W.B);
			if (W.B->dead) { sp = 0; break; } // any-hit exit, as above
			W.lim = W.B->cut;
		}
	}
	limf = f_up(W.lim);
	return next;
}

// the exit marker of a mesh visit: Trimesh::intersectLocal's winner goes back to the scene level as
// ONE candidate, and the walk returns to world space
template <int FIXED_MODE> RT_COLD void walk_mesh_exit(const DScene& S, Walk& W, RayF& F, const DNode2f*& nodesf, float& limf, int& level, int sp)
{
	const int mode = FIXED_MODE >= 0 ? FIXED_MODE : W.mode;
	const MeshVisit& V = W.V;
	TravClosest& C = *W.C;
	if (mode == TRAV_CLOSEST && V.have) {
		const double t = V.best_t / V.len;
		if (!C.have || t < C.best_t || (t == C.best_t && V.rank < C.best_rank)) {
			C.have = true;
			C.best_t = t;
			C.best_rank = V.rank;
			C.best_obj = V.obj;
			C.best_fpos = V.item_off + V.best_pos;
		}
	}
	W.o = W.ow;
	W.d = W.dw;
	// the world-space inverse again (recomputed: a walk rarely returns), unless nothing is left to test with it
	if (sp > 0) make_inv(W.I, F, W.o, W.d, W.prune ? S.scene_slack : 0.0, S.scene_slack * 1e9);
	W.nodes = S.snodes2;
	nodesf = S.snodes2f;
	level = 0;
	W.lim = mode == TRAV_CLOSEST ? (C.have ? C.best_t : RT_T_INF) : (mode == TRAV_SHADOW ? W.B->cut : W.Y->prune());
	limf = f_up(W.lim);
}

// FIXED_MODE >= 0 compiles the walk for one query kind only (the wavefront's stage kernels).
// budget > 0: give up once that many tree nodes have been visited (noticed at the next stack pop) and return false - the
// caller hands the ray to the warp-cooperative walk (rt_coop.cuh).  A few rays visit thousands of nodes (a zero direction
// component makes the reference's slab test accept every box along that axis, bbox.cc:43) and a
// warp waits for its slowest lane; the result of an abandoned walk is not used.
// All ways out of the loop are in the pop section: a second exit inside the node or leaf loops costs
// the warp its reconvergence points (profiles/r1_variants.txt: 3x slower).
template <bool PRUNE, int FIXED_MODE = -1>
RT_HD bool traverse(const DScene& S, const d3& ow, const d3& dw, const int mode_arg, TravClosest& C, SBuf& B, int budget = 0, const DNode2f* top = nullptr, AnyHit* Y = nullptr)
{
	const int mode = FIXED_MODE >= 0 ? FIXED_MODE : mode_arg;
	RT_WORK_SCOPE(mode);
	const bool prune = PRUNE || mode != TRAV_CLOSEST;
	Walk W;
	W.ow = ow;
	W.dw = dw;
	W.C = &C;
	W.B = &B;
	W.Y = Y;
	W.mode = mode;
	W.prune = prune;
	StackEnt stk[RT_STACK];
	RayF F;
	float limf;
	const DNode2f* nodesf = S.snodes2f;
	int level = 0;
	int sp = 0;
	int cur = walk_begin(S, W, F, limf);
	if (cur == RT_REF_NONE) return true;
	if (RT_TOP_K > 0 && top && cur >= 0) cur = RT_REF_TOP; // the root is entry 0 of the staged table
	bool done = true;
	const bool limited = budget > 0;
	for (;;) {
		// ---- 1. internal nodes: test both children, go near, push far ---------------------------
#pragma unroll 1
		while (cur >= 0) {
			const bool staged = RT_TOP_K > 0 && (cur & RT_REF_TOP) != 0;
			const DNode2f nf = staged ? top[cur & (RT_REF_TOP - 1)] : ld_node2f(nodesf + cur);
			RT_CNT(nodes, mode);
			budget--;
			float e0, e1;
			int r0 = boxf(nf.lo0, nf.hi0, F, e0);
			int r1 = boxf(nf.lo1, nf.hi1, F, e1);
			if (r0 == 2) { const StackEnt r = walk_resolve(W, staged ? nf.pad[0] : cur, 0); r0 = r.ref; e0 = r.e; }
			if (r1 == 2) { const StackEnt r = walk_resolve(W, staged ? nf.pad[0] : cur, 1); r1 = r.ref; e1 = r.e; }
			bool h0 = r0 != 0, h1 = r1 != 0;
			if (prune) { h0 = h0 && !(e0 > limf); h1 = h1 && !(e1 > limf); }
			if (h0 && h1) {
				const bool swap = prune && e1 < e0;
				stk[sp].ref = swap ? nf.c0 : nf.c1;
				stk[sp].e = swap ? e0 : e1;
				sp++;
				cur = swap ? nf.c1 : nf.c0;
			} else cur = h0 ? nf.c0 : (h1 ? nf.c1 : RT_REF_NONE);
		}
		// ---- 2. a leaf --------------------------------------------------------------------------
		if (cur != RT_REF_NONE) {
			const int first = leaf_first(cur), count = leaf_count(cur);
			cur = RT_REF_NONE;
			if (level == 1) {
				walk_leaf_tris<FIXED_MODE>(S, W, first, count, limf, sp);
			} else {
				// float images of the candidates' boxes first (32 bytes each): what is certainly missed, or
				// certainly entered beyond the limit, never reaches the double-precision code
				// (bit i: candidate i is pending; bit 8 + i: its box is CERTAINLY hit, the double test of the box is not needed)
				unsigned pend = 0;
#pragma unroll 1
				for (int i = 0; i < count; i++) {
					RT_CNT(objs, mode);
					const DItemBox& ib = S.sboxf[first + i];
					float e;
					const int r = boxf(ib.lo, ib.hi, F, e);
					if (r == 0 || (r == 1 && prune && e > limf)) continue;
					pend |= (r == 1 ? 0x101u : 1u) << i;
				}
				if (pend) cur = walk_leaf_items<FIXED_MODE>(S, W, first, count, pend, F, nodesf, limf, level, sp, stk);
			}
		}
		// ---- 3. pop -----------------------------------------------------------------------------
		if (cur == RT_REF_NONE) {
#pragma unroll 1
			for (;;) {
				if (sp == 0) break;
				if (limited && budget <= 0) { done = false; break; }
				sp--;
				const StackEnt a = stk[sp];
				if (a.ref == RT_REF_MESH_EXIT) {
					walk_mesh_exit<FIXED_MODE>(S, W, F, nodesf, limf, level, sp);
					continue;
				}
				if (prune && a.e > limf) continue;
				cur = a.ref;
				break;
			}
			if (cur == RT_REF_NONE) break;
		}
	}
	return done;
}

// Scene::intersect.  On return out.obj/out.fpos/out.t identify the winner; finalize_hit fills the rest.
template <bool PRUNE> RT_HD bool closest_hit(const DScene& S, const d3& o, const d3& d, Hit& out)
{
	TravClosest C;
	C.have = false;
	C.best_t = RT_T_INF;
	C.best_rank = 0x7fffffff;
	C.best_obj = -1;
	C.best_fpos = -1;
	SBuf unused;
	unused.n = 0;
	unused.cut = RT_T_INF;
	traverse<PRUNE>(S, o, d, TRAV_CLOSEST, C, unused);
	out.obj = -1;
	out.face = -1;
	out.fpos = -1;
	out.t = 1000.0; // scene.cpp:174-175
	if (!C.have) return false;
	out.obj = C.best_obj;
	out.fpos = C.best_fpos;
	out.t = C.best_t;
	return true;
}


// ---- Light::srsAttenuation's walk over the sorted hits (reference scene/light.cpp:28-52) --------
struct ShadowState {
	d3 sattn;    // running attenuation
	d3 cur;      // r2l position, advanced hit by hit (light.cpp:38)
	double last_t;
	double limit; // hits beyond this cannot change the result (past a point light)
	int point;
};

// AreaLightRect::impact / AreaLightCirc::impact (light.cpp:105-111, 138-146): where the ray meets the
// light's plane; the disc variants answer (0,0,0) outside the radius
RT_HD d3 area_impact(const rt_light& Lt, const d3& p, const d3& dir)
{
	const d3 ori = ld3(Lt.orientation), lpos = ld3(Lt.position);
	double t = dot(ori, dir);
	t = dot(lpos - p, ori) / t;
	const d3 at = p + t * dir;
	if (Lt.type == RT_LIGHT_AREA_RECT) return at;
	if (dot(at - lpos, at - lpos) < (Lt.radius * Lt.radius)) return at;
	return mk3(0.0, 0.0, 0.0);
}

RT_COLD void shadow_begin(ShadowState& st, SBuf& B, const rt_light& Lt, const d3& pos)
{
	st.point = Lt.type != RT_LIGHT_DIRECTIONAL;
	st.limit = RT_T_INF;
	// area and spot lights: the limit check is against the ray's impact point on the light's plane
	// (light.cpp:91-99), no distance bound is used for culling and no any-hit exit is taken
	B.sure_t = Lt.type == RT_LIGHT_DIRECTIONAL ? RT_T_INF : -1.0;
	B.dead = false;
	if (Lt.type == RT_LIGHT_POINT) {
		const double D = length(ld3(Lt.position) - pos);
		st.limit = D * (1.0 + 1e-9) + 1e-9;
		B.sure_t = 0.49 * D;
	}
	st.sattn = mk3(1.0, 1.0, 1.0);
	st.last_t = 0.0;
	st.cur = pos;
	B.floor_t = -1.0;
	B.floor_key = 0;
}
RT_HD void shadow_pass_reset(const ShadowState& st, SBuf& B)
{
	B.n = 0;
	B.cut = st.limit;
	B.min_dropped = RT_T_INF;
}
// Consumes one collection pass.  Returns true when the attenuation is final (in st.sattn); false
// when another pass (continuing after B.floor) is needed.
RT_COLD bool shadow_walk(ShadowState& st, SBuf& B, const rt_light& Lt, const d3& dir, const rt_params& prm)
{
	if (B.dead) { st.sattn = mk3(0.0, 0.0, 0.0); return true; }
	const bool overflow = B.min_dropped < RT_T_INF && B.min_dropped <= B.cut;
	const bool aterm = prm.aterm_thresh > 0.0;
	const d3 lpos = ld3(Lt.position);
	for (int k = 0; k < B.n; k++) {
		const SHit& h = B.h[k];
		if (!(h.t <= B.cut)) break;
		const double t = h.t - st.last_t;
		st.last_t = h.t;
		const bool inside = (h.flags & 1) != 0;
		st.cur = st.cur + t * dir;
		// curr/next media: inside ? (m_in -> air) : (air -> m_in); overlapping-objects mode off
		if (Lt.type == RT_LIGHT_POINT) { // PointLight::sattnLimitCheck (light.cpp:65-71)
			if (dot(lpos - (st.cur + t * dir), dir) <= 0) return true;
		} else if (st.point) { // AreaLight::sattnLimitCheck (light.cpp:91-99) with impact() of the Rect / Circ subclass
			if (dot(area_impact(Lt, st.cur, dir) - (st.cur + t * dir), dir) <= 0) return true;
		}
		const bool next_trans = inside ? true : (h.flags & 2) != 0;
		if (!next_trans) { st.sattn = mk3(0.0, 0.0, 0.0); return true; }
		if (aterm && dot(st.sattn, st.sattn) < prm.aterm_thresh * prm.aterm_thresh) { st.sattn = mk3(0.0, 0.0, 0.0); return true; }
		const d3 kt = inside ? h.kt : mk3(1.0, 1.0, 1.0);
		st.sattn = st.sattn * vpow(kt, t);
		// once the attenuation is exactly 0 every later exit of the reference's loop returns 0
		if (st.sattn.x == 0.0 && st.sattn.y == 0.0 && st.sattn.z == 0.0) return true;
		B.floor_t = h.t;
		B.floor_key = h.key;
	}
	if (!overflow && !(B.cut < st.limit)) return true; // the list is exhausted
	if (!overflow) { // walked everything up to a cut that did not end the walk (delta == 0 case)
		B.floor_t = B.cut;
		B.floor_key = ~0ull;
	}
	return false;
}

// Light::srsAttenuation for point and directional lights
RT_HD d3 shadow_attenuation(const DScene& S, const rt_light& Lt, const d3& pos, const d3& dir, const rt_params& prm)
{
	ShadowState st;
	SBuf B;
	TravClosest unused;
	unused.have = false;
	shadow_begin(st, B, Lt, pos);
	for (int pass = 0; pass < 1 << 20; pass++) {
		shadow_pass_reset(st, B);
		traverse<true>(S, pos, dir, TRAV_SHADOW, unused, B);
		if (shadow_walk(st, B, Lt, dir, prm)) break;
	}
	return st.sattn;
}

} // namespace rtd
