// rt_shadow.cuh - Light::shadowAttenuation / srsAttenuation (reference scene/light.cpp:16-53)
// without materialising and sorting every hit along the ray.
//
// Reference: Scene::intersectList gathers ALL hits of ALL candidates (each primitive's list
// variant), std::sort by t, then walks them: per hit k, delta = t_k - t_{k-1}; inside = N.dir > 0;
// the ray origin advances by delta; point lights stop (returning the running attenuation) once
// dot(L - (origin + delta*dir), dir) <= 0; an entered non-transmissive surface returns 0; else the
// attenuation is multiplied by kt_of_the_medium_just_crossed ^ delta.
//
// Device re-design: the walk only ever looks at hits in increasing t and stops at the first
// "terminator", so one traversal pass keeps the K nearest hits beyond a floor key in a sorted
// per-thread buffer and lowers a cut distance whenever a terminator is found (entered opaque
// surface, or an exit from a kt=0 medium); subtrees beyond the cut are skipped.  Hits past a point
// light are irrelevant (end of list and the limit check both return the running attenuation).
// If more than K hits matter, or the walk survives a cut, another pass continues from the last
// processed key.  Ties in t are ordered by candidate rank (what the reference's insertion-sort
// sized std::sort does for short lists).
#pragma once
#include "rt_material.cuh"

namespace rtd {

#define RT_SHADOW_K 8

struct SHit {
	double t;
	unsigned long long key; // candidate order: scene rank << 32 | (mesh item pos or sub-hit index)
	d3 kt;
	int flags; // bit0 inside, bit1 trans, bit2 terminator
};

struct SBuf {
	SHit h[RT_SHADOW_K];
	int n;
	double cut;         // hits beyond this cannot influence the result of this pass
	double floor_t;     // only hits with (t,key) > (floor_t,floor_key) are collected
	unsigned long long floor_key;
	double min_dropped; // smallest t that was dropped for lack of space
};

RT_HD bool key_less(double ta, unsigned long long ka, double tb, unsigned long long kb) { return ta < tb || (ta == tb && ka < kb); }

RT_HD void sbuf_add(SBuf& B, double t, unsigned long long key, bool inside, bool trans, const d3& kt)
{
	if (!(t <= B.cut)) return;
	if (!key_less(B.floor_t, B.floor_key, t, key)) return;
	// a hit that ends the walk whenever it is reached: entering a non-transmissive surface returns
	// 0; leaving a medium with kt = 0 zeroes the attenuation (unless delta is exactly 0, which the
	// walk detects and answers with another pass)
	const bool term = inside ? (kt.x == 0.0 && kt.y == 0.0 && kt.z == 0.0) : !trans;
	int pos = B.n;
	while (pos > 0 && key_less(t, key, B.h[pos - 1].t, B.h[pos - 1].key)) pos--;
	if (B.n == RT_SHADOW_K) {
		if (pos == RT_SHADOW_K) { if (t < B.min_dropped) B.min_dropped = t; if (term && t < B.cut) B.cut = t; return; }
		if (B.h[RT_SHADOW_K - 1].t < B.min_dropped) B.min_dropped = B.h[RT_SHADOW_K - 1].t;
		B.n--;
	}
	for (int i = B.n; i > pos; i--) B.h[i] = B.h[i - 1];
	B.h[pos].t = t;
	B.h[pos].key = key;
	B.h[pos].kt = kt;
	B.h[pos].flags = (inside ? 1 : 0) | (trans ? 2 : 0) | (term ? 4 : 0);
	B.n++;
	if (term && t < B.cut) B.cut = t;
}

// ---- list variants of the primitives (all hits) ---------------------------------------------
// emit(t_local, N_local, uv) for every hit the reference's intersectLocalList reports.
template <typename F> RT_HD void sphere_list(d3 o, d3 d, F&& emit)
{
	// Sphere::intersectLocalList (reference SceneObjects/Sphere.cpp:42-72)
	d = normalize(d);
	d3 v = -o;
	double b = dot(v, d);
	double disc = b * b - dot(v, v) + 1;
	if (disc < 0.0) return;
	disc = sqrt(disc);
	double t1 = b - disc, t2 = b + disc;
	d2 uv; uv.x = uv.y = 0.0;
	if (t1 > RT_RAY_EPSILON) emit(t1, normalize(o + t1 * d), uv, 0);
	if (t2 > RT_RAY_EPSILON) emit(t2, normalize(o + t2 * d), uv, 1);
}
template <typename F> RT_HD void box_list(d3 o, d3 d, F&& emit)
{
	// Box::intersectLocalList (reference SceneObjects/Box.cpp:65-97); the bump normal is applied by the caller
	for (int it = 0; it < 6; it++) {
		const int m0 = it % 3, m1 = (it + 1) % 3, m2 = (it + 2) % 3;
		const double dm = comp(d, m0);
		if (dm == 0) continue;
		const double t = ((it / 3) - 0.5 - comp(o, m0)) / dm;
		if (t < RT_RAY_EPSILON) continue;
		const double x = comp(o, m1) + t * comp(d, m1);
		const double y = comp(o, m2) + t * comp(d, m2);
		if (x <= 0.5 && x >= -0.5 && y <= 0.5 && y >= -0.5) {
			d3 P = o + t * d;
			const int lo = m1 < m2 ? m1 : m2, hi = m1 < m2 ? m2 : m1;
			d2 uv;
			uv.x = 0.5 + ((it < 3) ? -1.0 : 1.0) * comp(P, lo);
			uv.y = 0.5 + comp(P, hi);
			d3 N = (it < 3) ? mk3(-(double)(it == 0), -(double)(it == 1), -(double)(it == 2))
			                : mk3((double)(it == 3), (double)(it == 4), (double)(it == 5));
			emit(t, N, uv, it);
		}
	}
}
template <typename F> RT_HD void cylinder_list(d3 o, d3 d, F&& emit)
{
	// Cylinder::intersectCapsList then intersectBodyList (reference SceneObjects/Cylinder.cpp:155-263)
	d2 uv; uv.x = uv.y = 0.0;
	const double pz = o.z, dz = d.z;
	if (!(0.0 == dz)) {
		double t1, t2;
		if (dz > 0.0) { t1 = (-pz) / dz; t2 = (1.0 - pz) / dz; }
		else { t1 = (1.0 - pz) / dz; t2 = (-pz) / dz; }
		if (t1 >= RT_RAY_EPSILON) {
			d3 p = o + t1 * d;
			if ((p.x * p.x + p.y * p.y) <= 1.0) emit(t1, (dz > 0.0) ? mk3(0.0, 0.0, -1.0) : mk3(0.0, 0.0, 1.0), uv, 0);
		}
		if (t2 >= RT_RAY_EPSILON) {
			d3 p = o + t2 * d;
			if ((p.x * p.x + p.y * p.y) <= 1.0) emit(t2, (dz > 0.0) ? mk3(0.0, 0.0, 1.0) : mk3(0.0, 0.0, -1.0), uv, 1);
		}
	}
	double x0 = o.x, y0 = o.y, x1 = d.x, y1 = d.y;
	double a = x1 * x1 + y1 * y1;
	double b = 2.0 * (x0 * x1 + y0 * y1);
	double c = x0 * x0 + y0 * y0 - 1.0;
	if (0.0 == a) return;
	double disc = b * b - 4.0 * a * c;
	if (disc < 0.0) return;
	disc = sqrt(disc);
	double t1 = (-b - disc) / (2.0 * a);
	double t2 = (-b + disc) / (2.0 * a);
	if (t1 > RT_RAY_EPSILON) {
		d3 P = o + t1 * d;
		if (P.z >= 0.0 && P.z <= 1.0) emit(t1, normalize(mk3(P.x, P.y, 0.0)), uv, 2);
	}
	if (t2 > RT_RAY_EPSILON) {
		d3 P = o + t2 * d;
		if (P.z >= 0.0 && P.z <= 1.0) emit(t2, normalize(mk3(P.x, P.y, 0.0)), uv, 3);
	}
}
template <typename F> RT_HD void cone_list(d3 R0, d3 Rd, const double* prm, int capped, F&& emit)
{
	// Cone::intersectLocalList (reference SceneObjects/Cone.cpp:108-210)
	const double height = prm[0], b_radius = prm[1], t_radius = prm[2], beta_squared = prm[3], gamma = prm[4];
	d2 uv; uv.x = uv.y = 0.0;
	double pz = R0.z, dz = Rd.z;
	double a = Rd.x * Rd.x + Rd.y * Rd.y - beta_squared * Rd.z * Rd.z;
	if (a == 0.0) return;
	double b = 2 * (R0.x * Rd.x + R0.y * Rd.y - beta_squared * ((R0.z + gamma) * Rd.z));
	double c = -beta_squared * (gamma + R0.z) * (gamma + R0.z) + R0.x * R0.x + R0.y * R0.y;
	double disc = b * b - 4 * a * c;
	double theRoot = RT_RAY_EPSILON;
	if (disc <= 0) return;
	disc = sqrt(disc);
	double nearRoot = (-b + disc) / (2 * a);
	double farRoot = (-b - disc) / (2 * a);
	d3 Pn = R0 + nearRoot * Rd;
	bool nearGood = !(Pn.z < 0 || Pn.z > height);
	if (nearGood && (nearRoot > theRoot)) {
		theRoot = nearRoot;
		d3 P = R0 + theRoot * Rd;
		d3 normal = mk3(P.x, P.y, -2.0 * beta_squared * (P.z + gamma));
		if (!capped && dot(normal, Rd) > 0) normal = -normal;
		emit(theRoot, normalize(normal), uv, 0);
	}
	if (farRoot != nearRoot) {
		d3 Pf = R0 + farRoot * Rd;
		bool farGood = !(Pf.z < 0 || Pf.z > height);
		if (farGood && farRoot > RT_RAY_EPSILON) {
			theRoot = farRoot;
			d3 P = R0 + theRoot * Rd;
			d3 normal = mk3(P.x, P.y, -2.0 * beta_squared * (P.z + gamma));
			if (!capped && dot(normal, Rd) > 0) normal = -normal;
			emit(theRoot, normalize(normal), uv, 1);
		}
	}
	double t1 = (-pz) / dz;
	double t2 = (height - pz) / dz;
	if (capped) {
		d3 p = R0 + t1 * Rd;
		if (p.x * p.x + p.y * p.y <= b_radius * b_radius) {
			if (t1 > RT_RAY_EPSILON) emit(t1, normalize((dz > 0.0) ? mk3(0.0, 0.0, -1.0) : mk3(0.0, 0.0, 1.0)), uv, 2);
		}
		d3 q = R0 + t2 * Rd;
		if (q.x * q.x + q.y * q.y <= t_radius * t_radius) {
			if (t2 > RT_RAY_EPSILON) emit(t2, normalize((dz > 0.0) ? mk3(0.0, 0.0, 1.0) : mk3(0.0, 0.0, -1.0)), uv, 3);
		}
	}
}

// Geometry::intersectList for one scene-level candidate (reference scene/scene.cpp:40-64)
RT_HD void object_list(const DScene& S, int obj, unsigned rank, const d3& o, const d3& d, const d3& sl, SBuf& B, int* stk, double* stke, int sp0)
{
	const rt_object& ob = S.objs[obj];
	double e;
	if (!box_hit(ob.bbox, ob.bbox + 3, o, d, sl, e)) return;
	if (e > B.cut) return;
	const LocalRay L = to_local(ob, o, d);
	if (ob.type != RT_OBJ_TRIMESH) {
		const rt_material& m = S.mats[ob.material];
		auto emit = [&](double tl, d3 Nl, d2 uv, int sub) {
			const double t = tl / L.len;
			if (!(t <= B.cut)) return;
			const d3 N = world_normal(S, ob, Nl, uv);
			const bool inside = dot(N, d) > 0;
			sbuf_add(B, t, ((unsigned long long)rank << 32) | (unsigned)sub, inside, m.trans != 0, mat_param(S, m, RT_KT, uv));
		};
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
		return;
	}
	// Trimesh::intersectLocalList (reference SceneObjects/trimesh.cpp:97-105): every face the
	// per-mesh tree offers that the ray hits
	const rt_mesh& M = S.meshes[ob.mesh];
	const rt_node* nodes = S.mnodes + M.node_off;
	const DTri* tris = S.tris + M.item_off;
	const d3 lsl = slack_vec(L.d, S.mesh_slack[ob.mesh]);
	double le;
	if (!box_hit(nodes[0].bmin, nodes[0].bmax, L.o, L.d, lsl, le)) return;
	int sp = sp0, node = 0;
	for (;;) {
		const double lcut = B.cut < RT_T_INF ? (B.cut * L.len) * (1.0 + 1e-9) + 1e-300 : RT_T_INF;
		const rt_node& n = nodes[node];
		bool descend = false;
		if (n.right < 0) {
			for (int i = 0; i < n.count; i++) {
				const int pos = n.first + i;
				double tl;
				d3 bary;
				if (!tri_hit<true>(tris[pos], L.o, L.d, tl, bary)) continue;
				const double t = tl / L.len;
				if (!(t <= B.cut)) continue;
				Hit h;
				h.obj = obj;
				h.fpos = (int)M.item_off + pos;
				h.face = S.mitems[h.fpos];
				h.bary = bary;
				h.uv.x = h.uv.y = 0.0;
				const rt_face& f = S.faces[M.face_off + h.face];
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
				d3 kt;
				bool trans;
				eval_kt(S, h, kt, trans);
				sbuf_add(B, t, ((unsigned long long)rank << 32) | (unsigned)pos, dot(N, d) > 0, trans, kt);
			}
		} else {
			const int c0 = node + 1, c1 = n.right;
			double e0, e1;
			bool h0 = box_hit(nodes[c0].bmin, nodes[c0].bmax, L.o, L.d, lsl, e0) && !(e0 > lcut);
			bool h1 = box_hit(nodes[c1].bmin, nodes[c1].bmax, L.o, L.d, lsl, e1) && !(e1 > lcut);
			if (h0 && h1) {
				const bool swap = e1 < e0;
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
			if (sp == sp0) { done = true; break; }
			sp--;
			node = stk[sp];
			const double lc = B.cut < RT_T_INF ? (B.cut * L.len) * (1.0 + 1e-9) + 1e-300 : RT_T_INF;
			if (!(stke[sp] > lc)) break;
		}
		if (done) return;
	}
}

// One collection pass over the scene-level tree (Scene::intersectList, scene.cpp:182-197).
RT_HD void shadow_pass(const DScene& S, const d3& o, const d3& d, SBuf& B)
{
	int stk[RT_STACK];
	double stke[RT_STACK];
	const rt_node* nodes = S.snodes;
	const d3 sl = slack_vec(d, S.scene_slack);
	double e;
	if (!box_hit(nodes[0].bmin, nodes[0].bmax, o, d, sl, e)) return;
	int sp = 0, node = 0;
	for (;;) {
		const rt_node& n = nodes[node];
		bool descend = false;
		if (n.right < 0) {
			for (int i = 0; i < n.count; i++) {
				const int rank = n.first + i;
				object_list(S, S.sitems[rank], (unsigned)rank, o, d, sl, B, stk, stke, sp);
			}
		} else {
			const int c0 = node + 1, c1 = n.right;
			double e0, e1;
			bool h0 = box_hit(nodes[c0].bmin, nodes[c0].bmax, o, d, sl, e0) && !(e0 > B.cut);
			bool h1 = box_hit(nodes[c1].bmin, nodes[c1].bmax, o, d, sl, e1) && !(e1 > B.cut);
			if (h0 && h1) {
				const bool swap = e1 < e0;
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
			if (!(stke[sp] > B.cut)) break;
		}
		if (done) return;
	}
}

// Light::srsAttenuation for point and directional lights (overlapping-objects mode off: the
// medium outside every surface is `air`, kt = 1, transmissive; reference material.cpp:17-21).
RT_HD d3 shadow_attenuation(const DScene& S, const rt_light& Lt, const d3& pos, const d3& dir, const rt_params& prm)
{
	const bool point = Lt.type != RT_LIGHT_DIRECTIONAL;
	const d3 lpos = ld3(Lt.position);
	double limit = RT_T_INF;
	if (point) {
		const double D = length(lpos - pos);
		limit = D * (1.0 + 1e-9) + 1e-9;
	}
	d3 sattn = mk3(1.0, 1.0, 1.0);
	double last_t = 0.0;
	d3 cur = pos; // r2l position, advanced hit by hit (light.cpp:38)
	const bool aterm = prm.aterm_thresh > 0.0;
	SBuf B;
	B.floor_t = -1.0;
	B.floor_key = 0;
	for (int pass = 0; pass < 1 << 20; pass++) {
		B.n = 0;
		B.cut = limit;
		B.min_dropped = RT_T_INF;
		shadow_pass(S, pos, dir, B);
		const bool overflow = B.min_dropped < RT_T_INF && B.min_dropped <= B.cut;
		for (int k = 0; k < B.n; k++) {
			const SHit& h = B.h[k];
			if (!(h.t <= B.cut)) break;
			const double t = h.t - last_t;
			last_t = h.t;
			const bool inside = (h.flags & 1) != 0;
			cur = cur + t * dir;
			// curr/next media: inside ? (m_in -> air) : (air -> m_in)
			if (point) { // PointLight::sattnLimitCheck (light.cpp:65-71)
				if (dot(lpos - (cur + t * dir), dir) <= 0) return sattn;
			}
			const bool next_trans = inside ? true : (h.flags & 2) != 0;
			if (!next_trans) return mk3(0.0, 0.0, 0.0);
			if (aterm && dot(sattn, sattn) < prm.aterm_thresh * prm.aterm_thresh) return mk3(0.0, 0.0, 0.0);
			const d3 kt = inside ? h.kt : mk3(1.0, 1.0, 1.0);
			sattn = sattn * vpow(kt, t);
			// once the attenuation is exactly 0 every later exit of the reference's loop returns 0
			if (sattn.x == 0.0 && sattn.y == 0.0 && sattn.z == 0.0) return sattn;
			B.floor_t = h.t;
			B.floor_key = h.key;
		}
		if (!overflow && !(B.cut < limit)) return sattn; // the list is exhausted
		if (!overflow) { // walked everything up to a cut that did not end the walk (delta == 0 case)
			B.floor_t = B.cut;
			B.floor_key = ~0ull;
		}
	}
	return sattn;
}

} // namespace rtd
