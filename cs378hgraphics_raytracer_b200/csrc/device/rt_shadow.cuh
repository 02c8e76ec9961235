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
	double sure_t;      // an entered opaque surface at t <= sure_t makes the result exactly 0 (see sbuf_add)
	bool dead;          // such a surface was found: the walk can stop, srsAttenuation returns dvec3(0)
};

RT_HD bool key_less(double ta, unsigned long long ka, double tb, unsigned long long kb) { return ta < tb || (ta == tb && ka < kb); }

RT_COLD void sbuf_add(SBuf& B, double t, unsigned long long key, bool inside, bool trans, const d3& kt)
{
	if (!(t <= B.cut)) return;
	if (!key_less(B.floor_t, B.floor_key, t, key)) return;
	// a hit that ends the walk whenever it is reached: entering a non-transmissive surface returns
	// 0; leaving a medium with kt = 0 zeroes the attenuation (unless delta is exactly 0, which the
	// walk detects and answers with another pass)
	const bool term = inside ? (kt.x == 0.0 && kt.y == 0.0 && kt.z == 0.0) : !trans;
	// Any-hit exit.  srsAttenuation's loop can only return something other than 0 before it reaches
	// an entered opaque surface through the light's limit check (light.cpp:45), which for a point
	// light at distance D fires at hit k when D - t_k - (t_k - t_{k-1}) <= 0, i.e. never while
	// t_k <= D/2, and never for a directional light (light.cpp:60).  Every other way out of the loop
	// before that surface (light.cpp:47-48) returns 0, and so does the surface itself.  So a
	// non-transmissive entry at t <= sure_t (0.49 D, or infinity) settles the result whatever else
	// the ray meets: no need to collect the hits in front of it.
	if (term && !inside && t <= B.sure_t) { B.dead = true; B.cut = t; return; }
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

} // namespace rtd
