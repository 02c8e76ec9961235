// rt_prims.cuh - bounding-box slab test and the analytic primitives, closest-hit variants.
// Each function restates the reference member it cites, comparison for comparison.
#pragma once
#include "rt_scene.cuh"

namespace rtd {

// BoundingBox::intersect (reference ray/src/scene/bbox.cc:33-70).  The per-axis early returns
// are monotone (tMin only grows, tMax only shrinks) so testing once after the loop is identical.
// `entry` receives tMin lowered by the ray's pruning slack: a conservative entry distance used
// ONLY to cull subtrees that cannot beat the current closest hit; box membership itself is the
// reference's decision (see DESIGN.md "pruning without changing the winner").
//
// box_hit_exact: the six quotients are correctly rounded (div_by, rt_math.cuh), i.e. the
// reference's own numbers.
//
// Zero direction components.  The reference skips an axis with d == 0 altogether (bbox.cc:43), so
// a ray of the centre pixel column (d.x == 0) "touches" every box whose other two slabs it crosses,
// wherever the box lies along x: tens of thousands of nodes per ray, all of whose items then fail
// their own intersection test.  A PRUNED walk (I.guard >= 0: set for rays that have a zero component) may refuse such a box when the origin
// lies outside its slab by more than the guard distance (1e-9 of the tree's extent): every item
// under the box lies inside it, the ray never comes closer to it than the guard along that axis,
// and a hit needs the intersection point inside the item up to rounding (1e-16 relative), so the
// refused subtree holds no candidate the reference would have accepted - the same argument as for
// distance pruning: the box test only filters candidates, the winner is chosen among the items.
RT_COLD bool box_hit_exact(const double* bmin, const double* bmax, const d3& o, const d3& d, const RayInv& I, double& entry)
{
	double tMin = -1.0e308, tMax = 1.0e308;
#define RT_SLAB(O, D, Y, LO, HI)                                            \
	if (D != 0.0) {                                                         \
		double t1 = div_by(LO - O, D, Y, I.span());                           \
		double t2 = div_by(HI - O, D, Y, I.span());                           \
		if (t1 > t2) { double tt = t1; t1 = t2; t2 = tt; }                  \
		if (t1 > tMin) tMin = t1;                                           \
		if (t2 < tMax) tMax = t2;                                           \
	}
	RT_SLAB(o.x, d.x, I.y.x, bmin[0], bmax[0])
	RT_SLAB(o.y, d.y, I.y.y, bmin[1], bmax[1])
	RT_SLAB(o.z, d.z, I.y.z, bmin[2], bmax[2])
#undef RT_SLAB
	entry = tMin - I.slmax;
	if (tMin > tMax) return false;
	if (tMax < RT_RAY_EPSILON) return false;
	return true;
}

// box_hit: the same decision from ONE multiplication per quotient.  q~ = a * RN(1/d) differs from
// the correctly rounded a/d by less than 3.4e-16 |q| (two roundings), and so do tMin~ = max of the
// near quotients and tMax~ = min of the far ones from the exact tMin, tMax (max/min of perturbed
// values move by at most the perturbation of the element that ends up selected, and an element
// x <= T satisfies x + e|x| <= T + e|T|).  Hence with s = 1e-14 (|tMin~| + |tMax~|) (30x the bound):
//     tMax~ - tMin~ < -s                   =>  exact tMin > tMax       : miss
//     tMax~ < EPS - 1e-20                  =>  exact tMax < EPS        : miss
//     tMax~ - tMin~ > s, tMax~ > EPS+1e-20 =>  neither                 : hit
// and only the sliver in between (grazing rays, boxes touching the origin) takes box_hit_exact.
// Non-finite or denormal intermediates fail all three tests and fall through to it as well.
// the zero-component refusal described above; only rays that have such a component get here
RT_COLD bool zero_axis_outside(const double* bmin, const double* bmax, const d3& o, const d3& d, double g)
{
	if (d.x == 0.0 && (o.x < bmin[0] - g || o.x > bmax[0] + g)) return true;
	if (d.y == 0.0 && (o.y < bmin[1] - g || o.y > bmax[1] + g)) return true;
	if (d.z == 0.0 && (o.z < bmin[2] - g || o.z > bmax[2] + g)) return true;
	return false;
}

RT_HD bool box_hit(const double* bmin, const double* bmax, const d3& o, const d3& d, const RayInv& I, double& entry)
{
	if (I.fma_ok()) {
		double tn = -1.0e308, tf = 1.0e308;
#define RT_SLAB(O, D, Y, LO, HI)                                            \
		if (D != 0.0) {                                                     \
			const double q1 = (LO - O) * Y, q2 = (HI - O) * Y;              \
			const double nr = q2 < q1 ? q2 : q1, fr = q2 < q1 ? q1 : q2;    \
			if (nr > tn) tn = nr;                                           \
			if (fr < tf) tf = fr;                                           \
		}
		RT_SLAB(o.x, d.x, I.y.x, bmin[0], bmax[0])
		RT_SLAB(o.y, d.y, I.y.y, bmin[1], bmax[1])
		RT_SLAB(o.z, d.z, I.y.z, bmin[2], bmax[2])
#undef RT_SLAB
		const double s = 1e-14 * (fabs(tn) + fabs(tf)) + 1e-290;
		const double gap = tf - tn;
		entry = tn - I.slmax;
		if (gap < -s || tf < RT_RAY_EPSILON - 1e-20) return false;
		if (gap > s && tf > RT_RAY_EPSILON + 1e-20) return true;
	}
	return box_hit_exact(bmin, bmax, o, d, I, entry);
}

// Tree nodes (never the per-object bounding boxes, whose pre-test stays the reference's): box_hit
// plus the zero-component refusal.  Such rays have F.amax = +inf, so all their node tests come here.
#ifndef RT_NODE_ATTR
#define RT_NODE_ATTR RT_COLD // out of line: the hot loop keeps only the float test (profiles/r1_variants.txt)
#endif
RT_NODE_ATTR bool box_hit_node(const double* bmin, const double* bmax, const d3& o, const d3& d, const RayInv& I, double& entry)
{
#ifndef RT_NO_ZERO_GUARD
	if (I.guard >= 0.0f && zero_axis_outside(bmin, bmax, o, d, (double)I.guard)) { entry = 0.0; return false; }
#endif
	return box_hit(bmin, bmax, o, d, I, entry);
}

// Float slab test of one box against the ray's single-precision image (RayF, rt_math.cuh): 6 FMAs,
// min / max, and a three-way answer by the margins derived there:
//   0 = the reference's double test certainly answers "miss"
//   1 = it certainly answers "hit"; e = a conservative entry distance (never above the exact tMin
//       minus the pruning slack: tn - s lies below the exact tMin by more than the rounding of the
//       two float subtractions that follow)
//   2 = ambiguous (grazing rays, boxes touching the origin, zero direction components, NaN): the
//       caller runs the double test on the double box
RT_HD int boxf(const float* lo, const float* hi, const RayF& F, float& e)
{
	const float qx1 = fmaf(lo[0], F.yx, F.px), qx2 = fmaf(hi[0], F.yx, F.px);
	const float qy1 = fmaf(lo[1], F.yy, F.py), qy2 = fmaf(hi[1], F.yy, F.py);
	const float qz1 = fmaf(lo[2], F.yz, F.pz), qz2 = fmaf(hi[2], F.yz, F.pz);
	const float tn = fmaxf(fmaxf(fminf(qx1, qx2), fminf(qy1, qy2)), fminf(qz1, qz2));
	const float tf = fminf(fminf(fmaxf(qx1, qx2), fmaxf(qy1, qy2)), fmaxf(qz1, qz2));
	const float s = fmaf(0x1p-20f, fabsf(tn) + fabsf(tf), 2.0f * F.amax);
	const float gap = tf - tn;
	e = (tn - s) - F.slf;
	if (gap < -s || tf + s < (float)RT_RAY_EPSILON) return 0;
	if (gap > s && tf - s > (float)RT_RAY_EPSILON * 1.0001f) return 1;
	return 2;
}

// The double-precision decision for an ambiguous box of a tree node, entry distance rounded down to float.
RT_COLD int boxf_resolve(const double* lo, const double* hi, const d3& o, const d3& d, const RayInv& I, float& e)
{
	double ed;
	const bool h = box_hit_node(lo, hi, o, d, I, ed);
	e = h ? f_dn(ed) : 0.0f;
	return h ? 1 : 0;
}

// Both children of a node with double entry distances (the warp-cooperative walk, rt_coop.cuh).
RT_HD void box2_hit(const DNode2f& nf, const DNode2& n, const d3& o, const d3& d, const RayInv& I, const RayF& F, bool& h0, bool& h1, double& e0, double& e1)
{
	float f0, f1;
	int r0 = boxf(nf.lo0, nf.hi0, F, f0), r1 = boxf(nf.lo1, nf.hi1, F, f1);
	if (r0 == 2) r0 = boxf_resolve(n.lo0, n.hi0, o, d, I, f0);
	if (r1 == 2) r1 = boxf_resolve(n.lo1, n.hi1, o, d, I, f1);
	h0 = r0 != 0;
	h1 = r1 != 0;
	e0 = (double)f0;
	e1 = (double)f1;
}

// Object bounding boxes (Geometry::intersect's pre-test, scene.cpp:15): true when the float test
// proves that the reference's double test answers "miss", or - pruned walks - that the box is entered
// beyond `limf` (a float >= the walk's limit); anything else - hit, ambiguous, zero direction
// components (amax = inf), NaN - is false and the caller runs the double test on the object's own box.
RT_HD bool box_miss_f(const DItemBox& b, const RayF& F, float limf = INFINITY)
{
	float e;
	const int r = boxf(b.lo, b.hi, F, e);
	return r == 0 || (r == 1 && e > limf);
}

// ---- object-space float pre-test of a scene-level candidate (DPrimF, rt_scene.cuh) ------------------------------------------
// The same function builds the record on the host (rt_derive.hpp) and on the device (rt_derive_dev.cuh): plain double
// arithmetic, no contraction, so the two images are byte-identical.
RT_HD void derive_primf(const rt_object& ob, const DItemBox& b, DPrimF& p)
{
	for (int k = 0; k < 9; k++) p.m[k] = 0.0f;
	p.lc[0] = p.lc[1] = p.lc[2] = 0.0f;
	p.kind = RT_PF_NONE;
	p.pad[0] = p.pad[1] = p.pad[2] = 0;
	int kind;
	double az = 0.0; // centre of the shape's own bounding box (on the z axis)
	switch (ob.type) {
	case RT_OBJ_BOX: kind = RT_PF_BOX; break;           // [-0.5, 0.5]^3 (Box.cpp)
	case RT_OBJ_SPHERE: kind = RT_PF_SPHERE; break;     // radius 1 (Sphere.cpp)
	case RT_OBJ_SQUARE: kind = RT_PF_SQUARE; break;     // [-0.5, 0.5]^2 in z = 0 (Square.cpp)
	case RT_OBJ_CYLINDER: kind = RT_PF_CYLINDER; az = 0.5; break; // radius 1, z in [0, 1], capped (Cylinder.cpp)
	default: return; // trimeshes are entered, cones keep the double test (their root bookkeeping is followed literally)
	}
	// inverse transform: element (row r, column c) = inv[4 c + r] (rt_math.cuh: xform_point)
	double M[9], n[3];
	for (int r = 0; r < 3; r++) {
		for (int c = 0; c < 3; c++) M[3 * r + c] = ob.inv[4 * c + r];
		n[r] = (M[3 * r] * M[3 * r] + M[3 * r + 1] * M[3 * r + 1]) + M[3 * r + 2] * M[3 * r + 2];
	}
	// rotation x uniform scale only: rows of equal length, orthogonal (to 1e-6).  Then |m v| = s |v| for every v, no component
	// of m v is a small difference of large terms beyond what |v| allows, and the error bounds in prim_miss_f hold.
	const double mx = n[0] > n[1] ? (n[0] > n[2] ? n[0] : n[2]) : (n[1] > n[2] ? n[1] : n[2]);
	const double mn = n[0] < n[1] ? (n[0] < n[2] ? n[0] : n[2]) : (n[1] < n[2] ? n[1] : n[2]);
	if (!(mn > 1e-60) || !(mx < 1e60) || !(mx - mn <= 1e-6 * mx)) return;
	for (int r = 0; r < 3; r++) {
		const int q = (r + 1) % 3;
		const double dp = (M[3 * r] * M[3 * q] + M[3 * r + 1] * M[3 * q + 1]) + M[3 * r + 2] * M[3 * q + 2];
		if (!(dp <= 1e-6 * mx && dp >= -1e-6 * mx)) return;
	}
	// reference point: the centre of the item's float box, as prim_miss_f recomputes it
	const float cf[3] = {0.5f * (b.lo[0] + b.hi[0]), 0.5f * (b.lo[1] + b.hi[1]), 0.5f * (b.lo[2] + b.hi[2])};
	const d3 lc = xform_point(ob.inv, mk3((double)cf[0], (double)cf[1], (double)cf[2]));
	const double l[3] = {lc.x, lc.y, lc.z - az};
	for (int c = 0; c < 3; c++)
		if (!(l[c] < 4.0 && l[c] > -4.0)) return; // the box centre of a unit-sized shape lies within it; anything else is left alone
	for (int k = 0; k < 9; k++) p.m[k] = (float)M[k];
	for (int c = 0; c < 3; c++) p.lc[c] = (float)l[c];
	p.kind = kind;
}

// true = the LINE of the ray certainly misses the shape, so Geometry::intersect (scene.cpp:13-40: world -> object ray,
// intersectLocal) certainly answers "no hit" whatever t range it looks at.  o, d = the world ray (double), b = the item's box.
//  * q = o + t0 d, t0 ~ where the ray passes the box centre: a point of the ray NEAR the object, computed in double precision,
//    so that the float numbers below are small (object-sized) whatever the distance of the ray's origin;
//  * l = m (q - c) + lc and e = m d: the line in object space, centred on the shape's bounding box; with |l|_1 < 8 (checked;
//    a ray through the box of a unit-sized shape passes its centre within 4.5) their components are within 6e-6 and
//    4e-7 |e|_1 of the exact ones (four float roundings per component on terms of that size, rows of m of equal length),
//    a cross product component (l x e)_k within 1e-5 |e|_1;
//  * box / square / cylinder: separating axes e x u_k of a line and a box with half extents h (the cylinder's bounding box):
//    |(l x e)_k| > h_i |e_j| + h_j |e_i|; h is widened by 2e-3 and an axis is used only if |e_i| + |e_j| > 0.05 |e|_1, which
//    makes the widening worth >= 1e-4 |e|_1: ten times the error above (the primitives themselves accept nothing outside
//    their shape: Box.cpp / Square.cpp / Cylinder.cpp compare against the exact extents);
//  * sphere / cylinder: distance of the line from the centre / the axis against the radius widened the same way.
// NaN, infinities, zero directions: every comparison is written so that they answer "false" (no certain miss).
RT_HD bool prim_miss_f(const DPrimF& p, const DItemBox& b, const d3& o, const d3& d)
{
	if (p.kind == RT_PF_NONE) return false;
	const float cx = 0.5f * (b.lo[0] + b.hi[0]), cy = 0.5f * (b.lo[1] + b.hi[1]), cz = 0.5f * (b.lo[2] + b.hi[2]);
	const float dx = (float)d.x, dy = (float)d.y, dz = (float)d.z;
	const float t0 = ((cx - (float)o.x) * dx + (cy - (float)o.y) * dy + (cz - (float)o.z) * dz) / (dx * dx + dy * dy + dz * dz);
	if (!(t0 < 1e30f && t0 > -1e30f)) return false;
	const double t = (double)t0;
	const float qx = (float)((o.x + t * d.x) - (double)cx), qy = (float)((o.y + t * d.y) - (double)cy), qz = (float)((o.z + t * d.z) - (double)cz);
	const float lx = (p.m[0] * qx + p.m[1] * qy + p.m[2] * qz) + p.lc[0];
	const float ly = (p.m[3] * qx + p.m[4] * qy + p.m[5] * qz) + p.lc[1];
	const float lz = (p.m[6] * qx + p.m[7] * qy + p.m[8] * qz) + p.lc[2];
	const float ex = p.m[0] * dx + p.m[1] * dy + p.m[2] * dz;
	const float ey = p.m[3] * dx + p.m[4] * dy + p.m[5] * dz;
	const float ez = p.m[6] * dx + p.m[7] * dy + p.m[8] * dz;
	if (!(fabsf(lx) + fabsf(ly) + fabsf(lz) < 8.0f)) return false;
	const float ax = fabsf(ex), ay = fabsf(ey), az = fabsf(ez);
	const float en = ax + ay + az;
	if (!(en > 1e-30f && en < 1e30f)) return false;
	const float kx = ly * ez - lz * ey, ky = lz * ex - lx * ez, kz = lx * ey - ly * ex; // l x e
	const float w = 2e-3f;
	if (p.kind == RT_PF_SPHERE) {
		const float r = 1.0f + w;
		return (kx * kx + ky * ky) + kz * kz > (r * r) * ((ex * ex + ey * ey) + ez * ez);
	}
	const bool round_xy = p.kind == RT_PF_CYLINDER;
	const float hx = (round_xy ? 1.0f : 0.5f) + w, hy = hx, hz = (p.kind == RT_PF_SQUARE ? 0.0f : 0.5f) + w;
	const float thr = 0.05f * en;
	if (ay + az > thr && fabsf(kx) > hy * az + hz * ay) return true;
	if (az + ax > thr && fabsf(ky) > hz * ax + hx * az) return true;
	if (ax + ay > thr && fabsf(kz) > hx * ay + hy * ax) return true;
	if (round_xy) {
		const float exy = ex * ex + ey * ey;
		if (exy > (thr * thr) && kz * kz > (hx * hx) * exy) return true;
	}
	return false;
}

// What a primitive's intersectLocal leaves in the isect (object space).
struct LocalHit {
	double t;
	d3 N;
	d2 uv;
};

// Sphere::intersectLocal (reference SceneObjects/Sphere.cpp:9-40).  It re-normalises the already
// normalised direction (:10), which can move it by an ulp, so that is reproduced.
template <bool FULL> RT_HD bool sphere_hit(d3 o, d3 d, LocalHit& h)
{
	d = normalize(d);
	d3 v = -o;
	double b = dot(v, d);
	double disc = b * b - dot(v, v) + 1;
	if (disc < 0.0) return false;
	disc = sqrt(disc);
	double t2 = b + disc;
	if (t2 <= RT_RAY_EPSILON) return false;
	double t1 = b - disc;
	double t = (t1 > RT_RAY_EPSILON) ? t1 : t2;
	h.t = t;
	if (FULL) h.N = normalize(o + t * d);
	return true;
}

// Box::intersectLocal (reference SceneObjects/Box.cpp:11-64).  The bump-mapped normal
// (Box::computeNormal :99-108) needs the material and is applied by the caller from `face`.
template <bool FULL> RT_HD bool box_prim_hit(d3 o, d3 d, LocalHit& h, int& face)
{
	double bestT = 1e100;
	int best = -1;
#pragma unroll
	for (int it = 0; it < 6; it++) {
		const int m0 = it % 3, m1 = (it + 1) % 3, m2 = (it + 2) % 3;
		const double dm = comp(d, m0);
		if (dm == 0) continue;
		const double t = ((it / 3) - 0.5 - comp(o, m0)) / dm;
		if (t < RT_RAY_EPSILON || t > bestT) continue;
		const double x = comp(o, m1) + t * comp(d, m1);
		const double y = comp(o, m2) + t * comp(d, m2);
		if (x <= 0.5 && x >= -0.5 && y <= 0.5 && y >= -0.5) {
			if (bestT > t) { bestT = t; best = it; }
		}
	}
	if (best < 0) return false;
	h.t = bestT;
	face = best;
	if (FULL) {
		d3 P = o + bestT * d;
		const int i1 = (best + 1) % 3, i2 = (best + 2) % 3;
		const int lo = i1 < i2 ? i1 : i2, hi = i1 < i2 ? i2 : i1;
		if (best < 3) { h.uv.x = 0.5 - comp(P, lo); h.uv.y = 0.5 + comp(P, hi); }
		else { h.uv.x = 0.5 + comp(P, lo); h.uv.y = 0.5 + comp(P, hi); }
		if (best < 3) h.N = mk3(-(double)(best == 0), -(double)(best == 1), -(double)(best == 2));
		else h.N = mk3((double)(best == 3), (double)(best == 4), (double)(best == 5));
	}
	return true;
}

// Square::intersectLocal (reference SceneObjects/Square.cpp:9-46)
template <bool FULL> RT_HD bool square_hit(d3 o, d3 d, LocalHit& h)
{
	if (d.z == 0.0) return false;
	double t = -o.z / d.z;
	if (t <= RT_RAY_EPSILON) return false;
	d3 P = o + t * d;
	if (P.x < -0.5 || P.x > 0.5) return false;
	if (P.y < -0.5 || P.y > 0.5) return false;
	h.t = t;
	if (FULL) {
		h.N = (d.z > 0.0) ? mk3(0.0, 0.0, -1.0) : mk3(0.0, 0.0, 1.0);
		h.uv.x = P.x + 0.5;
		h.uv.y = P.y + 0.5;
	}
	return true;
}

// Cylinder::intersectBody (reference SceneObjects/Cylinder.cpp:30-93); capped is always true
// (Cylinder.h:11) so the normal flip for the uncapped case never runs.
template <bool FULL> RT_HD bool cyl_body(d3 o, d3 d, LocalHit& h)
{
	double x0 = o.x, y0 = o.y, x1 = d.x, y1 = d.y;
	double a = x1 * x1 + y1 * y1;
	double b = 2.0 * (x0 * x1 + y0 * y1);
	double c = x0 * x0 + y0 * y0 - 1.0;
	if (0.0 == a) return false;
	double disc = b * b - 4.0 * a * c;
	if (disc < 0.0) return false;
	disc = sqrt(disc);
	double t2 = (-b + disc) / (2.0 * a);
	if (t2 <= RT_RAY_EPSILON) return false;
	double t1 = (-b - disc) / (2.0 * a);
	if (t1 > RT_RAY_EPSILON) {
		d3 P = o + t1 * d;
		if (P.z >= 0.0 && P.z <= 1.0) {
			h.t = t1;
			if (FULL) h.N = normalize(mk3(P.x, P.y, 0.0));
			return true;
		}
	}
	d3 P = o + t2 * d;
	if (P.z >= 0.0 && P.z <= 1.0) {
		h.t = t2;
		if (FULL) h.N = normalize(mk3(P.x, P.y, 0.0));
		return true;
	}
	return false;
}

// Cylinder::intersectCaps (reference SceneObjects/Cylinder.cpp:95-153)
template <bool FULL> RT_HD bool cyl_caps(d3 o, d3 d, LocalHit& h)
{
	double pz = o.z, dz = d.z;
	if (0.0 == dz) return false;
	double t1, t2;
	if (dz > 0.0) { t1 = (-pz) / dz; t2 = (1.0 - pz) / dz; }
	else { t1 = (1.0 - pz) / dz; t2 = (-pz) / dz; }
	if (t2 < RT_RAY_EPSILON) return false;
	if (t1 >= RT_RAY_EPSILON) {
		d3 p = o + t1 * d;
		if ((p.x * p.x + p.y * p.y) <= 1.0) {
			h.t = t1;
			if (FULL) h.N = (dz > 0.0) ? mk3(0.0, 0.0, -1.0) : mk3(0.0, 0.0, 1.0);
			return true;
		}
	}
	d3 p = o + t2 * d;
	if ((p.x * p.x + p.y * p.y) <= 1.0) {
		h.t = t2;
		if (FULL) h.N = (dz > 0.0) ? mk3(0.0, 0.0, 1.0) : mk3(0.0, 0.0, -1.0);
		return true;
	}
	return false;
}

// Cylinder::intersectLocal (reference SceneObjects/Cylinder.cpp:8-28): caps first, the body
// replaces them only when strictly closer.
template <bool FULL> RT_HD bool cylinder_hit(d3 o, d3 d, LocalHit& h)
{
	if (cyl_caps<FULL>(o, d, h)) {
		LocalHit b;
		if (cyl_body<FULL>(o, d, b)) {
			if (b.t < h.t) { h.t = b.t; if (FULL) h.N = b.N; }
		}
		return true;
	}
	return cyl_body<FULL>(o, d, h);
}

// Cone::intersectLocal (reference SceneObjects/Cone.cpp:7-107), root bookkeeping kept literal.
// prm = {height, b_radius, t_radius, beta_squared, gamma}
template <bool FULL> RT_HD bool cone_hit(d3 R0, d3 Rd, const double* prm, int capped, LocalHit& h)
{
	const double height = prm[0], b_radius = prm[1], t_radius = prm[2], beta_squared = prm[3], gamma = prm[4];
	d3 normal = mk3(0.0, 0.0, 0.0);
	double pz = R0.z, dz = Rd.z;
	double a = Rd.x * Rd.x + Rd.y * Rd.y - beta_squared * Rd.z * Rd.z;
	if (a == 0.0) return false;
	double b = 2 * (R0.x * Rd.x + R0.y * Rd.y - beta_squared * ((R0.z + gamma) * Rd.z));
	double c = -beta_squared * (gamma + R0.z) * (gamma + R0.z) + R0.x * R0.x + R0.y * R0.y;
	double disc = b * b - 4 * a * c;
	double theRoot = RT_RAY_EPSILON;
	if (disc <= 0) return false;
	disc = sqrt(disc);
	double nearRoot = (-b + disc) / (2 * a);
	double farRoot = (-b - disc) / (2 * a);
	d3 Pn = R0 + nearRoot * Rd;
	bool nearGood = !(Pn.z < 0 || Pn.z > height);
	if (nearGood && (nearRoot > theRoot)) {
		theRoot = nearRoot;
		d3 P = R0 + theRoot * Rd;
		normal = mk3(P.x, P.y, -2.0 * beta_squared * (P.z + gamma));
	}
	d3 Pf = R0 + farRoot * Rd;
	bool farGood = !(Pf.z < 0 || Pf.z > height);
	if (farGood && ((nearGood && farRoot < theRoot) || farRoot > RT_RAY_EPSILON)) {
		theRoot = farRoot;
		d3 P = R0 + theRoot * Rd;
		normal = mk3(P.x, P.y, -2.0 * beta_squared * (P.z + gamma));
	}
	if (!capped && dot(normal, Rd) > 0) normal = -normal;
	double t1 = (-pz) / dz;
	double t2 = (height - pz) / dz;
	if (capped) {
		d3 p = R0 + t1 * Rd;
		if (p.x * p.x + p.y * p.y <= b_radius * b_radius) {
			if (t1 < theRoot && t1 > RT_RAY_EPSILON) {
				theRoot = t1;
				normal = (dz > 0.0) ? mk3(0.0, 0.0, -1.0) : mk3(0.0, 0.0, 1.0);
			}
		}
		d3 q = R0 + t2 * Rd;
		if (q.x * q.x + q.y * q.y <= t_radius * t_radius) {
			if (t2 < theRoot && t2 > RT_RAY_EPSILON) {
				theRoot = t2;
				normal = (dz > 0.0) ? mk3(0.0, 0.0, 1.0) : mk3(0.0, 0.0, -1.0);
			}
		}
	}
	if (theRoot <= RT_RAY_EPSILON) return false;
	h.t = theRoot;
	if (FULL) h.N = normalize(normal);
	return true;
}

} // namespace rtd
