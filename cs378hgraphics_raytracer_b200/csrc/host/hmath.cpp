// hmath.cpp - matrix inverse and transform constructors in glm 0.9.8.4's evaluation order
// (func_matrix.inl compute_inverse<mat3/mat4>, gtc/matrix_transform.inl translate/scale/rotate).
// These run once per transform node at load time (reference scene/scene.h:119-129).
#include "hmath.hpp"

namespace hm {

M3 inverse(const M3& m)
{
	const double a = m[0].x, b = m[0].y, c = m[0].z; // column 0
	const double d = m[1].x, e = m[1].y, f = m[1].z; // column 1
	const double g = m[2].x, h = m[2].y, i = m[2].z; // column 2
	const double det = a * (e * i - h * f) - d * (b * i - h * c) + g * (b * f - e * c);
	const double r = 1.0 / det;
	M3 o;
	o[0].x = (e * i - h * f) * r;
	o[1].x = -(d * i - g * f) * r;
	o[2].x = (d * h - g * e) * r;
	o[0].y = -(b * i - h * c) * r;
	o[1].y = (a * i - g * c) * r;
	o[2].y = -(a * h - g * b) * r;
	o[0].z = (b * f - e * c) * r;
	o[1].z = -(a * f - d * c) * r;
	o[2].z = (a * e - d * b) * r;
	return o;
}

// 2x2 sub-determinant of rows (r0,r1) taken from columns (c0,c1): m[c0][r0]*m[c1][r1] - m[c1][r0]*m[c0][r1]
static inline double sub2(const M4& m, int c0, int r0, int c1, int r1) { return m[c0][r0] * m[c1][r1] - m[c1][r0] * m[c0][r1]; }

M4 inverse(const M4& m)
{
	// the 18 distinct 2x2 minors of columns 1..3 / rows pairs, grouped as glm does (Fac0..Fac5)
	const double s00 = sub2(m, 2, 2, 3, 3), s02 = sub2(m, 1, 2, 3, 3), s03 = sub2(m, 1, 2, 2, 3);
	const double s04 = sub2(m, 2, 1, 3, 3), s06 = sub2(m, 1, 1, 3, 3), s07 = sub2(m, 1, 1, 2, 3);
	const double s08 = sub2(m, 2, 1, 3, 2), s10 = sub2(m, 1, 1, 3, 2), s11 = sub2(m, 1, 1, 2, 2);
	const double s12 = sub2(m, 2, 0, 3, 3), s14 = sub2(m, 1, 0, 3, 3), s15 = sub2(m, 1, 0, 2, 3);
	const double s16 = sub2(m, 2, 0, 3, 2), s18 = sub2(m, 1, 0, 3, 2), s19 = sub2(m, 1, 0, 2, 2);
	const double s20 = sub2(m, 2, 0, 3, 1), s22 = sub2(m, 1, 0, 3, 1), s23 = sub2(m, 1, 0, 2, 1);
	const V4 f0(s00, s00, s02, s03), f1(s04, s04, s06, s07), f2(s08, s08, s10, s11);
	const V4 f3(s12, s12, s14, s15), f4(s16, s16, s18, s19), f5(s20, s20, s22, s23);
	const V4 v0(m[1].x, m[0].x, m[0].x, m[0].x);
	const V4 v1(m[1].y, m[0].y, m[0].y, m[0].y);
	const V4 v2(m[1].z, m[0].z, m[0].z, m[0].z);
	const V4 v3(m[1].w, m[0].w, m[0].w, m[0].w);
	const V4 i0 = (v1 * f0 - v2 * f1) + v3 * f2;
	const V4 i1 = (v0 * f0 - v2 * f3) + v3 * f4;
	const V4 i2 = (v0 * f1 - v1 * f3) + v3 * f5;
	const V4 i3 = (v0 * f2 - v1 * f4) + v2 * f5;
	const V4 sa(+1, -1, +1, -1), sb(-1, +1, -1, +1);
	M4 adj(i0 * sa, i1 * sb, i2 * sa, i3 * sb);
	const V4 row0(adj[0].x, adj[1].x, adj[2].x, adj[3].x);
	const V4 d = m[0] * row0;
	const double det = (d.x + d.y) + (d.z + d.w);
	const double r = 1.0 / det;
	return M4(adj[0] * r, adj[1] * r, adj[2] * r, adj[3] * r);
}

M4 translation(V3 v)
{
	M4 id, r;
	r[3] = ((id[0] * v.x + id[1] * v.y) + id[2] * v.z) + id[3];
	return r;
}

M4 scaling(V3 v)
{
	M4 id, r;
	r[0] = id[0] * v.x;
	r[1] = id[1] * v.y;
	r[2] = id[2] * v.z;
	r[3] = id[3];
	return r;
}

M4 rotation(double angle, V3 axis_in)
{
	const double c = std::cos(angle), s = std::sin(angle);
	const V3 ax = normalize(axis_in);
	const V3 t = (1.0 - c) * ax;
	double R[3][3];
	R[0][0] = c + t.x * ax.x;
	R[0][1] = t.x * ax.y + s * ax.z;
	R[0][2] = t.x * ax.z - s * ax.y;
	R[1][0] = t.y * ax.x - s * ax.z;
	R[1][1] = c + t.y * ax.y;
	R[1][2] = t.y * ax.z + s * ax.x;
	R[2][0] = t.z * ax.x + s * ax.y;
	R[2][1] = t.z * ax.y - s * ax.x;
	R[2][2] = c + t.z * ax.z;
	M4 id, r;
	for (int k = 0; k < 3; k++) r[k] = (id[0] * R[k][0] + id[1] * R[k][1]) + id[2] * R[k][2];
	r[3] = id[3];
	return r;
}

} // namespace hm
