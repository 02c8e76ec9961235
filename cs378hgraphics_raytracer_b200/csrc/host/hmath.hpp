// hmath.hpp - host-side double-precision vector/matrix arithmetic for the scene loader.
//
// The reference does all scene setup with glm 0.9.8.4 (reference ray/cmake/glm.cmake:9-17):
// transforms (parser/Parser.cpp:265,287,317,340), their inverses (scene/scene.h:127-128), face
// normals (SceneObjects/trimesh.h:114-126), bounding boxes (scene/scene.cpp:78-116).  For the
// flattened scene to be bit-identical to the one the reference builds, every operation here uses
// glm 0.9.8.4's evaluation order.  Compile with -ffp-contract=off.
#pragma once
#include <cmath>

namespace hm {

struct V3 {
	double x, y, z;
	V3() : x(0), y(0), z(0) {}
	V3(double a, double b, double c) : x(a), y(b), z(c) {}
	double& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
	double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct V4 {
	double x, y, z, w;
	V4() : x(0), y(0), z(0), w(0) {}
	V4(double a, double b, double c, double d) : x(a), y(b), z(c), w(d) {}
	double& operator[](int i) { return i == 0 ? x : (i == 1 ? y : (i == 2 ? z : w)); }
	double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : (i == 2 ? z : w)); }
};

inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, double s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(double s, V3 a) { return V3(s * a.x, s * a.y, s * a.z); }
inline V3 operator/(V3 a, double s) { return V3(a.x / s, a.y / s, a.z / s); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline V4 operator+(V4 a, V4 b) { return V4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline V4 operator-(V4 a, V4 b) { return V4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline V4 operator*(V4 a, V4 b) { return V4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline V4 operator*(V4 a, double s) { return V4(a.x * s, a.y * s, a.z * s, a.w * s); }

// glm: dot = (x*x' + y*y') + z*z'
inline double dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline double length(V3 a) { return std::sqrt(dot(a, a)); }
// glm: normalize = v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x)
inline V3 normalize(V3 a) { return a * (1.0 / std::sqrt(dot(a, a))); }
inline V3 cross(V3 a, V3 b) { return V3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
// glm: min(x,y) = (y < x) ? y : x ; max(x,y) = (x < y) ? y : x
inline double gmin(double x, double y) { return (y < x) ? y : x; }
inline double gmax(double x, double y) { return (x < y) ? y : x; }
inline V3 gmin(V3 a, V3 b) { return V3(gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z)); }
inline V3 gmax(V3 a, V3 b) { return V3(gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z)); }
inline V4 gmin(V4 a, V4 b) { return V4(gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z), gmin(a.w, b.w)); }
inline V4 gmax(V4 a, V4 b) { return V4(gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z), gmax(a.w, b.w)); }

// column-major: c[col][row]
struct M3 {
	V3 c[3];
	M3() { c[0] = V3(1, 0, 0); c[1] = V3(0, 1, 0); c[2] = V3(0, 0, 1); }
	M3(V3 a, V3 b, V3 d) { c[0] = a; c[1] = b; c[2] = d; }
	V3& operator[](int i) { return c[i]; }
	const V3& operator[](int i) const { return c[i]; }
};
struct M4 {
	V4 c[4];
	M4() { c[0] = V4(1, 0, 0, 0); c[1] = V4(0, 1, 0, 0); c[2] = V4(0, 0, 1, 0); c[3] = V4(0, 0, 0, 1); }
	M4(V4 a, V4 b, V4 d, V4 e) { c[0] = a; c[1] = b; c[2] = d; c[3] = e; }
	V4& operator[](int i) { return c[i]; }
	const V4& operator[](int i) const { return c[i]; }
};

inline V3 mul(const M3& m, V3 v)
{
	return V3(m[0].x * v.x + m[1].x * v.y + m[2].x * v.z,
	          m[0].y * v.x + m[1].y * v.y + m[2].y * v.z,
	          m[0].z * v.x + m[1].z * v.y + m[2].z * v.z);
}
// glm type_mat4x4.inl: (m0*v0 + m1*v1) + (m2*v2 + m3*v3)
inline V4 mul(const M4& m, V4 v)
{
	V4 a0 = m[0] * v.x + m[1] * v.y;
	V4 a1 = m[2] * v.z + m[3] * v.w;
	return a0 + a1;
}
// glm: column j = ((a0*b[j][0] + a1*b[j][1]) + a2*b[j][2]) + a3*b[j][3]
inline M4 mul(const M4& a, const M4& b)
{
	M4 r;
	for (int j = 0; j < 4; j++) r[j] = ((a[0] * b[j].x + a[1] * b[j].y) + a[2] * b[j].z) + a[3] * b[j].w;
	return r;
}
inline M3 transpose(const M3& m)
{
	return M3(V3(m[0].x, m[1].x, m[2].x), V3(m[0].y, m[1].y, m[2].y), V3(m[0].z, m[1].z, m[2].z));
}
inline M4 transpose(const M4& m)
{
	return M4(V4(m[0].x, m[1].x, m[2].x, m[3].x), V4(m[0].y, m[1].y, m[2].y, m[3].y),
	          V4(m[0].z, m[1].z, m[2].z, m[3].z), V4(m[0].w, m[1].w, m[2].w, m[3].w));
}
inline M3 upper3(const M4& m) { return M3(V3(m[0].x, m[0].y, m[0].z), V3(m[1].x, m[1].y, m[1].z), V3(m[2].x, m[2].y, m[2].z)); }

M3 inverse(const M3& m);
M4 inverse(const M4& m);
M4 translation(V3 v);
M4 scaling(V3 v);
M4 rotation(double angle, V3 axis);

} // namespace hm
