// glm-compatible shim for building the UNMODIFIED reference sources as the oracle.
//
// TEST INFRASTRUCTURE ONLY (oracle/): the reference needs glm 0.9.8.4, which is an
// un-vendored submodule (reference ray/cmake/glm.cmake:9-17, .SUBMODULES.json) and is
// absent here.  This header is a from-scratch restatement of the subset of the glm API
// the reference's CLI path uses, written to glm 0.9.8.4's published evaluation orders
// (func_geometric.inl, func_matrix.inl, type_mat4x4.inl, gtc/matrix_transform.inl):
//   - vectors default-construct to zero, matrices to identity (0.9.8 semantics)
//   - dot(vec3) = (x+y)+z, dot(vec4) = (x+y)+(z+w)
//   - normalize(v) = v * (1/sqrt(dot(v,v)))
//   - mat4*vec4 = (m0*v0 + m1*v1) + (m2*v2 + m3*v3)
//   - inverse() = cofactor formulas of func_matrix.inl
// No reference test pins glm's results, so parity is "unpinned at the glm boundary":
// this shim DEFINES the arithmetic for the oracle; the product's host math
// (csrc/host/hmath.hpp) is an independent second implementation and
// tests/test_arena_parity.py proves the two agree bit-for-bit on every scene.
#pragma once
#include <cmath>
#include <cstddef>
#include <ostream>
#include <numeric>

namespace glm {

template <int N> struct dvec;

template <> struct dvec<2> {
	double x, y;
	dvec() : x(0), y(0) {}
	explicit dvec(double s) : x(s), y(s) {}
	dvec(double a, double b) : x(a), y(b) {}
	double& operator[](int i) { return (&x)[i]; }
	const double& operator[](int i) const { return (&x)[i]; }
};
template <> struct dvec<4>;
template <> struct dvec<3> {
	double x, y, z;
	dvec() : x(0), y(0), z(0) {}
	explicit dvec(double s) : x(s), y(s), z(s) {}
	dvec(double a, double b, double c) : x(a), y(b), z(c) {}
	explicit dvec(const dvec<4>& v);
	double& operator[](int i) { return (&x)[i]; }
	const double& operator[](int i) const { return (&x)[i]; }
};
template <> struct dvec<4> {
	double x, y, z, w;
	dvec() : x(0), y(0), z(0), w(0) {}
	explicit dvec(double s) : x(s), y(s), z(s), w(s) {}
	dvec(double a, double b, double c, double d) : x(a), y(b), z(c), w(d) {}
	dvec(const dvec<3>& v, double d) : x(v.x), y(v.y), z(v.z), w(d) {}
	double& operator[](int i) { return (&x)[i]; }
	const double& operator[](int i) const { return (&x)[i]; }
};
inline dvec<3>::dvec(const dvec<4>& v) : x(v.x), y(v.y), z(v.z) {}

typedef dvec<2> dvec2;
typedef dvec<3> dvec3;
typedef dvec<4> dvec4;

// ---- component-wise arithmetic ------------------------------------------------------
#define GLMSHIM_VEC_OP(OP)                                                               \
	template <int N> inline dvec<N> operator OP(const dvec<N>& a, const dvec<N>& b) {     \
		dvec<N> r;                                                                        \
		for (int i = 0; i < N; i++) r[i] = a[i] OP b[i];                                  \
		return r;                                                                         \
	}                                                                                     \
	template <int N> inline dvec<N> operator OP(const dvec<N>& a, double s) {             \
		dvec<N> r;                                                                        \
		for (int i = 0; i < N; i++) r[i] = a[i] OP s;                                     \
		return r;                                                                         \
	}                                                                                     \
	template <int N> inline dvec<N> operator OP(double s, const dvec<N>& a) {             \
		dvec<N> r;                                                                        \
		for (int i = 0; i < N; i++) r[i] = s OP a[i];                                     \
		return r;                                                                         \
	}                                                                                     \
	template <int N> inline dvec<N>& operator OP##=(dvec<N>& a, const dvec<N>& b) {       \
		for (int i = 0; i < N; i++) a[i] OP## = b[i];                                     \
		return a;                                                                         \
	}                                                                                     \
	template <int N> inline dvec<N>& operator OP##=(dvec<N>& a, double s) {               \
		for (int i = 0; i < N; i++) a[i] OP## = s;                                        \
		return a;                                                                         \
	}
GLMSHIM_VEC_OP(+)
GLMSHIM_VEC_OP(-)
GLMSHIM_VEC_OP(*)
GLMSHIM_VEC_OP(/)
#undef GLMSHIM_VEC_OP

template <int N> inline dvec<N> operator-(const dvec<N>& a) {
	dvec<N> r;
	for (int i = 0; i < N; i++) r[i] = -a[i];
	return r;
}
template <int N> inline bool operator==(const dvec<N>& a, const dvec<N>& b) {
	for (int i = 0; i < N; i++)
		if (!(a[i] == b[i])) return false;
	return true;
}
template <int N> inline bool operator!=(const dvec<N>& a, const dvec<N>& b) { return !(a == b); }

// ---- geometric ----------------------------------------------------------------------
inline double dot(const dvec2& a, const dvec2& b) { dvec2 t = a * b; return t.x + t.y; }
inline double dot(const dvec3& a, const dvec3& b) { dvec3 t = a * b; return (t.x + t.y) + t.z; }
inline double dot(const dvec4& a, const dvec4& b) { dvec4 t = a * b; return (t.x + t.y) + (t.z + t.w); }
template <int N> inline double length(const dvec<N>& v) { return std::sqrt(dot(v, v)); }
template <int N> inline double distance(const dvec<N>& a, const dvec<N>& b) { return length(b - a); }
template <int N> inline dvec<N> normalize(const dvec<N>& v) { return v * (1.0 / std::sqrt(dot(v, v))); }
inline dvec3 cross(const dvec3& x, const dvec3& y) {
	return dvec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}

// ---- common -------------------------------------------------------------------------
inline double min(double x, double y) { return (y < x) ? y : x; }
inline double max(double x, double y) { return (x < y) ? y : x; }
inline double abs(double x) { return std::fabs(x); }
inline double clamp(double x, double lo, double hi) { return min(max(x, lo), hi); }
inline double sqrt(double x) { return std::sqrt(x); }
inline double cos(double x) { return std::cos(x); }
inline double sin(double x) { return std::sin(x); }
inline double tan(double x) { return std::tan(x); }
inline double pow(double x, double y) { return std::pow(x, y); }

#define GLMSHIM_VEC_FN2(FN)                                                              \
	template <int N> inline dvec<N> FN(const dvec<N>& a, const dvec<N>& b) {              \
		dvec<N> r;                                                                        \
		for (int i = 0; i < N; i++) r[i] = FN(a[i], b[i]);                                \
		return r;                                                                         \
	}                                                                                     \
	template <int N> inline dvec<N> FN(const dvec<N>& a, double b) {                      \
		dvec<N> r;                                                                        \
		for (int i = 0; i < N; i++) r[i] = FN(a[i], b);                                   \
		return r;                                                                         \
	}
GLMSHIM_VEC_FN2(min)
GLMSHIM_VEC_FN2(max)
GLMSHIM_VEC_FN2(pow)
#undef GLMSHIM_VEC_FN2
template <int N> inline dvec<N> abs(const dvec<N>& a) {
	dvec<N> r;
	for (int i = 0; i < N; i++) r[i] = std::fabs(a[i]);
	return r;
}
template <int N> inline dvec<N> clamp(const dvec<N>& a, double lo, double hi) {
	dvec<N> r;
	for (int i = 0; i < N; i++) r[i] = clamp(a[i], lo, hi);
	return r;
}

// ---- matrices (column-major, m[col][row]) -------------------------------------------
template <int C, int R> struct dmat {
	dvec<R> c[C];
	dmat() {
		for (int i = 0; i < C; i++)
			for (int j = 0; j < R; j++) c[i][j] = (i == j) ? 1.0 : 0.0;
	}
	explicit dmat(double s) {
		for (int i = 0; i < C; i++)
			for (int j = 0; j < R; j++) c[i][j] = (i == j) ? s : 0.0;
	}
	dmat(const dvec<R>& a, const dvec<R>& b) { c[0] = a; c[1] = b; }
	dmat(const dvec<R>& a, const dvec<R>& b, const dvec<R>& d) { c[0] = a; c[1] = b; c[2] = d; }
	dmat(const dvec<R>& a, const dvec<R>& b, const dvec<R>& d, const dvec<R>& e) { c[0] = a; c[1] = b; c[2] = d; c[3] = e; }
	template <int C2, int R2> explicit dmat(const dmat<C2, R2>& m) {
		for (int i = 0; i < C; i++)
			for (int j = 0; j < R; j++) c[i][j] = (i < C2 && j < R2) ? m[i][j] : ((i == j) ? 1.0 : 0.0);
	}
	dvec<R>& operator[](int i) { return c[i]; }
	const dvec<R>& operator[](int i) const { return c[i]; }
};
typedef dmat<3, 3> dmat3;
typedef dmat<3, 3> dmat3x3;
typedef dmat<4, 4> dmat4;
typedef dmat<4, 4> dmat4x4;
typedef dmat<2, 3> dmat2x3;

inline dvec3 operator*(const dmat3& m, const dvec3& v) {
	return dvec3(m[0][0] * v.x + m[1][0] * v.y + m[2][0] * v.z,
	             m[0][1] * v.x + m[1][1] * v.y + m[2][1] * v.z,
	             m[0][2] * v.x + m[1][2] * v.y + m[2][2] * v.z);
}
inline dvec4 operator*(const dmat4& m, const dvec4& v) {
	dvec4 Mul0 = m[0] * v[0], Mul1 = m[1] * v[1];
	dvec4 Add0 = Mul0 + Mul1;
	dvec4 Mul2 = m[2] * v[2], Mul3 = m[3] * v[3];
	dvec4 Add1 = Mul2 + Mul3;
	return Add0 + Add1;
}
inline dvec3 operator*(const dmat2x3& m, const dvec2& v) {
	return dvec3(m[0][0] * v.x + m[1][0] * v.y, m[0][1] * v.x + m[1][1] * v.y, m[0][2] * v.x + m[1][2] * v.y);
}
inline dmat4 operator*(const dmat4& a, const dmat4& b) {
	dmat4 r;
	for (int j = 0; j < 4; j++) r[j] = a[0] * b[j][0] + a[1] * b[j][1] + a[2] * b[j][2] + a[3] * b[j][3];
	return r;
}
inline dmat4 operator*(const dmat4& m, double s) { return dmat4(m[0] * s, m[1] * s, m[2] * s, m[3] * s); }

template <int C, int R> inline dmat<R, C> transpose(const dmat<C, R>& m) {
	dmat<R, C> r;
	for (int i = 0; i < C; i++)
		for (int j = 0; j < R; j++) r[j][i] = m[i][j];
	return r;
}

inline dmat3 inverse(const dmat3& m) {
	double OneOverDeterminant = 1.0 / (+m[0][0] * (m[1][1] * m[2][2] - m[2][1] * m[1][2])
	                                   - m[1][0] * (m[0][1] * m[2][2] - m[2][1] * m[0][2])
	                                   + m[2][0] * (m[0][1] * m[1][2] - m[1][1] * m[0][2]));
	dmat3 I;
	I[0][0] = +(m[1][1] * m[2][2] - m[2][1] * m[1][2]) * OneOverDeterminant;
	I[1][0] = -(m[1][0] * m[2][2] - m[2][0] * m[1][2]) * OneOverDeterminant;
	I[2][0] = +(m[1][0] * m[2][1] - m[2][0] * m[1][1]) * OneOverDeterminant;
	I[0][1] = -(m[0][1] * m[2][2] - m[2][1] * m[0][2]) * OneOverDeterminant;
	I[1][1] = +(m[0][0] * m[2][2] - m[2][0] * m[0][2]) * OneOverDeterminant;
	I[2][1] = -(m[0][0] * m[2][1] - m[2][0] * m[0][1]) * OneOverDeterminant;
	I[0][2] = +(m[0][1] * m[1][2] - m[1][1] * m[0][2]) * OneOverDeterminant;
	I[1][2] = -(m[0][0] * m[1][2] - m[1][0] * m[0][2]) * OneOverDeterminant;
	I[2][2] = +(m[0][0] * m[1][1] - m[1][0] * m[0][1]) * OneOverDeterminant;
	return I;
}

inline dmat4 inverse(const dmat4& m) {
	double Coef00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
	double Coef02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
	double Coef03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
	double Coef04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
	double Coef06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
	double Coef07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
	double Coef08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
	double Coef10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
	double Coef11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
	double Coef12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
	double Coef14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
	double Coef15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
	double Coef16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
	double Coef18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
	double Coef19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
	double Coef20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
	double Coef22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
	double Coef23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];
	dvec4 Fac0(Coef00, Coef00, Coef02, Coef03);
	dvec4 Fac1(Coef04, Coef04, Coef06, Coef07);
	dvec4 Fac2(Coef08, Coef08, Coef10, Coef11);
	dvec4 Fac3(Coef12, Coef12, Coef14, Coef15);
	dvec4 Fac4(Coef16, Coef16, Coef18, Coef19);
	dvec4 Fac5(Coef20, Coef20, Coef22, Coef23);
	dvec4 Vec0(m[1][0], m[0][0], m[0][0], m[0][0]);
	dvec4 Vec1(m[1][1], m[0][1], m[0][1], m[0][1]);
	dvec4 Vec2(m[1][2], m[0][2], m[0][2], m[0][2]);
	dvec4 Vec3(m[1][3], m[0][3], m[0][3], m[0][3]);
	dvec4 Inv0(Vec1 * Fac0 - Vec2 * Fac1 + Vec3 * Fac2);
	dvec4 Inv1(Vec0 * Fac0 - Vec2 * Fac3 + Vec3 * Fac4);
	dvec4 Inv2(Vec0 * Fac1 - Vec1 * Fac3 + Vec3 * Fac5);
	dvec4 Inv3(Vec0 * Fac2 - Vec1 * Fac4 + Vec2 * Fac5);
	dvec4 SignA(+1, -1, +1, -1);
	dvec4 SignB(-1, +1, -1, +1);
	dmat4 Inverse(Inv0 * SignA, Inv1 * SignB, Inv2 * SignA, Inv3 * SignB);
	dvec4 Row0(Inverse[0][0], Inverse[1][0], Inverse[2][0], Inverse[3][0]);
	dvec4 Dot0(m[0] * Row0);
	double Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
	double OneOverDeterminant = 1.0 / Dot1;
	return Inverse * OneOverDeterminant;
}

// ---- gtc/matrix_transform + gtx/transform -------------------------------------------
inline dmat4 translate(const dmat4& m, const dvec3& v) {
	dmat4 Result(m);
	Result[3] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2] + m[3];
	return Result;
}
inline dmat4 scale(const dmat4& m, const dvec3& v) {
	dmat4 Result;
	Result[0] = m[0] * v[0];
	Result[1] = m[1] * v[1];
	Result[2] = m[2] * v[2];
	Result[3] = m[3];
	return Result;
}
inline dmat4 rotate(const dmat4& m, double angle, const dvec3& v) {
	const double a = angle;
	const double c = std::cos(a);
	const double s = std::sin(a);
	dvec3 axis(normalize(v));
	dvec3 temp((1.0 - c) * axis);
	dmat4 Rotate;
	Rotate[0][0] = c + temp[0] * axis[0];
	Rotate[0][1] = temp[0] * axis[1] + s * axis[2];
	Rotate[0][2] = temp[0] * axis[2] - s * axis[1];
	Rotate[1][0] = temp[1] * axis[0] - s * axis[2];
	Rotate[1][1] = c + temp[1] * axis[1];
	Rotate[1][2] = temp[1] * axis[2] + s * axis[0];
	Rotate[2][0] = temp[2] * axis[0] + s * axis[1];
	Rotate[2][1] = temp[2] * axis[1] - s * axis[0];
	Rotate[2][2] = c + temp[2] * axis[2];
	dmat4 Result;
	Result[0] = m[0] * Rotate[0][0] + m[1] * Rotate[0][1] + m[2] * Rotate[0][2];
	Result[1] = m[0] * Rotate[1][0] + m[1] * Rotate[1][1] + m[2] * Rotate[1][2];
	Result[2] = m[0] * Rotate[2][0] + m[1] * Rotate[2][1] + m[2] * Rotate[2][2];
	Result[3] = m[3];
	return Result;
}
inline dmat4 translate(const dvec3& v) { return translate(dmat4(1.0), v); }
inline dmat4 scale(const dvec3& v) { return scale(dmat4(1.0), v); }
inline dmat4 rotate(double angle, const dvec3& v) { return rotate(dmat4(1.0), angle, v); }

// ---- gtx/io ---------------------------------------------------------------------------
template <int N> inline std::ostream& operator<<(std::ostream& os, const dvec<N>& v) {
	os << "(";
	for (int i = 0; i < N; i++) os << (i ? ", " : "") << v[i];
	return os << ")";
}

} // namespace glm
