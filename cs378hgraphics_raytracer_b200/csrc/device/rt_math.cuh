// rt_math.cuh - FP64 vector helpers for the device path, in glm 0.9.8.4's evaluation order.
//
// Everything on the hit-decision path of the reference is double precision glm arithmetic built
// for baseline x86-64 (no FMA contraction).  To select the same hit primitives bit for bit, the
// device code spells every expression term by term in the same association order (SURVEY.md
// Appendix B) and this translation unit is compiled with -fmad=false, IEEE division and sqrt
// (nvcc defaults --prec-div=true --prec-sqrt=true for FP64).
//
// RT_HD functions also compile as plain C++ (g++ -ffp-contract=off) so that tests can single-step
// the same per-ray code on the CPU next to the reference (tests/hostsim); the product library
// never contains that host build.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define RT_HD __host__ __device__ __forceinline__
#define RT_HDN __host__ __device__ __noinline__
#else
#define RT_HD inline
#define RT_HDN inline
#endif
// RT_COLD: code outside the box/triangle loop.  The trace kernel is instruction-fetch bound when
// everything is inlined (profiles/r1_v2_*: 47k SASS instructions, 52-80% stall_no_inst), so the
// rarely executed parts are real calls and the tree-walk loop stays within the 32 KB L1.5 I-cache.
#if defined(__CUDACC__) && !defined(RT_INLINE_ALL)
#define RT_COLD __host__ __device__ __noinline__
#else
#define RT_COLD RT_HD
#endif

namespace rtd {

struct d3 {
	double x, y, z;
};
struct d2 {
	double x, y;
};

RT_HD d3 mk3(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_HD d3 ld3(const double* p) { return mk3(p[0], p[1], p[2]); }
RT_HD d3 operator+(d3 a, d3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD d3 operator-(d3 a, d3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_HD d3 operator*(d3 a, d3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
RT_HD d3 operator*(d3 a, double s) { return mk3(a.x * s, a.y * s, a.z * s); }
RT_HD d3 operator*(double s, d3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
RT_HD d3 operator-(d3 a) { return mk3(-a.x, -a.y, -a.z); }
RT_HD double comp(const d3& v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

RT_HD double dot(d3 a, d3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// one copy each of the IEEE sqrt / division expansions (~50 SASS instructions apiece)
RT_COLD double dsqrt(double x) { return sqrt(x); }
RT_COLD double inv_sqrt(double x) { return 1.0 / sqrt(x); }
RT_COLD double ddiv(double a, double b) { return a / b; }
RT_HD double length(d3 a) { return dsqrt(dot(a, a)); }
RT_HD d3 normalize(d3 a) { return a * inv_sqrt(dot(a, a)); }
RT_HD d3 cross(d3 a, d3 b) { return mk3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
RT_HD double gmin(double x, double y) { return (y < x) ? y : x; }
RT_HD double gmax(double x, double y) { return (x < y) ? y : x; }
RT_HD double gclamp(double x, double lo, double hi) { return gmin(gmax(x, lo), hi); }
RT_HD d3 gclamp(d3 v, double lo, double hi) { return mk3(gclamp(v.x, lo, hi), gclamp(v.y, lo, hi), gclamp(v.z, lo, hi)); }
RT_HD d3 vpow(d3 b, double e) { return mk3(pow(b.x, e), pow(b.y, e), pow(b.z, e)); }

// reference scene.h:57-62 on a column-major dmat4 m[col*4+row]: (x,y,z,1) in, xyz out;
// glm mat4*vec4 = (m0*v0 + m1*v1) + (m2*v2 + m3*v3)
RT_HD d3 xform_point(const double* m, d3 v)
{
	return mk3((m[0] * v.x + m[4] * v.y) + (m[8] * v.z + m[12] * 1.0),
	           (m[1] * v.x + m[5] * v.y) + (m[9] * v.z + m[13] * 1.0),
	           (m[2] * v.x + m[6] * v.y) + (m[10] * v.z + m[14] * 1.0));
}
// glm mat3*vec3 on a column-major dmat3 m[col*3+row]
RT_HD d3 mul33(const double* m, d3 v)
{
	return mk3(m[0] * v.x + m[3] * v.y + m[6] * v.z,
	           m[1] * v.x + m[4] * v.y + m[7] * v.z,
	           m[2] * v.x + m[5] * v.y + m[8] * v.z);
}

// ---- IEEE division by a per-ray constant, without the ~25-instruction DDIV sequence --------------
// The slab test divides six numbers per box by the three direction components.  With y = 1.0/d
// (one real division per axis per ray) Markstein's correction recovers the CORRECTLY ROUNDED
// quotient a/d from FMAs: q0 = a*y is within 2 ulp; r0 = a - q0*d is exact in an FMA; q1 = q0 + r0*y
// is faithful; one more step (r1, q2) is the correctly rounded a/d (Markstein 1990; Muller et al.,
// Handbook of Floating-Point Arithmetic, "division with an FMA").  The proof needs no overflow or
// underflow in the intermediates, hence the range guards; outside them (never seen in practice) the
// real division runs.  tests/test_gpu_parity.py::test_division_selftest compares 2^28 quotients
// against the hardware sequence bit for bit.
struct RayInv {
	d3 y;          // 1/d per axis (IEEE)
	double slmax;  // pruning slack: max over the axes with d != 0 of slack/|d|
	float guard;   // one register for two rarely-used facts:
	               //   NaN  - some |d| lies outside [2^-300, 2^300]: every quotient is a real division
	               //   >= 0 - pruned walk of a ray with a zero direction component: boxes farther than
	               //          this from the origin along that axis are refused (rt_prims.cuh)
	               //   < 0  - neither
	RT_HD bool fma_ok() const { return guard == guard; }
	// width of the exponent window of `a` served by div_by's FMA path
	RT_HD unsigned span() const { return fma_ok() ? 1201u : 0u; }
};
// Single-precision image of the ray for the box PRE-test (rt_prims.cuh: boxf).  A slab quotient is
// computed as ONE fused multiply-add  q~ = RN32(b~ * y~ + p~)  with b~ = RN32(box coordinate),
// y~ = RN32(1/d_i), p~ = RN32(-(o_i * (1/d_i))) (product in double, rounded once).  With u = 2^-24:
//     q~ - (b - o_i)/d_i  =  y (b (e_b + e_y) - o_i e_p) + e_q q~,   |e_*| <= u (1 + 2^-20)
// so |q~ - q| < 2u (E + |o_i|) |y_i| + u |q|  (E = largest |coordinate| of the tree's root box; the
// reference's own double rounding, 2^-53 relative, is far below this).  amax carries the absolute part
// with a 3x margin (6u max_i (E + |o_i|)|y_i|); the relative part is covered by the 2^-20 (|tn| + |tf|)
// term of the test (16x).  amax is +inf when a direction component is zero or anything is out of float
// range, which makes every box "ambiguous" and hands the decision to the double-precision test.
RT_HD void make_inv(RayInv& I, const d3& d, double slack);
struct RayF {
	float yx, yy, yz; // RN32(1/d)
	float px, py, pz; // RN32(-(o * (1/d)))
	float amax;
	float slf;        // RayInv::slmax rounded up: the pruning slack of the float entry distances
};
// smallest float >= x / largest float <= x (x finite or +-inf)
RT_HD float f_up(double x)
{
#ifdef __CUDA_ARCH__
	return __double2float_ru(x);
#else
	float f = (float)x;
	if ((double)f < x) f = nextafterf(f, INFINITY);
	return f;
#endif
}
RT_HD float f_dn(double x)
{
#ifdef __CUDA_ARCH__
	return __double2float_rd(x);
#else
	float f = (float)x;
	if ((double)f > x) f = nextafterf(f, -INFINITY);
	return f;
#endif
}
RT_HD bool div_range_ok(double d) { const double a = fabs(d); return d == 0.0 || (a >= 0x1p-300 && a <= 0x1p300); }
RT_HD void make_rayf(RayF& F, const d3& o, const d3& d, const d3& y, double extent, double slmax)
{
	F.yx = (float)y.x; F.yy = (float)y.y; F.yz = (float)y.z;
	F.px = (float)(-(o.x * y.x)); F.py = (float)(-(o.y * y.y)); F.pz = (float)(-(o.z * y.z));
	const double ax = (extent + fabs(o.x)) * fabs(y.x), ay = (extent + fabs(o.y)) * fabs(y.y), az = (extent + fabs(o.z)) * fabs(y.z);
	double m = ax > ay ? ax : ay;
	m = az > m ? az : m;
	const bool ok = d.x != 0.0 && d.y != 0.0 && d.z != 0.0 && fabs(y.x) <= 1e15 && fabs(y.y) <= 1e15 && fabs(y.z) <= 1e15 &&
	                extent + fabs(o.x) <= 1e15 && extent + fabs(o.y) <= 1e15 && extent + fabs(o.z) <= 1e15;
	F.amax = ok ? (float)(m * 0x1.8p-22) * 1.0001f + 1e-30f : INFINITY; // 6u m, rounded up
	F.slf = f_up(slmax);
}
RT_COLD void make_inv(RayInv& I, RayF& F, const d3& o, const d3& d, double slack, double extent)
{
	make_inv(I, d, slack);
	make_rayf(F, o, d, I.y, extent, I.slmax);
}
RT_HD void make_inv(RayInv& I, const d3& d, double slack)
{
	I.y = mk3(1.0 / d.x, 1.0 / d.y, 1.0 / d.z);
	const d3 sl = mk3(slack * fabs(I.y.x), slack * fabs(I.y.y), slack * fabs(I.y.z));
	I.slmax = 0.0;
	if (d.x != 0.0) I.slmax = sl.x;
	if (d.y != 0.0 && sl.y > I.slmax) I.slmax = sl.y;
	if (d.z != 0.0 && sl.z > I.slmax) I.slmax = sl.z;
	const bool zero = d.x == 0.0 || d.y == 0.0 || d.z == 0.0;
	I.guard = zero && slack > 0.0 && slack < 1e30 ? (float)slack * 1.0001f + 1e-37f : -1.0f;
	if (!(div_range_ok(d.x) && div_range_ok(d.y) && div_range_ok(d.z))) I.guard = NAN; // no FMA path, no refusal either
}
RT_HD unsigned biased_exp(double a)
{
#ifdef __CUDA_ARCH__
	return ((unsigned)__double2hiint(a) >> 20) & 0x7ffu;
#else
	uint64_t u;
	memcpy(&u, &a, 8);
	return (unsigned)(u >> 52) & 0x7ffu;
#endif
}
RT_HD double div_by(double a, double d, double y, unsigned span)
{
	// FMA path for 2^-600 <= |a| < 2^601 (biased exponent 423..1623) and for a == 0
	if (biased_exp(a) - 423u >= span) {
		if (span == 0u || a != 0.0) return ddiv(a, d);
	}
	double q = a * y;
	double r = fma(-q, d, a);
	q = fma(r, y, q);
	r = fma(-q, d, a);
	return fma(r, y, q);
}

// hammersley(k, N) (reference util.cpp:3-11): radical inverse base 2; the second coordinate is
// always 0 because n has been shifted down to 0 by the time it is used.
RT_HD double radical_inverse2(int n)
{
	double mul = 0.5, result = 0.0;
	while (n > 0) {
		result += (n % 2) ? mul : 0;
		n /= 2;
		mul /= 2.0;
	}
	return result;
}

#define RT_RAY_EPSILON 0.00000001
#define RT_EPS32 (RT_RAY_EPSILON / 32)

} // namespace rtd
