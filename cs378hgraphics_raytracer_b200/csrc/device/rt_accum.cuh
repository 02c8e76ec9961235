// rt_accum.cuh - order-independent colour accumulators.
//
// The reference adds a sample's colour up the recursion (RayTracer.cpp:98-170: shade + kr * reflected + kt * refracted);
// the device adds every ray's weighted term to its sample the moment the term is known, from whichever thread holds it.
// FP64 atomics would make the sum depend on the order threads retire in (two renderings of one frame differing in a last
// bit now and then).  Here each colour channel is a fixed-point number held in two signed 64-bit limbs,
//
//     value = (limb[1] * 2^44 + limb[0]) * 2^-68
//
// and a term is split into its two limb parts and added with integer atomics (no return value: RED.ADD.64).  Integer
// addition is associative and commutative, so the limbs - and every frame byte derived from them - are the same for any
// launch order, queue order, number of partitions or GPUs.  Reading a channel rounds the exact limb sum to the nearest
// double once.  Twice the atomics of a plain FP64 add: k_wave_shade therefore STORES the terms that need no atomic at all
// (each camera ray's own, see acc_read3) - A/B in profiles/r2_variants.txt.
//
//  - limb 1 holds each term rounded to a multiple of 2^-24 and limb 0 the rest (|.| <= 2^-25, 44 bits); the other 20
//    bits of each limb are head room, so 2^19 terms of one sample cannot overflow a limb whatever their signs
//    (a depth-10 sample with both children everywhere and 16 lights has 2^11 * 17 < 2^16 terms);
//  - a term with |x| >= 2^-16 (its last mantissa bit is >= 2^-68) is held exactly: a sample with ONE term - flat
//    emissive / ambient colour, where authored constants such as 0.6 sit exactly on a byte boundary (255 * 0.6 = 153) -
//    reads back as that very double, and a sum of such terms is the correctly rounded exact sum;
//  - smaller terms are rounded (to nearest, ties to even) to a multiple of 2^-68 = 3.4e-21, five orders of magnitude
//    below the last bit of any value that can move a byte;
//  - |x| >= 2^38 is clamped to +-2^38 (the sample is white or black either way: sample_value clamps to [0, 1]) and a NaN
//    term adds nothing (the reference converts a NaN colour to a byte with an undefined cast, RayTracer.cpp:388-394).
#pragma once
#include "rt_math.cuh"

namespace rtd {

#define RT_ACC_LIMBS 2            // per channel
#define RT_ACC_STRIDE 6           // long longs per sample (3 channels)
#define RT_ACC_BYTES 48           // per sample
#define RT_ACC_LOW_BITS 44        // limb 1 counts multiples of 2^(44 - 68) = 2^-24
#define RT_ACC_UNIT_EXP (-68)     // limb 0 counts multiples of 2^-68

// x -> (lo, hi) with x = (hi * 2^44 + lo) * 2^-68 (to the rounding stated above), |lo| <= 2^43.
// Six FP64 / conversion instructions: hi = x rounded to a multiple of 2^-24, counted in those units; what is left of x
// (exact: the remainder of rounding a double to a coarser grid is a double) counted in units of 2^-68.  Any (lo, hi) with
// the same total would do - the limbs are only ever added.
RT_HD void acc_split(double x, long long& lo, long long& hi)
{
	const double big = 274877906944.0; // 2^38
	if (x != x) x = 0.0;               // a NaN term adds nothing
	x = x > big ? big : (x < -big ? -big : x);
#ifdef __CUDA_ARCH__
	hi = __double2ll_rn(x * 16777216.0); // 2^24
	const double r = fma(-(double)hi, 1.0 / 16777216.0, x);
	lo = __double2ll_rn(r * 295147905179352825856.0); // 2^68
#else
	hi = llrint(x * 16777216.0);
	const double r = x - (double)hi * (1.0 / 16777216.0); // the product is exact, and so is the difference
	lo = llrint(r * 295147905179352825856.0);
#endif
}

// (lo, hi) -> the double nearest to (hi * 2^44 + lo) * 2^-68
RT_HD double acc_merge(long long lo, long long hi)
{
	// 128-bit two's complement sum in two 64-bit words: (hi << 44) + lo
	unsigned long long w0 = ((unsigned long long)hi << RT_ACC_LOW_BITS);
	long long w1 = hi >> (64 - RT_ACC_LOW_BITS); // arithmetic shift: sign extended
	const unsigned long long s0 = w0 + (unsigned long long)lo;
	w1 += (lo < 0 ? -1ll : 0ll) + (s0 < w0 ? 1ll : 0ll);
	w0 = s0;
	const bool neg = w1 < 0;
	unsigned long long a1 = (unsigned long long)w1, a0 = w0;
	if (neg) { // magnitude
		a0 = ~a0 + 1ull;
		a1 = ~a1 + (a0 == 0 ? 1ull : 0ull);
	}
	if ((a0 | a1) == 0) return 0.0;
	// the top 64 bits of the magnitude with everything below them folded into the last bit (sticky):
	// the conversion to double then rounds once, to nearest even, as if it had seen all 128 bits
	unsigned long long top;
	int shift; // magnitude ~= top * 2^shift
	if (a1 == 0) {
		top = a0;
		shift = 0;
	} else {
#ifdef __CUDA_ARCH__
		const int lz = __clzll((long long)a1);
#else
		const int lz = __builtin_clzll(a1);
#endif
		if (lz == 0) {
			top = a1 | (a0 != 0 ? 1ull : 0ull);
		} else {
			top = (a1 << lz) | (a0 >> (64 - lz));
			if ((a0 << lz) != 0) top |= 1ull;
		}
		shift = 64 - lz;
	}
	// * 2^(shift - 68): an exact power of two built from its exponent field (0 <= shift <= 64: always a normal number)
	const unsigned long long pw = (unsigned long long)(1023 + shift + RT_ACC_UNIT_EXP) << 52;
	double scale;
	memcpy(&scale, &pw, 8);
	const double v = (double)top * scale;
	return neg ? -v : v;
}

#ifdef __CUDACC__
__device__ __forceinline__ void acc_add(long long* ch, double x)
{
	long long lo, hi;
	acc_split(x, lo, hi);
	if (lo) atomicAdd((unsigned long long*)ch, (unsigned long long)lo);
	if (hi) atomicAdd((unsigned long long*)ch + 1, (unsigned long long)hi);
}
#endif
RT_HD double acc_read(const long long* ch) { return acc_merge(ch[0], ch[1]); }
// One sample's colour.  `base` (optional, 3 doubles per camera-ray slot, `lenses` slots per sample) holds the terms that were
// STORED rather than added - each camera ray's own term: the slot is its alone, so the term needs no atomic - and they join
// the limbs here, as more integer additions.
RT_HD d3 acc_read3(const long long* acc, const double* base, int lenses, long long sample)
{
	const long long* a = acc + RT_ACC_STRIDE * sample;
	double v[3];
	for (int c = 0; c < 3; c++) {
		unsigned long long l0 = (unsigned long long)a[2 * c], l1 = (unsigned long long)a[2 * c + 1];
		if (base)
			for (int l = 0; l < lenses; l++) {
				long long lo, hi;
				acc_split(base[3 * (sample * lenses + l) + c], lo, hi);
				l0 += (unsigned long long)lo;
				l1 += (unsigned long long)hi;
			}
		v[c] = acc_merge((long long)l0, (long long)l1);
	}
	return mk3(v[0], v[1], v[2]);
}

} // namespace rtd
