// bs_math.cuh — device atan2 / sin / cos that follow Java's StrictMath specification (the public fdlibm 5.3
// algorithms) operation for operation.  The reference calls Math.atan2/cos/sin at
//   F/alg/feature/orientation/impl/ImplOrientationAverageGradientIntegral.java:101,126
//   F/alg/feature/orientation/impl/ImplOrientationSlidingWindowIntegral.java:100,187
//   F/alg/feature/describe/DescribePointSurf.java:191
// CUDA's own atan2()/sincos() differ from fdlibm in the last ulp, and a 1-ulp change of the angle can move a descriptor
// sample across an (int) truncation boundary, so the product carries its own implementation.
// Every step is a single IEEE-754 double operation; this TU is compiled with -fmad=false so nothing is contracted.
#pragma once
#include <cstdint>

namespace bs {

__device__ __forceinline__ int dhi(double x) { return __double2hiint(x); }
__device__ __forceinline__ unsigned dlo(double x) { return (unsigned)__double2loint(x); }
__device__ __forceinline__ double dmake(int hi, unsigned lo) { return __hiloint2double(hi, (int)lo); }

static __constant__ double k_atan_hi[4] = {4.63647609000806093515e-01, 7.85398163397448278999e-01,
                                           9.82793723247329054082e-01, 1.57079632679489655800e+00};
static __constant__ double k_atan_lo[4] = {2.26987774529616870924e-17, 3.06161699786838301793e-17,
                                           1.39033110312309984516e-17, 6.12323399573676603587e-17};
// high words of n*pi/2 for n = 1..32
static __constant__ int k_npio2_hw[32] = {0x3FF921FB, 0x400921FB, 0x4012D97C, 0x401921FB, 0x401F6A7A, 0x4022D97C, 0x4025FDBB,
                                          0x402921FB, 0x402C463A, 0x402F6A7A, 0x4031475C, 0x4032D97C, 0x40346B9C, 0x4035FDBB,
                                          0x40378FDB, 0x403921FB, 0x403AB41B, 0x403C463A, 0x403DD85A, 0x403F6A7A, 0x40407E4C,
                                          0x4041475C, 0x4042106C, 0x4042D97C, 0x4043A28C, 0x40446B9C, 0x404534AC, 0x4045FDBB,
                                          0x4046C6CB, 0x40478FDB, 0x404858EB, 0x404921FB};

// atan(x): argument reduction to one of four break points, then an odd/even split degree-11 polynomial in x^2
__device__ inline double sm_atan(double x) {
	const double* hi_tab = k_atan_hi;
	const double* lo_tab = k_atan_lo;
	const double c0 = 3.33333333333329318027e-01, c1 = -1.99999999998764832476e-01, c2 = 1.42857142725034663711e-01,
	             c3 = -1.11111104054623557880e-01, c4 = 9.09088713343650656196e-02, c5 = -7.69187620504482999495e-02,
	             c6 = 6.66107313738753120669e-02, c7 = -5.83357013379057348645e-02, c8 = 4.97687799461593236017e-02,
	             c9 = -3.65315727442169155270e-02, c10 = 1.62858201153657823623e-02;
	const int hx = dhi(x);
	const int ax = hx & 0x7fffffff;
	int seg;
	if (ax >= 0x44100000) {   // |x| >= 2^66, inf, nan
		if (ax > 0x7ff00000 || (ax == 0x7ff00000 && dlo(x) != 0)) return x + x;
		return hx > 0 ? hi_tab[3] + lo_tab[3] : -hi_tab[3] - lo_tab[3];
	}
	if (ax < 0x3fdc0000) {    // |x| < 7/16
		if (ax < 0x3e200000) return x;
		seg = -1;
	} else {
		x = fabs(x);
		if (ax < 0x3ff30000) {
			if (ax < 0x3fe60000) { seg = 0; x = (2.0 * x - 1.0) / (2.0 + x); }
			else { seg = 1; x = (x - 1.0) / (x + 1.0); }
		} else {
			if (ax < 0x40038000) { seg = 2; x = (x - 1.5) / (1.0 + 1.5 * x); }
			else { seg = 3; x = -1.0 / x; }
		}
	}
	const double z = x * x;
	const double w = z * z;
	const double even = z * (c0 + w * (c2 + w * (c4 + w * (c6 + w * (c8 + w * c10)))));
	const double odd = w * (c1 + w * (c3 + w * (c5 + w * (c7 + w * c9))));
	if (seg < 0) return x - x * (even + odd);
	const double r = hi_tab[seg] - ((x * (even + odd) - lo_tab[seg]) - x);
	return hx < 0 ? -r : r;
}

__device__ inline double sm_atan2(double y, double x) {
	const double tiny = 1.0e-300, quarter_pi = 7.8539816339744827900E-01, half_pi = 1.5707963267948965580E+00,
	             pi = 3.1415926535897931160E+00, pi_tail = 1.2246467991473531772E-16;
	const int hx = dhi(x), hy = dhi(y);
	const int ax = hx & 0x7fffffff, ay = hy & 0x7fffffff;
	const unsigned lx = dlo(x), ly = dlo(y);
	if (((unsigned)ax | ((lx | (0u - lx)) >> 31)) > 0x7ff00000u || ((unsigned)ay | ((ly | (0u - ly)) >> 31)) > 0x7ff00000u)
		return x + y;
	if ((((unsigned)(hx - 0x3ff00000)) | lx) == 0) return sm_atan(y);
	const int quad = ((hy >> 31) & 1) | ((hx >> 30) & 2);
	if (((unsigned)ay | ly) == 0) {
		if (quad < 2) return y;
		return quad == 2 ? pi + tiny : -pi - tiny;
	}
	if (((unsigned)ax | lx) == 0) return hy < 0 ? -half_pi - tiny : half_pi + tiny;
	if (ax == 0x7ff00000) {
		if (ay == 0x7ff00000) {
			if (quad == 0) return quarter_pi + tiny;
			if (quad == 1) return -quarter_pi - tiny;
			if (quad == 2) return 3.0 * quarter_pi + tiny;
			return -3.0 * quarter_pi - tiny;
		}
		if (quad == 0) return 0.0;
		if (quad == 1) return -0.0;
		return quad == 2 ? pi + tiny : -pi - tiny;
	}
	if (ay == 0x7ff00000) return hy < 0 ? -half_pi - tiny : half_pi + tiny;
	const int k = (ay - ax) >> 20;
	double z;
	if (k > 60) z = half_pi + 0.5 * pi_tail;
	else if (hx < 0 && k < -60) z = 0.0;
	else z = sm_atan(fabs(y / x));
	if (quad == 0) return z;
	if (quad == 1) return dmake(dhi(z) ^ (int)0x80000000, dlo(z));
	if (quad == 2) return pi - (z - pi_tail);
	return (z - pi_tail) - pi;
}

// sin/cos kernels on [-pi/4, pi/4] with a tail term
__device__ inline double sm_ksin(double x, double tail, int have_tail) {
	const double s1 = -1.66666666666666324348e-01, s2 = 8.33333333332248946124e-03, s3 = -1.98412698298579493134e-04,
	             s4 = 2.75573137070700676789e-06, s5 = -2.50507602534068634195e-08, s6 = 1.58969099521155010221e-10;
	if ((dhi(x) & 0x7fffffff) < 0x3e400000) { if ((int)x == 0) return x; }
	const double z = x * x, v = z * x;
	const double r = s2 + z * (s3 + z * (s4 + z * (s5 + z * s6)));
	if (!have_tail) return x + v * (s1 + z * r);
	return x - ((z * (0.5 * tail - v * r) - tail) - v * s1);
}

__device__ inline double sm_kcos(double x, double tail) {
	const double k1 = 4.16666666666666019037e-02, k2 = -1.38888888888741095749e-03, k3 = 2.48015872894767294178e-05,
	             k4 = -2.75573143513906633035e-07, k5 = 2.08757232129817482790e-09, k6 = -1.13596475577881948265e-11;
	const int ax = dhi(x) & 0x7fffffff;
	if (ax < 0x3e400000) { if ((int)x == 0) return 1.0; }
	const double z = x * x;
	const double r = z * (k1 + z * (k2 + z * (k3 + z * (k4 + z * (k5 + z * k6)))));
	if (ax < 0x3FD33333) return 1.0 - (0.5 * z - (z * r - x * tail));
	const double qx = (ax > 0x3fe90000) ? 0.28125 : dmake(ax - 0x00200000, 0u);
	const double hz = 0.5 * z - qx, a = 1.0 - qx;
	return a - (hz - (z * r - x * tail));
}

// x - n*pi/2 as head+tail for |x| <= 2^19*pi/2 (orientation angles are atan2 results, |x| <= pi)
__device__ inline int sm_reduce_pio2(double x, double& head, double& tail) {
	const double inv_pio2 = 6.36619772367581382433e-01, p1 = 1.57079632673412561417e+00, p1t = 6.07710050650619224932e-11,
	             p2 = 6.07710050630396597660e-11, p2t = 2.02226624879595063154e-21, p3 = 2.02226624871116645580e-21,
	             p3t = 8.47842766036889956997e-32;
	const int hx = dhi(x), ax = hx & 0x7fffffff;
	if (ax <= 0x3fe921fb) { head = x; tail = 0; return 0; }
	if (ax < 0x4002d97c) {
		if (hx > 0) {
			double z = x - p1;
			if (ax != 0x3ff921fb) { head = z - p1t; tail = (z - head) - p1t; }
			else { z -= p2; head = z - p2t; tail = (z - head) - p2t; }
			return 1;
		}
		double z = x + p1;
		if (ax != 0x3ff921fb) { head = z + p1t; tail = (z - head) + p1t; }
		else { z += p2; head = z + p2t; tail = (z - head) + p2t; }
		return -1;
	}
	double t = fabs(x);
	const int n = (int)(t * inv_pio2 + 0.5);
	const double fn = (double)n;
	double r = t - fn * p1;
	double w = fn * p1t;
	// when |x| is within a high word of a multiple of pi/2 the first subtraction cancels badly: extend the reduction
	bool near_multiple = true;
	if (n < 32) near_multiple = (ax == k_npio2_hw[n - 1]);
	head = r - w;
	if (near_multiple) {
		const int j = ax >> 20;
		int lost = j - ((dhi(head) >> 20) & 0x7ff);
		if (lost > 16) {
			t = r; w = fn * p2; r = t - w; w = fn * p2t - ((t - r) - w); head = r - w;
			lost = j - ((dhi(head) >> 20) & 0x7ff);
			if (lost > 49) { t = r; w = fn * p3; r = t - w; w = fn * p3t - ((t - r) - w); head = r - w; }
		}
	}
	tail = (r - head) - w;
	if (hx < 0) { head = -head; tail = -tail; return -n; }
	return n;
}

// Arguments beyond 2^19*pi/2 need Payne-Hanek reduction, which this path never reaches (angles come from atan2);
// CUDA's sin/cos are used there.
__device__ inline double sm_sin(double x) {
	const int ax = dhi(x) & 0x7fffffff;
	if (ax <= 0x3fe921fb) return sm_ksin(x, 0.0, 0);
	if (ax >= 0x7ff00000) return x - x;
	if (ax > 0x413921fb) return sin(x);
	double h, t;
	const int n = sm_reduce_pio2(x, h, t) & 3;
	if (n == 0) return sm_ksin(h, t, 1);
	if (n == 1) return sm_kcos(h, t);
	if (n == 2) return -sm_ksin(h, t, 1);
	return -sm_kcos(h, t);
}

__device__ inline double sm_cos(double x) {
	const int ax = dhi(x) & 0x7fffffff;
	if (ax <= 0x3fe921fb) return sm_kcos(x, 0.0);
	if (ax >= 0x7ff00000) return x - x;
	if (ax > 0x413921fb) return cos(x);
	double h, t;
	const int n = sm_reduce_pio2(x, h, t) & 3;
	if (n == 0) return sm_kcos(h, t);
	if (n == 1) return -sm_ksin(h, t, 1);
	if (n == 2) return -sm_kcos(h, t);
	return sm_ksin(h, t, 1);
}

}  // namespace bs
