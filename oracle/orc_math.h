// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the shipped product path; only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// orc_math.h: atan2 / sin / cos exactly as Java's StrictMath specifies them (the freely distributable
// "fdlibm 5.3" algorithms: e_atan2.c, s_atan.c, s_sin.c, s_cos.c, k_sin.c, k_cos.c, e_rem_pio2.c).
// The reference calls Math.atan2 / Math.cos / Math.sin at
//   ImplOrientationAverageGradientIntegral.java:101,126   (atan2 of the summed gradient)
//   ImplOrientationSlidingWindowIntegral.java:100,187     (atan2 per sample and of the best window)
//   DescribePointSurf.java:191                            (cos/sin of the orientation)
// Java allows Math.* to differ from StrictMath.* by 1 ulp; StrictMath is the specified algorithm, so it
// is what is restated here.  Every operation is a plain IEEE-754 double op (build with -ffp-contract=off),
// which lets the CUDA side (boofcv_b200/csrc/bs_math.cuh, written independently, -fmad=false) be
// bit-identical to this file.  Checked against glibc to <= 1 ulp in tests/test_oracle_math.py.
#pragma once
#include <cstdint>
#include <cstring>
#include <cmath>

namespace orc {

static inline int32_t hi_word(double x) { uint64_t u; std::memcpy(&u, &x, 8); return (int32_t)(u >> 32); }
static inline uint32_t lo_word(double x) { uint64_t u; std::memcpy(&u, &x, 8); return (uint32_t)u; }
static inline double from_words(int32_t hi, uint32_t lo) {
	uint64_t u = ((uint64_t)(uint32_t)hi << 32) | lo; double d; std::memcpy(&d, &u, 8); return d;
}

// ---- s_atan.c ----
static inline double fd_atan(double x) {
	static const double atanhi[4] = {4.63647609000806093515e-01, 7.85398163397448278999e-01,
	                                 9.82793723247329054082e-01, 1.57079632679489655800e+00};
	static const double atanlo[4] = {2.26987774529616870924e-17, 3.06161699786838301793e-17,
	                                 1.39033110312309984516e-17, 6.12323399573676603587e-17};
	static const double aT[11] = {3.33333333333329318027e-01, -1.99999999998764832476e-01,
	                              1.42857142725034663711e-01, -1.11111104054623557880e-01,
	                              9.09088713343650656196e-02, -7.69187620504482999495e-02,
	                              6.66107313738753120669e-02, -5.83357013379057348645e-02,
	                              4.97687799461593236017e-02, -3.65315727442169155270e-02,
	                              1.62858201153657823623e-02};
	int32_t hx = hi_word(x), ix = hx & 0x7fffffff;
	int id;
	if (ix >= 0x44100000) {  // |x| >= 2^66
		if (ix > 0x7ff00000 || (ix == 0x7ff00000 && lo_word(x) != 0)) return x + x;  // NaN
		return hx > 0 ? atanhi[3] + atanlo[3] : -atanhi[3] - atanlo[3];
	}
	if (ix < 0x3fdc0000) {      // |x| < 0.4375
		if (ix < 0x3e200000) return x;  // |x| < 2^-29
		id = -1;
	} else {
		x = std::fabs(x);
		if (ix < 0x3ff30000) {          // |x| < 1.1875
			if (ix < 0x3fe60000) { id = 0; x = (2.0 * x - 1.0) / (2.0 + x); }
			else                 { id = 1; x = (x - 1.0) / (x + 1.0); }
		} else {
			if (ix < 0x40038000) { id = 2; x = (x - 1.5) / (1.0 + 1.5 * x); }
			else                 { id = 3; x = -1.0 / x; }
		}
	}
	double z = x * x;
	double w = z * z;
	double s1 = z * (aT[0] + w * (aT[2] + w * (aT[4] + w * (aT[6] + w * (aT[8] + w * aT[10])))));
	double s2 = w * (aT[1] + w * (aT[3] + w * (aT[5] + w * (aT[7] + w * aT[9]))));
	if (id < 0) return x - x * (s1 + s2);
	z = atanhi[id] - ((x * (s1 + s2) - atanlo[id]) - x);
	return hx < 0 ? -z : z;
}

// ---- e_atan2.c ----
static inline double fd_atan2(double y, double x) {
	const double tiny = 1.0e-300, pi_o_4 = 7.8539816339744827900E-01, pi_o_2 = 1.5707963267948965580E+00,
	             pi = 3.1415926535897931160E+00, pi_lo = 1.2246467991473531772E-16;
	int32_t hx = hi_word(x), ix = hx & 0x7fffffff; uint32_t lx = lo_word(x);
	int32_t hy = hi_word(y), iy = hy & 0x7fffffff; uint32_t ly = lo_word(y);
	if (((uint32_t)ix | ((lx | (0u - lx)) >> 31)) > 0x7ff00000u ||
	    ((uint32_t)iy | ((ly | (0u - ly)) >> 31)) > 0x7ff00000u) return x + y;  // NaN
	if (((uint32_t)(hx - 0x3ff00000) | lx) == 0) return fd_atan(y);            // x == 1.0
	int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);                                 // 2*sign(x)+sign(y)
	if (((uint32_t)iy | ly) == 0) {                                              // y == 0
		switch (m) { case 0: case 1: return y; case 2: return pi + tiny; default: return -pi - tiny; }
	}
	if (((uint32_t)ix | lx) == 0) return hy < 0 ? -pi_o_2 - tiny : pi_o_2 + tiny;  // x == 0
	if (ix == 0x7ff00000) {                                                      // x is INF
		if (iy == 0x7ff00000) {
			switch (m) { case 0: return pi_o_4 + tiny; case 1: return -pi_o_4 - tiny;
			             case 2: return 3.0 * pi_o_4 + tiny; default: return -3.0 * pi_o_4 - tiny; }
		} else {
			switch (m) { case 0: return 0.0; case 1: return -0.0; case 2: return pi + tiny; default: return -pi - tiny; }
		}
	}
	if (iy == 0x7ff00000) return hy < 0 ? -pi_o_2 - tiny : pi_o_2 + tiny;       // y is INF
	int k = (iy - ix) >> 20;
	double z;
	if (k > 60) z = pi_o_2 + 0.5 * pi_lo;
	else if (hx < 0 && k < -60) z = 0.0;
	else z = fd_atan(std::fabs(y / x));
	switch (m) {
		case 0: return z;
		case 1: return from_words(hi_word(z) ^ (int32_t)0x80000000, lo_word(z));
		case 2: return pi - (z - pi_lo);
		default: return (z - pi_lo) - pi;
	}
}

// ---- k_sin.c / k_cos.c ----
static inline double fd_kernel_sin(double x, double y, int iy) {
	const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
	             S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
	             S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
	int32_t ix = hi_word(x) & 0x7fffffff;
	if (ix < 0x3e400000) { if ((int)x == 0) return x; }
	double z = x * x, v = z * x;
	double r = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)));
	if (iy == 0) return x + v * (S1 + z * r);
	return x - ((z * (0.5 * y - v * r) - y) - v * S1);
}

static inline double fd_kernel_cos(double x, double y) {
	const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
	             C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
	             C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
	int32_t ix = hi_word(x) & 0x7fffffff;
	if (ix < 0x3e400000) { if ((int)x == 0) return 1.0; }
	double z = x * x;
	double r = z * (C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6)))));
	if (ix < 0x3FD33333) return 1.0 - (0.5 * z - (z * r - x * y));
	double qx = (ix > 0x3fe90000) ? 0.28125 : from_words(ix - 0x00200000, 0);
	double hz = 0.5 * z - qx, a = 1.0 - qx;
	return a - (hz - (z * r - x * y));
}

// ---- e_rem_pio2.c, small and medium ranges (|x| <= 2^19*pi/2); returns n, y[0]+y[1] = x - n*pi/2 ----
static inline int fd_rem_pio2(double x, double* y) {
	static const int32_t npio2_hw[32] = {
	    0x3FF921FB, 0x400921FB, 0x4012D97C, 0x401921FB, 0x401F6A7A, 0x4022D97C, 0x4025FDBB, 0x402921FB,
	    0x402C463A, 0x402F6A7A, 0x4031475C, 0x4032D97C, 0x40346B9C, 0x4035FDBB, 0x40378FDB, 0x403921FB,
	    0x403AB41B, 0x403C463A, 0x403DD85A, 0x403F6A7A, 0x40407E4C, 0x4041475C, 0x4042106C, 0x4042D97C,
	    0x4043A28C, 0x40446B9C, 0x404534AC, 0x4045FDBB, 0x4046C6CB, 0x40478FDB, 0x404858EB, 0x404921FB};
	const double invpio2 = 6.36619772367581382433e-01, pio2_1 = 1.57079632673412561417e+00,
	             pio2_1t = 6.07710050650619224932e-11, pio2_2 = 6.07710050630396597660e-11,
	             pio2_2t = 2.02226624879595063154e-21, pio2_3 = 2.02226624871116645580e-21,
	             pio2_3t = 8.47842766036889956997e-32;
	int32_t hx = hi_word(x), ix = hx & 0x7fffffff;
	if (ix <= 0x3fe921fb) { y[0] = x; y[1] = 0; return 0; }
	if (ix < 0x4002d97c) {  // |x| < 3pi/4
		if (hx > 0) {
			double z = x - pio2_1;
			if (ix != 0x3ff921fb) { y[0] = z - pio2_1t; y[1] = (z - y[0]) - pio2_1t; }
			else { z -= pio2_2; y[0] = z - pio2_2t; y[1] = (z - y[0]) - pio2_2t; }
			return 1;
		} else {
			double z = x + pio2_1;
			if (ix != 0x3ff921fb) { y[0] = z + pio2_1t; y[1] = (z - y[0]) + pio2_1t; }
			else { z += pio2_2; y[0] = z + pio2_2t; y[1] = (z - y[0]) + pio2_2t; }
			return -1;
		}
	}
	// medium size; callers guarantee ix <= 0x413921fb
	double t = std::fabs(x);
	int n = (int)(t * invpio2 + 0.5);
	double fn = (double)n;
	double r = t - fn * pio2_1;
	double w = fn * pio2_1t;
	if (n < 32 && ix != npio2_hw[n - 1]) {
		y[0] = r - w;
	} else {
		int j = ix >> 20;
		y[0] = r - w;
		int i = j - ((hi_word(y[0]) >> 20) & 0x7ff);
		if (i > 16) {
			t = r; w = fn * pio2_2; r = t - w; w = fn * pio2_2t - ((t - r) - w); y[0] = r - w;
			i = j - ((hi_word(y[0]) >> 20) & 0x7ff);
			if (i > 49) { t = r; w = fn * pio2_3; r = t - w; w = fn * pio2_3t - ((t - r) - w); y[0] = r - w; }
		}
	}
	y[1] = (r - y[0]) - w;
	if (hx < 0) { y[0] = -y[0]; y[1] = -y[1]; return -n; }
	return n;
}

// ---- s_sin.c / s_cos.c.  Beyond 2^19*pi/2 (never reached on this path: arguments are atan2 results)
// the huge-argument Payne-Hanek branch is not restated; libm is used there. ----
static inline double fd_sin(double x) {
	int32_t ix = hi_word(x) & 0x7fffffff;
	if (ix <= 0x3fe921fb) return fd_kernel_sin(x, 0.0, 0);
	if (ix >= 0x7ff00000) return x - x;
	if (ix > 0x413921fb) return std::sin(x);
	double y[2]; int n = fd_rem_pio2(x, y);
	switch (n & 3) {
		case 0: return fd_kernel_sin(y[0], y[1], 1);
		case 1: return fd_kernel_cos(y[0], y[1]);
		case 2: return -fd_kernel_sin(y[0], y[1], 1);
		default: return -fd_kernel_cos(y[0], y[1]);
	}
}

static inline double fd_cos(double x) {
	int32_t ix = hi_word(x) & 0x7fffffff;
	if (ix <= 0x3fe921fb) return fd_kernel_cos(x, 0.0);
	if (ix >= 0x7ff00000) return x - x;
	if (ix > 0x413921fb) return std::cos(x);
	double y[2]; int n = fd_rem_pio2(x, y);
	switch (n & 3) {
		case 0: return fd_kernel_cos(y[0], y[1]);
		case 1: return -fd_kernel_sin(y[0], y[1], 1);
		case 2: return -fd_kernel_cos(y[0], y[1]);
		default: return fd_kernel_sin(y[0], y[1], 1);
	}
}

}  // namespace orc
