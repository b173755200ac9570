// bs_box.cuh — integral-image box sums and the sparse gradient operator, in the reference's evaluation order.
//   ImplIntegralImageOps.block_unsafe / block_zero / convolveSparse
//     (IP/alg/transform/ii/impl/ImplIntegralImageOps.java:172-214 F32, :216-258 S32)
//   SparseIntegralGradient_NoBorder_F32/_I32.compute (IP/alg/transform/ii/impl/SparseIntegralGradient_NoBorder_F32.java:46-76)
//   SparseScaleGradient.isInBounds (main/boofcv-types/src/main/java/boofcv/struct/sparse/SparseScaleGradient.java:47-50)
#pragma once
#include <cstdint>

namespace bs {

template <typename T> struct Arith;
template <> struct Arith<float> {
	static __device__ __forceinline__ float sum4(float br, float tr, float bl, float tl) { return br - tr - bl + tl; }
	static __device__ __forceinline__ float sub(float a, float b) { return a - b; }
};
template <> struct Arith<int32_t> {
	static __device__ __forceinline__ int32_t sum4(int32_t br, int32_t tr, int32_t bl, int32_t tl) {
		return (int32_t)((uint32_t)br - (uint32_t)tr - (uint32_t)bl + (uint32_t)tl);
	}
	static __device__ __forceinline__ int32_t sub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
};

// block_unsafe: corners known to be inside the image
template <typename T>
__device__ __forceinline__ T box_in(const T* __restrict__ ii, int W, int x0, int y0, int x1, int y1) {
	const T* r0 = ii + (long)y0 * W;
	const T* r1 = ii + (long)y1 * W;
	return Arith<T>::sum4(__ldg(r1 + x1), __ldg(r0 + x1), __ldg(r1 + x0), __ldg(r0 + x0));
}
// block_zero: clamp to the last row/column, zero where a coordinate is negative
template <typename T>
__device__ __forceinline__ T box_zero(const T* __restrict__ ii, int W, int H, int x0, int y0, int x1, int y1) {
	x0 = min(x0, W - 1); y0 = min(y0, H - 1); x1 = min(x1, W - 1); y1 = min(y1, H - 1);
	T br = 0, tr = 0, bl = 0, tl = 0;
	if (x1 >= 0 && y1 >= 0) br = __ldg(ii + (long)y1 * W + x1);
	if (y0 >= 0 && x1 >= 0) tr = __ldg(ii + (long)y0 * W + x1);
	if (x0 >= 0 && y1 >= 0) bl = __ldg(ii + (long)y1 * W + x0);
	if (x0 >= 0 && y0 >= 0) tl = __ldg(ii + (long)y0 * W + x0);
	return Arith<T>::sum4(br, tr, bl, tl);
}

// Sparse gradient of kernel radius r at (x,y): 12 integral reads; (0,0) when the kernel leaves the image
// (SparseGradientSafe.compute, main/boofcv-types/src/main/java/boofcv/struct/sparse/SparseGradientSafe.java:56-61).
// The reference also has an unchecked path that it only takes after SurfDescribeOps.isInside proved every sample is
// inside; the check is kept here on both paths because a GPU thread must never read outside the allocation.
// Index arithmetic is 32-bit on purpose (one IMAD.WIDE per load instead of 64-bit multiplies and adds: ncu's source page
// had the address code at 12 % of describe_kernel's instructions): W*H < 2^31 is checked where the image enters the library.
template <typename T>
__device__ __forceinline__ void sparse_gradient(const T* __restrict__ ii, int W, int H, int r, int x, int y, T& dx, T& dy) {
	const int x0 = x - r - 1, y0 = y - r - 1;
	if ((x0 | y0) < 0 || x + r >= W || y + r >= H) { dx = 0; dy = 0; return; }
	const int w = 2 * r + 1;
	const int a = y0 * W + x0;
	const int b = a + r * W;
	const int c = b + W;
	const int d = c + r * W;
	const T p0 = __ldg(ii + a), p1 = __ldg(ii + (a + r)), p2 = __ldg(ii + (a + r + 1)), p3 = __ldg(ii + (a + w));
	const T p11 = __ldg(ii + b), p4 = __ldg(ii + (b + w));
	const T p10 = __ldg(ii + c), p5 = __ldg(ii + (c + w));
	const T p9 = __ldg(ii + d), p8 = __ldg(ii + (d + r)), p7 = __ldg(ii + (d + r + 1)), p6 = __ldg(ii + (d + w));
	const T left = Arith<T>::sum4(p8, p9, p1, p0);
	const T right = Arith<T>::sum4(p6, p7, p3, p2);
	const T top = Arith<T>::sum4(p4, p11, p3, p0);
	const T bottom = Arith<T>::sum4(p6, p9, p5, p10);
	dx = Arith<T>::sub(right, left);      // the reference widens these to double at the call site
	dy = Arith<T>::sub(bottom, top);
}

// The Haar wavelet of ConfigSurfDescribe.useHaar (SparseIntegralHaar_NoBorder_F32.java:37-73, _I32): radius r, width 2r, eight
// integral reads; bounds x0 = y0 = -r, x1 = y1 = r (SparseScaleGradient.isInBounds); (0,0) outside, like SparseGradientSafe.
template <typename T>
__device__ __forceinline__ void sparse_haar(const T* __restrict__ ii, int W, int H, int r, int x, int y, T& dx, T& dy) {
	const int x0 = x - r, y0 = y - r;
	if ((x0 | y0) < 0 || x + r >= W || y + r >= H) { dx = 0; dy = 0; return; }
	const int w = 2 * r;
	const int a = y0 * W + x0, b = y * W + x0, c = (y + r) * W + x0;
	const T p0 = __ldg(ii + a), p1 = __ldg(ii + (a + r)), p2 = __ldg(ii + (a + w));
	const T p3 = __ldg(ii + b), p5 = __ldg(ii + (b + w));
	const T p6 = __ldg(ii + c), p7 = __ldg(ii + (c + r)), p8 = __ldg(ii + (c + w));
	const T left = Arith<T>::sum4(p7, p1, p6, p0);
	const T right = Arith<T>::sum4(p8, p2, p7, p1);
	const T top = Arith<T>::sum4(p5, p2, p3, p0);
	const T bottom = Arith<T>::sum4(p8, p5, p6, p3);
	dx = Arith<T>::sub(right, left);
	dy = Arith<T>::sub(bottom, top);
}

// SparseIntegralGradient_NoBorder.setWidth (IP/alg/transform/ii/SparseIntegralGradient_NoBorder.java:41-47)
__device__ __forceinline__ int gradient_radius(double width) {
	int r = ((int)(width + 0.5)) / 2;
	if (r <= 0) r = 1;
	return r;
}

}  // namespace bs
