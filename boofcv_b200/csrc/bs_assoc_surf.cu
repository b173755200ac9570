// bs_assoc_surf.cu — AssociateSurfBasic: features are split by the sign of the Laplacian, the two classes are associated
// independently and the indices are mapped back to the callers' lists.
// Reference: AssociateSurfBasic.sort (F/alg/feature/associate/AssociateSurfBasic.java:129-143: stable split, the Helper keeps
// the index in the input list), associate (:83-122: positive lists first, then negative; matches and unassociated sources
// are reported with their global indices).
//
// Device side: a stable partition of the index range by `white` (one CTA, ballot/popc scan — lists are at most a few
// hundred thousand entries), a row gather of the descriptors into the two contiguous class matrices the greedy kernels
// take, and a scatter that rewrites the per-class pairs[] with global indices.
#include "bs_common.cuh"

namespace bs {

constexpr int PART_THREADS = 1024;

// idx[0 .. n_white) = indices with white != 0 in ascending order, idx[n_white .. n) = the others in ascending order;
// counts[0] = n_white
__global__ void __launch_bounds__(PART_THREADS) partition_white_kernel(const uint8_t* __restrict__ white, int n, int32_t* __restrict__ idx,
                                                                       int32_t* __restrict__ counts) {
	__shared__ int warp_tot[PART_THREADS / 32];
	__shared__ int base_sh;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	// pass 0 counts the white entries, passes 1 and 2 place the two classes
	int n_white = 0;
	for (int pass = 0; pass < 3; pass++) {
		if (tid == 0) base_sh = pass == 2 ? n_white : 0;
		__syncthreads();
		for (int i0 = 0; i0 < n; i0 += PART_THREADS) {
			const int i = i0 + tid;
			const bool w = i < n && white[i] != 0;
			const bool mine = i < n && (pass == 2 ? !w : w);
			const unsigned bal = __ballot_sync(0xffffffffu, mine);
			if (lane == 0) warp_tot[warp] = __popc(bal);
			__syncthreads();
			int before = 0, total = 0;
			for (int k = 0; k < PART_THREADS / 32; k++) {
				const int t = warp_tot[k];
				before += k < warp ? t : 0;
				total += t;
			}
			const int base = base_sh;
			if (mine && pass > 0) idx[base + before + __popc(bal & ((1u << lane) - 1u))] = i;
			__syncthreads();
			if (tid == 0) base_sh = base + total;
		}
		__syncthreads();
		if (pass == 0) n_white = base_sh;
		__syncthreads();
	}
	if (tid == 0) counts[0] = n_white;
}

// out[r][:] = in[idx[r]][:]   (dof doubles per row; one warp per row)
__global__ void gather_rows_kernel(const double* __restrict__ in, const int32_t* __restrict__ idx, int n, int dof, double* __restrict__ out) {
	const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (r >= n) return;
	const double* src = in + (long)idx[r] * dof;
	double* dst = out + (long)r * dof;
	for (int k = threadIdx.x & 31; k < dof; k += 32) dst[k] = src[k];
}

// class-local result -> global: pairs[src_idx[r]] = dst_idx[pairs_c[r]] (or -1), fit likewise
__global__ void scatter_pairs_kernel(const int32_t* __restrict__ pairs_c, const double* __restrict__ fit_c, int n_c,
                                     const int32_t* __restrict__ src_idx, const int32_t* __restrict__ dst_idx, int32_t* __restrict__ pairs,
                                     double* __restrict__ fit) {
	const int r = blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= n_c) return;
	const int p = pairs_c[r];
	const int g = src_idx[r];
	pairs[g] = p >= 0 ? dst_idx[p] : -1;
	fit[g] = fit_c[r];
}

void launch_partition_white(const uint8_t* d_white, int n, int32_t* d_idx, int32_t* d_count, cudaStream_t st) {
	BS_LAUNCH(partition_white_kernel, 1, PART_THREADS, 0, st, d_white, n, d_idx, d_count);
}
void launch_gather_rows(const double* d_in, const int32_t* d_idx, int n, int dof, double* d_out, cudaStream_t st) {
	if (n <= 0) return;
	BS_LAUNCH(gather_rows_kernel, div_up(n, 8), 256, 0, st, d_in, d_idx, n, dof, d_out);
}
void launch_scatter_pairs(const int32_t* d_pairs_c, const double* d_fit_c, int n_c, const int32_t* d_src_idx, const int32_t* d_dst_idx,
                          int32_t* d_pairs, double* d_fit, cudaStream_t st) {
	if (n_c <= 0) return;
	BS_LAUNCH(scatter_pairs_kernel, div_up(n_c, 256), 256, 0, st, d_pairs_c, d_fit_c, n_c, d_src_idx, d_dst_idx, d_pairs, d_fit);
}

}  // namespace bs
