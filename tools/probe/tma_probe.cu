// Dev probe (not part of the library): semantics of cp.async.bulk.tensor.3d on a float tensor — negative coordinates,
// odd box sizes, element strides.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o tma_probe tma_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(const __grid_constant__ CUtensorMap tm, int c0, int c1, int c2, int words, uint32_t tx_bytes, float* out, int* status) {
	extern __shared__ __align__(128) unsigned char raw[];
	float* buf = reinterpret_cast<float*>(raw);
	uint64_t* bar = reinterpret_cast<uint64_t*>(raw + ((words * 4 + 127) & ~127));
	if (threadIdx.x == 0) {
		for (int i = 0; i < words; i++) buf[i] = -777.0f;
		asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(bar)));
		asm volatile("fence.mbarrier_init.release.cluster;");
		asm volatile("fence.proxy.async.shared::cta;");
		asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(tx_bytes));
		asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(s32(buf)),
		             "l"(reinterpret_cast<uint64_t>(&tm)), "r"(s32(bar)), "r"(c0), "r"(c1), "r"(c2)
		             : "memory");
		int ok = 0;
		for (int spin = 0; spin < 2000000 && !ok; spin++) {
			uint32_t done;
			asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(s32(bar)) : "memory");
			ok = done;
		}
		*status = ok;
		for (int i = 0; i < words; i++) out[i] = buf[i];
	}
}

typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                        const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
	const int only = argc > 1 ? atoi(argv[1]) : -1;
	const int W = 400, H = 300, F = 2;
	std::vector<float> h((size_t)W * H * F);
	for (int f = 0; f < F; f++) for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) h[((size_t)f * H + y) * W + x] = f * 1000000.f + y * 1000.f + x;
	float* d; cudaMalloc(&d, h.size() * 4); cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
	void* p = nullptr; cudaDriverEntryPointQueryResult q;
	cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
	Enc enc = (Enc)p;
	struct Case { int bw, bh, estr, c0, c1, c2; const char* name; } cases[] = {
	    {92, 43, 1, 0, 0, 0, "box 92x43 at origin"},      {92, 43, 1, -14, -14, 1, "box 92x43 at (-14,-14), frame 1"},
	    {92, 43, 1, 0, 0, 1, "frame 1 only"},              {92, 43, 1, -14, 0, 0, "x=-14 only"},
	    {92, 43, 1, -16, 0, 0, "x=-16 only"},              {92, 43, 1, 0, -14, 0, "y=-14 only"},
	    {64, 8, 1, 4, 4, 0, "box 64x8"},                   {120, 8, 2, 3, 2, 0, "box 120x8 element stride 2 at x=3"},
	    {120, 8, 2, -5, -1, 0, "stride 2 at x=-5,y=-1"},   {92, 43, 1, 380, 290, 0, "box crossing the right/bottom edge"}};
	float* dout; int* dst; cudaMalloc(&dout, 1 << 20); cudaMalloc(&dst, 4);
	int idx = -1;
	for (auto& c : cases) {
		idx++;
		if (only >= 0 && idx != only) continue;
		CUtensorMap tm;
		cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)F};
		cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
		cuuint32_t box[3] = {(cuuint32_t)c.bw, (cuuint32_t)c.bh, 1};
		cuuint32_t es[3] = {(cuuint32_t)c.estr, 1, 1};
		CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
		                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
		const int loaded_w = (c.bw + c.estr - 1) / c.estr;
		const int words = loaded_w * c.bh;
		printf("== %s: encode=%d, expecting %d x %d words\n", c.name, (int)r, loaded_w, c.bh);
		if (r != CUDA_SUCCESS) continue;
		cudaMemset(dst, 0xff, 4);
		probe<<<1, 32, ((words * 4 + 127) & ~127) + 64>>>(tm, c.c0, c.c1, c.c2, words, (uint32_t)words * 4, dout, dst);
		cudaError_t e = cudaDeviceSynchronize();
		int st = -5; std::vector<float> o(words);
		if (e == cudaSuccess) { cudaMemcpy(&st, dst, 4, cudaMemcpyDeviceToHost); cudaMemcpy(o.data(), dout, words * 4, cudaMemcpyDeviceToHost); }
		printf("   sync=%s completed=%d\n", cudaGetErrorString(e), st);
		if (e != cudaSuccess) return 1;
		printf("   row0: %.0f %.0f %.0f %.0f ... row1: %.0f %.0f | last row: %.0f %.0f\n", o[0], o[1], o[2], o[3], o[loaded_w], o[loaded_w + 1],
		       o[(size_t)loaded_w * (c.bh - 1)], o[(size_t)loaded_w * (c.bh - 1) + 1]);
		int row_first = -1; for (int i = 0; i < words; i++) if (o[i] != 0.f && o[i] != -777.f) { row_first = i; break; }
		printf("   first non-zero word %d (row %d col %d) value %.0f; untouched words: ", row_first, row_first / loaded_w, row_first % loaded_w, row_first >= 0 ? o[row_first] : 0.f);
		int unt = 0; for (int i = 0; i < words; i++) unt += o[i] == -777.f; printf("%d\n", unt);
	}
	return 0;
}
