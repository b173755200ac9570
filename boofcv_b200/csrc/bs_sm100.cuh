// bs_sm100.cuh — thin inline-PTX wrappers for the sm_100a-only machinery used by the association filter:
// mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld) and the UMMA descriptors.
// Bit layouts follow the PTX ISA tcgen05 chapter (instruction descriptor, shared-memory matrix descriptor).
#pragma once
#include <cstdint>
#include <cuda.h>

namespace bs {
namespace sm100 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Bounded wait: a protocol bug must not hang the GPU (the box is shared), it traps instead.  The loop is kept small on purpose
// (no unrolling, no back-off ladder): try_wait already suspends the thread for a hardware time slice, and the first version of this
// wait — 64 unrolled polls before a nanosleep — made the filter kernel's epilogue loop 40 KB of code (12 % of its stall samples
// were instruction fetch).
__device__ __forceinline__ bool mbar_try_wait(uint32_t addr, uint32_t parity) {
	uint32_t done;
	asm volatile(
	    "{\n\t.reg .pred p;\n\t"
	    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
	    "selp.u32 %0, 1, 0, p;\n\t}"
	    : "=r"(done)
	    : "r"(addr), "r"(parity)
	    : "memory");
	return done != 0;
}
__device__ __forceinline__ void mbar_wait_addr(uint32_t addr, uint32_t parity) {
	uint32_t spin = 0;
#pragma unroll 1
	while (!mbar_try_wait(addr, parity))
		if (++spin == (1u << 24)) __trap();
}
// The same wait with a suspend-time hint: a failing poll may keep the warp asleep for up to `ns` (the hardware wakes it when the
// phase completes; ptxas emits NANOSLEEP.SYNCS), so a warp that waits long executes a handful of polls instead of one every ~60
// cycles — the counter arithmetic of the poll loop runs on the ALU pipe the filter's epilogue is short of.
__device__ __forceinline__ void mbar_wait_addr_sleepy(uint32_t addr, uint32_t parity, uint32_t ns) {
	uint32_t spin = 0;
#pragma unroll 1
	while (true) {
		uint32_t done;
		asm volatile(
		    "{\n\t.reg .pred p;\n\t"
		    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
		    "selp.u32 %0, 1, 0, p;\n\t}"
		    : "=r"(done)
		    : "r"(addr), "r"(parity), "r"(ns)
		    : "memory");
		if (done) break;
		if (++spin == (1u << 22)) __trap();
	}
}
__device__ __forceinline__ void mbar_arrive_addr(uint32_t addr) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
	uint32_t v;
	asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
	return v;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
	const uint32_t addr = smem_u32(bar);
	uint32_t spin = 0;
#pragma unroll 1
	while (!mbar_try_wait(addr, parity))
		if (++spin == (1u << 24)) __trap();
}

// One lane of the (converged) warp.  A branch on this predicate is the form ptxas recognises as single-thread code: tcgen05.mma /
// commit operands then go to uniform registers directly.  Behind `if (lane == 0)` every UTCHMMA is wrapped in an ELECT / R2UR /
// VOTEU loop instead (measured with tools/probe/tc_probe.cu: 112-124 cycles per 128x128x16 MMA against 73 behind elect.sync).
__device__ __forceinline__ uint32_t elect_one() {
	uint32_t pred = 0;
	asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, %1;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred) : "r"(0xffffffffu));
	return pred;
}

// ---------------------------------------------------------------- TMA
// plain 1-D bulk copy global -> shared (16-byte aligned, size a multiple of 16), completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
	             "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(smem_u32(bar))
	             : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
	asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tiled load global -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* m, int c0, int c1, uint64_t* bar) {
	asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
	             "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
	             : "memory");
}

// 3-D tiled load (x, y, frame); out-of-range coordinates are filled with zeros
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* m, int c0, int c1, int c2, uint64_t* bar) {
	asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
	                 smem_u32(dst)),
	             "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
	             : "memory");
}

// ---------------------------------------------------------------- tcgen05 / TMEM
// all 32 lanes of ONE warp; the TMEM base address is written to *slot (shared memory)
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* slot) {
	asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "n"(COLS) : "memory");
	asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
	asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T, bf16 inputs, fp32 accumulate; one thread issues for the CTA
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
	asm volatile(
	    "{\n\t.reg .pred p;\n\t"
	    "setp.ne.b32 p, %4, 0;\n\t"
	    "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
	    "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
	    : "memory");
}
// the same with the A operand in TMEM (lane = row, 32-bit column j = elements k = 2j | 2j+1 of a 16-bit type; a K = 16 step is 8
// columns): no shared-memory traffic for A.  Measured: 73 cycles per 128x128x16 against 97 with both operands in shared memory.
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
	asm volatile(
	    "{\n\t.reg .pred p;\n\t"
	    "setp.ne.b32 p, %4, 0;\n\t"
	    "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
	    "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
	    : "memory");
}
// arrive on `bar` when every tcgen05.mma issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 lanes x 32 consecutive columns: thread l of warp w reads TMEM lane 32*(w%4)+l, columns [col, col+32).
// tmem_ld32_issue starts the load; the registers are valid only after tmem_ld_wait on the same array (the wait takes the
// registers as in/out operands so that the compiler cannot move their consumers above it).
#define BS_R32(m) m(0), m(1), m(2), m(3), m(4), m(5), m(6), m(7), m(8), m(9), m(10), m(11), m(12), m(13), m(14), m(15), m(16), m(17), \
	m(18), m(19), m(20), m(21), m(22), m(23), m(24), m(25), m(26), m(27), m(28), m(29), m(30), m(31)
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t (&r)[32]) {
#define BS_OUT(k) "=r"(r[k])
	asm volatile(
	    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
	    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
	    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
	    : BS_R32(BS_OUT)
	    : "r"(taddr)
	    : "memory");
#undef BS_OUT
}
// 32 lanes x 16 consecutive columns into the first sixteen registers of a 32-register array (the ragged end of a 112-column accumulator)
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[32]) {
	asm volatile(
	    "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
	    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
	    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
	      "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
	    : "r"(taddr)
	    : "memory");
}
// Packed loads of 16-bit accumulators (kind::f16 with an F16 accumulator keeps one element per 32-bit TMEM column): register j =
// {column 2j (low half), column 2j + 1 (high half)}.  64 / 32 / 16 columns into 32 / 16 / 8 registers (tools/probe/f16acc_probe.cu).
__device__ __forceinline__ void tmem_ld64p_issue(uint32_t taddr, uint32_t (&r)[32]) {
#define BS_OUT(k) "=r"(r[k])
	asm volatile(
	    "tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
	    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
	    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
	    : BS_R32(BS_OUT)
	    : "r"(taddr)
	    : "memory");
#undef BS_OUT
}
__device__ __forceinline__ void tmem_ld32p_issue(uint32_t taddr, uint32_t (&r)[16]) {
	asm volatile(
	    "tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 "
	    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
	    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
	      "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
	    : "r"(taddr)
	    : "memory");
}
__device__ __forceinline__ void tmem_ld16p_issue(uint32_t taddr, uint32_t (&r)[8]) {
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.pack::16b.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
	             : "r"(taddr)
	             : "memory");
}
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&r)[16]) {
	asm volatile("tcgen05.wait::ld.sync.aligned;"
	             : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]),
	               "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])::"memory");
}
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&r)[8]) {
	asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])::"memory");
}
// packed fp16 pair maximum (HMNMX2: one issue slot per two elements) and the halves of a pair as floats (exact)
__device__ __forceinline__ uint32_t hmax2(uint32_t a, uint32_t b) {
	uint32_t d;
	asm("max.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
	return d;
}
__device__ __forceinline__ float h2lo(uint32_t v) {
	float f;
	asm("{\n\t.reg .b16 lo, hi;\n\tmov.b32 {lo, hi}, %1;\n\tcvt.f32.f16 %0, lo;\n\t}" : "=f"(f) : "r"(v));
	return f;
}
__device__ __forceinline__ float h2hi(uint32_t v) {
	float f;
	asm("{\n\t.reg .b16 lo, hi;\n\tmov.b32 {lo, hi}, %1;\n\tcvt.f32.f16 %0, hi;\n\t}" : "=f"(f) : "r"(v));
	return f;
}
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&r)[32]) {
#define BS_INOUT(k) "+r"(r[k])
	asm volatile("tcgen05.wait::ld.sync.aligned;" : BS_R32(BS_INOUT)::"memory");
#undef BS_INOUT
}

// registers -> TMEM, same shape as tmem_ld32_issue; returns once the store has completed
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
#define BS_IN(k) "r"(r[k])
	asm volatile(
	    "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
	    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
	    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31};" ::BS_R32(BS_IN),
	    "r"(taddr)
	    : "memory");
#undef BS_IN
	asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// packed fp32 pair arithmetic and the three-input minimum of sm_100
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
	unsigned long long ra, rb, rc, rd;
	asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
	asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
	asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
	asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
	float2 d;
	asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
	return d;
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
	float4 v;
	asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
	return v;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
	float d;
	asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
	return d;
}
__device__ __forceinline__ float fmin3(float a, float b, float c) {
	float d;
	asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
	return d;
}

// ---------------------------------------------------------------- UMMA descriptors
// Shared-memory matrix descriptor, K-major operand stored as rows of 128 bytes with the 128-byte swizzle (what a TMA box
// of 64 bf16 x rows with CU_TENSOR_MAP_SWIZZLE_128B writes; the region must be 1024-byte aligned):
//   bits [0,14)  start address >> 4          bits [16,30) leading byte offset >> 4 (unused for swizzled K-major: 1)
//   bits [32,46) stride byte offset >> 4 (8 rows * 128 B = 1024 -> 64)      bits [46,48) version = 1 (sm_100)
//   bits [61,64) layout type: 2 = SWIZZLE_128B
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
	return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)64 << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// Instruction descriptor for kind::f16: D = F32 (bits 4-5 = 1), A = B = BF16 (bits 7-9, 10-12 = 1), both K-major (bits 15,16 = 0),
// N >> 3 at bits 17-22, M >> 4 at bits 24-28
// the same for D = F16, A = B = F16 (all format fields 0)
__host__ __device__ constexpr uint32_t umma_idesc_f16_f16acc(int M, int N) { return ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24); }
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
	return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

}  // namespace sm100
}  // namespace bs
