// tcgen05 / TMEM / mbarrier building blocks shared by the network kernels (hn_network_tc.cu: inference) and the training MLP
// kernels (hn_train_tc.cu: forward with stored activations, backward dX, split-K dW). sm_100a only.
//
// Shared-memory operand layouts (no swizzle; CUTLASS cute/atom/mma_traits_sm100.hpp "LayoutType::INTERLEAVE"), both built from
// "planes" [c][row][8 halfs] of 16-byte rows:
//   K-major  operand (rows = M or N index, planes = 8-wide chunks of K): descriptor LBO = plane bytes (next K chunk), SBO = 128
//            (next 8-row group); one K = 16 instruction starts at plane 2 * kk.
//   MN-major operand (rows = K index, planes = 8-wide chunks of M or N): LBO = 128 (next 8 K rows), SBO = plane bytes (next 8
//            M/N); one K = 16 instruction starts 256 bytes further down every plane. Instruction descriptor bits 15 / 16 select it.
#pragma once
#include "hn_internal.h"
#include <cuda_fp16.h>

namespace hn {
namespace tc {

constexpr int TILE = 128;                 // rows per CTA tile = TMEM lanes
constexpr uint32_t A_PLANE = TILE * 16;   // bytes of one K-chunk plane of the A operand
constexpr uint32_t A_BYTES = 8 * A_PLANE; // K up to 64
constexpr uint32_t TMEM_COLS = 64;        // fp32 accumulator 128 x 64

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
	uint32_t done;
	do {
		asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
			: "=r"(done) : "r"(bar), "r"(parity) : "memory");
	} while (!done);
}
// ---- proxies / tcgen05 ordering
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

// ---- TMEM allocation (one warp, all lanes converged)
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
	asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(dst_smem), "r"(cols) : "memory");
	asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
	asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(taddr), "r"(cols) : "memory");
}

// ---- descriptors (cute/arch/mma_sm100_desc.hpp: SmemDescriptor, InstrDescriptor)
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
	return (uint64_t)((saddr >> 4) & 0x3FFFu)          // start address, bits [0,14)
		| ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) // leading byte offset: next 16-byte K chunk
		| ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) // stride byte offset: next 8-row group
		| (1ull << 46);                                  // descriptor version 1 (Blackwell); swizzle none, base offset 0
}
__host__ __device__ constexpr uint32_t instr_desc(uint32_t M, uint32_t N) {
	// c_format F32 (bits 4-5) | a,b format F16 (0) | a,b K-major (0) | N >> 3 (bits 17-22) | M >> 4 (bits 24-28)
	return (1u << 4) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]^T, issued by ONE thread
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
		:: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" :: "r"(bar) : "memory");
}

// ---- TMEM -> registers: 32 lanes x 32 bit, N consecutive columns; thread t of warp w reads lane 32*(w%4)+t
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
		: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
		  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
		: "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&r)[4]) {
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];\n"
		: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}

// fp32 pair -> packed fp16 pair (lo = a, hi = b), with ReLU fused into the conversion
__device__ __forceinline__ uint32_t pack_relu(uint32_t a_bits, uint32_t b_bits) {
	uint32_t d;
	asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;\n" : "=r"(d) : "f"(__uint_as_float(b_bits)), "f"(__uint_as_float(a_bits)));
	return d;
}
__device__ __forceinline__ uint32_t pack_plain(uint32_t a_bits, uint32_t b_bits) {
	uint32_t d;
	asm("cvt.rn.f16x2.f32 %0, %1, %2;\n" : "=r"(d) : "f"(__uint_as_float(b_bits)), "f"(__uint_as_float(a_bits)));
	return d;
}


// ---- weights: global row-major [out][in] -> canonical B planes [k/8][out][8 halfs]
__device__ __forceinline__ uint32_t stage_layer(const __half* __restrict__ w, int N, int K, uint8_t* dst) {
	const int chunks = K / 8;
	for (int i = threadIdx.x; i < N * chunks; i += blockDim.x) {
		const int n = i / chunks, c = i - n * chunks;
		*reinterpret_cast<uint4*>(dst + (size_t)c * N * 16 + n * 16) = __ldg(reinterpret_cast<const uint4*>(w + (size_t)n * K) + c);
	}
	return (uint32_t)(N * K * 2);
}
__host__ __device__ inline uint32_t mlp_bytes(int hidden) { return (64 * 32 + (hidden - 1) * 64 * 64 + 16 * 64) * 2; }

// one layer = K/16 instructions into the same accumulator, then a commit that arrives on `bar` when they have all retired
__device__ __forceinline__ void issue_layer(uint32_t d_tmem, uint32_t a_s, uint32_t w_s, int K, int N, uint32_t bar) {
	const uint32_t idesc = instr_desc(TILE, (uint32_t)N);
	const uint32_t w_plane = (uint32_t)N * 16u;
	for (int kk = 0; kk < K / 16; ++kk) {
		const uint64_t ad = smem_desc(a_s + (uint32_t)kk * 2u * A_PLANE, A_PLANE, 128u);
		const uint64_t bd = smem_desc(w_s + (uint32_t)kk * 2u * w_plane, w_plane, 128u);
		umma_f16(d_tmem, ad, bd, idesc, kk > 0 ? 1u : 0u);
	}
	umma_commit(bar);
}

// 64 fp32 accumulator columns of this thread's row -> ReLU -> fp16 -> the 8 K-chunks of the next A operand
__device__ __forceinline__ void hidden_epilogue(uint32_t t_row, uint8_t* a_row) {
	#pragma unroll
	for (int h = 0; h < 2; ++h) {
		uint32_t v0[16], v1[16];
		tmem_ld16(t_row + h * 32, v0);
		tmem_ld16(t_row + h * 32 + 16, v1);
		tc_wait_ld();
		uint4 c0 = make_uint4(pack_relu(v0[0], v0[1]), pack_relu(v0[2], v0[3]), pack_relu(v0[4], v0[5]), pack_relu(v0[6], v0[7]));
		uint4 c1 = make_uint4(pack_relu(v0[8], v0[9]), pack_relu(v0[10], v0[11]), pack_relu(v0[12], v0[13]), pack_relu(v0[14], v0[15]));
		uint4 c2 = make_uint4(pack_relu(v1[0], v1[1]), pack_relu(v1[2], v1[3]), pack_relu(v1[4], v1[5]), pack_relu(v1[6], v1[7]));
		uint4 c3 = make_uint4(pack_relu(v1[8], v1[9]), pack_relu(v1[10], v1[11]), pack_relu(v1[12], v1[13]), pack_relu(v1[14], v1[15]));
		*reinterpret_cast<uint4*>(a_row + (4 * h + 0) * A_PLANE) = c0;
		*reinterpret_cast<uint4*>(a_row + (4 * h + 1) * A_PLANE) = c1;
		*reinterpret_cast<uint4*>(a_row + (4 * h + 2) * A_PLANE) = c2;
		*reinterpret_cast<uint4*>(a_row + (4 * h + 3) * A_PLANE) = c3;
	}
}

// G independent tile groups of 128 threads share one copy of the weights and one TMEM allocation (64 columns each); a
// group synchronises on its own named barrier and its own mbarrier, so the groups of a CTA drift apart freely and one
// group's gather overlaps the other groups' MMA rounds.
__device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, %1;\n" :: "r"(g + 1), "r"(TILE) : "memory"); }

} // namespace tc
} // namespace hn
