// a-21: the MLP part of the training step on the 5th-generation tensor cores - the reference's kernel_mlp_fused forward with stored
// activations, kernel_mlp_fused_backward (tcnn/src/fully_fused_mlp.cu:150-259,307-311) and its split-K weight gradients
// (:775-829) - as three hand-written tcgen05 / TMEM kernels (no cuBLAS, no separate ReLU / copy / combine launches):
//
//   train_forward_tc_kernel   tile = 128 samples. enc [n][32] -> density MLP -> [16 density outputs | 16 SH] -> colour MLP, every
//                             hidden activation (post-ReLU, fp16) and both outputs stored to global memory for the backward pass.
//   train_backward_tc_kernel  dL/d(out) [n][16] -> colour MLP top-down -> + dL/d(sigma) on density channel 0 -> density MLP top-down
//                             -> dL/d(enc) [n][32]. B operands are the TRANSPOSED weights (staged once per CTA), the ReLU mask of the
//                             stored activation is applied in the TMEM epilogue, every pre-activation gradient is stored (fp16).
//   train_dw_tc_kernel        dW = dY^T X for all layers, reduction over the samples: both operands MN-major straight from the
//                             row-major [sample][feature] matrices, fp32 accumulation in TMEM across a CTA's slice of the batch
//                             (split-K over CTAs), one red.global.add per weight and CTA into the fp32 gradient buffer.
//
// Buffers follow hn_train.cu: act[i] = forward activations (enc | density hidden... | density out | rgb in | rgb hidden... | out),
// gact[i] = gradient w.r.t. the pre-activation value of act[i] (same shapes).
#include "hn_tc.cuh"
#include "hn_network.cuh"
#include "hn_train_tc.h"
#include <map>
#include <mutex>

namespace hn {
namespace tc {

// 64 fp32 accumulator columns of this thread's row -> ReLU -> fp16 -> the 8 K-chunks of the next A operand AND the activation row
// in global memory (128 contiguous bytes per thread)
__device__ __forceinline__ void hidden_epilogue_store(uint32_t t_row, uint8_t* a_row, __half* g_row /* nullptr: row beyond n */) {
	#pragma unroll
	for (int h = 0; h < 2; ++h) {
		uint32_t v0[16], v1[16];
		tmem_ld16(t_row + h * 32, v0);
		tmem_ld16(t_row + h * 32 + 16, v1);
		tc_wait_ld();
		const uint4 c0 = make_uint4(pack_relu(v0[0], v0[1]), pack_relu(v0[2], v0[3]), pack_relu(v0[4], v0[5]), pack_relu(v0[6], v0[7]));
		const uint4 c1 = make_uint4(pack_relu(v0[8], v0[9]), pack_relu(v0[10], v0[11]), pack_relu(v0[12], v0[13]), pack_relu(v0[14], v0[15]));
		const uint4 c2 = make_uint4(pack_relu(v1[0], v1[1]), pack_relu(v1[2], v1[3]), pack_relu(v1[4], v1[5]), pack_relu(v1[6], v1[7]));
		const uint4 c3 = make_uint4(pack_relu(v1[8], v1[9]), pack_relu(v1[10], v1[11]), pack_relu(v1[12], v1[13]), pack_relu(v1[14], v1[15]));
		*reinterpret_cast<uint4*>(a_row + (4 * h + 0) * A_PLANE) = c0;
		*reinterpret_cast<uint4*>(a_row + (4 * h + 1) * A_PLANE) = c1;
		*reinterpret_cast<uint4*>(a_row + (4 * h + 2) * A_PLANE) = c2;
		*reinterpret_cast<uint4*>(a_row + (4 * h + 3) * A_PLANE) = c3;
		if (g_row) {
			uint4* g = reinterpret_cast<uint4*>(g_row) + 4 * h;
			g[0] = c0; g[1] = c1; g[2] = c2; g[3] = c3;
		}
	}
}

template <int G>
__global__ void __launch_bounds__(TILE * G, 2) train_forward_tc_kernel(const __grid_constant__ TrainMlpArgs a) {
	extern __shared__ __align__(1024) uint8_t smem[];
	const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
	const int grp = warp_u >> 2;
	const int lane = threadIdx.x & 31, warp = warp_u & 3, tid = warp * 32 + lane;
	uint8_t* a_s = smem + (size_t)grp * A_BYTES;
	uint8_t* wd_s = smem + (size_t)G * A_BYTES;
	uint8_t* wr_s = wd_s + mlp_bytes(a.density_hidden);
	uint64_t* bar_p = reinterpret_cast<uint64_t*>(wr_s + mlp_bytes(a.rgb_hidden)) + grp;
	uint32_t* tmem_p = reinterpret_cast<uint32_t*>(reinterpret_cast<uint64_t*>(wr_s + mlp_bytes(a.rgb_hidden)) + G);
	{
		uint8_t* d = wd_s;
		const __half* w = a.w_density;
		d += stage_layer(w, 64, 32, d); w += 64 * 32;
		for (int l = 1; l < a.density_hidden; ++l) { d += stage_layer(w, 64, 64, d); w += 64 * 64; }
		stage_layer(w, 16, 64, d);
		d = wr_s; w = a.w_rgb;
		d += stage_layer(w, 64, 32, d); w += 64 * 32;
		for (int l = 1; l < a.rgb_hidden; ++l) { d += stage_layer(w, 64, 64, d); w += 64 * 64; }
		stage_layer(w, 16, 64, d);
	}
	const uint32_t bar = smem_addr(bar_p);
	if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
	if (threadIdx.x < 32) tmem_alloc(smem_addr(tmem_p), TMEM_COLS * G);
	fence_proxy_async();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem_base = *tmem_p + (uint32_t)grp * TMEM_COLS;
	const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16);
	const uint32_t a_addr = smem_addr(a_s), wd_addr = smem_addr(wd_s), wr_addr = smem_addr(wr_s);
	uint8_t* a_row = a_s + tid * 16;
	uint32_t phase = 0;
	const uint32_t n_tiles = (a.n + TILE - 1) / TILE;

	#define HN_ROUND(K, N, W_ADDR) do { \
		fence_proxy_async(); tc_fence_before(); group_sync(grp); \
		if (tid == 0) { tc_fence_after(); issue_layer(tmem_base, a_addr, (W_ADDR), (K), (N), bar); } \
	} while (0)
	#define HN_WAIT() do { mbar_wait(bar, phase); phase ^= 1u; tc_fence_after(); } while (0)

	for (uint32_t tile = blockIdx.x * G + grp; tile < n_tiles; tile += gridDim.x * G) {
		const size_t row = (size_t)tile * TILE + tid;
		const bool valid = row < a.n;
		{ // A <- enc row
			const uint4* e = reinterpret_cast<const uint4*>(a.enc + row * 32);
			#pragma unroll
			for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(a_row + c * A_PLANE) = valid ? __ldg(e + c) : make_uint4(0u, 0u, 0u, 0u);
		}
		uint32_t w_addr = wd_addr;
		HN_ROUND(32, 64, w_addr); w_addr += 64 * 32 * 2;
		for (int l = 1; l < a.density_hidden; ++l) {
			HN_WAIT(); hidden_epilogue_store(t_row, a_row, valid ? a.hid_d[l - 1] + row * 64 : nullptr);
			HN_ROUND(64, 64, w_addr); w_addr += 64 * 64 * 2;
		}
		HN_WAIT(); hidden_epilogue_store(t_row, a_row, valid ? a.hid_d[a.density_hidden - 1] + row * 64 : nullptr);
		HN_ROUND(64, 16, w_addr);
		{ // density output (16, no activation) -> d_out, rgb_in[:, :16], A planes 0..1; SH half of rgb_in -> A planes 2..3
			uint4 s0 = make_uint4(0u, 0u, 0u, 0u), s1 = s0;
			if (valid) { const uint4* sh = reinterpret_cast<const uint4*>(a.rgb_in + row * 32 + 16); s0 = __ldg(sh); s1 = __ldg(sh + 1); }
			HN_WAIT();
			uint32_t v[16];
			tmem_ld16(t_row, v);
			tc_wait_ld();
			const uint4 d0 = make_uint4(pack_plain(v[0], v[1]), pack_plain(v[2], v[3]), pack_plain(v[4], v[5]), pack_plain(v[6], v[7]));
			const uint4 d1 = make_uint4(pack_plain(v[8], v[9]), pack_plain(v[10], v[11]), pack_plain(v[12], v[13]), pack_plain(v[14], v[15]));
			*reinterpret_cast<uint4*>(a_row + 0 * A_PLANE) = d0;
			*reinterpret_cast<uint4*>(a_row + 1 * A_PLANE) = d1;
			*reinterpret_cast<uint4*>(a_row + 2 * A_PLANE) = s0;
			*reinterpret_cast<uint4*>(a_row + 3 * A_PLANE) = s1;
			if (valid) {
				uint4* o = reinterpret_cast<uint4*>(a.d_out + row * 16); o[0] = d0; o[1] = d1;
				uint4* r = reinterpret_cast<uint4*>(a.rgb_in + row * 32); r[0] = d0; r[1] = d1;
			}
		}
		w_addr = wr_addr;
		HN_ROUND(32, 64, w_addr); w_addr += 64 * 32 * 2;
		for (int l = 1; l < a.rgb_hidden; ++l) {
			HN_WAIT(); hidden_epilogue_store(t_row, a_row, valid ? a.hid_r[l - 1] + row * 64 : nullptr);
			HN_ROUND(64, 64, w_addr); w_addr += 64 * 64 * 2;
		}
		HN_WAIT(); hidden_epilogue_store(t_row, a_row, valid ? a.hid_r[a.rgb_hidden - 1] + row * 64 : nullptr);
		HN_ROUND(64, 16, w_addr);
		HN_WAIT();
		{
			uint32_t v[16];
			tmem_ld16(t_row, v);
			tc_wait_ld();
			if (valid) {
				uint4* o = reinterpret_cast<uint4*>(a.out + row * 16);
				o[0] = make_uint4(pack_plain(v[0], v[1]), pack_plain(v[2], v[3]), pack_plain(v[4], v[5]), pack_plain(v[6], v[7]));
				o[1] = make_uint4(pack_plain(v[8], v[9]), pack_plain(v[10], v[11]), pack_plain(v[12], v[13]), pack_plain(v[14], v[15]));
			}
			tc_fence_before(); // the next tile's first MMA overwrites these TMEM columns
		}
	}
	#undef HN_ROUND
	#undef HN_WAIT
	tc_fence_before();
	__syncthreads();
	if (threadIdx.x < 32) tmem_dealloc(*tmem_p, TMEM_COLS * G);
}

// ---- backward dX -------------------------------------------------------------------------------------------------------------------
// weights for dX = dY W: B operand = W^T, i.e. rows = input index n, K = output index k: planes [k/8][n][8 halfs], element j of
// plane c, row n = W[8 c + j][n]. One-time gather per CTA (<= 28 KB).
__device__ __forceinline__ uint32_t stage_layer_T(const __half* __restrict__ w, int N_out, int K_in, uint8_t* dst) {
	const int chunks = N_out / 8;
	for (int i = threadIdx.x; i < K_in * chunks; i += blockDim.x) {
		const int c = i / K_in, n = i - c * K_in;
		__half v[8];
		#pragma unroll
		for (int j = 0; j < 8; ++j) v[j] = w[(size_t)(8 * c + j) * K_in + n];
		*reinterpret_cast<uint4*>(dst + (size_t)c * K_in * 16 + n * 16) = *reinterpret_cast<const uint4*>(v);
	}
	return (uint32_t)(N_out * K_in * 2);
}

// 64 accumulator columns = dL/d(post-ReLU activation) of this row; times the ReLU mask of the stored forward activation -> fp16 ->
// A operand of the next (lower) layer and the gradient row in global memory
__device__ __forceinline__ void masked_epilogue_store(uint32_t t_row, uint8_t* a_row, const __half* act_row, __half* g_row) {
	const __half2 zero = __float2half2_rn(0.f);
	#pragma unroll
	for (int h = 0; h < 2; ++h) {
		uint4 m[4];
		#pragma unroll
		for (int q = 0; q < 4; ++q) m[q] = act_row ? __ldg(reinterpret_cast<const uint4*>(act_row) + 4 * h + q) : make_uint4(0u, 0u, 0u, 0u);
		uint32_t v0[16], v1[16];
		tmem_ld16(t_row + h * 32, v0);
		tmem_ld16(t_row + h * 32 + 16, v1);
		tc_wait_ld();
		uint4 c[4];
		#pragma unroll
		for (int q = 0; q < 4; ++q) {
			const uint32_t* v = q < 2 ? v0 + 8 * q : v1 + 8 * (q - 2);
			uint32_t p[4] = {pack_plain(v[0], v[1]), pack_plain(v[2], v[3]), pack_plain(v[4], v[5]), pack_plain(v[6], v[7])};
			const uint32_t mm[4] = {m[q].x, m[q].y, m[q].z, m[q].w};
			#pragma unroll
			for (int e = 0; e < 4; ++e) {
				const __half2 g = *reinterpret_cast<const __half2*>(&p[e]);
				const __half2 keep = __hgt2(*reinterpret_cast<const __half2*>(&mm[e]), zero); // 1.0 where the activation was > 0
				const __half2 r = __hmul2(g, keep);
				p[e] = *reinterpret_cast<const uint32_t*>(&r);
			}
			c[q] = make_uint4(p[0], p[1], p[2], p[3]);
		}
		#pragma unroll
		for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4*>(a_row + (4 * h + q) * A_PLANE) = c[q];
		if (g_row) {
			uint4* g = reinterpret_cast<uint4*>(g_row) + 4 * h;
			g[0] = c[0]; g[1] = c[1]; g[2] = c[2]; g[3] = c[3];
		}
	}
}

template <int G>
__global__ void __launch_bounds__(TILE * G, 2) train_backward_tc_kernel(const __grid_constant__ TrainMlpArgs a) {
	extern __shared__ __align__(1024) uint8_t smem[];
	const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
	const int grp = warp_u >> 2;
	const int lane = threadIdx.x & 31, warp = warp_u & 3, tid = warp * 32 + lane;
	uint8_t* a_s = smem + (size_t)grp * A_BYTES;
	uint8_t* wd_s = smem + (size_t)G * A_BYTES;
	uint8_t* wr_s = wd_s + mlp_bytes(a.density_hidden);
	uint64_t* bar_p = reinterpret_cast<uint64_t*>(wr_s + mlp_bytes(a.rgb_hidden)) + grp;
	uint32_t* tmem_p = reinterpret_cast<uint32_t*>(reinterpret_cast<uint64_t*>(wr_s + mlp_bytes(a.rgb_hidden)) + G);
	{ // transposed weights, same order and sizes as the forward staging: [first 64x32 | hidden 64x64 ... | last 16x64]
		uint8_t* d = wd_s;
		const __half* w = a.w_density;
		d += stage_layer_T(w, 64, 32, d); w += 64 * 32;
		for (int l = 1; l < a.density_hidden; ++l) { d += stage_layer_T(w, 64, 64, d); w += 64 * 64; }
		stage_layer_T(w, 16, 64, d);
		d = wr_s; w = a.w_rgb;
		d += stage_layer_T(w, 64, 32, d); w += 64 * 32;
		for (int l = 1; l < a.rgb_hidden; ++l) { d += stage_layer_T(w, 64, 64, d); w += 64 * 64; }
		stage_layer_T(w, 16, 64, d);
	}
	const uint32_t bar = smem_addr(bar_p);
	if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
	if (threadIdx.x < 32) tmem_alloc(smem_addr(tmem_p), TMEM_COLS * G);
	fence_proxy_async();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem_base = *tmem_p + (uint32_t)grp * TMEM_COLS;
	const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16);
	const uint32_t a_addr = smem_addr(a_s), wd_addr = smem_addr(wd_s), wr_addr = smem_addr(wr_s);
	uint8_t* a_row = a_s + tid * 16;
	uint32_t phase = 0;
	const uint32_t n_tiles = (a.n + TILE - 1) / TILE;
	// byte offsets of the layers inside a staged MLP
	const uint32_t off_last_d = (uint32_t)(64 * 32 + (a.density_hidden - 1) * 64 * 64) * 2u;
	const uint32_t off_last_r = (uint32_t)(64 * 32 + (a.rgb_hidden - 1) * 64 * 64) * 2u;

	#define HN_ROUND(K, N, W_ADDR) do { \
		fence_proxy_async(); tc_fence_before(); group_sync(grp); \
		if (tid == 0) { tc_fence_after(); issue_layer(tmem_base, a_addr, (W_ADDR), (K), (N), bar); } \
	} while (0)
	#define HN_WAIT() do { mbar_wait(bar, phase); phase ^= 1u; tc_fence_after(); } while (0)

	for (uint32_t tile = blockIdx.x * G + grp; tile < n_tiles; tile += gridDim.x * G) {
		const size_t row = (size_t)tile * TILE + tid;
		const bool valid = row < a.n;
		{ // A <- dL/d(out) row (16 wide)
			const uint4* g = reinterpret_cast<const uint4*>(a.g_out + row * 16);
			*reinterpret_cast<uint4*>(a_row + 0 * A_PLANE) = valid ? __ldg(g) : make_uint4(0u, 0u, 0u, 0u);
			*reinterpret_cast<uint4*>(a_row + 1 * A_PLANE) = valid ? __ldg(g + 1) : make_uint4(0u, 0u, 0u, 0u);
		}
		// ---- colour MLP, top-down
		HN_ROUND(16, 64, wr_addr + off_last_r);
		for (int l = a.rgb_hidden - 1; l >= 1; --l) {
			HN_WAIT(); masked_epilogue_store(t_row, a_row, valid ? a.hid_r[l] + row * 64 : nullptr, valid ? a.g_hid_r[l] + row * 64 : nullptr);
			HN_ROUND(64, 64, wr_addr + (uint32_t)(64 * 32 + (l - 1) * 64 * 64) * 2u);
		}
		HN_WAIT(); masked_epilogue_store(t_row, a_row, valid ? a.hid_r[0] + row * 64 : nullptr, valid ? a.g_hid_r[0] + row * 64 : nullptr);
		HN_ROUND(64, 32, wr_addr);
		{ // dL/d(rgb_in)[:, :16] + dL/d(sigma) on channel 0 = dL/d(density out); the SH half has no parameters behind it
			const float ds = valid ? __ldg(a.d_sigma + row) : 0.f;
			HN_WAIT();
			uint32_t v[16];
			tmem_ld16(t_row, v);
			tc_wait_ld();
			v[0] = __float_as_uint(__uint_as_float(v[0]) + ds);
			const uint4 d0 = make_uint4(pack_plain(v[0], v[1]), pack_plain(v[2], v[3]), pack_plain(v[4], v[5]), pack_plain(v[6], v[7]));
			const uint4 d1 = make_uint4(pack_plain(v[8], v[9]), pack_plain(v[10], v[11]), pack_plain(v[12], v[13]), pack_plain(v[14], v[15]));
			*reinterpret_cast<uint4*>(a_row + 0 * A_PLANE) = d0;
			*reinterpret_cast<uint4*>(a_row + 1 * A_PLANE) = d1;
			if (valid) { uint4* o = reinterpret_cast<uint4*>(a.g_dout + row * 16); o[0] = d0; o[1] = d1; }
		}
		// ---- density MLP, top-down
		HN_ROUND(16, 64, wd_addr + off_last_d);
		for (int l = a.density_hidden - 1; l >= 1; --l) {
			HN_WAIT(); masked_epilogue_store(t_row, a_row, valid ? a.hid_d[l] + row * 64 : nullptr, valid ? a.g_hid_d[l] + row * 64 : nullptr);
			HN_ROUND(64, 64, wd_addr + (uint32_t)(64 * 32 + (l - 1) * 64 * 64) * 2u);
		}
		HN_WAIT(); masked_epilogue_store(t_row, a_row, valid ? a.hid_d[0] + row * 64 : nullptr, valid ? a.g_hid_d[0] + row * 64 : nullptr);
		HN_ROUND(64, 32, wd_addr);
		HN_WAIT();
		{ // dL/d(enc) [32]
			uint32_t v0[16], v1[16];
			tmem_ld16(t_row, v0);
			tmem_ld16(t_row + 16, v1);
			tc_wait_ld();
			if (valid) {
				uint4* o = reinterpret_cast<uint4*>(a.g_enc + row * 32);
				o[0] = make_uint4(pack_plain(v0[0], v0[1]), pack_plain(v0[2], v0[3]), pack_plain(v0[4], v0[5]), pack_plain(v0[6], v0[7]));
				o[1] = make_uint4(pack_plain(v0[8], v0[9]), pack_plain(v0[10], v0[11]), pack_plain(v0[12], v0[13]), pack_plain(v0[14], v0[15]));
				o[2] = make_uint4(pack_plain(v1[0], v1[1]), pack_plain(v1[2], v1[3]), pack_plain(v1[4], v1[5]), pack_plain(v1[6], v1[7]));
				o[3] = make_uint4(pack_plain(v1[8], v1[9]), pack_plain(v1[10], v1[11]), pack_plain(v1[12], v1[13]), pack_plain(v1[14], v1[15]));
			}
			tc_fence_before();
		}
	}
	#undef HN_ROUND
	#undef HN_WAIT
	tc_fence_before();
	__syncthreads();
	if (threadIdx.x < 32) tmem_dealloc(*tmem_p, TMEM_COLS * G);
}

// ---- weight gradients ----------------------------------------------------------------------------------------------------------------
// dW = dY^T X, a reduction over the samples. One item = (128-sample tile, layer): D[128 x nb] (+)= A^T B with A = a 64-wide
// [sample][feature] matrix (zero-padded to M = 128) and B = an nb-wide one (nb = 16 / 32 / 64), BOTH MN-major: a 16-byte chunk of a
// row-major row is already a row of a canonical core matrix, so the tiles are copied from global memory as they lie (cp.async).
//   layers with 64 outputs: A = dY [n][64], B = X [n][in]   -> D[out][in]
//   layers with 16 outputs: A = X [n][64],  B = dY [n][16]  -> D[in][out] (written transposed)
// Every layer keeps its own accumulator columns in TMEM for the whole kernel; at the end each CTA adds its partial sums into the
// fp32 gradient buffer (split-K over CTAs).
__host__ __device__ constexpr uint32_t instr_desc_mn(uint32_t M, uint32_t N) {
	// as instr_desc(), with a_major (bit 15) and b_major (bit 16) = MN-major
	return (1u << 4) | (1u << 15) | (1u << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void* src, bool valid) {
	const uint32_t bytes = valid ? 16u : 0u; // src-size 0: the 16 destination bytes are zero-filled
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" :: "r"(dst_smem), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" :: "n"(N) : "memory"); }

constexpr uint32_t DW_THREADS = 256;
constexpr uint32_t DW_A_BYTES = 16 * A_PLANE; // 128 (padded) MN x 128 samples
constexpr uint32_t DW_B_BYTES = 8 * A_PLANE;
constexpr uint32_t DW_STAGE = DW_A_BYTES + DW_B_BYTES;
constexpr uint32_t DW_TMEM_COLS = 512;

__global__ void __launch_bounds__(DW_THREADS, 1) train_dw_tc_kernel(const __grid_constant__ DwArgs a) {
	extern __shared__ __align__(1024) uint8_t smem[];
	uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * DW_STAGE); // [0], [1]: stage free again; [2]: everything done
	uint32_t* tmem_p = reinterpret_cast<uint32_t*>(bars + 3);
	const uint32_t tid = threadIdx.x;
	// the upper 8 planes of both A stages (M rows 64..127) stay zero for the whole kernel
	for (uint32_t i = tid; i < 2 * (8 * A_PLANE / 16); i += DW_THREADS) {
		const uint32_t st = i / (8 * A_PLANE / 16), k = i % (8 * A_PLANE / 16);
		*reinterpret_cast<uint4*>(smem + st * DW_STAGE + 8 * A_PLANE + k * 16) = make_uint4(0u, 0u, 0u, 0u);
	}
	if (tid == 0) { mbar_init(smem_addr(bars), 1); mbar_init(smem_addr(bars + 1), 1); mbar_init(smem_addr(bars + 2), 1); fence_barrier_init(); }
	if (tid < 32) tmem_alloc(smem_addr(tmem_p), DW_TMEM_COLS);
	fence_proxy_async();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem_base = *tmem_p;
	const uint32_t n_tiles = (a.n + TILE - 1) / TILE;
	const uint32_t my_tiles = blockIdx.x < n_tiles ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
	const uint32_t n_items = my_tiles * (uint32_t)a.n_layers;

	auto load_item = [&](uint32_t item) {
		const uint32_t tile = blockIdx.x + (item / (uint32_t)a.n_layers) * gridDim.x;
		const DwLayer& L = a.layer[item % (uint32_t)a.n_layers];
		const uint32_t st = smem_addr(smem + (item & 1u) * DW_STAGE);
		const size_t row0 = (size_t)tile * TILE;
		// A: 8 planes x 128 rows of 16 B; chunk i -> (row = i / 8, plane = i % 8): a thread's global reads follow the row
		for (uint32_t i = tid; i < 8u * TILE; i += DW_THREADS) {
			const uint32_t r = i >> 3, c = i & 7u;
			const bool ok = row0 + r < a.n;
			cp_async16(st + c * A_PLANE + r * 16, L.a + (ok ? (row0 + r) * 64 + c * 8 : 0), ok);
		}
		const uint32_t pb = (uint32_t)L.nb >> 3; // planes of B
		for (uint32_t i = tid; i < pb * TILE; i += DW_THREADS) {
			const uint32_t r = i / pb, c = i % pb;
			const bool ok = row0 + r < a.n;
			cp_async16(st + DW_A_BYTES + c * A_PLANE + r * 16, L.b + (ok ? (row0 + r) * (size_t)L.nb + c * 8 : 0), ok);
		}
		cp_async_commit();
	};

	uint32_t ph[2] = {0u, 0u};
	if (n_items) load_item(0);
	for (uint32_t i = 0; i < n_items; ++i) {
		if (i + 1 < n_items) {
			const uint32_t s = (i + 1) & 1u;
			if (i >= 1) { mbar_wait(smem_addr(bars + s), ph[s]); ph[s] ^= 1u; } // the MMAs of item i - 1 have read that stage
			load_item(i + 1);
			cp_async_wait<1>(); // item i has landed (item i + 1 may still be in flight)
		} else {
			cp_async_wait<0>();
		}
		fence_proxy_async();
		__syncthreads();
		if (tid == 0) {
			tc_fence_after();
			const DwLayer& L = a.layer[i % (uint32_t)a.n_layers];
			const uint32_t st = smem_addr(smem + (i & 1u) * DW_STAGE);
			const uint32_t idesc = instr_desc_mn(TILE, (uint32_t)L.nb);
			const bool first_tile = i < (uint32_t)a.n_layers;
			#pragma unroll
			for (uint32_t kk = 0; kk < 8; ++kk) { // 16 samples per instruction = two 8-row K groups = 256 bytes down every plane
				const uint64_t ad = smem_desc(st + kk * 256u, 128u, A_PLANE);
				const uint64_t bd = smem_desc(st + DW_A_BYTES + kk * 256u, 128u, A_PLANE);
				umma_f16(tmem_base + (uint32_t)L.col, ad, bd, idesc, (first_tile && kk == 0) ? 0u : 1u);
			}
			umma_commit(smem_addr(bars + (i & 1u)));
			if (i + 1 == n_items) umma_commit(smem_addr(bars + 2));
		}
		// no barrier here: the next iteration's loads go to the OTHER stage after waiting for its own MMAs
	}
	if (n_items) {
		mbar_wait(smem_addr(bars + 2), 0u);
		tc_fence_after();
		const uint32_t warp = tid >> 5, lane = tid & 31u;
		if (warp < 2) { // accumulator rows 0..63 are the real ones (rows 64..127 multiply the zero planes)
			const uint32_t m = warp * 32 + lane;
			const uint32_t t_row = tmem_base + ((warp * 32u) << 16);
			for (int l = 0; l < a.n_layers; ++l) {
				const DwLayer& L = a.layer[l];
				for (int c = 0; c < L.nb; c += 16) {
					uint32_t v[16];
					tmem_ld16(t_row + (uint32_t)(L.col + c), v);
					tc_wait_ld();
					#pragma unroll
					for (int j = 0; j < 16; ++j) {
						const float g = __uint_as_float(v[j]);
						float* dst = L.transposed ? L.gw + (size_t)(c + j) * 64 + m : L.gw + (size_t)m * L.nb + (c + j);
						atomicAdd(dst, g);
					}
				}
			}
		}
	}
	tc_fence_before();
	__syncthreads();
	if (tid < 32) tmem_dealloc(tmem_base, DW_TMEM_COLS);
}

} // namespace tc

namespace {
// cudaFuncAttributeMaxDynamicSharedMemorySize once per (kernel, device)
template <typename K>
void raise_smem(K kernel, size_t smem) {
	static std::mutex mu;
	static std::map<std::pair<const void*, int>, size_t> raised;
	int dev = 0;
	cudaGetDevice(&dev);
	std::lock_guard<std::mutex> lock(mu);
	size_t& r = raised[std::make_pair((const void*)kernel, dev)];
	if (r < smem) { cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); r = smem; }
}
int sm_count_now() {
	int dev = 0, n = 0;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
	return n;
}
} // namespace

void launch_train_mlp_forward(const tc::TrainMlpArgs& a, cudaStream_t s) {
	if (!a.n) return;
	constexpr int G = 4;
	const size_t smem = (size_t)G * tc::A_BYTES + tc::mlp_bytes(a.density_hidden) + tc::mlp_bytes(a.rgb_hidden) + 8 * G + 16;
	const uint32_t tiles = (a.n + tc::TILE - 1) / tc::TILE, ctas = (tiles + G - 1) / G, persistent = (uint32_t)sm_count_now() * 2u;
	raise_smem(tc::train_forward_tc_kernel<G>, smem);
	tc::train_forward_tc_kernel<G><<<ctas < persistent ? ctas : persistent, tc::TILE * G, smem, s>>>(a);
}
void launch_train_mlp_backward(const tc::TrainMlpArgs& a, cudaStream_t s) {
	if (!a.n) return;
	constexpr int G = 4;
	const size_t smem = (size_t)G * tc::A_BYTES + tc::mlp_bytes(a.density_hidden) + tc::mlp_bytes(a.rgb_hidden) + 8 * G + 16;
	const uint32_t tiles = (a.n + tc::TILE - 1) / tc::TILE, ctas = (tiles + G - 1) / G, persistent = (uint32_t)sm_count_now() * 2u;
	raise_smem(tc::train_backward_tc_kernel<G>, smem);
	tc::train_backward_tc_kernel<G><<<ctas < persistent ? ctas : persistent, tc::TILE * G, smem, s>>>(a);
}
void launch_train_dw(const tc::DwArgs& a, cudaStream_t s) {
	if (!a.n || !a.n_layers) return;
	const size_t smem = 2 * (size_t)tc::DW_STAGE + 64;
	const uint32_t tiles = (a.n + tc::TILE - 1) / tc::TILE, sms = (uint32_t)sm_count_now();
	raise_smem(tc::train_dw_tc_kernel, smem);
	tc::train_dw_tc_kernel<<<tiles < sms ? tiles : sms, tc::DW_THREADS, smem, s>>>(a);
}

} // namespace hn
