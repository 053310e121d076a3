// Shared pieces of the network evaluation kernels (hn_network.cu: warp-MMA path, hn_network_tc.cu: tcgen05/TMEM path).
#pragma once
#include "hn_internal.h"
#include "hn_shade.cuh"

namespace hn {

struct EvalArgs {
	GridLevels g;
	const __half* table;
	const __half* w_density;
	const __half* w_rgb;
	int density_hidden, rgb_hidden;
	float train_lo[3], train_diag[3];
	const float4* samples;          // [n_chunks][32] (x,y,z,dt) un-warped
	const uint32_t* sample_counts;  // [n_chunks] valid samples per chunk (nullptr = all 32)
	const uint32_t* alive_list;     // [n_chunks] ray id per chunk (nullptr with dirs_override)
	const float4* d_mw;             // ray directions by ray id
	const float* dirs_override;     // [n_chunks][3] (operator entry point)
	const uint32_t* n_chunks_ptr;   // device-side chunk count (nullptr -> n_chunks)
	uint32_t n_chunks;
	uint2* out;                     // [n_chunks][32] (r,g | b,sigma) fp16 raw network output (unfused path)
	// fused compositing (tcgen05 kernel only): the warp that evaluated a ray chunk also composites it
	int fused;
	int density_only;               // tcgen05 kernel: stop after the density MLP, out = (0, 0 | 0, sigma) (light rays, a-16)
	cudaTextureObject_t tex16, tex8; // the whole parameter buffer as 16-byte / 8-byte texels (TEX-pipe gather of some levels)
	uint32_t tex_base;              // grid table offset inside the parameter buffer, in 8-byte texels
	CompArgs comp;
};

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
	__half2 h = __floats2half2_rn(a, b);
	return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- hash grid: one sample, all levels. F = features per level (2 or 4), L = 32 / F levels -> 32 halfs in `enc`.
// The 8 corners are fetched as 4 x-pairs: when the table slots of (x, y, z) and (x+1, y, z) are the two halves of one
// aligned 2-entry block (always for even x: the hash only XORs x in, dense rows are contiguous) ONE load brings both, which
// removes a quarter of the gather instructions and of the L1 wavefronts. Values and the order of the fp16 FMAs are those
// of tcnn's kernel_grid (grid.h:94-168), so the result stays bit-identical to the oracle.
template <int F, uint32_t TEXMASK = 0, int NDENSE = -1>
__device__ __forceinline__ void encode_sample(const GridLevels& g, const __half* __restrict__ table, float px, float py, float pz, uint32_t (&enc)[16],
	cudaTextureObject_t tex16 = 0, cudaTextureObject_t tex8 = 0, uint32_t tex_base = 0) {
	constexpr int L = 32 / F;
	#pragma unroll
	for (int l = 0; l < L; ++l) {
		const float scale = g.scale[l];
		const uint32_t res = g.res[l], size = g.size[l], mask = g.mask[l];
		const __half* __restrict__ lvl = table + (size_t)g.offset[l] * F;
		// pos_fract (common_device.h:842-854)
		float fx = fmaf(scale, px, 0.5f), fy = fmaf(scale, py, 0.5f), fz = fmaf(scale, pz, 0.5f);
		float flx = floorf(fx), fly = floorf(fy), flz = floorf(fz);
		uint32_t gx = (uint32_t)(int)flx, gy = (uint32_t)(int)fly, gz = (uint32_t)(int)flz;
		fx -= flx; fy -= fly; fz -= flz;
		float wx[2] = {__fsub_rn(1.0f, fx), fx}, wy[2] = {__fsub_rn(1.0f, fy), fy}, wz[2] = {__fsub_rn(1.0f, fz), fz};
		// NDENSE >= 0: the launcher has checked that exactly the first NDENSE levels are dense and every hashed level has a
		// power-of-two table, so the level kind and the index reduction are compile-time per unrolled level
		const bool hashed = NDENSE >= 0 ? (l >= NDENSE) : (g.hashed[l] != 0);
		// grid_index (common_device.h:690-707): dense x + y*R + z*R^2, or coherent prime hash; then % hashmap_size
		// (a power-of-two size is a mask; a dense index below the size needs no reduction at all)
		uint32_t idx[8];
		#pragma unroll
		for (int c = 0; c < 8; ++c) {
			uint32_t x = gx + (c & 1), y = gy + ((c >> 1) & 1), z = gz + ((c >> 2) & 1);
			uint32_t i = hashed ? (x ^ (y * 2654435761u) ^ (z * 805459861u)) : (x + y * res + z * res * res);
			if (NDENSE >= 0) idx[c] = (l >= NDENSE) ? (i & mask) : (i < size ? i : i % size);
			else idx[c] = mask ? (i & mask) : (i < size ? i : i % size);
		}
		if (F == 4) {
			uint2 v[8];
			#pragma unroll
			for (int c = 0; c < 8; c += 2) {
				const bool paired = (idx[c] ^ idx[c + 1]) == 1u;
				const bool via_tex = (TEXMASK >> (4 * l + (c >> 1))) & 1u; // compile-time per unrolled (level, corner pair): this gather rides the TEX pipe
				const uint32_t tbase = tex_base + g.offset[l];
				const uint4 q = via_tex ? tex1Dfetch<uint4>(tex16, (int)((tbase + idx[c]) >> 1)) : __ldg(reinterpret_cast<const uint4*>(lvl) + (idx[c] >> 1));
				const bool odd = idx[c] & 1u;
				v[c] = odd ? make_uint2(q.z, q.w) : make_uint2(q.x, q.y);
				v[c + 1] = odd ? make_uint2(q.x, q.y) : make_uint2(q.z, q.w);
				if (!paired) v[c + 1] = via_tex ? tex1Dfetch<uint2>(tex8, (int)(tbase + idx[c + 1])) : __ldg(reinterpret_cast<const uint2*>(lvl) + idx[c + 1]);
			}
			__half2 a0 = __float2half2_rn(0.f), a1 = a0;
			#pragma unroll
			for (int c = 0; c < 8; ++c) {
				float w = __fmul_rn(__fmul_rn(__fmul_rn(1.0f, wx[c & 1]), wy[(c >> 1) & 1]), wz[(c >> 2) & 1]);
				__half2 w2 = __float2half2_rn(w);
				a0 = __hfma2(w2, *reinterpret_cast<__half2*>(&v[c].x), a0);
				a1 = __hfma2(w2, *reinterpret_cast<__half2*>(&v[c].y), a1);
			}
			enc[2 * l] = *reinterpret_cast<uint32_t*>(&a0);
			enc[2 * l + 1] = *reinterpret_cast<uint32_t*>(&a1);
		} else {
			uint32_t v[8];
			#pragma unroll
			for (int c = 0; c < 8; c += 2) {
				const bool paired = (idx[c] ^ idx[c + 1]) == 1u;
				const uint2 q = __ldg(reinterpret_cast<const uint2*>(lvl) + (idx[c] >> 1));
				const bool odd = idx[c] & 1u;
				v[c] = odd ? q.y : q.x;
				v[c + 1] = odd ? q.x : q.y;
				if (!paired) v[c + 1] = __ldg(reinterpret_cast<const uint32_t*>(lvl) + idx[c + 1]);
			}
			__half2 a0 = __float2half2_rn(0.f);
			#pragma unroll
			for (int c = 0; c < 8; ++c) {
				float w = __fmul_rn(__fmul_rn(__fmul_rn(1.0f, wx[c & 1]), wy[(c >> 1) & 1]), wz[(c >> 2) & 1]);
				a0 = __hfma2(__float2half2_rn(w), *reinterpret_cast<__half2*>(&v[c]), a0);
			}
			enc[l] = *reinterpret_cast<uint32_t*>(&a0);
		}
	}
}

// ---- degree-4 real spherical harmonics (common_device.h:340-366), input = unit direction components
__device__ __forceinline__ void sh4(float x, float y, float z, float (&o)[16]) {
	float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
	o[0] = 0.28209479177387814f;
	o[1] = -0.48860251190291987f * y;
	o[2] = 0.48860251190291987f * z;
	o[3] = -0.48860251190291987f * x;
	o[4] = 1.0925484305920792f * xy;
	o[5] = -1.0925484305920792f * yz;
	o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
	o[7] = -1.0925484305920792f * xz;
	o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
	o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
	o[10] = 2.8906114426405538f * xy * z;
	o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
	o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
	o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
	o[14] = 1.4453057213202769f * z * (x2 - y2);
	o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
}

void launch_eval_tc(const EvalArgs& a, uint32_t F, uint32_t max_chunks, cudaStream_t s); // hn_network_tc.cu

} // namespace hn
