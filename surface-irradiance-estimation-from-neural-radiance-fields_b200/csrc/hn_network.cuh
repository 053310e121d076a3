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
	int unit_box;                   // the training box is the unit cube: warp_position is the identity, the three IEEE divisions are skipped
	const uint4* ray_sh;            // [ray][2] fp16 SH coefficients of the ray direction, computed once per frame (nullptr: per sample)
	int tex_split;                  // option eval_texsplit: which fine-level fetches use the TEX pipe (see launch_eval_tc)
	int groups;                     // option eval_groups: tile groups per CTA of the tcgen05 kernel (4 = 512 threads x 64 registers, 3 = 384 x 80)
	int gather_variant;             // option eval_gather: 1 (default) / 2 = encode_sample_pow2 with / without paired fetches on hashed levels, 0 = encode_sample
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

// ---- hash grid, F = 4, every level with a power-of-two table (configs/nerf/base.json: dense 16^3 / 32^3 / 64^3 and hashed 2^19).
// Same values and the same fp16 FMA order as encode_sample<4> (and so as tcnn's kernel_grid), with the per-sample instruction
// count cut: (i) index reduction is ONE and-mask for dense levels too (no compare / branch / modulo per corner), (ii) whether the
// two x-corners of a pair share one aligned 16-byte block depends only on the parity of the cell's x (the hash XORs x in, dense
// rows are contiguous and every resolution is even), so the pairing test is one predicate per level instead of one per pair; on
// dense levels that parity also says which half of the block is the x-corner (two selects per pair instead of four), (iii) the
// two weights of an x-pair are rounded to fp16 by one cvt and fed to HFMA2 as .H0_H0 / .H1_H1 operands, (iv) no multiplication
// by 1. PAIRH = false fetches the 8 corners of a hashed level as 8 independent 8-byte texels: no selects and no pairing logic at
// all, at the price of 8 instead of ~6 texture requests per level.
// The four features of a level go straight to the caller's A-operand row (a_row + (l / 2) * plane_bytes + (l % 2) * 8): nothing of
// a finished level stays in registers, which leaves them to the loads of the next levels.
template <uint32_t TEXMASK, int NDENSE, bool PAIRH>
__device__ __forceinline__ void encode_sample_pow2(const GridLevels& g, const __half* __restrict__ table, float px, float py, float pz, uint8_t* a_row, uint32_t plane_bytes,
	cudaTextureObject_t tex16, cudaTextureObject_t tex8, uint32_t tex_base) {
	#pragma unroll
	for (int l = 0; l < 8; ++l) {
		const float scale = g.scale[l];
		const uint32_t mask = g.mask[l];
		// pos_fract (common_device.h:842-854)
		float fx = fmaf(scale, px, 0.5f), fy = fmaf(scale, py, 0.5f), fz = fmaf(scale, pz, 0.5f);
		const float flx = floorf(fx), fly = floorf(fy), flz = floorf(fz);
		const uint32_t gx = (uint32_t)(int)flx, gy = (uint32_t)(int)fly, gz = (uint32_t)(int)flz;
		fx -= flx; fy -= fly; fz -= flz;
		const float wx0 = __fsub_rn(1.0f, fx), wy0 = __fsub_rn(1.0f, fy), wz0 = __fsub_rn(1.0f, fz);
		// (y, z) part of the index of the four x-pairs: p = yb + 2 zb
		uint32_t yz[4];
		if (l >= NDENSE) { // coherent prime hash (common_device.h:651-707)
			const uint32_t hy0 = gy * 2654435761u, hz0 = gz * 805459861u, hy1 = hy0 + 2654435761u, hz1 = hz0 + 805459861u;
			yz[0] = hy0 ^ hz0; yz[1] = hy1 ^ hz0; yz[2] = hy0 ^ hz1; yz[3] = hy1 ^ hz1;
		} else {           // dense: x + y R + z R^2
			const uint32_t res = g.res[l];
			const uint32_t ry0 = gy * res, rz0 = gz * res * res, ry1 = ry0 + res, rz1 = rz0 + res * res;
			yz[0] = ry0 + rz0; yz[1] = ry1 + rz0; yz[2] = ry0 + rz1; yz[3] = ry1 + rz1;
		}
		const float wxy[4] = {__fmul_rn(wx0, wy0), __fmul_rn(fx, wy0), __fmul_rn(wx0, fy), __fmul_rn(fx, fy)}; // [xb + 2 yb]
		const bool odd = gx & 1u;
		const uint32_t tbase = tex_base + g.offset[l];
		const __half* __restrict__ lvl = table + (size_t)g.offset[l] * 4;
		__half2 a0 = __float2half2_rn(0.f), a1 = a0;
		#pragma unroll
		for (int p = 0; p < 4; ++p) {
			const uint32_t i0 = (l >= NDENSE ? (gx ^ yz[p]) : (gx + yz[p])) & mask;
			const uint32_t i1 = (l >= NDENSE ? ((gx + 1u) ^ yz[p]) : (gx + 1u + yz[p])) & mask;
			const bool via_tex = (TEXMASK >> (4 * l + p)) & 1u;
			uint2 v0, v1;
			if (l >= NDENSE && !PAIRH) {
				v0 = via_tex ? tex1Dfetch<uint2>(tex8, (int)(tbase + i0)) : __ldg(reinterpret_cast<const uint2*>(lvl) + i0);
				v1 = via_tex ? tex1Dfetch<uint2>(tex8, (int)(tbase + i1)) : __ldg(reinterpret_cast<const uint2*>(lvl) + i1);
			} else {
				const uint4 q = via_tex ? tex1Dfetch<uint4>(tex16, (int)((tbase + i0) >> 1)) : __ldg(reinterpret_cast<const uint4*>(lvl) + (i0 >> 1));
				// which half of the block is corner x: the parity of i0 - on a dense level that is the parity of the cell's x
				const bool hi0 = l >= NDENSE ? (i0 & 1u) != 0u : odd;
				v0 = hi0 ? make_uint2(q.z, q.w) : make_uint2(q.x, q.y);
				v1 = (l >= NDENSE && hi0) ? make_uint2(q.x, q.y) : make_uint2(q.z, q.w);
				if (odd) v1 = via_tex ? tex1Dfetch<uint2>(tex8, (int)(tbase + i1)) : __ldg(reinterpret_cast<const uint2*>(lvl) + i1);
			}
			const float wz = (p & 2) ? fz : wz0;
			const __half2 w2 = __floats2half2_rn(__fmul_rn(wxy[(p & 1) * 2], wz), __fmul_rn(wxy[(p & 1) * 2 + 1], wz));
			const __half2 wl = __low2half2(w2), wh = __high2half2(w2);
			a0 = __hfma2(wl, *reinterpret_cast<const __half2*>(&v0.x), a0);
			a1 = __hfma2(wl, *reinterpret_cast<const __half2*>(&v0.y), a1);
			a0 = __hfma2(wh, *reinterpret_cast<const __half2*>(&v1.x), a0);
			a1 = __hfma2(wh, *reinterpret_cast<const __half2*>(&v1.y), a1);
		}
		*reinterpret_cast<uint2*>(a_row + (l >> 1) * plane_bytes + (l & 1) * 8) = make_uint2(*reinterpret_cast<uint32_t*>(&a0), *reinterpret_cast<uint32_t*>(&a1));
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
void launch_ray_sh(const uint32_t* list, const uint32_t* n_ptr, uint32_t n_max, const float4* d_mw, uint4* out, cudaStream_t s); // hn_network_tc.cu

} // namespace hn
