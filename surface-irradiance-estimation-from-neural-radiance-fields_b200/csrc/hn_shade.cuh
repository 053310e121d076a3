// Activations and colour-space helpers shared by the compositing code (hn_composite.cu and the fused epilogue of
// hn_network_tc.cu): nerf_device.cuh:203-263 (network_to_rgb/density), common_device.cuh srgb<->linear.
#pragma once
#include "hn_internal.h"

namespace hn {

#define HN_MIN_STEP (1.73205080757f / 1024.0f)

__device__ __forceinline__ float act_rgb(float v, int act) {
	switch (act) {
		case 1: return fmaxf(v, 0.0f);
		case 2: return __fdividef(1.0f, 1.0f + __expf(-v));
		case 3: return __expf(fminf(fmaxf(v, -10.0f), 10.0f));
		default: return v;
	}
}
__device__ __forceinline__ float act_density(float v, int act) {
	switch (act) {
		case 1: return fmaxf(v, 0.0f);
		case 2: return 1.0f / (1.0f + __expf(-v));
		case 3: return __expf(v);
		default: return v;
	}
}
__device__ __forceinline__ float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
__device__ __forceinline__ float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

// Everything the per-chunk compositing needs (shared by composite_ray_kernel and the fused epilogue of the tcgen05 kernel)
struct CompArgs {
	const uint32_t* alive_list;   // [n_chunks] ray id per list position
	const float4* samples;        // [n_chunks][32] (x,y,z,dt)
	float train_lo[3], train_diag[3];
	float cam_fwd[3], cam_pos[3];
	float min_transmittance;
	int rgb_act, density_act, linear_colors, depth_test, depth_clip;
	int render_mode;              // ERenderMode (common.h:58-72): 0 AO, 1..4 Shade, 5 Normals, 6 Positions, 7 Depth, 9 Cost, 12 EncodingVis
	float depth_scale;            // 1 / dataset scale (src/testbed_nerf.cu:2121)
	RaySoA rays;
	float4* frame;
	float* depth_buffer;
	uint32_t* survive;            // [n_chunks] 1 = the ray continues into the next pass
	uint32_t* counters;
};

// (d) Compositing of ONE ray chunk by ONE warp: lane j holds sample j of the chunk (warp-segmented scan;
// composite_kernel_nerf src/testbed_nerf.cu:528-735, final blend
// shade_kernel_nerf_geometry src/testbed_geometry_training.cu:1826-1871). Transmittance = warp product scan, early-out
// sample by ballot, colour/alpha by warp reductions, depth = warp arg-max of the weights. The raw network output is
// rounded to fp16 first, so the numbers equal the unfused path's. Lane 0 owns the ray state. Returns samples composited.
// The ray state and the lane's sample are fetched by composite_prefetch() right after the last layer has been issued,
// so their global-memory latency hides behind that MMA round instead of extending the tile.
struct RayPre { uint32_t r, n_steps; float4 rgba; float depth, max_weight; float4 s; float4 o_t; };
__device__ __forceinline__ RayPre composite_prefetch(const CompArgs& a, uint32_t chunk, uint32_t count, int lane) {
	RayPre p;
	p.r = a.alive_list[chunk];
	p.rgba = a.rays.rgba[p.r];
	p.depth = a.rays.depth[p.r];
	p.max_weight = a.rays.d_mw[p.r].w;
	p.n_steps = a.rays.n_steps[p.r];
	p.o_t = a.render_mode == 7 ? a.rays.o_t[p.r] : make_float4(0.f, 0.f, 0.f, 0.f);
	p.s = make_float4(0.f, 0.f, 0.f, 0.f);
	if ((uint32_t)lane < count) p.s = __ldg(a.samples + (size_t)chunk * CHUNK + lane);
	return p;
}
// PLAIN = the common frame fixed at compile time (ERenderMode::Shade, logistic colour, exponential density - configs/*/base.json): the
// run-time switches on activation and render mode cost ~15 % of the instructions of a kernel that is bound by its issue slots.
template <bool PLAIN = false>
__device__ __forceinline__ uint32_t composite_chunk(const CompArgs& a, const RayPre& pre, uint32_t chunk, uint32_t count, int lane, uint2 o) {
	const int render_mode = PLAIN ? 1 : a.render_mode, rgb_act = PLAIN ? 2 : a.rgb_act, density_act = PLAIN ? 3 : a.density_act;
	const uint32_t r = pre.r;
	float4 rgba = pre.rgba;
	float depth = pre.depth;
	float max_weight = pre.max_weight;
	uint32_t n_steps = pre.n_steps;
	bool terminated = false;
	uint32_t composited = 0;
	if (count > 0) {
		const bool has = (uint32_t)lane < count;
		float alpha = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, zview = 0.f;
		if (has) {
			const float4 s = pre.s;
			const __half2 rg = *reinterpret_cast<const __half2*>(&o.x), bs = *reinterpret_cast<const __half2*>(&o.y);
			const float sigma = act_density(__high2float(bs), density_act);
			// the reference sends dt and the position through warp/unwarp (network input); that round trip moves them by at most an
			// ulp, far below the compositing tolerance, and costs five IEEE divisions per sample: use the marched values directly
			alpha = 1.f - __expf(-sigma * s.w);
			cr = act_rgb(__low2float(rg), rgb_act); cg = act_rgb(__high2float(rg), rgb_act); cb = act_rgb(__low2float(bs), rgb_act);
			zview = a.cam_fwd[0] * (s.x - a.cam_pos[0]) + a.cam_fwd[1] * (s.y - a.cam_pos[1]) + a.cam_fwd[2] * (s.z - a.cam_pos[2]);
			// per-sample colour of the visualisation modes (composite_kernel_nerf, src/testbed_nerf.cu:686-702)
			if (render_mode == 6) { cr = (s.x - 0.5f) / 2.0f + 0.5f; cg = (s.y - 0.5f) / 2.0f + 0.5f; cb = (s.z - 0.5f) / 2.0f + 0.5f; }
			else if (render_mode == 12) { cr = (s.x - a.train_lo[0]) / a.train_diag[0]; cg = (s.y - a.train_lo[1]) / a.train_diag[1]; cb = (s.z - a.train_lo[2]) / a.train_diag[2]; }
			else if (render_mode == 7) { cr = cg = cb = (a.cam_fwd[0] * (s.x - pre.o_t.x) + a.cam_fwd[1] * (s.y - pre.o_t.y) + a.cam_fwd[2] * (s.z - pre.o_t.z)) * a.depth_scale; }
			else if (render_mode == 0) { cr = cg = cb = alpha; }
			else if (render_mode == 5) { cr = __low2float(rg); cg = __high2float(rg); cb = __low2float(bs); } // unit normal written by normals_kernel (:688-693)
		}
		float P = 1.f - alpha; // inclusive product scan of (1 - alpha)
		#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const float up = __shfl_up_sync(0xffffffffu, P, d);
			if (lane >= d) P *= up;
		}
		float Pex = __shfl_up_sync(0xffffffffu, P, 1);
		if (lane == 0) Pex = 1.f;
		const float T_in = 1.f - rgba.w;
		const float weight = alpha * (T_in * Pex);
		const float A_after = 1.f - T_in * P;
		const uint32_t term_mask = __ballot_sync(0xffffffffu, has && (A_after > (1.0f - a.min_transmittance)));
		const bool term = term_mask != 0;
		const uint32_t last = term ? (uint32_t)(__ffs(term_mask) - 1) : count - 1;
		const bool use = has && (uint32_t)lane <= last;
		float sw = use ? weight : 0.f, sr = use ? cr * weight : 0.f, sg = use ? cg * weight : 0.f, sb = use ? cb * weight : 0.f;
		// Arg-max of the weights: one warp-wide float max (redux.sync.max.f32 = CREDUX on sm_100a), then the lowest lane that holds it (ties go
		// to the earlier sample). The four sums share one butterfly: after the first two exchanges a lane carries ONE of the four totals (lanes
		// 0-7 alpha, 8-15 red, 16-23 green, 24-31 blue), three more exchanges finish it and four broadcasts hand every lane all four - 10
		// shuffles instead of 30, in a kernel that is bound by its issue slots.
		float bw = use ? weight : -1.f;
		uint32_t bj;
		{
			float wmax;
			asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(wmax) : "f"(bw));
			bj = (uint32_t)(__ffs(__ballot_sync(0xffffffffu, bw == wmax)) - 1) & 31u;
			bw = wmax;
		}
		{
			const bool hi16 = (lane & 16) != 0, hi8 = (lane & 8) != 0;
			const float x0 = (hi16 ? sg : sw) + __shfl_xor_sync(0xffffffffu, hi16 ? sw : sg, 16);
			const float x1 = (hi16 ? sb : sr) + __shfl_xor_sync(0xffffffffu, hi16 ? sr : sb, 16);
			float y = (hi8 ? x1 : x0) + __shfl_xor_sync(0xffffffffu, hi8 ? x0 : x1, 8);
			y += __shfl_xor_sync(0xffffffffu, y, 4);
			y += __shfl_xor_sync(0xffffffffu, y, 2);
			y += __shfl_xor_sync(0xffffffffu, y, 1);
			sw = __shfl_sync(0xffffffffu, y, 0); sr = __shfl_sync(0xffffffffu, y, 8);
			sg = __shfl_sync(0xffffffffu, y, 16); sb = __shfl_sync(0xffffffffu, y, 24);
		}
		const float zbest = __shfl_sync(0xffffffffu, zview, bj);
		rgba.x += sr; rgba.y += sg; rgba.z += sb; rgba.w += sw;
		if (bw > max_weight) { max_weight = bw; depth = zbest; }
		n_steps += last + 1;
		composited = last + 1;
		if (term) {
			const float inv = rgba.w;
			rgba.x /= inv; rgba.y /= inv; rgba.z /= inv; rgba.w /= inv;
			terminated = true;
		}
	}
	const bool alive_next = !terminated && count == CHUNK;
	// every lane holds the same ray state here (the reductions are butterflies); the sRGB -> linear conversion of a finished
	// ray (three powf, ~50 instructions each) is spread over lanes 0..2 so the warp issues it once instead of three times
	const bool finished = !alive_next && rgba.w > 0.001f; // compact_kernel_nerf keeps finished rays with alpha > 0.001
	float4 tmp = rgba;
	const bool shade_mode = render_mode >= 1 && render_mode <= 4; // only Shade (and Slice) accumulate in linear colours (shade_kernel_nerf :1391-1394)
	if (render_mode == 9) { const float col = (float)n_steps / 128.0f; tmp = make_float4(col, col, col, 1.0f); } // Cost (:1381-1383)
	if (render_mode == 5) { // Normals (:1379-1381): (0.5 n + 0.5) * alpha
		const float len = sqrtf(rgba.x * rgba.x + rgba.y * rgba.y + rgba.z * rgba.z);
		tmp.x = (0.5f * (rgba.x / len) + 0.5f) * rgba.w; tmp.y = (0.5f * (rgba.y / len) + 0.5f) * rgba.w; tmp.z = (0.5f * (rgba.z / len) + 0.5f) * rgba.w;
	}
	if (finished && !a.linear_colors && shade_mode) {
		float c = lane == 0 ? rgba.x : (lane == 1 ? rgba.y : rgba.z);
		c = srgb_to_linear(c);
		tmp.x = __shfl_sync(0xffffffffu, c, 0); tmp.y = __shfl_sync(0xffffffffu, c, 1); tmp.z = __shfl_sync(0xffffffffu, c, 2);
	}
	if (lane == 0) {
		a.rays.rgba[r] = rgba;
		a.rays.depth[r] = depth;
		a.rays.d_mw[r].w = max_weight;
		a.rays.n_steps[r] = n_steps;
		a.survive[chunk] = alive_next ? 1u : 0u;
		if (finished) {
			const uint32_t pix = a.rays.pixel[r];
			if (!(a.depth_test && !a.depth_clip && depth > a.depth_buffer[pix])) { // NeRF behind the mesh: drop (per-pixel rule)
				const float4 fb = a.frame[pix];
				const float ia = 1.0f - tmp.w;
				a.frame[pix] = make_float4(tmp.x + fb.x * ia, tmp.y + fb.y * ia, tmp.z + fb.z * ia, tmp.w + fb.w * ia);
				if (tmp.w > 0.2f) a.depth_buffer[pix] = depth;
			}
		}
	}
	return lane == 0 ? composited : 0u;
}

CompArgs make_comp_args(const FrameArgs& fa, RaySoA rays, const uint32_t* alive_list, const float4* samples, float4* frame, float* depth_buffer,
	uint32_t* survive, uint32_t* counters); // hn_composite.cu
void launch_composite_rays(const CompArgs& a, const uint32_t* n_alive_ptr, const uint32_t* sample_counts, const uint2* net_out, uint32_t max_rays, cudaStream_t s);

} // namespace hn
