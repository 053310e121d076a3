// (d) Alpha compositing as a warp-segmented scan + warp-level compaction of live rays + final depth-merged blend.
// A warp owns 32 consecutive live rays (lane k keeps ray k's state in registers, loaded/stored coalesced). For each of its
// rays in turn the 32 lanes take the ray's 32 samples of this pass: transmittance is a warp product scan, the early-out
// sample is found with a ballot, colour/alpha are warp reductions, depth is a warp arg-max. Survivors are appended to the
// next pass's list with ONE atomic per warp; rays that finish are blended straight into the frame buffer with the
// mesh depth test (composite_kernel_nerf src/testbed_nerf.cu:528-735, compact_kernel_nerf :1403-1426,
// shade_kernel_nerf_geometry src/testbed_geometry_training.cu:1826-1871).
#include "hn_shade.cuh"

namespace hn {

__global__ void __launch_bounds__(256) composite_kernel(FrameArgs a, RaySoA rays, const uint32_t* __restrict__ alive_in, const uint32_t* __restrict__ n_alive_ptr,
	uint32_t* __restrict__ alive_out, uint32_t* __restrict__ n_alive_out, uint32_t* __restrict__ counters, const float4* __restrict__ samples,
	const uint32_t* __restrict__ sample_counts, const uint2* __restrict__ net_out, float4* __restrict__ frame, float* __restrict__ depth_buffer) {
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const uint32_t n_alive = *n_alive_ptr;
	const uint32_t p0 = warp_global * 32u;
	if (p0 >= n_alive) return;
	const uint32_t p = p0 + lane;
	const bool valid = p < n_alive;

	uint32_t r = 0, count = 0, n_steps = 0;
	float4 rgba = make_float4(0.f, 0.f, 0.f, 0.f);
	float depth = 0.f, max_weight = 0.f;
	if (valid) {
		r = alive_in[p];
		count = sample_counts[p];
		rgba = rays.rgba[r];
		depth = rays.depth[r];
		max_weight = rays.d_mw[r].w;
		n_steps = rays.n_steps[r];
	}
	bool terminated = false;
	const float cfx = a.cam1[6], cfy = a.cam1[7], cfz = a.cam1[8];
	const float cpx = a.cam1[9], cpy = a.cam1[10], cpz = a.cam1[11];
	const float dt_range = __fsub_rn(__fmul_rn(HN_MIN_STEP, 128.0f), HN_MIN_STEP); // unwarp_dt range, nerf_device.cuh:306-314
	uint32_t composited_total = 0;

	const uint32_t n_here = min(32u, n_alive - p0);
	for (uint32_t rr = 0; rr < n_here; ++rr) {
		const uint32_t cnt = __shfl_sync(0xffffffffu, count, rr);
		if (cnt == 0) continue;
		const float A_in = __shfl_sync(0xffffffffu, rgba.w, rr);
		const float mw_in = __shfl_sync(0xffffffffu, max_weight, rr);
		const bool has = lane < cnt;
		float alpha = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, zview = 0.f;
		if (has) {
			const size_t si = (size_t)(p0 + rr) * CHUNK + lane;
			const uint2 o = __ldg(net_out + si);
			const float4 s = __ldg(samples + si);
			const __half2 rg = *reinterpret_cast<const __half2*>(&o.x), bs = *reinterpret_cast<const __half2*>(&o.y);
			const float sigma = act_density(__high2float(bs), a.density_act);
			// warp_dt then unwarp_dt, as the value travels through the network input in the reference
			const float wdt = __fdiv_rn(__fsub_rn(s.w, HN_MIN_STEP), dt_range);
			const float dt = __fadd_rn(__fmul_rn(wdt, dt_range), HN_MIN_STEP);
			alpha = 1.f - __expf(-sigma * dt);
			cr = act_rgb(__low2float(rg), a.rgb_act); cg = act_rgb(__high2float(rg), a.rgb_act); cb = act_rgb(__low2float(bs), a.rgb_act);
			// unwarp_position(warp_position(pos)) (nerf_device.cuh:265-279), view depth along the camera forward axis
			const float lx = a.train_aabb[0], ly = a.train_aabb[1], lz = a.train_aabb[2];
			const float dx = a.train_aabb[3] - lx, dy = a.train_aabb[4] - ly, dz = a.train_aabb[5] - lz;
			const float ux = __fadd_rn(lx, __fmul_rn(__fdiv_rn(__fsub_rn(s.x, lx), dx), dx));
			const float uy = __fadd_rn(ly, __fmul_rn(__fdiv_rn(__fsub_rn(s.y, ly), dy), dy));
			const float uz = __fadd_rn(lz, __fmul_rn(__fdiv_rn(__fsub_rn(s.z, lz), dz), dz));
			zview = cfx * (ux - cpx) + cfy * (uy - cpy) + cfz * (uz - cpz);
		}
		// inclusive product scan of (1 - alpha): P_j = prod_{k<=j}
		float P = 1.f - alpha;
		#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			float up = __shfl_up_sync(0xffffffffu, P, d);
			if (lane >= (uint32_t)d) P *= up;
		}
		float Pex = __shfl_up_sync(0xffffffffu, P, 1);
		if (lane == 0) Pex = 1.f;
		const float T_in = 1.f - A_in;
		const float weight = alpha * (T_in * Pex);
		const float A_after = 1.f - T_in * P;
		// first sample whose accumulated alpha exceeds 1 - min_transmittance ends the ray (:716-719)
		const uint32_t term_mask = __ballot_sync(0xffffffffu, has && (A_after > (1.0f - a.min_transmittance)));
		const bool term = term_mask != 0;
		const uint32_t last = term ? (uint32_t)(__ffs(term_mask) - 1) : cnt - 1;
		const bool use = has && lane <= last;
		float sw = use ? weight : 0.f, sr = use ? cr * weight : 0.f, sg = use ? cg * weight : 0.f, sb = use ? cb * weight : 0.f;
		// arg-max of the weights (earliest sample wins ties, strict > against the carried max, :711-714)
		float bw = use ? weight : -1.f;
		uint32_t bj = lane;
		#pragma unroll
		for (int d = 16; d > 0; d >>= 1) {
			sw += __shfl_xor_sync(0xffffffffu, sw, d);
			sr += __shfl_xor_sync(0xffffffffu, sr, d);
			sg += __shfl_xor_sync(0xffffffffu, sg, d);
			sb += __shfl_xor_sync(0xffffffffu, sb, d);
			float ow = __shfl_xor_sync(0xffffffffu, bw, d);
			uint32_t oj = __shfl_xor_sync(0xffffffffu, bj, d);
			if (ow > bw || (ow == bw && oj < bj)) { bw = ow; bj = oj; }
		}
		const float zbest = __shfl_sync(0xffffffffu, zview, bj);
		if (lane == rr) {
			rgba.x += sr; rgba.y += sg; rgba.z += sb; rgba.w += sw;
			if (bw > mw_in) { max_weight = bw; depth = zbest; }
			n_steps += last + 1;
			composited_total = last + 1;
			if (term) {
				const float inv = rgba.w;
				rgba.x /= inv; rgba.y /= inv; rgba.z /= inv; rgba.w /= inv;
				terminated = true;
			}
		}
	}

	// ---- state write-back, compaction of survivors (one atomic per warp), blend of finished rays
	const bool alive_next = valid && !terminated && count == CHUNK;
	if (valid) {
		rays.rgba[r] = rgba;
		rays.depth[r] = depth;
		rays.d_mw[r].w = max_weight;
		rays.n_steps[r] = n_steps;
	}
	const uint32_t amask = __ballot_sync(0xffffffffu, alive_next);
	if (amask) {
		uint32_t base = 0;
		const int leader = __ffs(amask) - 1;
		if ((int)lane == leader) base = atomicAdd(n_alive_out, __popc(amask));
		base = __shfl_sync(0xffffffffu, base, leader);
		if (alive_next) alive_out[base + __popc(amask & ((1u << lane) - 1u))] = r;
	}
	uint32_t ct = composited_total;
	#pragma unroll
	for (int d = 16; d > 0; d >>= 1) ct += __shfl_xor_sync(0xffffffffu, ct, d);
	if (lane == 0 && ct) atomicAdd((unsigned long long*)&counters[CNT_COMPOSITED_LO], (unsigned long long)ct);

	if (valid && !alive_next && rgba.w > 0.001f) { // compact_kernel_nerf keeps finished rays with alpha > 0.001
		const uint32_t pix = rays.pixel[r];
		if (a.depth_test && !a.depth_clip && depth > depth_buffer[pix]) return; // NeRF behind the mesh: drop (per-pixel rule)
		float4 tmp = rgba;
		if (!a.linear_colors) { tmp.x = srgb_to_linear(tmp.x); tmp.y = srgb_to_linear(tmp.y); tmp.z = srgb_to_linear(tmp.z); }
		const float4 fb = frame[pix];
		const float ia = 1.0f - tmp.w;
		frame[pix] = make_float4(tmp.x + fb.x * ia, tmp.y + fb.y * ia, tmp.z + fb.z * ia, tmp.w + fb.w * ia);
		if (tmp.w > 0.2f) depth_buffer[pix] = depth;
	}
}

// render_buffer.cu:228-262 (Linear colour space): running mean over spp
__global__ void accumulate_kernel(const float4* __restrict__ frame, float4* __restrict__ accum, uint32_t n, float sample_count) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 c = frame[i], t = sample_count > 0.f ? accum[i] : make_float4(0.f, 0.f, 0.f, 0.f);
	float inv = sample_count + 1.f;
	t.x = (t.x * sample_count + c.x) / inv; t.y = (t.y * sample_count + c.y) / inv;
	t.z = (t.z * sample_count + c.z) / inv; t.w = (t.w * sample_count + c.w) / inv;
	accum[i] = t;
}
// tonemap curves (ETonemapCurve: 0 Identity, 1 ACES, 2 Hable, 3 Reinhard; src/render_buffer.cu:264-321)
__device__ __forceinline__ void tonemap_curve(float& r, float& g, float& b, int curve) {
	if (curve == 0) return;
	r = fmaxf(r, 0.0f); g = fmaxf(g, 0.0f); b = fmaxf(b, 0.0f);
	float k0, k1, k2, k3, k4, k5;
	if (curve == 1) {
		k0 = 0.6f * 0.6f * 2.51f; k1 = 0.6f * 0.03f; k2 = 0.0f; k3 = 0.6f * 0.6f * 2.43f; k4 = 0.6f * 0.59f; k5 = 0.14f;
	} else if (curve == 2) {
		const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
		k0 = A * F - A * E; k1 = C * B * F - B * E; k2 = 0.0f; k3 = A * F; k4 = B * F; k5 = D * F * F;
		const float W = 11.2f;
		const float nom = k0 * (W * W) + k1 * W + k2, denom = k3 * (W * W) + k4 * W + k5;
		const float white_scale = denom / nom;
		k0 = 4.0f * k0 * white_scale; k1 = 2.0f * k1 * white_scale; k2 = k2 * white_scale; k3 = 4.0f * k3; k4 = 2.0f * k4;
	} else {
		const float Y = 0.2126f * r + 0.7152f * g + 0.0722f * b;
		const float s = 1.f / (Y + 1.0f);
		r *= s; g *= s; b *= s;
		return;
	}
	r = (r * r * k0 + k1 * r + k2) / (k3 * (r * r) + k4 * r + k5);
	g = (g * g * k0 + k1 * g + k2) / (k3 * (g * g) + k4 * g + k5);
	b = (b * b * k0 + k1 * b + k2) / (k3 * (b * b) + k4 * b + k5);
}
// render_buffer.cu:529-560 (accumulation in linear colours): background blend, exposure, curve, optional sRGB, optional alpha
// un-multiply, optional clamp; linear device buffer instead of a surface
__global__ void tonemap_kernel(const float4* __restrict__ accum, float4* __restrict__ out, uint32_t n, int to_srgb, float exposure_scale, float4 bg,
	int curve, int clamp_output, int unmultiply_alpha) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 c = accum[i];
	float w = (1.f - c.w) * bg.w;
	c.x += srgb_to_linear(bg.x) * w; c.y += srgb_to_linear(bg.y) * w; c.z += srgb_to_linear(bg.z) * w; c.w += w;
	c.x *= exposure_scale; c.y *= exposure_scale; c.z *= exposure_scale;
	tonemap_curve(c.x, c.y, c.z, curve);
	if (to_srgb) { c.x = linear_to_srgb(c.x); c.y = linear_to_srgb(c.y); c.z = linear_to_srgb(c.z); }
	if (unmultiply_alpha && c.w > 0.0f) { c.x /= c.w; c.y /= c.w; c.z /= c.w; }
	if (clamp_output) { c.x = fminf(fmaxf(c.x, 0.f), 1.f); c.y = fminf(fmaxf(c.y, 0.f), 1.f); c.z = fminf(fmaxf(c.z, 0.f), 1.f); c.w = fminf(fmaxf(c.w, 0.f), 1.f); }
	out[i] = c;
}

// tile (un)packing for the multi-GPU gather: packed = [owned tile][tile_h*tile_w] x (rgba, depth)
__global__ void pack_tiles_kernel(int w, int h, int rank, int world, int tw, int th, float4* frame, float* depth, float* packed, int unpack) {
	uint32_t tiles_x = (w + tw - 1) / tw, tiles_y = (h + th - 1) / th;
	uint32_t n_tiles = tiles_x * tiles_y;
	uint32_t per_tile = tw * th;
	uint32_t local_tile = blockIdx.x;
	uint32_t tile = local_tile * world + rank;
	if (tile >= n_tiles) return;
	uint32_t tx = tile % tiles_x, ty = tile / tiles_x;
	for (uint32_t k = threadIdx.x; k < per_tile; k += blockDim.x) {
		uint32_t x = tx * tw + k % tw, y = ty * th + k / tw;
		if (x >= (uint32_t)w || y >= (uint32_t)h) continue;
		float* dst = packed + ((size_t)local_tile * per_tile + k) * 5;
		uint32_t pix = x + w * y;
		if (unpack) {
			frame[pix] = make_float4(dst[0], dst[1], dst[2], dst[3]);
			depth[pix] = dst[4];
		} else {
			float4 c = frame[pix];
			dst[0] = c.x; dst[1] = c.y; dst[2] = c.z; dst[3] = c.w; dst[4] = depth[pix];
		}
	}
}

// Multi-GPU frame gather without a collective: a rank copies the tiles it owns from its local frame / depth straight into the ROOT
// rank's buffers, mapped into this process through CUDA IPC - plain stores to peer memory over NVLink / NVSwitch, one CTA per
// owned tile, 16-byte coalesced rows (1 KB of colour + 256 B of depth per 64-pixel tile row). Replaces pack -> all_gather (every
// rank receiving all 41.5 MB) -> one unpack launch per peer.
__global__ void __launch_bounds__(256) scatter_tiles_kernel(int w, int h, int rank, int world, int tw, int th, const float4* __restrict__ frame, const float* __restrict__ depth,
	float4* __restrict__ frame_remote, float* __restrict__ depth_remote) {
	const uint32_t tiles_x = (w + tw - 1) / tw, tiles_y = (h + th - 1) / th;
	const uint32_t tile = blockIdx.x * world + rank;
	if (tile >= tiles_x * tiles_y) return;
	const uint32_t tx = tile % tiles_x, ty = tile / tiles_x;
	for (uint32_t k = threadIdx.x; k < (uint32_t)(tw * th); k += blockDim.x) {
		const uint32_t x = tx * tw + k % tw, y = ty * th + k / tw;
		if (x >= (uint32_t)w || y >= (uint32_t)h) continue;
		const uint32_t pix = x + w * y;
		frame_remote[pix] = frame[pix];
		if (depth_remote) depth_remote[pix] = depth[pix];
	}
}
void launch_scatter_tiles(int w, int h, int rank, int world, int tw, int th, const float4* frame, const float* depth, float4* frame_remote, float* depth_remote, cudaStream_t s) {
	const uint32_t tiles = ((w + tw - 1) / tw) * ((h + th - 1) / th);
	const uint32_t mine = (tiles + world - 1 - rank) / world;
	if (!mine) return;
	scatter_tiles_kernel<<<mine, 256, 0, s>>>(w, h, rank, world, tw, th, frame, depth, frame_remote, depth_remote);
}

CompArgs make_comp_args(const FrameArgs& fa, RaySoA rays, const uint32_t* alive_list, const float4* samples, float4* frame, float* depth_buffer,
	uint32_t* survive, uint32_t* counters) {
	CompArgs c;
	c.alive_list = alive_list; c.samples = samples;
	for (int i = 0; i < 3; ++i) {
		c.train_lo[i] = fa.train_aabb[i]; c.train_diag[i] = fa.train_aabb[3 + i] - fa.train_aabb[i];
		c.cam_fwd[i] = fa.cam1[6 + i]; c.cam_pos[i] = fa.cam1[9 + i];
	}
	c.min_transmittance = fa.min_transmittance; c.rgb_act = fa.rgb_act; c.density_act = fa.density_act; c.linear_colors = fa.linear_colors;
	c.depth_test = fa.depth_test; c.depth_clip = fa.depth_clip;
	c.render_mode = fa.render_mode; c.depth_scale = fa.depth_scale;
	c.rays = rays; c.frame = frame; c.depth_buffer = depth_buffer; c.survive = survive; c.counters = counters;
	return c;
}

// One warp per live ray (lane j = sample j of this pass's chunk): every load of the kernel is issued up front, so the
// kernel runs at memory latency x waves instead of 32 dependent round trips per warp (composite_kernel above keeps 32 rays
// per warp and is latency-bound by its serial ray loop). Survivors are flagged; compact_kernel builds the next list.
template <bool PLAIN>
__global__ void __launch_bounds__(512) composite_ray_kernel(const __grid_constant__ CompArgs a, const uint32_t* __restrict__ n_alive_ptr,
	const uint32_t* __restrict__ sample_counts, const uint2* __restrict__ net_out, uint32_t max_rays) {
	__shared__ uint32_t block_total;
	if (threadIdx.x == 0) block_total = 0;
	__syncthreads();
	const uint32_t chunk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (chunk < max_rays) { // warp-uniform; max_rays = the host's upper bound of the live count, the size every buffer below was grown to
		// The kernel is bound by its chain of dependent loads (live count -> sample count -> sample / network output, list slot -> ray state),
		// not by bandwidth: everything indexed by the chunk alone is requested at once, whether or not the chunk turns out to be live or the
		// lane to hold a sample (stale rows of an earlier pass sit there: read, never used). Only the ray state waits for the list slot.
		// (volatile asm: the compiler otherwise sinks these loads back behind the branches that use them)
		uint2 o_raw; float4 s_raw; uint32_t count;
		asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(o_raw.x), "=r"(o_raw.y) : "l"(net_out + (size_t)chunk * CHUNK + lane));
		asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(s_raw.x), "=f"(s_raw.y), "=f"(s_raw.z), "=f"(s_raw.w) : "l"(a.samples + (size_t)chunk * CHUNK + lane));
		asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(count) : "l"(sample_counts + chunk));
		if (chunk < *n_alive_ptr) {
			RayPre pre;
			pre.r = a.alive_list[chunk];
			pre.rgba = a.rays.rgba[pre.r];
			pre.depth = a.rays.depth[pre.r];
			pre.max_weight = a.rays.d_mw[pre.r].w;
			pre.n_steps = a.rays.n_steps[pre.r];
			pre.o_t = (!PLAIN && a.render_mode == 7) ? a.rays.o_t[pre.r] : make_float4(0.f, 0.f, 0.f, 0.f);
			const bool has = (uint32_t)lane < count;
			pre.s = has ? s_raw : make_float4(0.f, 0.f, 0.f, 0.f);
			const uint32_t n = composite_chunk<PLAIN>(a, pre, chunk, count, lane, has ? o_raw : make_uint2(0u, 0u));
			if (lane == 0 && n) atomicAdd(&block_total, n);
		}
	}
	__syncthreads();
	if (threadIdx.x == 0 && block_total) atomicAdd((unsigned long long*)&a.counters[CNT_COMPOSITED_LO], (unsigned long long)block_total);
}
void launch_composite_rays(const CompArgs& a, const uint32_t* n_alive_ptr, const uint32_t* sample_counts, const uint2* net_out, uint32_t max_rays, cudaStream_t s) {
	if (!max_rays) return;
	const bool plain = a.render_mode >= 1 && a.render_mode <= 4 && a.rgb_act == 2 && a.density_act == 3;
	if (plain) composite_ray_kernel<true><<<(max_rays + 15) / 16, 512, 0, s>>>(a, n_alive_ptr, sample_counts, net_out, max_rays);
	else composite_ray_kernel<false><<<(max_rays + 15) / 16, 512, 0, s>>>(a, n_alive_ptr, sample_counts, net_out, max_rays);
}

// ERenderMode::Slice, second half (compute_nerf_rgba with depth 0.01, src/testbed_nerf.cu:408-421, then shade_kernel_nerf :1361-1401):
// alpha = clamp(1 - exp(-0.01 sigma)), colour premultiplied, taken to linear like Shade; the depth buffer is left alone
__global__ void __launch_bounds__(256) slice_shade_kernel(uint32_t n, const uint32_t* __restrict__ pixels, const uint2* __restrict__ net_out, int rgb_act, int density_act,
	int linear_colors, float4* __restrict__ frame) {
	const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= n) return;
	const uint32_t idx = pixels[p];
	if (idx == 0xffffffffu) return;
	const uint2 o = __ldg(net_out + (size_t)p * CHUNK);
	const __half2 rg = *reinterpret_cast<const __half2*>(&o.x), bs = *reinterpret_cast<const __half2*>(&o.y);
	const float alpha = fminf(fmaxf(1.f - __expf(-act_density(__high2float(bs), density_act) * 0.01f), 0.0f), 1.0f);
	float r = act_rgb(__low2float(rg), rgb_act) * alpha, g = act_rgb(__high2float(rg), rgb_act) * alpha, b = act_rgb(__low2float(bs), rgb_act) * alpha;
	if (!linear_colors) { r = srgb_to_linear(r); g = srgb_to_linear(g); b = srgb_to_linear(b); }
	const float4 fb = frame[idx];
	const float ia = 1.0f - alpha;
	frame[idx] = make_float4(r + fb.x * ia, g + fb.y * ia, b + fb.z * ia, alpha + fb.w * ia);
}
void launch_slice_shade(uint32_t n, const uint32_t* pixels, const uint2* net_out, int rgb_act, int density_act, int linear_colors, float4* frame, cudaStream_t s) {
	if (n) slice_shade_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, pixels, net_out, rgb_act, density_act, linear_colors, frame);
}
// the blit of a foveated frame (Testbed::blit_texture + its fragment shader, src/testbed.cu:2600-2640, without the GL y flip): output
// pixel centre uv -> unwarp(uv) -> bilinear fetch (GL_LINEAR, clamp to edge) of the rendered, warped frame and depth
struct FovCoeffs { float c[2][12]; };
__device__ __forceinline__ float fovc_unwarp(const float* c, float y) { // FoveationPiecewiseQuadratic::unwarp, common_device.cuh:210-219
	y = fminf(fmaxf(y, 0.0f), 1.0f);
	if (y < c[10]) return (sqrtf(-4 * c[0] * c[2] + 4 * c[0] * y + c[1] * c[1]) - c[1]) / (2 * c[0]);
	if (y > c[11]) return (sqrtf(-4 * c[5] * c[7] + 4 * c[5] * y + c[6] * c[6]) - c[6]) / (2 * c[5]);
	return (y - c[4]) / c[3];
}
__global__ void __launch_bounds__(256) unwarp_foveated_kernel(FovCoeffs f, int w_in, int h_in, const float4* __restrict__ frame_in, const float* __restrict__ depth_in,
	int w_out, int h_out, float4* __restrict__ frame_out, float* __restrict__ depth_out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= (uint32_t)w_out * (uint32_t)h_out) return;
	const int x = (int)(i % (uint32_t)w_out), y = (int)(i / (uint32_t)w_out);
	const float u = fovc_unwarp(f.c[0], ((float)x + 0.5f) / (float)w_out), v = fovc_unwarp(f.c[1], ((float)y + 0.5f) / (float)h_out);
	// GL_LINEAR: texel centres at (k + 0.5) / N
	const float sx = u * (float)w_in - 0.5f, sy = v * (float)h_in - 0.5f;
	const float fx0 = floorf(sx), fy0 = floorf(sy);
	const float wx = sx - fx0, wy = sy - fy0;
	const int x0 = min(max((int)fx0, 0), w_in - 1), x1 = min(max((int)fx0 + 1, 0), w_in - 1), y0 = min(max((int)fy0, 0), h_in - 1), y1 = min(max((int)fy0 + 1, 0), h_in - 1);
	const float w00 = (1 - wx) * (1 - wy), w10 = wx * (1 - wy), w01 = (1 - wx) * wy, w11 = wx * wy;
	const float4 a = frame_in[x0 + y0 * w_in], b = frame_in[x1 + y0 * w_in], c = frame_in[x0 + y1 * w_in], d = frame_in[x1 + y1 * w_in];
	frame_out[i] = make_float4(w00 * a.x + w10 * b.x + w01 * c.x + w11 * d.x, w00 * a.y + w10 * b.y + w01 * c.y + w11 * d.y,
		w00 * a.z + w10 * b.z + w01 * c.z + w11 * d.z, w00 * a.w + w10 * b.w + w01 * c.w + w11 * d.w);
	if (depth_in && depth_out)
		depth_out[i] = w00 * depth_in[x0 + y0 * w_in] + w10 * depth_in[x1 + y0 * w_in] + w01 * depth_in[x0 + y1 * w_in] + w11 * depth_in[x1 + y1 * w_in];
}
void launch_unwarp_foveated(const float coeffs[24], int w_in, int h_in, const float4* frame_in, const float* depth_in, int w_out, int h_out, float4* frame_out, float* depth_out, cudaStream_t s) {
	FovCoeffs f;
	for (int i = 0; i < 24; ++i) f.c[i / 12][i % 12] = coeffs[i];
	const uint32_t n = (uint32_t)w_out * (uint32_t)h_out;
	if (n) unwarp_foveated_kernel<<<(n + 255) / 256, 256, 0, s>>>(f, w_in, h_in, frame_in, depth_in, w_out, h_out, frame_out, depth_out);
}

// a-16 light rays: one warp per ray chunk, lane j = sample j. T_j = T_in * prod_{k<=j} exp(-sigma_k dt_k) (warp product scan);
// the ray ends at the first sample with T_j < min_transmittance (oracle nerf_transmittance) or when it leaves the volume.
__global__ void __launch_bounds__(512) transmittance_kernel(FrameArgs a, RaySoA rays, const uint32_t* __restrict__ alive_list,
	const uint32_t* __restrict__ n_alive_ptr, const float4* __restrict__ samples, const uint32_t* __restrict__ sample_counts,
	const uint2* __restrict__ net_out, uint32_t* __restrict__ survive, uint32_t* __restrict__ counters, uint32_t max_rays) {
	__shared__ uint32_t block_total;
	if (threadIdx.x == 0) block_total = 0;
	__syncthreads();
	const uint32_t chunk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31u;
	uint2 o = make_uint2(0u, 0u); float dt_in = 0.0f; uint32_t count = 0;
	if (chunk < max_rays) { // as composite_ray_kernel: what the chunk index alone addresses is requested before the dependent loads
		const size_t si = (size_t)chunk * CHUNK + lane;
		asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(o.x), "=r"(o.y) : "l"(net_out + si));
		asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(dt_in) : "l"(&samples[si].w));
		asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(count) : "l"(sample_counts + chunk));
	}
	if (chunk < max_rays && chunk < *n_alive_ptr) { // warp-uniform
		const uint32_t r = alive_list[chunk];
		const float T_in = rays.rgba[r].w;
		float f = 1.0f;
		if (lane < count) {
			const __half2 bs = *reinterpret_cast<const __half2*>(&o.y);
			const float sigma = act_density(__high2float(bs), a.density_act);
			const float dt_range = __fsub_rn(__fmul_rn(HN_MIN_STEP, 128.0f), HN_MIN_STEP);
			const float dt = __fadd_rn(__fmul_rn(__fdiv_rn(__fsub_rn(dt_in, HN_MIN_STEP), dt_range), dt_range), HN_MIN_STEP);
			f = __expf(-sigma * dt);
		}
		float P = f;
		#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const float up = __shfl_up_sync(0xffffffffu, P, d);
			if (lane >= (uint32_t)d) P *= up;
		}
		const float T = T_in * P;
		const uint32_t term_mask = __ballot_sync(0xffffffffu, lane < count && T < a.min_transmittance);
		const uint32_t last = term_mask ? (uint32_t)(__ffs(term_mask) - 1) : (count ? count - 1 : 0u);
		const float T_out = __shfl_sync(0xffffffffu, T, last);
		if (lane == 0) {
			if (count) { rays.rgba[r].w = T_out; rays.n_steps[r] += last + 1; atomicAdd(&block_total, last + 1); }
			survive[chunk] = (!term_mask && count == CHUNK) ? 1u : 0u;
		}
	}
	__syncthreads();
	if (threadIdx.x == 0 && block_total) atomicAdd((unsigned long long*)&counters[CNT_COMPOSITED_LO], (unsigned long long)block_total);
}
void launch_transmittance(const FrameArgs& a, RaySoA rays, const uint32_t* alive_list, const uint32_t* n_alive_ptr, const float4* samples,
	const uint32_t* sample_counts, const uint2* net_out, uint32_t* survive, uint32_t* counters, uint32_t max_rays, cudaStream_t s) {
	if (!max_rays) return;
	transmittance_kernel<<<(max_rays + 15) / 16, 512, 0, s>>>(a, rays, alive_list, n_alive_ptr, samples, sample_counts, net_out, survive, counters, max_rays);
}

// next pass's live-ray list from the survive flags of the fused network+composite kernel (one atomic per warp)
__global__ void __launch_bounds__(256) compact_kernel(const uint32_t* __restrict__ alive_in, const uint32_t* __restrict__ n_alive_ptr,
	const uint32_t* __restrict__ survive, uint32_t* __restrict__ alive_out, uint32_t* __restrict__ n_alive_out) {
	const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t lane = threadIdx.x & 31u;
	const bool keep = p < *n_alive_ptr && survive[p] != 0u;
	const uint32_t mask = __ballot_sync(0xffffffffu, keep);
	if (!mask) return;
	uint32_t base = 0;
	const int leader = __ffs(mask) - 1;
	if ((int)lane == leader) base = atomicAdd(n_alive_out, __popc(mask));
	base = __shfl_sync(0xffffffffu, base, leader);
	if (keep) alive_out[base + __popc(mask & ((1u << lane) - 1u))] = alive_in[p];
}
void launch_compact(const uint32_t* alive_in, const uint32_t* n_alive_ptr, const uint32_t* survive, uint32_t* alive_out, uint32_t* n_alive_out, uint32_t max_rays, cudaStream_t s) {
	if (!max_rays) return;
	compact_kernel<<<(max_rays + 255) / 256, 256, 0, s>>>(alive_in, n_alive_ptr, survive, alive_out, n_alive_out);
}

void launch_composite(const FrameArgs& a, RaySoA rays, const uint32_t* alive_in, const uint32_t* n_alive_ptr, uint32_t* alive_out,
	uint32_t* counters, const float4* samples, const uint32_t* sample_counts, const uint2* net_out,
	float4* frame, float* depth_buffer, uint32_t max_rays, cudaStream_t s) {
	if (!max_rays) return;
	// n_alive_ptr points at one of counters[CNT_ALIVE_A/B]; the other one receives the survivors
	uint32_t* n_out = (n_alive_ptr == counters + CNT_ALIVE_A) ? counters + CNT_ALIVE_B : counters + CNT_ALIVE_A;
	uint32_t warps = (max_rays + 31) / 32;
	composite_kernel<<<(warps + 7) / 8, 256, 0, s>>>(a, rays, alive_in, n_alive_ptr, alive_out, n_out, counters, samples, sample_counts, net_out, frame, depth_buffer);
}
void launch_accumulate(const float4* frame, float4* accum, uint32_t n, uint32_t spp_index, cudaStream_t s) {
	accumulate_kernel<<<(n + 255) / 256, 256, 0, s>>>(frame, accum, n, (float)spp_index);
}
void launch_tonemap(const float4* accum, float4* out, uint32_t n, int to_srgb, float exposure, const float* bg, int curve, int clamp_output, int unmultiply_alpha, cudaStream_t s) {
	tonemap_kernel<<<(n + 255) / 256, 256, 0, s>>>(accum, out, n, to_srgb, exp2f(exposure), make_float4(bg[0], bg[1], bg[2], bg[3]), curve, clamp_output, unmultiply_alpha);
}
void launch_pack_tiles(int w, int h, int rank, int world, int tw, int th, const float4* frame, const float* depth, float* packed, bool unpack, cudaStream_t s) {
	uint32_t tiles = ((w + tw - 1) / tw) * ((h + th - 1) / th);
	uint32_t mine = (tiles + world - 1 - rank) / world;
	if (!mine) return;
	pack_tiles_kernel<<<mine, 256, 0, s>>>(w, h, rank, world, tw, th, (float4*)frame, (float*)depth, packed, unpack ? 1 : 0);
}

} // namespace hn
