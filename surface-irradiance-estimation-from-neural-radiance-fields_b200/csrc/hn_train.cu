// a-21 NeRF training step for sm_100a: ray sampling from an in-memory dataset, occupancy-DDA sample generation, forward
// pass with stored fp16 activations, per-ray two-pass compositing loss, backward through both MLPs and the hash grid, Adam.
//   Testbed::train_nerf / train_nerf_step       src/testbed_nerf.cu:2929-3075, 3183-3430
//   generate_training_samples_nerf              src/testbed_nerf.cu:737-890       (train_march_kernel, hn_march.cu)
//   compute_loss_kernel_train_nerf              src/testbed_nerf.cu:893-1212      (train_loss_kernel)
//   kernel_grid_backward                        tcnn grid.h:214-320               (grid_backward_kernel: one 16-byte vector atomic per corner, fp32)
//   MLP forward/backward + weight gradients     tcnn fully_fused_mlp.cu:150-259,775-829 (plain GEMMs: cuBLAS, fp16 in / fp32 accumulate;
//                                               the reference runs CUTLASS split-K GEMMs for the same weight gradients)
//   Trainer::training_step, adam_step, Ema, ExponentialDecay   tcnn trainer.h:163-190, optimizers/adam.h:47-118, ema.h, exponential_decay.h
// Layout: activations [samples][width] fp16 row-major, one buffer per layer; gradients fp32 [n_params] in the layout of the
// parameter blob [density MLP | rgb MLP | hash grid]. Round-1 scope: the MLP GEMMs are library calls, everything else is
// hand-written; a fused tcgen05 forward/backward is the next step.
#include "hn_context.h"
#include "hn_train_tc.h"
#include "hn_network.cuh"
#include "hn_shade.cuh"
#include "hn_device_math.cuh"
#include <cublas_v2.h>
#include <algorithm>
#include <cmath>

namespace hn {

#define HN_CUBLAS(expr) do { cublasStatus_t s__ = (expr); if (s__ != CUBLAS_STATUS_SUCCESS) throw std::runtime_error(std::string(#expr) + " failed: cuBLAS status " + std::to_string((int)s__)); } while (0)

// ---- pcg32 (O'Neill; tcnn/dependencies/pcg32): the generator the reference samples training rays with ----
struct Pcg32 {
	uint64_t state, inc;
	__host__ __device__ Pcg32(uint64_t initstate, uint64_t initseq = 1u) {
		state = 0u; inc = (initseq << 1u) | 1u;
		next_uint(); state += initstate; next_uint();
	}
	__host__ __device__ uint32_t next_uint() {
		uint64_t old = state;
		state = old * 0x5851f42d4c957f2dULL + inc;
		uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
		uint32_t rot = (uint32_t)(old >> 59u);
		return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
	}
	__host__ __device__ float next_float() {
		union { uint32_t u; float f; } x;
		x.u = (next_uint() >> 9) | 0x3f800000u;
		return x.f - 1.0f;
	}
	__host__ __device__ void advance(uint64_t delta) { // skip-ahead in O(log delta)
		uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
		while (delta > 0) {
			if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
			cur_plus = (cur_mult + 1) * cur_plus;
			cur_mult *= cur_mult;
			delta /= 2;
		}
		state = acc_mult * state + acc_plus;
	}
};

struct TrainDataDev {
	const float4* images; const float* cams;
	uint32_t n_images; int w, h;
	const ImageMeta* meta; // [n_images] focal length, principal point and lens of every image (TrainingImageMetadata)
};

// ---- error map (src/testbed_nerf.cu:1080-1106, 1776-1832, 2986-3075; nerf_device.cuh:536-633) ----
// Per training image a coarse map of the mean loss of the rays that landed there (bilinear deposit), turned every
// n_steps_between_error_map_updates steps into CDFs (x given y, y, image) that can steer the choice of image and pixel.
// depth supervision (src/testbed_nerf.cu:1063-1067,1149-1154): per-ray target distance (<= 0 or nullptr: none), lambda, ELossType of the depth term;
// ray origins and the training box un-warp the sample positions for distance(pos, ray_o)
struct DepthSup {
	const float* target; const float* ray_o;
	float lambda; int loss_type;
	float lo[3], diag[3];
};
struct ErrorMapDev {
	float* data;                  // [n_images][res_y][res_x] accumulated loss, nullptr = off
	int res[2];
	const float* cdf_x_cond_y;    // nullptr: pixels uniformly (nerf.training.sample_focal_plane_proportional_to_error off / no CDF yet)
	const float* cdf_y;
	const float* cdf_img;         // nullptr: image by ray index (sample_image_proportional_to_error off / no CDF yet)
	int cdf_res[2];
};
// common.h:242-264
__host__ __device__ inline uint32_t cdf_binary_search(float val, const float* __restrict__ data, uint32_t length) {
	if (length == 0) return 0;
	uint32_t first = 0, count = length;
	while (count > 0) {
		const uint32_t step = count / 2, it = first + step;
		if (data[it] < val) { first = it + 1; count -= step + 1; } else count = step;
	}
	return first < length - 1 ? first : length - 1;
}
// image_idx (nerf_device.cuh:617-639)
__host__ __device__ inline uint32_t train_image_idx(uint32_t base_idx, uint32_t n_rays, uint32_t n_images, const float* __restrict__ cdf, float& pdf) {
	pdf = 1.0f;
	if (cdf) {
		const float sample = hn_ld_random_val(base_idx, 0xdeadbeefu);
		const uint32_t img = cdf_binary_search(sample, cdf, n_images);
		const float prev = img > 0 ? cdf[img - 1] : 0.0f;
		pdf = (cdf[img] - prev) * (float)n_images;
		return img;
	}
	return ((base_idx * n_images) / n_rays) % n_images; // 32-bit arithmetic as in the reference (wraps for base_idx * n_images >= 2^32)
}
// sample_cdf_2d (nerf_device.cuh:538-568): half of the samples stay uniform (UNIFORM_SAMPLING_FRACTION)
__host__ __device__ inline void train_sample_cdf_2d(float& sx, float& sy, uint32_t img, int res_x, int res_y, const float* __restrict__ cdf_x_cond_y, const float* __restrict__ cdf_y, float& pdf) {
	if (sx < 0.5f) { sx /= 0.5f; return; }
	sx = (sx - 0.5f) / (1.0f - 0.5f);
	cdf_y += (size_t)img * res_y;
	const uint32_t y = cdf_binary_search(sy, cdf_y, (uint32_t)res_y);
	float prev = y > 0 ? cdf_y[y - 1] : 0.0f;
	const float pmf_y = cdf_y[y] - prev;
	sy = (sy - prev) / pmf_y;
	cdf_x_cond_y += (size_t)img * res_y * res_x + (size_t)y * res_x;
	const uint32_t x = cdf_binary_search(sx, cdf_x_cond_y, (uint32_t)res_x);
	prev = x > 0 ? cdf_x_cond_y[x - 1] : 0.0f;
	const float pmf_x = cdf_x_cond_y[x] - prev;
	sx = (sx - prev) / pmf_x;
	pdf = pmf_x * pmf_y * (float)(res_x * res_y);
	sx = ((float)x + sx) / (float)res_x; sy = ((float)y + sy) / (float)res_y;
}
// compute_loss_kernel_train_nerf :1080-1106 (the clamp of the cell index against the IMAGE resolution - 2 is the reference's)
__device__ __forceinline__ void error_map_deposit(float* __restrict__ map, int res_x, int res_y, int img_w, int img_h, uint32_t img, float u, float v, float val) {
	const float px = fminf(fmaxf(u * (float)res_x - 0.5f, 0.0f), (float)res_x - (1.0f + 1e-4f));
	const float py = fminf(fmaxf(v * (float)res_y - 0.5f, 0.0f), (float)res_y - (1.0f + 1e-4f));
	const int ix = (int)px, iy = (int)py;
	const float wx = px - (float)ix, wy = py - (float)iy;
	const int x = min(max(ix, 0), img_w - 2), y = min(max(iy, 0), img_h - 2);
	float* m = map + (size_t)img * res_x * res_y;
	atomicAdd(m + y * res_x + x, (1 - wx) * (1 - wy) * val);
	atomicAdd(m + y * res_x + x + 1, wx * (1 - wy) * val);
	atomicAdd(m + (y + 1) * res_x + x, (1 - wx) * wy * val);
	atomicAdd(m + (y + 1) * res_x + x + 1, wx * wy * val);
}
// construct_cdf_2d / construct_cdf_1d (:1776-1832): one thread per (row, image) / per image, sequential sums (same order -> same bits)
__global__ void construct_cdf_2d_kernel(uint32_t n_images, uint32_t height, uint32_t width, const float* __restrict__ data, float* __restrict__ cdf_x_cond_y, float* __restrict__ cdf_y) {
	const uint32_t y = threadIdx.x + blockIdx.x * blockDim.x, img = threadIdx.y + blockIdx.y * blockDim.y;
	if (y >= height || img >= n_images) return;
	const size_t off = ((size_t)img * height + y) * width;
	data += off; cdf_x_cond_y += off;
	float cum = 0;
	for (uint32_t x = 0; x < width; ++x) { cum += data[x] + 1e-10f; cdf_x_cond_y[x] = cum; }
	cdf_y[img * height + y] = cum;
	const float norm = __frcp_rn(cum);
	for (uint32_t x = 0; x < width; ++x) cdf_x_cond_y[x] = (1.0f - 0.01f) * cdf_x_cond_y[x] * norm + 0.01f * (float)(x + 1) / (float)width;
}
__global__ void construct_cdf_1d_kernel(uint32_t n_images, uint32_t height, float* __restrict__ cdf_y, float* __restrict__ cdf_img) {
	const uint32_t img = threadIdx.x + blockIdx.x * blockDim.x;
	if (img >= n_images) return;
	cdf_y += (size_t)img * height;
	float cum = 0;
	for (uint32_t y = 0; y < height; ++y) { cum += cdf_y[y]; cdf_y[y] = cum; }
	cdf_img[img] = cum;
	const float norm = __frcp_rn(cum);
	for (uint32_t y = 0; y < height; ++y) cdf_y[y] = (1.0f - 0.01f) * cdf_y[y] * norm + 0.01f * (float)(y + 1) / (float)height;
}

// Training rays from the dataset (train_nerf_step ray selection, src/testbed_nerf.cu:789-846 + :1007-1037): ray i of the
// global stream belongs to image (i * n_images) / n_rays_total, pixel uniformly random (snapped to the pixel centre),
// random background colour, target = linear_to_srgb(rgb + (1 - a) * bg) (Linear image colour space, sRGB training space).
// Start t = AABB entry advanced by a random fraction of a step (:817-821).
__global__ void train_raygen_kernel(TrainDataDev D, FrameArgs a, uint32_t n_rays, uint32_t ray_offset, uint32_t n_rays_total, uint32_t seed, uint32_t step,
	int random_bg, int srgb_space, float bg_r, float bg_g, float bg_b, float* __restrict__ ray_o, float* __restrict__ ray_d, float* __restrict__ ray_t,
	float* __restrict__ ray_target, float* __restrict__ ray_bg, ErrorMapDev E, uint32_t* __restrict__ ray_img, float* __restrict__ ray_uv, float* __restrict__ ray_pdf,
	const float* __restrict__ depth_images, float* __restrict__ ray_target_depth) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_rays) return;
	const uint32_t gi = i + ray_offset;
	Pcg32 rng((uint64_t)seed + ((uint64_t)step << 32));
	rng.advance((uint64_t)gi * 16u); // N_MAX_RANDOM_SAMPLES_PER_RAY
	// image and pixel: by ray index / uniform, or from the error-map CDFs (image_idx, nerf_random_image_pos_training)
	float img_pdf = 1.0f, uv_pdf = 1.0f;
	const uint32_t img = train_image_idx(gi, n_rays_total, D.n_images, E.cdf_img, img_pdf);
	float u = rng.next_float(), v = rng.next_float();
	if (E.cdf_x_cond_y) train_sample_cdf_2d(u, v, img, E.cdf_res[0], E.cdf_res[1], E.cdf_x_cond_y, E.cdf_y, uv_pdf);
	int px = min(max((int)(u * (float)D.w), 0), D.w - 1), py = min(max((int)(v * (float)D.h), 0), D.h - 1);
	const float uu = ((float)px + 0.5f) / (float)D.w, vv = ((float)py + 0.5f) / (float)D.h;
	const float jitter = rng.next_float();
	float bg[3] = {bg_r, bg_g, bg_b};
	if (random_bg) { bg[0] = rng.next_float(); bg[1] = rng.next_float(); bg[2] = rng.next_float(); }
	const float4 tex = __ldg(D.images + ((size_t)img * D.h + py) * D.w + px);
	const float* c = D.cams + 12 * img;
	// uv_to_ray with the image's own focal length, principal point and lens (generate_training_samples_nerf, src/testbed_nerf.cu:806-808)
	const ImageMeta m = D.meta[img];
	V3 dc;
	const bool valid_ray = lens_direction(m.lens_mode, m.lens_params, uu, vv, m.principal[0], m.principal[1], D.w, D.h, m.focal[0], m.focal[1], dc);
	if (!valid_ray) dc = mk(0.f, 0.f, 1.f);
	float d0 = c[0] * dc.x + c[3] * dc.y + c[6] * dc.z, d1 = c[1] * dc.x + c[4] * dc.y + c[7] * dc.z, d2 = c[2] * dc.x + c[5] * dc.y + c[8] * dc.z;
	const float len2 = d0 * d0 + d1 * d1 + d2 * d2;
	const float inv = rsqrtf(len2);
	// target distance of the depth term: |un-normalised ray direction| x the image's depth at the pixel (:1065), -1 = none
	if (ray_target_depth) ray_target_depth[i] = depth_images ? sqrtf(len2) * __ldg(depth_images + ((size_t)img * D.h + py) * D.w + px) : -1.0f;
	d0 *= inv; d1 *= inv; d2 *= inv;
	ray_o[3 * i] = c[9]; ray_o[3 * i + 1] = c[10]; ray_o[3 * i + 2] = c[11];
	ray_d[3 * i] = d0; ray_d[3 * i + 1] = d1; ray_d[3 * i + 2] = d2;
	ray_img[i] = img; ray_uv[2 * i] = uu; ray_uv[2 * i + 1] = vv; ray_pdf[i] = img_pdf * uv_pdf;
	// slab test against the render box
	float tmin = -1e30f, tmax = 1e30f;
	const float oo[3] = {c[9], c[10], c[11]}, dd[3] = {d0, d1, d2};
	#pragma unroll
	for (int k = 0; k < 3; ++k) {
		const float t0 = (a.render_aabb[k] - oo[k]) / dd[k], t1 = (a.render_aabb[3 + k] - oo[k]) / dd[k];
		tmin = fmaxf(tmin, fminf(t0, t1)); tmax = fminf(tmax, fmaxf(t0, t1));
	}
	// startt = advance_n_steps(tmin, cone_angle, random) (:833): a random fraction of the LOCAL step, which grows with t when the cone angle is > 0 (aabb_scale > 1)
	ray_t[i] = (tmin <= tmax && valid_ray) ? advance_n_steps(fmaxf(tmin, 0.0f), a.cone_angle, jitter) : 3.0e4f; // beyond MAX_DEPTH: the march finds nothing (:812-815 drops invalid rays)
	#pragma unroll
	for (int k = 0; k < 3; ++k) {
		const float texc = k == 0 ? tex.x : (k == 1 ? tex.y : tex.z);
		// :1023-1037: linear images are composited over the background, then taken to sRGB; sRGB (byte) images are blended in sRGB
		ray_target[3 * i + k] = srgb_space ? linear_to_srgb(texc) * tex.w + (1.0f - tex.w) * linear_to_srgb(bg[k]) : linear_to_srgb(texc * tex.w + (1.0f - tex.w) * bg[k]);
		ray_bg[3 * i + k] = linear_to_srgb(bg[k]);
	}
}

// hash-grid features (tcnn fp16 arithmetic, as inference) -> enc [n][32]; SH of the ray direction -> rgb_in[:, 16:32]
template <int F>
__global__ void train_encode_kernel(GridLevels g, const __half* __restrict__ table, uint32_t n, const float4* __restrict__ coords,
	const uint32_t* __restrict__ sample_ray, const float* __restrict__ ray_d, __half* __restrict__ enc, __half* __restrict__ rgb_in) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float4 c = __ldg(coords + i);
	uint32_t e[16];
	encode_sample<F>(g, table, c.x, c.y, c.z, e);
	uint4* dst = reinterpret_cast<uint4*>(enc + (size_t)i * 32);
	#pragma unroll
	for (int k = 0; k < 4; ++k) dst[k] = make_uint4(e[4 * k], e[4 * k + 1], e[4 * k + 2], e[4 * k + 3]);
	const uint32_t r = sample_ray[i];
	float dx = ray_d[3 * r], dy = ray_d[3 * r + 1], dz = ray_d[3 * r + 2];
	dx = __fmul_rn(__fadd_rn(dx, 1.0f), 0.5f) * 2.f - 1.f; dy = __fmul_rn(__fadd_rn(dy, 1.0f), 0.5f) * 2.f - 1.f; dz = __fmul_rn(__fadd_rn(dz, 1.0f), 0.5f) * 2.f - 1.f;
	float sh[16];
	sh4(dx, dy, dz, sh);
	uint4* d2 = reinterpret_cast<uint4*>(rgb_in + (size_t)i * 32 + 16);
	d2[0] = make_uint4(pack_h2(sh[0], sh[1]), pack_h2(sh[2], sh[3]), pack_h2(sh[4], sh[5]), pack_h2(sh[6], sh[7]));
	d2[1] = make_uint4(pack_h2(sh[8], sh[9]), pack_h2(sh[10], sh[11]), pack_h2(sh[12], sh[13]), pack_h2(sh[14], sh[15]));
}

__global__ void relu_kernel(__half2* __restrict__ x, size_t n2) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n2) x[i] = __hmax2(x[i], __float2half2_rn(0.f));
}
// gradient of a hidden layer: zero where the forward activation was clamped
__global__ void relu_backward_kernel(__half2* __restrict__ dx, const __half2* __restrict__ act, size_t n2) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n2) return;
	const __half2 a = act[i]; __half2 d = dx[i];
	if (!(__low2float(a) > 0.f)) d = __halves2half2(__float2half(0.f), __high2half(d));
	if (!(__high2float(a) > 0.f)) d = __halves2half2(__low2half(d), __float2half(0.f));
	dx[i] = d;
}
// density-MLP output (16) -> first half of the colour MLP input
__global__ void copy_density_out_kernel(const __half* __restrict__ d_out, __half* __restrict__ rgb_in, uint32_t n) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint4* s = reinterpret_cast<const uint4*>(d_out + (size_t)i * 16);
	uint4* d = reinterpret_cast<uint4*>(rgb_in + (size_t)i * 32);
	d[0] = s[0]; d[1] = s[1];
}
// dL/d(density out) = dL/d(rgb_in)[:, :16] + dL/dsigma on channel 0
__global__ void combine_density_grad_kernel(const __half* __restrict__ d_rgb_in, const float* __restrict__ d_sigma, __half* __restrict__ d_dout, uint32_t n) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const __half* s = d_rgb_in + (size_t)i * 32;
	__half* d = d_dout + (size_t)i * 16;
	d[0] = __float2half(__half2float(s[0]) + d_sigma[i]);
	#pragma unroll
	for (int k = 1; k < 16; ++k) d[k] = s[k];
}

// loss_and_gradient (nerf_device.cuh:74-142,640-655) per channel. ids: 0 L2, 1 Huber(0.1)/5, 2 RelativeL2, 3 L1, 4 Mape, 5 Smape, 6 LogL1
__device__ __forceinline__ float loss_grad(int loss_type, float target, float pred, float& grad) {
	const float diff = pred - target;
	switch (loss_type) {
		case 1: {
			const float alpha = 0.1f, ad = fabsf(diff);
			const float l = ad > alpha ? (ad - 0.5f * alpha) : 0.5f / alpha * diff * diff;
			grad = (ad > alpha ? (diff > 0 ? 1.0f : -1.0f) : (diff / alpha)) / 5.0f;
			return l / 5.0f;
		}
		case 2: { const float denom = pred * pred + 1e-2f; grad = 2.0f * diff / denom; return diff * diff / denom; }
		case 3: grad = copysignf(1.0f, diff); return fabsf(diff);
		case 4: { const float denom = fabsf(pred) + 1e-2f; grad = copysignf(1.0f / denom, diff); return fabsf(diff) / denom; }
		case 5: { const float denom = 0.5f * (fabsf(pred) + fabsf(target)) + 1e-2f; grad = copysignf(1.0f / denom, diff); return fabsf(diff) / denom; }
		case 6: { const float div = fabsf(diff) + 1.0f; grad = copysignf(1.0f / div, diff); return logf(div); }
		default: grad = 2.0f * diff; return diff * diff;
	}
}
__device__ __forceinline__ float rgb_derivative(float v, int act) {
	switch (act) {
		case 1: return v > 0.0f ? 1.0f : 0.0f;
		case 2: { const float d = 1.0f / (1.0f + __expf(-v)); return d * (1 - d); }
		case 3: return __expf(fminf(fmaxf(v, -10.0f), 10.0f));
		default: return 1.0f;
	}
}
__device__ __forceinline__ float density_derivative(float v, int act) {
	switch (act) {
		case 1: return v > 0.0f ? 1.0f : 0.0f;
		case 2: { const float d = 1.0f / (1.0f + __expf(-v)); return d * (1 - d); }
		case 3: return __expf(fminf(fmaxf(v, -15.0f), 15.0f));
		default: return 1.0f;
	}
}

__device__ __forceinline__ float warp_incl_sum(float v, uint32_t lane) {
	#pragma unroll
	for (int s = 1; s < 32; s <<= 1) { const float up = __shfl_up_sync(0xffffffffu, v, s); if (lane >= (uint32_t)s) v += up; }
	return v;
}
__device__ __forceinline__ float warp_incl_prod(float v, uint32_t lane) {
	#pragma unroll
	for (int s = 1; s < 32; s <<= 1) { const float up = __shfl_up_sync(0xffffffffu, v, s); if (lane >= (uint32_t)s) v *= up; }
	return v;
}

// compute_loss_kernel_train_nerf (src/testbed_nerf.cu:893-1212), ONE WARP per ray (the reference: one thread per ray; at
// ~5 000 rays per step that leaves a B200 with 40 blocks): lane j takes sample 32c + j of chunk c. Pass 1 composites the ray
// colour (transmittance = warp product scan carried across chunks; the ray stops at the first sample whose incoming T is
// below 1e-4), pass 2 recomputes the weights and emits dL/d(network output) per sample from inclusive prefix sums of w*c
// (the "suffix" of the reference is ray colour minus that prefix). out/density_out: [samples][16] fp16. Writes
// d_out [samples][16] fp16 (rows 0..2) and d_sigma [samples] fp32, both scaled by loss_scale / n_rays_total; samples past the
// stop get zero gradients (the reference compacts them away). loss_acc += mean(loss_rgb) / n_rays_total.
__global__ void __launch_bounds__(128) train_loss_kernel(uint32_t n_rays, uint32_t n_rays_total, const uint32_t* __restrict__ numsteps_in, const uint32_t* __restrict__ base_in,
	const float4* __restrict__ coords, const __half* __restrict__ out, const __half* __restrict__ density_out, const float* __restrict__ ray_target,
	const float* __restrict__ ray_bg, int loss_type, float loss_scale, int rgb_act, int density_act, float l1_reg_density, __half* __restrict__ d_out, float* __restrict__ d_sigma,
	uint32_t* __restrict__ used_out, float* __restrict__ loss_acc, ErrorMapDev E, const uint32_t* __restrict__ ray_img, const float* __restrict__ ray_uv,
	const float* __restrict__ ray_pdf, int img_w, int img_h, DepthSup Z) {
	const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31u;
	if (i >= n_rays) return; // warp-uniform
	const uint32_t numsteps = numsteps_in[i], base = base_in[i];
	const float dt_range = HN_MIN_STEP * 128.0f - HN_MIN_STEP;
	const uint32_t n_chunks = (numsteps + 31u) / 32u;
	// ---- pass 1: ray colour
	float T = 1.f, rr = 0.f, rg = 0.f, rb = 0.f, depth_ray = 0.f;
	const bool use_depth = Z.target != nullptr && Z.lambda > 0.0f && Z.target[i] > 0.0f; // warp-uniform
	float ox = 0.f, oy = 0.f, oz = 0.f;
	if (use_depth) { ox = Z.ray_o[3 * i]; oy = Z.ray_o[3 * i + 1]; oz = Z.ray_o[3 * i + 2]; }
	uint32_t compacted = numsteps;
	for (uint32_t c = 0; c < n_chunks; ++c) {
		const uint32_t j = c * 32u + lane;
		const bool has = j < numsteps;
		float alpha = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, dist = 0.f;
		if (has) {
			const size_t s = base + j;
			const uint2 o = *reinterpret_cast<const uint2*>(out + s * 16);
			const __half2 h01 = *reinterpret_cast<const __half2*>(&o.x), h2x = *reinterpret_cast<const __half2*>(&o.y);
			const float4 cw = coords[s];
			if (use_depth) { const float qx = Z.lo[0] + cw.x * Z.diag[0] - ox, qy = Z.lo[1] + cw.y * Z.diag[1] - oy, qz = Z.lo[2] + cw.z * Z.diag[2] - oz; dist = sqrtf(qx * qx + qy * qy + qz * qz); }
			const float dt = (cw.w - HN_MIN_STEP) / dt_range * dt_range + HN_MIN_STEP;
			alpha = 1.f - __expf(-act_density(__half2float(density_out[s * 16]), density_act) * dt);
			cr = act_rgb(__low2float(h01), rgb_act); cg = act_rgb(__high2float(h01), rgb_act); cb = act_rgb(__low2float(h2x), rgb_act);
		}
		const float P = warp_incl_prod(1.f - alpha, lane);
		float Pex = __shfl_up_sync(0xffffffffu, P, 1);
		if (lane == 0) Pex = 1.f;
		const float T_before = T * Pex;
		const uint32_t stop = __ballot_sync(0xffffffffu, has && T_before < 1e-4f);
		const uint32_t last = stop ? (uint32_t)(__ffs(stop) - 1) : 32u; // samples [0, last) of this chunk are used
		const float w = (has && lane < last) ? alpha * T_before : 0.f;
		float sr = w * cr, sg = w * cg, sb = w * cb;
		#pragma unroll
		for (int dd = 16; dd > 0; dd >>= 1) { sr += __shfl_xor_sync(0xffffffffu, sr, dd); sg += __shfl_xor_sync(0xffffffffu, sg, dd); sb += __shfl_xor_sync(0xffffffffu, sb, dd); }
		rr += sr; rg += sg; rb += sb;
		if (use_depth) {
			float sd = w * dist;
			#pragma unroll
			for (int dd = 16; dd > 0; dd >>= 1) sd += __shfl_xor_sync(0xffffffffu, sd, dd);
			depth_ray += sd;
		}
		if (stop) { compacted = c * 32u + last; T = __shfl_sync(0xffffffffu, T_before, last); break; }
		T *= __shfl_sync(0xffffffffu, P, 31);
	}
	if (compacted == numsteps) { rr += T * ray_bg[3 * i]; rg += T * ray_bg[3 * i + 1]; rb += T * ray_bg[3 * i + 2]; }
	if (used_out && lane == 0) used_out[i] = compacted;
	float gr, gg, gb;
	const float l = loss_grad(loss_type, ray_target[3 * i], rr, gr) + loss_grad(loss_type, ray_target[3 * i + 1], rg, gg) + loss_grad(loss_type, ray_target[3 * i + 2], rb, gb);
	if (lane == 0 && numsteps) {
		// :1061-1106: the REPORTED loss is divided by the sampling pdf (the gradient is not: importance sampling is meant to re-weight the
		// objective); the same mean loss goes into the error map at the ray's image position
		const float mean_loss = l / 3.0f / (ray_pdf ? ray_pdf[i] : 1.0f);
		atomicAdd(loss_acc, mean_loss / (float)n_rays_total);
		if (E.data && ray_img) error_map_deposit(E.data, E.res[0], E.res[1], img_w, img_h, ray_img[i], ray_uv[2 * i], ray_uv[2 * i + 1], mean_loss);
	}
	float depth_loss_gradient = 0.0f;
	if (use_depth) { float g; loss_grad(Z.loss_type, Z.target[i], depth_ray, g); depth_loss_gradient = Z.lambda * g; }
	const float ls = loss_scale / (float)n_rays_total;
	const float l2_reg = rgb_act == 3 ? 1e-4f : 0.0f;
	// ---- pass 2: gradients
	float Tc = 1.f, pr = 0.f, pg = 0.f, pb = 0.f, pd = 0.f; // carried transmittance, colour prefix and depth prefix
	for (uint32_t c = 0; c < n_chunks; ++c) {
		const uint32_t j = c * 32u + lane;
		const bool has = j < numsteps;
		const size_t s = base + j;
		const bool used = j < compacted;
		float alpha = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, o0 = 0.f, o1 = 0.f, o2 = 0.f, sraw = 0.f, dt = 0.f, dist = 0.f;
		if (used) {
			const uint2 o = *reinterpret_cast<const uint2*>(out + s * 16);
			const __half2 h01 = *reinterpret_cast<const __half2*>(&o.x), h2x = *reinterpret_cast<const __half2*>(&o.y);
			o0 = __low2float(h01); o1 = __high2float(h01); o2 = __low2float(h2x); sraw = __half2float(density_out[s * 16]);
			const float4 cw = coords[s];
			if (use_depth) { const float qx = Z.lo[0] + cw.x * Z.diag[0] - ox, qy = Z.lo[1] + cw.y * Z.diag[1] - oy, qz = Z.lo[2] + cw.z * Z.diag[2] - oz; dist = sqrtf(qx * qx + qy * qy + qz * qz); }
			dt = (cw.w - HN_MIN_STEP) / dt_range * dt_range + HN_MIN_STEP;
			alpha = 1.f - __expf(-act_density(sraw, density_act) * dt);
			cr = act_rgb(o0, rgb_act); cg = act_rgb(o1, rgb_act); cb = act_rgb(o2, rgb_act);
		}
		const float P = warp_incl_prod(1.f - alpha, lane);
		float Pex = __shfl_up_sync(0xffffffffu, P, 1);
		if (lane == 0) Pex = 1.f;
		const float weight = alpha * (Tc * Pex);
		const float T_after = Tc * P;
		const float r2 = pr + warp_incl_sum(weight * cr, lane), g2 = pg + warp_incl_sum(weight * cg, lane), b2 = pb + warp_incl_sum(weight * cb, lane);
		const float d2 = use_depth ? pd + warp_incl_sum(weight * dist, lane) : 0.f;
		const float depth_supervision = use_depth ? depth_loss_gradient * (T_after * dist - (depth_ray - d2)) : 0.f;
		if (has) {
			uint4* dst = reinterpret_cast<uint4*>(d_out + s * 16);
			if (!used) { dst[0] = make_uint4(0, 0, 0, 0); dst[1] = make_uint4(0, 0, 0, 0); d_sigma[s] = 0.f; }
			else {
				const float sr = rr - r2, sg = rg - g2, sb = rb - b2; // suffix of the ray behind this sample
				const float d0 = ls * (weight * gr * rgb_derivative(o0, rgb_act) + fmaxf(0.0f, l2_reg * o0));
				const float d1 = ls * (weight * gg * rgb_derivative(o1, rgb_act) + fmaxf(0.0f, l2_reg * o1));
				const float d2 = ls * (weight * gb * rgb_derivative(o2, rgb_act) + fmaxf(0.0f, l2_reg * o2));
				dst[0] = make_uint4(pack_h2(d0, d1), pack_h2(d2, 0.f), 0, 0); dst[1] = make_uint4(0, 0, 0, 0);
				// + the reference's UNSCALED L1 density regulariser (src/testbed_nerf.cu:1113,1161-1164): pushes negative raw densities down
				// while the grid is nearly empty (mean density < NERF_MIN_OPTICAL_THICKNESS), which keeps such a grid from collapsing.
				// (its near-distance term is inert: near_distance is 0 on this path)
				d_sigma[s] = ls * density_derivative(sraw, density_act) * (dt * (gr * (T_after * cr - sr) + gg * (T_after * cg - sg) + gb * (T_after * cb - sb) + depth_supervision))
					+ (sraw < 0.0f ? -l1_reg_density : 0.0f);
			}
		}
		Tc *= __shfl_sync(0xffffffffu, P, 31);
		pr = __shfl_sync(0xffffffffu, r2, 31); pg = __shfl_sync(0xffffffffu, g2, 31); pb = __shfl_sync(0xffffffffu, b2, 31);
		if (use_depth) pd = __shfl_sync(0xffffffffu, d2, 31);
	}
}

// kernel_grid_backward (tcnn grid.h:214-320): thread = (sample, level); dL/dfeature scattered to the 8 corners with the
// trilinear weights. fp32 gradients; F = 4 uses ONE 16-byte vector atomic per corner (red.global.add.v4.f32, sm_90+).
template <int F>
__global__ void grid_backward_kernel(GridLevels g, uint32_t n, const float4* __restrict__ coords, const __half* __restrict__ d_enc, float* __restrict__ grid_grad) {
	constexpr uint32_t L = 32 / F;
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t i = tid / L, l = tid % L;
	if (i >= n) return;
	const __half* de = d_enc + (size_t)i * 32 + l * F;
	float gr[F];
	bool any = false;
	#pragma unroll
	for (int f = 0; f < F; ++f) { gr[f] = __half2float(de[f]); any |= gr[f] != 0.f; }
	if (!any) return;
	const float4 c = __ldg(coords + i);
	const float scale = g.scale[l];
	const uint32_t res = g.res[l], size = g.size[l], mask = g.mask[l];
	float fx = fmaf(scale, c.x, 0.5f), fy = fmaf(scale, c.y, 0.5f), fz = fmaf(scale, c.z, 0.5f);
	const float flx = floorf(fx), fly = floorf(fy), flz = floorf(fz);
	const uint32_t gx = (uint32_t)(int)flx, gy = (uint32_t)(int)fly, gz = (uint32_t)(int)flz;
	fx -= flx; fy -= fly; fz -= flz;
	float* lvl = grid_grad + (size_t)g.offset[l] * F;
	#pragma unroll
	for (int k = 0; k < 8; ++k) {
		const uint32_t x = gx + (k & 1), y = gy + ((k >> 1) & 1), z = gz + ((k >> 2) & 1);
		const uint32_t h = g.hashed[l] ? (x ^ (y * 2654435761u) ^ (z * 805459861u)) : (x + y * res + z * res * res);
		const uint32_t idx = mask ? (h & mask) : (h < size ? h : h % size);
		const float w = ((k & 1) ? fx : 1.f - fx) * ((k & 2) ? fy : 1.f - fy) * ((k & 4) ? fz : 1.f - fz);
		if (F == 4) atomicAdd(reinterpret_cast<float4*>(lvl + (size_t)idx * 4), make_float4(w * gr[0], w * gr[1], w * gr[2], w * gr[3]));
		else atomicAdd(reinterpret_cast<float2*>(lvl + (size_t)idx * 2), make_float2(w * gr[0], w * gr[1]));
	}
}

// adam_step (tcnn optimizers/adam.h:47-118) + Ema (ema.h) fused: hash-grid entries with a zero gradient are skipped and keep
// their own step counter; L2 regularisation only on the MLP ("matrix") weights.
__global__ void adam_kernel(uint32_t n, uint32_t n_matrix, float grad_scale, float inv_loss_scale, float lr, float beta1, float beta2, float eps, float l2_reg,
	int use_ema, float ema_decay, float ema_debias, float* __restrict__ master, __half* __restrict__ raw, __half* __restrict__ inference,
	const float* __restrict__ grads, float* __restrict__ m1, float* __restrict__ m2, float* __restrict__ ema, uint32_t* __restrict__ steps) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float w = master[i];
	float gradient = grads[i] * grad_scale * inv_loss_scale;
	const bool skip = i >= n_matrix && gradient == 0.f;
	if (!skip) {
		if (i < n_matrix) gradient += l2_reg * w;
		const float a = m1[i] = beta1 * m1[i] + (1 - beta1) * gradient;
		const float b = m2[i] = beta2 * m2[i] + (1 - beta2) * gradient * gradient;
		const uint32_t t = ++steps[i];
		const float l = lr * sqrtf(1 - powf(beta2, (float)t)) / (1 - powf(beta1, (float)t));
		w -= (l / (sqrtf(b) + eps)) * a;
		master[i] = w;
		raw[i] = __float2half(w);
	}
	if (use_ema) { // every parameter, every step (ema.h: tmp = decay * ema + (1 - decay) * w; weights_ema = tmp / (1 - decay^t))
		const float e = ema[i] = ema_decay * ema[i] + (1 - ema_decay) * w;
		inference[i] = __float2half(e / ema_debias);
	} else if (!skip) {
		inference[i] = __float2half(w);
	}
}
__global__ void half_to_float_kernel(const __half* __restrict__ in, float* __restrict__ out, size_t n) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = __half2float(in[i]);
}

// ---------------------------------------------------------------- occupancy grid from the live network
__device__ __forceinline__ uint32_t morton_invert(uint32_t x) {
	x = x & 0x49249249u;
	x = (x | (x >> 2)) & 0xc30c30c3u;
	x = (x | (x >> 4)) & 0x0f00f00fu;
	x = (x | (x >> 8)) & 0xff0000ffu;
	x = (x | (x >> 16)) & 0x0000ffffu;
	return x;
}
// generate_grid_samples_nerf_nonuniform (src/testbed_nerf.cu:185-217): sample i -> (cascade, cell, jittered position)
__global__ void grid_samples_kernel(uint32_t n, Pcg32 rng, uint32_t step, const float* __restrict__ grid_in, float4* __restrict__ out, uint32_t* __restrict__ indices,
	uint32_t n_cascades, float thresh) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	rng.advance((uint64_t)i * 4);
	const uint32_t level = (uint32_t)(rng.next_float() * n_cascades) % n_cascades;
	uint32_t idx = 0;
	for (uint32_t j = 0; j < 10; ++j) {
		idx = ((i + step * n) * 56924617u + j * 19349663u + 96925573u) % GRID_CELLS;
		idx += level * GRID_CELLS;
		if (grid_in[idx] > thresh) break;
	}
	const uint32_t pos_idx = idx % GRID_CELLS;
	const float x = (float)morton_invert(pos_idx), y = (float)morton_invert(pos_idx >> 1), z = (float)morton_invert(pos_idx >> 2);
	const float r0 = rng.next_float(), r1 = rng.next_float(), r2 = rng.next_float();
	const float sc = scalbnf(1.0f, (int)level);
	out[i] = make_float4(__fadd_rn(__fmul_rn(__fsub_rn(__fdiv_rn(__fadd_rn(x, r0), 128.0f), 0.5f), sc), 0.5f),
		__fadd_rn(__fmul_rn(__fsub_rn(__fdiv_rn(__fadd_rn(y, r1), 128.0f), 0.5f), sc), 0.5f),
		__fadd_rn(__fmul_rn(__fsub_rn(__fdiv_rn(__fadd_rn(z, r2), 128.0f), 0.5f), sc), 0.5f), HN_MIN_STEP);
	indices[i] = idx;
}
// splat_grid_samples_nerf_max_nearest_neighbor (:219-236): optical thickness of the smallest step, atomicMax on the bit pattern
__global__ void grid_splat_kernel(uint32_t n, const uint32_t* __restrict__ indices, const uint2* __restrict__ net_out, int density_act, float* __restrict__ tmp) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const __half2 bs = *reinterpret_cast<const __half2*>(&net_out[i].y);
	const float thickness = act_density(__high2float(bs), density_act) * HN_MIN_STEP;
	atomicMax(reinterpret_cast<uint32_t*>(tmp) + indices[i], __float_as_uint(thickness));
}
// ema_grid_samples_nerf (:258-282): maximum instead of a mean, so thin features switch a cell on at once
__global__ void grid_ema_kernel(uint32_t n, float decay, float* __restrict__ grid, const float* __restrict__ tmp) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float prev = grid[i];
	grid[i] = prev < 0.f ? prev : fmaxf(prev * decay, tmp[i]);
}
// mark_untrained_density_grid (:74-145), pinhole cameras: a cell no training view sees can never be trained -> -1
__global__ void mark_untrained_kernel(uint32_t n_elements, float* __restrict__ grid, TrainDataDev D) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	const uint32_t level = i / GRID_CELLS, pos_idx = i % GRID_CELLS;
	const float voxel = scalbnf(1.0f / 128.0f, (int)level), sc = scalbnf(1.0f, (int)level);
	const float px = ((float)morton_invert(pos_idx) / 128.0f - 0.5f) * sc + 0.5f, py = ((float)morton_invert(pos_idx >> 1) / 128.0f - 0.5f) * sc + 0.5f,
		pz = ((float)morton_invert(pos_idx >> 2) / 128.0f - 0.5f) * sc + 0.5f;
	bool seen = false;
	for (uint32_t j = 0; j < D.n_images && !seen; ++j) {
		const float* c = D.cams + 12 * j;
		const ImageMeta m = D.meta[j];
		// lenses without a forward mapping (f-theta) or that see everything (lat-long, equirectangular) count as seeing the cell (:115-120)
		if (m.lens_mode == 2 || m.lens_mode == 3 || m.lens_mode == 5) { seen = true; break; }
		for (uint32_t k = 0; k < 8 && !seen; ++k) {
			const float rx = px + ((k & 1) ? voxel : 0.f) - c[9], ry = py + ((k & 2) ? voxel : 0.f) - c[10], rz = pz + ((k & 4) ? voxel : 0.f) - c[11];
			const float len = sqrtf(rx * rx + ry * ry + rz * rz);
			const float cz = rx * c[6] + ry * c[7] + rz * c[8];
			if (cz / len < 1e-4f) continue;
			const float cx = rx * c[0] + ry * c[1] + rz * c[2], cy = rx * c[3] + ry * c[4] + rz * c[5];
			float ux = cx / cz, uy = cy / cz;
			if (m.lens_mode == 1 || m.lens_mode == 4) { // pos_to_uv adds the lens's distortion delta (common_device.cuh:521-531)
				float du, dv;
				if (m.lens_mode == 1) opencv_distortion_delta(m.lens_params, ux, uy, &du, &dv); else fisheye_distortion_delta(m.lens_params, ux, uy, &du, &dv);
				ux += du; uy += dv;
			}
			const float u = ux * m.focal[0] / (float)D.w + m.principal[0], v = uy * m.focal[1] / (float)D.h + m.principal[1];
			if (!(u > 0.f && v > 0.f && u < 1.f && v < 1.f)) continue;
			if (m.lens_mode == 1 || m.lens_mode == 4) {
				// pos_to_uv is not injective under lens distortion: the ray of that uv must point back at the corner (:131-137)
				V3 dc;
				lens_direction(m.lens_mode, m.lens_params, u, v, m.principal[0], m.principal[1], D.w, D.h, m.focal[0], m.focal[1], dc);
				float bx = c[0] * dc.x + c[3] * dc.y + c[6] * dc.z, by = c[1] * dc.x + c[4] * dc.y + c[7] * dc.z, bz = c[2] * dc.x + c[5] * dc.y + c[8] * dc.z;
				const float bl = sqrtf(bx * bx + by * by + bz * bz);
				bx = bx / bl - rx / len; by = by / bl - ry / len; bz = bz / bl - rz / len;
				if (!(sqrtf(bx * bx + by * by + bz * bz) < 1e-3f)) continue;
			}
			seen = true;
		}
	}
	grid[i] = seen ? 0.f : -1.f;
}

FrameArgs train_frame_args(hn_context* ctx); // hn_api.cpp: aabbs, occupancy bounds, cone angle, activations

namespace {

cublasHandle_t handle(hn_context* ctx) {
	if (!ctx->train.cublas) {
		cublasHandle_t h;
		HN_CUBLAS(cublasCreate(&h));
		HN_CUBLAS(cublasSetMathMode(h, CUBLAS_DEFAULT_MATH));
		ctx->train.cublas = h;
	}
	cublasHandle_t h = (cublasHandle_t)ctx->train.cublas;
	HN_CUBLAS(cublasSetStream(h, ctx->stream));
	return h;
}
// row-major views: Y[B][N] = X[B][K] W[N][K]^T ; dX[B][K] = dY[B][N] W[N][K] ; dW[N][K] (fp32) = dY[B][N]^T X[B][K]
void gemm_forward(cublasHandle_t h, uint32_t B, int N, int K, const __half* X, const __half* W, __half* Y) {
	const float one = 1.f, zero = 0.f;
	HN_CUBLAS(cublasGemmEx(h, CUBLAS_OP_T, CUBLAS_OP_N, N, (int)B, K, &one, W, CUDA_R_16F, K, X, CUDA_R_16F, K, &zero, Y, CUDA_R_16F, N, CUBLAS_COMPUTE_32F, CUBLAS_GEMM_DEFAULT));
}
void gemm_backward_input(cublasHandle_t h, uint32_t B, int N, int K, const __half* dY, const __half* W, __half* dX) {
	const float one = 1.f, zero = 0.f;
	HN_CUBLAS(cublasGemmEx(h, CUBLAS_OP_N, CUBLAS_OP_N, K, (int)B, N, &one, W, CUDA_R_16F, K, dY, CUDA_R_16F, N, &zero, dX, CUDA_R_16F, K, CUBLAS_COMPUTE_32F, CUBLAS_GEMM_DEFAULT));
}
void gemm_backward_weights(cublasHandle_t h, uint32_t B, int N, int K, const __half* dY, const __half* X, float* dW) {
	const float one = 1.f, zero = 0.f;
	HN_CUBLAS(cublasGemmEx(h, CUBLAS_OP_N, CUBLAS_OP_T, K, N, (int)B, &one, X, CUDA_R_16F, K, dY, CUDA_R_16F, N, &zero, dW, CUDA_R_32F, K, CUBLAS_COMPUTE_32F, CUBLAS_GEMM_DEFAULT));
}
inline uint32_t blocks(size_t n, uint32_t b = 256) { return (uint32_t)((n + b - 1) / b); }

struct Ev {
	cudaEvent_t a = nullptr, b = nullptr; bool on; cudaStream_t s;
	Ev(bool e, cudaStream_t st) : on(e), s(st) { if (on) { cudaEventCreate(&a); cudaEventCreate(&b); } }
	~Ev() { if (on) { cudaEventDestroy(a); cudaEventDestroy(b); } }
	void start() { if (on) cudaEventRecord(a, s); }
	void stop(float& acc) { if (on) { cudaEventRecord(b, s); cudaEventSynchronize(b); float ms = 0; cudaEventElapsedTime(&ms, a, b); acc = ms; } }
};

void ensure_batch(hn_context* ctx, uint32_t n_rays, uint32_t max_samples) {
	TrainState& t = ctx->train;
	t.ray_o.ensure((size_t)n_rays * 3); t.ray_d.ensure((size_t)n_rays * 3); t.ray_t.ensure(n_rays);
	t.ray_target.ensure((size_t)n_rays * 3); t.ray_bg.ensure((size_t)n_rays * 3);
	t.ray_numsteps.ensure(n_rays); t.ray_base.ensure(n_rays); t.counters.ensure(4); t.loss_acc.ensure(4);
	t.ray_img.ensure(n_rays); t.ray_uv.ensure((size_t)n_rays * 2); t.ray_pdf.ensure(n_rays); t.ray_target_depth.ensure(n_rays);
	t.coords.ensure(max_samples); t.sample_ray.ensure(max_samples);
	const uint32_t Hd = ctx->net.density_hidden, Hr = ctx->net.rgb_hidden;
	if (Hd + Hr + 4 > 12) throw std::runtime_error("training supports at most 8 hidden layers in total");
	for (uint32_t l = 0; l < Hd + Hr + 4; ++l) t.act[l].ensure((size_t)max_samples * 64);
	for (auto& d : t.dact) d.ensure((size_t)max_samples * 64);
}

// march (samples) -> forward -> loss -> backward into t.grads. Rays are already in t.ray_*.
void forward_backward(hn_context* ctx, uint32_t n_rays, uint32_t n_rays_total, uint32_t max_samples, int loss_type, float loss_scale, uint32_t* used_dev, const ErrorMapDev* emap = nullptr,
	const float* target_depth = nullptr, float depth_lambda = 0.0f, int depth_loss_type = 3) {
	TrainState& t = ctx->train;
	cudaStream_t s = ctx->stream;
	const NetCfg& net = ctx->net;
	const uint32_t Hd = net.density_hidden, Hr = net.rgb_hidden;
	const bool prof = ctx->opt.profile != 0;
	Ev ev(prof, s);
	FrameArgs a = train_frame_args(ctx);
	ev.start();
	HN_CUDA(cudaMemsetAsync(t.counters.p, 0, 4 * sizeof(uint32_t), s));
	HN_CUDA(cudaMemsetAsync(t.loss_acc.p, 0, 4 * sizeof(float), s));
	if ((size_t)n_rays * 1024u > t.march_staging.n) // the ray count drifts with the batch controller: grow in steps of 25 % + 4096 rows, never per step
		t.march_staging.ensure(((size_t)n_rays + n_rays / 4 + 4096u) * 1024u);
	launch_train_march(a, ctx->bitfield.p, n_rays, t.ray_o.p, t.ray_d.p, t.ray_t.p, max_samples, t.counters.p, t.ray_numsteps.p, t.ray_base.p, t.coords.p, t.sample_ray.p, t.march_staging.p, s);
	uint32_t n = 0;
	HN_CUDA(cudaMemcpyAsync(&n, t.counters.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	ev.stop(t.ms[0]);
	n = std::min(n, max_samples); // rays beyond the cap reserved slots but wrote nothing; their numsteps is 0
	t.n_samples_last = n; t.n_rays_last = n_rays;
	HN_CUDA(cudaMemsetAsync(t.grads.p, 0, net.n_params * sizeof(float), s));
	if (!n) return;
	cublasHandle_t h = handle(ctx);
	const __half* W = t.raw.p;
	// The GEMMs run on the sample count rounded up to 32 Ki rows (zero rows: no bias anywhere, so they stay zero through both
	// passes and add nothing to the weight gradients): the count changes every step, and a new GEMM shape costs cuBLAS a
	// heuristics search of ~2 ms per call on the host (measured: 26 ms per step for the 13 backward GEMMs).
	const uint32_t n_gemm = std::min<uint32_t>((n + 32767u) / 32768u * 32768u, (uint32_t)(t.act[0].n / 64));
	if (n_gemm > n) {
		HN_CUDA(cudaMemsetAsync(t.act[0].p + (size_t)n * 32, 0, (size_t)(n_gemm - n) * 32 * sizeof(__half), s));
		HN_CUDA(cudaMemsetAsync(t.act[Hd + 2].p + (size_t)n * 32, 0, (size_t)(n_gemm - n) * 32 * sizeof(__half), s));
		HN_CUDA(cudaMemsetAsync(t.dact[0].p + (size_t)n * 16, 0, (size_t)(n_gemm - n) * 16 * sizeof(__half), s));
	}
	__half* enc = t.act[0].p; __half* d_out = t.act[Hd + 1].p; __half* rgb_in = t.act[Hd + 2].p; __half* out = t.act[Hd + 3 + Hr].p;
	// ---- forward
	ev.start();
	if (net.F == 4) train_encode_kernel<4><<<blocks(n, 128), 128, 0, s>>>(ctx->levels, W + net.grid_offset, n, t.coords.p, t.sample_ray.p, t.ray_d.p, enc, rgb_in);
	else train_encode_kernel<2><<<blocks(n, 128), 128, 0, s>>>(ctx->levels, W + net.grid_offset, n, t.coords.p, t.sample_ray.p, t.ray_d.p, enc, rgb_in);
	const bool own_mlp = ctx->train_cfg.mlp_impl == 1 && Hd <= (uint32_t)tc::MAX_HIDDEN && Hr <= (uint32_t)tc::MAX_HIDDEN;
	tc::TrainMlpArgs M = {};
	if (own_mlp) {
		M.w_density = W + net.density_w_offset; M.w_rgb = W + net.rgb_w_offset;
		M.density_hidden = (int)Hd; M.rgb_hidden = (int)Hr; M.n = n;
		M.enc = enc; M.d_out = d_out; M.rgb_in = rgb_in; M.out = out;
		for (uint32_t l = 0; l < Hd; ++l) { M.hid_d[l] = t.act[1 + l].p; M.g_hid_d[l] = t.dact[8 + l].p; }
		for (uint32_t l = 0; l < Hr; ++l) { M.hid_r[l] = t.act[Hd + 3 + l].p; M.g_hid_r[l] = t.dact[4 + l].p; }
		M.g_out = t.dact[0].p; M.d_sigma = (const float*)t.dact[3].p; M.g_dout = t.dact[1].p; M.g_enc = t.dact[2].p;
	}
	auto mlp_forward = [&](const __half* w, uint32_t first_act, uint32_t hidden) {
		int K = 32;
		for (uint32_t l = 0; l < hidden; ++l) {
			gemm_forward(h, n_gemm, 64, K, t.act[first_act + l].p, w, t.act[first_act + l + 1].p);
			relu_kernel<<<blocks((size_t)n * 32), 256, 0, s>>>((__half2*)t.act[first_act + l + 1].p, (size_t)n * 32);
			w += 64 * K; K = 64;
		}
		gemm_forward(h, n_gemm, 16, 64, t.act[first_act + hidden].p, w, t.act[first_act + hidden + 1].p);
	};
	if (own_mlp) {
		launch_train_mlp_forward(M, s); // both MLPs, every activation stored, one launch
	} else {
		mlp_forward(W + net.density_w_offset, 0, Hd);
		copy_density_out_kernel<<<blocks(n), 256, 0, s>>>(d_out, rgb_in, n);
		mlp_forward(W + net.rgb_w_offset, Hd + 2, Hr);
	}
	ev.stop(t.ms[1]);
	// ---- loss and dL/d(output)
	DepthSup Z = {};
	if (target_depth && depth_lambda > 0.0f) {
		Z.target = target_depth; Z.ray_o = t.ray_o.p; Z.lambda = depth_lambda; Z.loss_type = depth_loss_type;
		for (int k = 0; k < 3; ++k) { Z.lo[k] = a.train_aabb[k]; Z.diag[k] = a.train_aabb[3 + k] - a.train_aabb[k]; }
	}
	ev.start();
	__half* g_out = t.dact[0].p;               // [n][16]
	float* d_sigma = (float*)t.dact[3].p;      // [n] fp32 (buffer holds n*64 halfs)
	train_loss_kernel<<<blocks((size_t)n_rays * 32, 128), 128, 0, s>>>(n_rays, n_rays_total, t.ray_numsteps.p, t.ray_base.p, t.coords.p, out, d_out, t.ray_target.p, t.ray_bg.p,
		loss_type, loss_scale, a.rgb_act, a.density_act, ctx->density_grid_mean < 0.01f ? 1e-4f : 0.0f, g_out, d_sigma, used_dev, t.loss_acc.p,
		emap ? *emap : ErrorMapDev{}, emap ? t.ray_img.p : nullptr, emap ? t.ray_uv.p : nullptr, emap ? t.ray_pdf.p : nullptr, t.img_w, t.img_h, Z);
	ev.stop(t.ms[2]);
	// ---- backward
	ev.start();
	// returns the gradient w.r.t. the MLP input [n][32] in dact[ret]
	auto mlp_backward = [&](const __half* w, float* gw, uint32_t first_act, uint32_t hidden, __half* g_top /*[n][16]*/, int buf_a, int buf_b) -> __half* {
		// weight offsets per layer
		std::vector<size_t> off(hidden + 1);
		size_t o = 0; int K = 32;
		for (uint32_t l = 0; l < hidden; ++l) { off[l] = o; o += (size_t)64 * K; K = 64; }
		off[hidden] = o;
		__half* g = g_top;
		int N = 16;
		int cur = buf_a;
		for (int l = (int)hidden; l >= 0; --l) {
			const int Kin = l == 0 ? 32 : 64;
			const __half* x = t.act[first_act + l].p;
			gemm_backward_weights(h, n_gemm, N, Kin, g, x, gw + off[l]);
			__half* gx = t.dact[cur].p;
			gemm_backward_input(h, n_gemm, N, Kin, g, w + off[l], gx);
			if (l > 0) relu_backward_kernel<<<blocks((size_t)n * 32), 256, 0, s>>>((__half2*)gx, (const __half2*)x, (size_t)n * 32);
			g = gx; N = 64;
			cur = (cur == buf_a) ? buf_b : buf_a;
		}
		return g;
	};
	__half* g_enc = nullptr;
	if (own_mlp) {
		launch_train_mlp_backward(M, s); // dX chain of both MLPs with ReLU masks and the sigma gradient folded in, one launch
		tc::DwArgs D = {};
		D.n = n;
		int col = 0;
		auto add = [&](const __half* a_mat, const __half* b_mat, int nb, int transposed, float* gw) {
			tc::DwLayer& L = D.layer[D.n_layers++];
			L.a = a_mat; L.b = b_mat; L.nb = nb; L.transposed = transposed; L.gw = gw; L.col = col;
			col += nb;
		};
		auto add_mlp = [&](float* gw, const __half* x0, __half* const* hid, __half* const* g_hid, uint32_t hidden, const __half* g_top) {
			add(g_hid[0], x0, 32, 0, gw); gw += 64 * 32;                                         // first layer: 64 x 32
			for (uint32_t l = 1; l < hidden; ++l) { add(g_hid[l], hid[l - 1], 64, 0, gw); gw += 64 * 64; } // hidden: 64 x 64
			add(hid[hidden - 1], g_top, 16, 1, gw);                                              // last: 16 x 64, accumulated as [in][out]
		};
		add_mlp(t.grads.p + net.density_w_offset, enc, M.hid_d, M.g_hid_d, Hd, M.g_dout);
		add_mlp(t.grads.p + net.rgb_w_offset, rgb_in, M.hid_r, M.g_hid_r, Hr, M.g_out);
		launch_train_dw(D, s);       // every weight gradient, split-K over the batch, one launch
		g_enc = M.g_enc;
	} else {
		__half* g_rgb_in = mlp_backward(W + net.rgb_w_offset, t.grads.p + net.rgb_w_offset, Hd + 2, Hr, g_out, 1, 2);
		__half* g_dout = t.dact[0].p; // dL/d(output) is consumed; reuse its buffer for [n][16]
		combine_density_grad_kernel<<<blocks(n), 256, 0, s>>>(g_rgb_in, d_sigma, g_dout, n);
		g_enc = mlp_backward(W + net.density_w_offset, t.grads.p + net.density_w_offset, 0, Hd, g_dout, 1, 2);
	}
	float* grid_grad = t.grads.p + net.grid_offset;
	if (net.F == 4) grid_backward_kernel<4><<<blocks((size_t)n * 8, 256), 256, 0, s>>>(ctx->levels, n, t.coords.p, g_enc, grid_grad);
	else grid_backward_kernel<2><<<blocks((size_t)n * 16, 256), 256, 0, s>>>(ctx->levels, n, t.coords.p, g_enc, grid_grad);
	ev.stop(t.ms[3]);
	HN_CUDA(cudaGetLastError());
}

// image CDF on the host from the per-image sums (:3052-3067): MIN_PMF = 0.1 of the mass stays uniform
void image_cdf_host(std::vector<float>& cdf /* in: unnormalised per-image mass, out: cdf */, std::vector<float>& pmf) {
	pmf = cdf;
	float cum = 0;
	for (float& f : cdf) { cum += f; f = cum; }
	const float norm = 1.0f / cum;
	const float n = (float)cdf.size();
	for (size_t i = 0; i < cdf.size(); ++i) {
		pmf[i] = (1.0f - 0.1f) * pmf[i] * norm + 0.1f / n;
		cdf[i] = (1.0f - 0.1f) * cdf[i] * norm + 0.1f * (float)(i + 1) / n;
	}
}
void launch_error_cdfs(uint32_t n_images, uint32_t height, uint32_t width, const float* data, float* cdf_x_cond_y, float* cdf_y, float* cdf_img, std::vector<float>& pmf, cudaStream_t s) {
	const dim3 threads(16, 8, 1), grid((height + 15) / 16, (n_images + 7) / 8, 1);
	construct_cdf_2d_kernel<<<grid, threads, 0, s>>>(n_images, height, width, data, cdf_x_cond_y, cdf_y);
	construct_cdf_1d_kernel<<<blocks(n_images), 256, 0, s>>>(n_images, height, cdf_y, cdf_img);
	std::vector<float> cdf(n_images);
	HN_CUDA(cudaMemcpyAsync(cdf.data(), cdf_img, n_images * sizeof(float), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	image_cdf_host(cdf, pmf);
	HN_CUDA(cudaMemcpyAsync(cdf_img, cdf.data(), n_images * sizeof(float), cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaStreamSynchronize(s)); // `cdf` lives on this stack frame
}
void build_error_cdfs(hn_context* ctx) {
	TrainState& t = ctx->train;
	t.cdf_res[0] = t.err_res[0]; t.cdf_res[1] = t.err_res[1];
	const size_t n_cells = (size_t)t.cdf_res[0] * t.cdf_res[1] * t.n_images;
	t.cdf_x_cond_y.ensure(n_cells); t.cdf_y.ensure((size_t)t.cdf_res[1] * t.n_images); t.cdf_img.ensure(t.n_images);
	launch_error_cdfs(t.n_images, (uint32_t)t.cdf_res[1], (uint32_t)t.cdf_res[0], t.error_map.p, t.cdf_x_cond_y.p, t.cdf_y.p, t.cdf_img.p, t.pmf_img, ctx->stream);
	t.cdf_valid = true;
}

void check_ready(hn_context* ctx, bool need_dataset) {
	if (!ctx->has_net || !ctx->has_params) throw std::runtime_error("training needs a network config and parameters");
	if (!ctx->has_bitfield) throw std::runtime_error("training needs an occupancy grid (hn_set_density_grid)");
	if (need_dataset && !ctx->train.has_dataset) throw std::runtime_error("no training dataset (hn_train_set_dataset)");
	if (!ctx->train.ready) train_init(ctx, ctx->train.seed);
}

} // namespace

void train_set_dataset(hn_context* ctx, const float* rgba, uint32_t n, int w, int h, const float* cams, const float focal[2], const float principal[2]) {
	if (!rgba || !cams || n == 0 || w < 2 || h < 2 || !(focal[0] > 0.f) || !(focal[1] > 0.f)) throw std::runtime_error("bad training dataset arguments");
	TrainState& t = ctx->train;
	const size_t px = (size_t)n * w * h;
	t.images.ensure(px); t.cams.ensure((size_t)n * 12);
	HN_CUDA(cudaMemcpy(t.images.p, rgba, px * sizeof(float4), cudaMemcpyHostToDevice));
	HN_CUDA(cudaMemcpy(t.cams.p, cams, (size_t)n * 12 * sizeof(float), cudaMemcpyHostToDevice));
	t.n_images = n; t.img_w = w; t.img_h = h;
	t.focal[0] = focal[0]; t.focal[1] = focal[1]; t.principal[0] = principal[0]; t.principal[1] = principal[1];
	{ // every image starts with the shared pinhole camera; hn_train_set_image_metadata / the loader refine it per image
		ImageMeta m = {};
		m.focal[0] = focal[0]; m.focal[1] = focal[1]; m.principal[0] = principal[0]; m.principal[1] = principal[1];
		t.meta_host.assign(n, m);
		t.meta.ensure(n);
		HN_CUDA(cudaMemcpy(t.meta.p, t.meta_host.data(), (size_t)n * sizeof(ImageMeta), cudaMemcpyHostToDevice));
	}
	t.has_dataset = true;
	t.has_depth = false;
	t.srgb_space = 0;
}

// metadata[img].depth of the reference's dataset: one float depth image per training image (z-depth in scene units, 0 = no measurement)
void train_set_depth(hn_context* ctx, const float* depth, uint32_t n, int w, int h) {
	TrainState& t = ctx->train;
	if (!depth) { t.has_depth = false; return; }
	if (!t.has_dataset || n != t.n_images || w != t.img_w || h != t.img_h) throw std::runtime_error("depth images: set the dataset first, one depth image per training image at its resolution");
	const size_t px = (size_t)n * w * h;
	t.depth.ensure(px);
	HN_CUDA(cudaMemcpy(t.depth.p, depth, px * sizeof(float), cudaMemcpyHostToDevice));
	t.has_depth = true;
}
// TrainingImageMetadata per image (nerf.h; read_focal_length / read_lens of the loader, src/nerf_loader.cu:175-271,678-696)
void train_set_image_metadata(hn_context* ctx, uint32_t n, const float* focal, const float* principal, const int32_t* lens_mode, const float* lens_params) {
	TrainState& t = ctx->train;
	if (!t.has_dataset || n != t.n_images) throw std::runtime_error("image metadata: set the dataset first, one entry per image");
	if (!focal) throw std::runtime_error("image metadata: focal lengths are required");
	for (uint32_t i = 0; i < n; ++i) {
		ImageMeta& m = t.meta_host[i];
		m.focal[0] = focal[2 * i]; m.focal[1] = focal[2 * i + 1];
		if (!(m.focal[0] > 0.f) || !(m.focal[1] > 0.f)) throw std::runtime_error("image metadata: focal lengths must be positive");
		if (principal) { m.principal[0] = principal[2 * i]; m.principal[1] = principal[2 * i + 1]; }
		if (lens_mode) {
			if (lens_mode[i] < 0 || lens_mode[i] > 5) throw std::runtime_error("image metadata: lens mode must be an ELensMode (0..5)");
			m.lens_mode = lens_mode[i];
		}
		if (lens_params) for (int k = 0; k < 7; ++k) m.lens_params[k] = lens_params[7 * i + k];
	}
	HN_CUDA(cudaMemcpy(t.meta.p, t.meta_host.data(), (size_t)n * sizeof(ImageMeta), cudaMemcpyHostToDevice));
}

void train_init(hn_context* ctx, uint32_t seed) {
	if (!ctx->has_net || !ctx->has_params) throw std::runtime_error("training needs a network config and parameters");
	TrainState& t = ctx->train;
	const size_t n = ctx->net.n_params;
	cudaStream_t s = ctx->stream;
	t.master.ensure(n); t.m1.ensure(n); t.m2.ensure(n); t.ema.ensure(n); t.grads.ensure(n); t.param_steps.ensure(n); t.raw.ensure(n);
	HN_CUDA(cudaMemcpyAsync(t.raw.p, ctx->params.p, n * sizeof(__half), cudaMemcpyDeviceToDevice, s));
	half_to_float_kernel<<<blocks(n), 256, 0, s>>>(t.raw.p, t.master.p, n);
	HN_CUDA(cudaMemsetAsync(t.m1.p, 0, n * sizeof(float), s)); HN_CUDA(cudaMemsetAsync(t.m2.p, 0, n * sizeof(float), s));
	HN_CUDA(cudaMemsetAsync(t.ema.p, 0, n * sizeof(float), s)); HN_CUDA(cudaMemsetAsync(t.grads.p, 0, n * sizeof(float), s));
	HN_CUDA(cudaMemsetAsync(t.param_steps.p, 0, n * sizeof(uint32_t), s));
	HN_CUDA(cudaStreamSynchronize(s));
	t.seed = seed; t.step = 0; t.ema_step = 0; t.rays_per_batch = 4096; t.loss_last = 0.f;
	t.n_steps_since_error_map_update = 0; t.n_steps_between_error_map_updates = 128; t.cdf_valid = false; // Testbed::reset_network, src/testbed.cu:3857-3860
	t.ready = true;
}

void train_accumulate(hn_context* ctx, uint32_t n_rays, float* loss_out) {
	if (ctx->train_cfg.auto_grid && ctx->train.has_dataset && ctx->has_net && ctx->has_params) {
		// Testbed::train (src/testbed.cu:4418-4440): training_prep_nerf every min(max(step / 16, 1), 16) steps; all cells for the
		// first 256 steps, then a quarter uniformly and a quarter among the occupied cells (:3433-3446)
		if (!ctx->train.ready) train_init(ctx, ctx->train.seed);
		TrainState& tt = ctx->train;
		const uint32_t skip = std::min(std::max(tt.step / 16u, 1u), 16u);
		if (tt.step % skip == 0) {
			const uint32_t nc = (uint32_t)ctx->opt.max_cascade + 1u;
			if (tt.step == 0) train_mark_untrained(ctx);
			if (tt.step < 256) train_update_density_grid(ctx, ctx->train_cfg.grid_decay, GRID_CELLS * nc, 0);
			else train_update_density_grid(ctx, ctx->train_cfg.grid_decay, GRID_CELLS / 4 * nc, GRID_CELLS / 4 * nc);
		}
	}
	check_ready(ctx, true);
	TrainState& t = ctx->train;
	const TrainCfg& c = ctx->train_cfg;
	cudaStream_t s = ctx->stream;
	const uint32_t world = (uint32_t)std::max(ctx->world, 1), rank = (uint32_t)ctx->rank;
	if (!n_rays) n_rays = t.rays_per_batch;
	const uint32_t max_samples = c.target_batch * 2u;
	ensure_batch(ctx, n_rays, max_samples);
	FrameArgs a = train_frame_args(ctx);
	TrainDataDev D = {t.images.p, t.cams.p, t.n_images, t.img_w, t.img_h, t.meta.p};
	const float* bg = ctx->opt.background_color;
	// error map: a fresh, zeroed map whenever a new accumulation period starts (train_nerf, src/testbed_nerf.cu:2986-2992); each rank keeps its own
	if (t.n_steps_since_error_map_update == 0) {
		const uint32_t n_samples_per_image = (uint32_t)(((uint64_t)t.n_steps_between_error_map_updates * t.rays_per_batch) / t.n_images);
		const int r = (int)(std::sqrt(std::sqrt((float)n_samples_per_image)) * 3.5f);
		t.err_res[0] = std::max(std::min(r, t.img_w), 2); t.err_res[1] = std::max(std::min(r, t.img_h), 2);
		const size_t n_cells = (size_t)t.err_res[0] * t.err_res[1] * t.n_images;
		t.error_map.ensure(n_cells);
		HN_CUDA(cudaMemsetAsync(t.error_map.p, 0, n_cells * sizeof(float), s));
	}
	ErrorMapDev E = {};
	E.data = t.error_map.p; E.res[0] = t.err_res[0]; E.res[1] = t.err_res[1];
	if (t.cdf_valid) {
		E.cdf_res[0] = t.cdf_res[0]; E.cdf_res[1] = t.cdf_res[1];
		if (c.sample_focal_plane_proportional_to_error) { E.cdf_x_cond_y = t.cdf_x_cond_y.p; E.cdf_y = t.cdf_y.p; }
		if (c.sample_image_proportional_to_error) E.cdf_img = t.cdf_img.p;
	}
	train_raygen_kernel<<<blocks(n_rays), 256, 0, s>>>(D, a, n_rays, rank * n_rays, n_rays * world, t.seed, t.step, c.random_bg, t.srgb_space, bg[0], bg[1], bg[2],
		t.ray_o.p, t.ray_d.p, t.ray_t.p, t.ray_target.p, t.ray_bg.p, E, t.ray_img.p, t.ray_uv.p, t.ray_pdf.p,
		(t.has_depth && c.depth_supervision_lambda > 0.0f) ? t.depth.p : nullptr, t.ray_target_depth.p);
	forward_backward(ctx, n_rays, n_rays * world, max_samples, c.loss_type, c.loss_scale, nullptr, &E,
		(t.has_depth && c.depth_supervision_lambda > 0.0f) ? t.ray_target_depth.p : nullptr, c.depth_supervision_lambda, c.depth_loss_type);
	// CDFs from the error map every n_steps_between_error_map_updates steps, then a 1.5x longer period (:3023-3075)
	if (++t.n_steps_since_error_map_update >= t.n_steps_between_error_map_updates) {
		build_error_cdfs(ctx);
		t.n_steps_since_error_map_update = 0;
		t.n_steps_between_error_map_updates = (uint32_t)(t.n_steps_between_error_map_updates * 1.5f);
	}
	float loss = 0.f;
	HN_CUDA(cudaMemcpyAsync(&loss, t.loss_acc.p, sizeof(float), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	t.loss_last = loss;
	if (loss_out) *loss_out = loss;
	// rays per batch follow the measured samples per ray towards the target batch size (src/testbed_nerf.cu:3049-3056)
	if (t.n_samples_last > 0) {
		const double per_ray = (double)t.n_samples_last / (double)n_rays;
		uint32_t next = (uint32_t)std::min<double>((double)c.target_batch / per_ray, 1u << 18); // the reference's cap (src/testbed_nerf.cu:2943-2944)
		next = std::max(128u, (next + 127u) / 128u * 128u);
		t.rays_per_batch = next;
	}
}

void train_apply(hn_context* ctx, float grad_scale) {
	check_ready(ctx, false);
	TrainState& t = ctx->train;
	const TrainCfg& c = ctx->train_cfg;
	cudaStream_t s = ctx->stream;
	Ev ev(ctx->opt.profile != 0, s);
	ev.start();
	// ExponentialDecay (exponential_decay.h): base lr scaled by decay_base^((step - start) / interval + 1) once step >= start
	float lr = c.lr;
	if (c.exp_decay && t.step >= c.decay_start && c.decay_interval > 0) lr *= std::pow(c.decay_base, (float)((t.step - c.decay_start) / c.decay_interval + 1));
	t.lr_last = lr;
	if (t.ema_step < (1u << 20)) ++t.ema_step;
	const float debias = c.ema ? 1.0f - std::pow(c.ema_decay, (float)t.ema_step) : 1.0f;
	const uint32_t n = (uint32_t)ctx->net.n_params, n_matrix = (uint32_t)ctx->net.grid_offset; // blob = [density MLP | rgb MLP | grid]
	adam_kernel<<<blocks(n), 256, 0, s>>>(n, n_matrix, grad_scale, 1.0f / c.loss_scale, lr, c.beta1, c.beta2, c.epsilon, c.l2_reg, c.ema ? 1 : 0, c.ema_decay, debias,
		t.master.p, t.raw.p, ctx->params.p, t.grads.p, t.m1.p, t.m2.p, t.ema.p, t.param_steps.p);
	ev.stop(t.ms[4]);
	++t.step;
	ctx->params_host.clear(); // the host copy (snapshot writer) is stale now; hn_save_snapshot reads the device copy back
	HN_CUDA(cudaGetLastError());
}

void train_op_gradients(hn_context* ctx, uint32_t n_rays, const float* o, const float* d, const float* tt, const float* targets, const float* bgs,
	int loss_type, float loss_scale, float* loss_out, float* grads_host, uint32_t* n_steps, const float* target_depth, float depth_lambda, int depth_loss_type) {
	check_ready(ctx, false);
	TrainState& t = ctx->train;
	cudaStream_t s = ctx->stream;
	if (!n_rays) throw std::runtime_error("no rays");
	const uint32_t max_samples = std::max<uint32_t>(n_rays * 64u, 1u << 16);
	ensure_batch(ctx, n_rays, max_samples);
	HN_CUDA(cudaMemcpyAsync(t.ray_o.p, o, (size_t)n_rays * 12, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemcpyAsync(t.ray_d.p, d, (size_t)n_rays * 12, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemcpyAsync(t.ray_t.p, tt, (size_t)n_rays * 4, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemcpyAsync(t.ray_target.p, targets, (size_t)n_rays * 12, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemcpyAsync(t.ray_bg.p, bgs, (size_t)n_rays * 12, cudaMemcpyHostToDevice, s));
	DevBuf<uint32_t> used; used.ensure(n_rays);
	HN_CUDA(cudaMemsetAsync(used.p, 0, (size_t)n_rays * 4, s));
	if (target_depth) HN_CUDA(cudaMemcpyAsync(t.ray_target_depth.p, target_depth, (size_t)n_rays * 4, cudaMemcpyHostToDevice, s));
	forward_backward(ctx, n_rays, n_rays, max_samples, loss_type, loss_scale, used.p, nullptr, target_depth ? t.ray_target_depth.p : nullptr, depth_lambda, depth_loss_type);
	if (loss_out) HN_CUDA(cudaMemcpyAsync(loss_out, t.loss_acc.p, sizeof(float), cudaMemcpyDeviceToHost, s));
	if (grads_host) HN_CUDA(cudaMemcpyAsync(grads_host, t.grads.p, ctx->net.n_params * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (n_steps) HN_CUDA(cudaMemcpyAsync(n_steps, used.p, (size_t)n_rays * 4, cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
}

// Testbed::update_density_grid_nerf (src/testbed_nerf.cu:2772-2861) + update_density_grid_mean_and_bitfield (:2863-2880)
void train_update_density_grid(hn_context* ctx, float decay, uint32_t n_uniform, uint32_t n_nonuniform) {
	if (!ctx->has_net || !ctx->has_params) throw std::runtime_error("density-grid update needs a network config and parameters");
	TrainState& t = ctx->train;
	cudaStream_t s = ctx->stream;
	const uint32_t n_cascades = (uint32_t)ctx->opt.max_cascade + 1u;
	const uint32_t n_elements = GRID_CELLS * n_cascades;
	if (ctx->density_grid.n < n_elements || ctx->grid_cascades != n_cascades) { // no grid yet (or another cascade count): start from "nothing known"
		ctx->density_grid.ensure(n_elements);
		HN_CUDA(cudaMemsetAsync(ctx->density_grid.p, 0, (size_t)n_elements * sizeof(float), s));
		ctx->grid_cascades = n_cascades;
	}
	const uint32_t n = n_uniform + n_nonuniform;
	if (!n) return;
	const uint32_t chunks = (n + CHUNK - 1) / CHUNK;
	t.grid_tmp.ensure(n_elements); t.grid_pos.ensure((size_t)chunks * CHUNK); t.grid_idx.ensure(n); t.grid_out.ensure((size_t)chunks * CHUNK);
	HN_CUDA(cudaMemsetAsync(t.grid_tmp.p, 0, (size_t)n_elements * sizeof(float), s));
	HN_CUDA(cudaMemsetAsync(t.grid_pos.p, 0, (size_t)chunks * CHUNK * sizeof(float4), s)); // padding samples sit at the box corner
	for (int pass = 0; pass < 2; ++pass) {
		const uint32_t cnt = pass == 0 ? n_uniform : n_nonuniform, off = pass == 0 ? 0u : n_uniform;
		Pcg32 rng(t.grid_seed);
		for (uint32_t k = 0; k < t.grid_rng_calls; ++k) rng.advance(1ull << 32);
		++t.grid_rng_calls; // density_grid_rng.advance() after every generate call, used or not (:2826,2839)
		if (cnt) grid_samples_kernel<<<blocks(cnt), 256, 0, s>>>(cnt, rng, t.grid_ema_step, ctx->density_grid.p, t.grid_pos.p + off, t.grid_idx.p + off, n_cascades,
			pass == 0 ? -0.01f : 0.01f);
	}
	FrameArgs a = train_frame_args(ctx);
	RaySoA none = {};
	const __half* weights = t.ready ? t.raw.p : ctx->params.p; // density(..., use_inference_params = false): the raw training weights
	launch_eval(ctx->net, ctx->levels, weights, a, t.grid_pos.p, nullptr, nullptr, none, nullptr, nullptr, chunks, t.grid_out.p, 1, nullptr, nullptr, nullptr, nullptr, s, true);
	grid_splat_kernel<<<blocks(n), 256, 0, s>>>(n, t.grid_idx.p, t.grid_out.p, a.density_act, t.grid_tmp.p);
	grid_ema_kernel<<<blocks(n_elements), 256, 0, s>>>(n_elements, decay, ctx->density_grid.p, t.grid_tmp.p);
	++t.grid_ema_step;
	HN_CUDA(cudaGetLastError());
	ctx->density_grid_host.clear(); // stale; the snapshot writer reads the device grid back
	refresh_bitfield_from_device_grid(ctx);
}

void train_mark_untrained(hn_context* ctx) {
	TrainState& t = ctx->train;
	if (!t.has_dataset) throw std::runtime_error("no training dataset (hn_train_set_dataset)");
	const uint32_t n_cascades = (uint32_t)ctx->opt.max_cascade + 1u, n_elements = GRID_CELLS * n_cascades;
	ctx->density_grid.ensure(n_elements);
	ctx->grid_cascades = n_cascades;
	TrainDataDev D = {t.images.p, t.cams.p, t.n_images, t.img_w, t.img_h, t.meta.p};
	mark_untrained_kernel<<<blocks(n_elements), 256, 0, ctx->stream>>>(n_elements, ctx->density_grid.p, D);
	t.grid_ema_step = 0;
	HN_CUDA(cudaGetLastError());
}

// Trainer state for snapshots (Trainer::serialize(true), tcnn trainer.h:275-288; AdamOptimizer::serialize, adam.h:278-299)
void train_export_state(hn_context* ctx, std::vector<float>& m1, std::vector<float>& m2, std::vector<uint32_t>& steps) {
	TrainState& t = ctx->train;
	if (!t.ready) throw std::runtime_error("trainer not initialised");
	const size_t n = ctx->net.n_params;
	m1.resize(n); m2.resize(n); steps.resize(n);
	HN_CUDA(cudaMemcpy(m1.data(), t.m1.p, n * sizeof(float), cudaMemcpyDeviceToHost));
	HN_CUDA(cudaMemcpy(m2.data(), t.m2.p, n * sizeof(float), cudaMemcpyDeviceToHost));
	HN_CUDA(cudaMemcpy(steps.data(), t.param_steps.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
}
// The snapshot's parameters are the inference (EMA) weights; as in Trainer::deserialize they become the training weights,
// and the EMA accumulator restarts from them (Ema::deserialize, ema.h:197-206).
void train_import_state(hn_context* ctx, uint32_t step, const float* m1, const float* m2, const uint32_t* steps, size_t n) {
	if (n != ctx->net.n_params) throw std::runtime_error("snapshot: optimizer state has the wrong number of parameters");
	train_init(ctx, ctx->train.seed);
	TrainState& t = ctx->train;
	HN_CUDA(cudaMemcpy(t.m1.p, m1, n * sizeof(float), cudaMemcpyHostToDevice));
	HN_CUDA(cudaMemcpy(t.m2.p, m2, n * sizeof(float), cudaMemcpyHostToDevice));
	if (steps) HN_CUDA(cudaMemcpy(t.param_steps.p, steps, n * sizeof(uint32_t), cudaMemcpyHostToDevice));
	HN_CUDA(cudaMemcpy(t.ema.p, t.master.p, n * sizeof(float), cudaMemcpyDeviceToDevice)); // accumulator = weights, debias 1
	t.step = step;
	t.ema_step = 1u << 20; // 1 - decay^step == 1: the restored accumulator is already debiased
}

// ---- error-map operators (test entry points: the same device functions the training kernels call) ----
namespace {
__global__ void op_sample_pixels_kernel(uint32_t n_rays, uint32_t ray_offset, uint32_t n_rays_total, uint32_t seed, uint32_t step, uint32_t n_images, ErrorMapDev E,
	uint32_t* __restrict__ img_out, float* __restrict__ uv_out, float* __restrict__ pdf_out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_rays) return;
	const uint32_t gi = i + ray_offset;
	Pcg32 rng((uint64_t)seed + ((uint64_t)step << 32));
	rng.advance((uint64_t)gi * 16u);
	float img_pdf = 1.0f, uv_pdf = 1.0f;
	const uint32_t img = train_image_idx(gi, n_rays_total, n_images, E.cdf_img, img_pdf);
	float u = rng.next_float(), v = rng.next_float();
	if (E.cdf_x_cond_y) train_sample_cdf_2d(u, v, img, E.cdf_res[0], E.cdf_res[1], E.cdf_x_cond_y, E.cdf_y, uv_pdf);
	img_out[i] = img; uv_out[2 * i] = u; uv_out[2 * i + 1] = v; pdf_out[i] = img_pdf * uv_pdf;
}
__global__ void op_deposit_kernel(uint32_t n, const uint32_t* __restrict__ img, const float* __restrict__ uv, const float* __restrict__ val, int res_x, int res_y, int img_w, int img_h, float* __restrict__ map) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) error_map_deposit(map, res_x, res_y, img_w, img_h, img[i], uv[2 * i], uv[2 * i + 1], val[i]);
}
template <typename T> struct Tmp { // scoped device buffer of the operator entry points
	T* p = nullptr;
	explicit Tmp(size_t n) { if (n) HN_CUDA(cudaMalloc((void**)&p, n * sizeof(T))); }
	~Tmp() { if (p) cudaFree(p); }
	Tmp(const Tmp&) = delete; Tmp& operator=(const Tmp&) = delete;
};
} // namespace

void train_op_error_map_cdf(hn_context* ctx, uint32_t n_images, uint32_t height, uint32_t width, const float* data, float* cdf_x_cond_y, float* cdf_y, float* cdf_img, float* pmf_img) {
	if (!n_images || !height || !width || !data) throw std::runtime_error("bad error-map arguments");
	cudaStream_t s = ctx->stream;
	const size_t n = (size_t)n_images * height * width;
	Tmp<float> d_data(n), d_x(n), d_y((size_t)n_images * height), d_i(n_images);
	HN_CUDA(cudaMemcpyAsync(d_data.p, data, n * sizeof(float), cudaMemcpyHostToDevice, s));
	std::vector<float> pmf;
	launch_error_cdfs(n_images, height, width, d_data.p, d_x.p, d_y.p, d_i.p, pmf, s);
	if (cdf_x_cond_y) HN_CUDA(cudaMemcpyAsync(cdf_x_cond_y, d_x.p, n * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (cdf_y) HN_CUDA(cudaMemcpyAsync(cdf_y, d_y.p, (size_t)n_images * height * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (cdf_img) HN_CUDA(cudaMemcpyAsync(cdf_img, d_i.p, n_images * sizeof(float), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	if (pmf_img) std::copy(pmf.begin(), pmf.end(), pmf_img);
	HN_CUDA(cudaGetLastError());
}
void train_op_sample_pixels(hn_context* ctx, uint32_t n_rays, uint32_t ray_offset, uint32_t n_rays_total, uint32_t seed, uint32_t step, uint32_t n_images, int cdf_res_x, int cdf_res_y,
	const float* cdf_x_cond_y, const float* cdf_y, const float* cdf_img, uint32_t* img, float* uv, float* pdf) {
	if (!n_rays || !n_images || !n_rays_total) throw std::runtime_error("bad sampling arguments");
	if ((cdf_x_cond_y != nullptr) != (cdf_y != nullptr)) throw std::runtime_error("cdf_x_cond_y and cdf_y come together");
	cudaStream_t s = ctx->stream;
	const size_t n = cdf_x_cond_y ? (size_t)n_images * cdf_res_x * cdf_res_y : 0;
	Tmp<float> d_x(n), d_y(cdf_y ? (size_t)n_images * cdf_res_y : 0), d_i(cdf_img ? n_images : 0), d_uv((size_t)n_rays * 2), d_pdf(n_rays);
	Tmp<uint32_t> d_img(n_rays);
	if (cdf_x_cond_y) {
		HN_CUDA(cudaMemcpyAsync(d_x.p, cdf_x_cond_y, n * sizeof(float), cudaMemcpyHostToDevice, s));
		HN_CUDA(cudaMemcpyAsync(d_y.p, cdf_y, (size_t)n_images * cdf_res_y * sizeof(float), cudaMemcpyHostToDevice, s));
	}
	if (cdf_img) HN_CUDA(cudaMemcpyAsync(d_i.p, cdf_img, n_images * sizeof(float), cudaMemcpyHostToDevice, s));
	ErrorMapDev E = {};
	E.cdf_x_cond_y = d_x.p; E.cdf_y = d_y.p; E.cdf_img = d_i.p; E.cdf_res[0] = cdf_res_x; E.cdf_res[1] = cdf_res_y;
	op_sample_pixels_kernel<<<blocks(n_rays), 256, 0, s>>>(n_rays, ray_offset, n_rays_total, seed, step, n_images, E, d_img.p, d_uv.p, d_pdf.p);
	HN_CUDA(cudaMemcpyAsync(img, d_img.p, (size_t)n_rays * 4, cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaMemcpyAsync(uv, d_uv.p, (size_t)n_rays * 8, cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaMemcpyAsync(pdf, d_pdf.p, (size_t)n_rays * 4, cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	HN_CUDA(cudaGetLastError());
}
void train_op_error_map_deposit(hn_context* ctx, uint32_t n, const uint32_t* img, const float* uv, const float* val, uint32_t n_images, int res_x, int res_y, int img_w, int img_h, float* map) {
	if (!n || !n_images || res_x < 2 || res_y < 2 || img_w < 2 || img_h < 2) throw std::runtime_error("bad error-map arguments");
	for (uint32_t i = 0; i < n; ++i) if (img[i] >= n_images) throw std::runtime_error("image index out of range");
	cudaStream_t s = ctx->stream;
	const size_t cells = (size_t)n_images * res_x * res_y;
	Tmp<float> d_map(cells), d_uv((size_t)n * 2), d_val(n);
	Tmp<uint32_t> d_img(n);
	HN_CUDA(cudaMemsetAsync(d_map.p, 0, cells * sizeof(float), s));
	HN_CUDA(cudaMemcpyAsync(d_img.p, img, (size_t)n * 4, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemcpyAsync(d_uv.p, uv, (size_t)n * 8, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemcpyAsync(d_val.p, val, (size_t)n * 4, cudaMemcpyHostToDevice, s));
	op_deposit_kernel<<<blocks(n), 256, 0, s>>>(n, d_img.p, d_uv.p, d_val.p, res_x, res_y, img_w, img_h, d_map.p);
	HN_CUDA(cudaMemcpyAsync(map, d_map.p, cells * sizeof(float), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	HN_CUDA(cudaGetLastError());
}
// m_nerf.training.error_map.{resolution, data} (the "Training view error" overlays of the reference GUI read these)
void train_get_error_map(hn_context* ctx, int res[2], uint32_t* n_images, float* data, size_t n_floats) {
	TrainState& t = ctx->train;
	res[0] = t.err_res[0]; res[1] = t.err_res[1];
	if (n_images) *n_images = t.n_images;
	if (!data) return;
	const size_t n = (size_t)t.err_res[0] * t.err_res[1] * t.n_images;
	if (!n || !t.error_map.p) throw std::runtime_error("no error map yet (train at least one step)");
	if (n_floats < n) throw std::runtime_error("error map buffer too small");
	HN_CUDA(cudaMemcpyAsync(data, t.error_map.p, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
}

void train_release(hn_context* ctx) {
	if (ctx->train.cublas) { cublasDestroy((cublasHandle_t)ctx->train.cublas); ctx->train.cublas = nullptr; }
}

} // namespace hn
