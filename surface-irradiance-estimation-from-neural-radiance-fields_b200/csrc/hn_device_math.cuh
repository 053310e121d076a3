// Device math for the IEEE-exact translation units (compiled with -fmad=false, no fast math): every expression is written
// in the operation order of the reference's host-callable functions so results match the CPU oracle bit for bit.
#pragma once
#include "hn_internal.h"
#include <cfloat>

namespace hn {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { return {x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, float s) { return {a.x - s, a.y - s, a.z - s}; }
__device__ __forceinline__ V3 operator+(V3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
__device__ __forceinline__ V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
// tcnn vec.h:425 dot: ((0 + a0*b0) + a1*b1) + a2*b2
__device__ __forceinline__ float dot3(V3 a, V3 b) { float r = 0.0f; r += a.x * b.x; r += a.y * b.y; r += a.z * b.z; return r; }
__device__ __forceinline__ float length3(V3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ V3 normalize3(V3 v) { float l = length3(v); if (l <= 0.0f) return {1.0f, 0.0f, 0.0f}; return v / l; }
__device__ __forceinline__ V3 cross3(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ float mixf(float a, float b, float c) { return a * (1.0f - c) + b * c; }
__device__ __forceinline__ V3 mix3(V3 a, V3 b, float c) { return a * (1.0f - c) + b * c; }
__device__ __forceinline__ V3 mix3(V3 a, V3 b, V3 c) { return {mixf(a.x, b.x, c.x), mixf(a.y, b.y, c.y), mixf(a.z, b.z, c.z)}; }
__device__ __forceinline__ V3 ld3(const float* f) { return {f[0], f[1], f[2]}; }

// Thread index -> pixel of THIS rank. One rank: row-major over the frame. Tile-sharded: the rank's own tiles are
// enumerated back to back (local tile k = global tile k*world + rank), so no thread is spent on a foreign pixel.
__device__ __forceinline__ bool owned_pixel(const FrameArgs& a, uint32_t i, uint32_t& x, uint32_t& y) {
	const uint32_t w = (uint32_t)a.w, h = (uint32_t)a.h;
	if (a.world <= 1) {
		if (i >= w * h) return false;
		x = i % w; y = i / w;
		return true;
	}
	const uint32_t tw = (uint32_t)a.tile_w, th = (uint32_t)a.tile_h, per_tile = tw * th;
	const uint32_t tiles_x = (w + tw - 1) / tw, tiles_y = (h + th - 1) / th;
	const uint32_t tile = (i / per_tile) * (uint32_t)a.world + (uint32_t)a.rank;
	if (tile >= tiles_x * tiles_y) return false;
	const uint32_t k = i % per_tile;
	x = (tile % tiles_x) * tw + k % tw;
	y = (tile / tiles_x) * th + k / tw;
	return x < w && y < h;
}
__host__ inline uint32_t owned_pixel_threads(const FrameArgs& a) {
	if (a.world <= 1) return (uint32_t)a.w * (uint32_t)a.h;
	const uint32_t tiles = (uint32_t)((a.w + a.tile_w - 1) / a.tile_w) * (uint32_t)((a.h + a.tile_h - 1) / a.tile_h);
	return ((tiles + a.world - 1 - a.rank) / a.world) * (uint32_t)(a.tile_w * a.tile_h);
}

struct Box { V3 lo, hi; };
__device__ __forceinline__ Box ldbox(const float* f) { return {{f[0], f[1], f[2]}, {f[3], f[4], f[5]}}; }
// bounding_box.cuh:313
__device__ __forceinline__ bool box_contains(const Box& b, V3 p) {
	return p.x >= b.lo.x && p.x <= b.hi.x && p.y >= b.lo.y && p.y <= b.hi.y && p.z >= b.lo.z && p.z <= b.hi.z;
}
// bounding_box.cuh:172-219: returns (tmin, tmax), (FLT_MAX, FLT_MAX) on a miss
__device__ __forceinline__ float2 box_ray_intersect(const Box& b, V3 pos, V3 dir) {
	float tmin = (b.lo.x - pos.x) / dir.x;
	float tmax = (b.hi.x - pos.x) / dir.x;
	if (tmin > tmax) { float t = tmin; tmin = tmax; tmax = t; }
	float tymin = (b.lo.y - pos.y) / dir.y;
	float tymax = (b.hi.y - pos.y) / dir.y;
	if (tymin > tymax) { float t = tymin; tymin = tymax; tymax = t; }
	if (tmin > tymax || tymin > tmax) return {FLT_MAX, FLT_MAX};
	if (tymin > tmin) tmin = tymin;
	if (tymax < tmax) tmax = tymax;
	float tzmin = (b.lo.z - pos.z) / dir.z;
	float tzmax = (b.hi.z - pos.z) / dir.z;
	if (tzmin > tzmax) { float t = tzmin; tzmin = tzmax; tzmax = t; }
	if (tmin > tzmax || tzmin > tmax) return {FLT_MAX, FLT_MAX};
	if (tzmin > tmin) tmin = tzmin;
	if (tzmax < tmax) tmax = tzmax;
	return {tmin, tmax};
}

// ---- Owen-scrambled Sobol, dimension 0 (random_val.cuh:207-345). sobol(index, 0) is the bit reversal of index. ----
__host__ __device__ __forceinline__ uint32_t hn_reverse_bits(uint32_t x) {
#ifdef __CUDA_ARCH__
	return __brev(x);
#else
	x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
	x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
	x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
	x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
	return (x >> 16) | (x << 16);
#endif
}
__host__ __device__ __forceinline__ uint32_t hn_laine_karras(uint32_t x, uint32_t seed) {
	x += seed;
	x ^= x * 0x6c50b47cu;
	x ^= x * 0xb82f1e52u;
	x ^= x * 0xc7afe638u;
	x ^= x * 0x8d22f6e6u;
	return x;
}
__host__ __device__ __forceinline__ uint32_t hn_owen(uint32_t x, uint32_t seed) { return hn_reverse_bits(hn_laine_karras(hn_reverse_bits(x), seed)); }
__host__ __device__ __forceinline__ uint32_t hn_hash_combine(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
__host__ __device__ __forceinline__ float hn_ld_random_val(uint32_t index, uint32_t seed) {
	index = hn_owen(index, seed);
	return (float)hn_owen(hn_reverse_bits(index), hn_hash_combine(seed, 0u)) * 2.3283064365386963e-10f; // 2^-32
}

// dimension 1 of the Sobol sequence: v_j = v_{j-1} ^ (v_{j-1} >> 1) (random_val.cuh:207-300)
__host__ __device__ __forceinline__ uint32_t hn_sobol_dim1(uint32_t index) {
	uint32_t X = 0, v = 0x80000000u;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		if ((index >> bit) & 1u) X ^= v;
		v ^= v >> 1;
	}
	return X;
}
__host__ __device__ __forceinline__ float2 hn_ld_random_val_2d(uint32_t index, uint32_t seed) {
	index = hn_owen(index, seed);
	float2 r;
	r.x = (float)hn_owen(hn_reverse_bits(index), hn_hash_combine(seed, 0u)) * 2.3283064365386963e-10f;
	r.y = (float)hn_owen(hn_sobol_dim1(index), hn_hash_combine(seed, 1u)) * 2.3283064365386963e-10f;
	return r;
}

#define HN_PI_F 3.14159265358979323846f

// ---- stepping space of the ray walk (nerf_device.cuh:316-368): linear for cone angle 0, logarithmic between the two break points otherwise
#define HN_STEPSIZE (1.73205080757f / 1024.0f)
#define HN_MAX_CONE_STEPSIZE (HN_STEPSIZE * 128 * 1024u / 128u)
__device__ __forceinline__ float to_stepping_space(float t, float cone) {
	if (cone <= 1e-5f) return t / HN_STEPSIZE;
	float log1p_c = logf(1.0f + cone);
	float a = (logf(HN_STEPSIZE) - logf(log1p_c)) / log1p_c;
	float b = (logf(HN_MAX_CONE_STEPSIZE) - logf(log1p_c)) / log1p_c;
	float at = expf(a * log1p_c), bt = expf(b * log1p_c);
	if (t <= at) return (t - at) / HN_STEPSIZE + a;
	if (t <= bt) return logf(t) / log1p_c;
	return (t - bt) / HN_MAX_CONE_STEPSIZE + b;
}
__device__ __forceinline__ float from_stepping_space(float n, float cone) {
	if (cone <= 1e-5f) return n * HN_STEPSIZE;
	float log1p_c = logf(1.0f + cone);
	float a = (logf(HN_STEPSIZE) - logf(log1p_c)) / log1p_c;
	float b = (logf(HN_MAX_CONE_STEPSIZE) - logf(log1p_c)) / log1p_c;
	float at = expf(a * log1p_c), bt = expf(b * log1p_c);
	if (n <= a) return (n - a) * HN_STEPSIZE + at;
	if (n <= b) return expf(n * log1p_c);
	return (n - b) * HN_MAX_CONE_STEPSIZE + bt;
}
__device__ __forceinline__ float advance_n_steps(float t, float cone, float n) { return from_stepping_space(to_stepping_space(t, cone) + n, cone); }
__device__ __forceinline__ float calc_dt(float t, float cone) { return advance_n_steps(t, cone, 1.0f) - t; }

// ---- camera models (common_device.cuh:249-483): lens undistortion, parallax shift, depth of field ----
// ELensMode (common.h:223-230): 0 Perspective, 1 OpenCV, 2 FTheta, 3 LatLong, 4 OpenCVFisheye, 5 Equirectangular
__device__ __forceinline__ void opencv_distortion_delta(const float* k, float u, float v, float* du, float* dv) {
	const float u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2;
	const float radial = k[0] * r2 + k[1] * r2 * r2;
	*du = u * radial + 2.0f * k[2] * uv + k[3] * (r2 + 2.0f * u2);
	*dv = v * radial + 2.0f * k[3] * uv + k[2] * (r2 + 2.0f * v2);
}
__device__ __forceinline__ void fisheye_distortion_delta(const float* k, float u, float v, float* du, float* dv) {
	const float r = sqrtf(u * u + v * v);
	if (r > (float)DBL_EPSILON) {
		const float theta = atanf(r), theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
		const float thetad = theta * (1.0f + k[0] * theta2 + k[1] * theta4 + k[2] * theta6 + k[3] * theta8);
		*du = u * thetad / r - u;
		*dv = v * thetad / r - v;
	} else { *du = 0.0f; *dv = 0.0f; }
}
// Newton iteration with central differences (iterative_lens_undistortion, common_device.cuh:281-321)
template <bool FISHEYE>
__device__ void iterative_undistortion(const float* k, float* u, float* v) {
	const float x0 = *u, y0 = *v;
	float x = x0, y = y0;
	for (uint32_t i = 0; i < 100; ++i) {
		const float step0 = fmaxf(FLT_EPSILON, fabsf(1e-6f * x)), step1 = fmaxf(FLT_EPSILON, fabsf(1e-6f * y));
		float dx, dy, bx0, by0, fx0, fy0, bx1, by1, fx1, fy1;
		if (FISHEYE) {
			fisheye_distortion_delta(k, x, y, &dx, &dy); fisheye_distortion_delta(k, x - step0, y, &bx0, &by0); fisheye_distortion_delta(k, x + step0, y, &fx0, &fy0);
			fisheye_distortion_delta(k, x, y - step1, &bx1, &by1); fisheye_distortion_delta(k, x, y + step1, &fx1, &fy1);
		} else {
			opencv_distortion_delta(k, x, y, &dx, &dy); opencv_distortion_delta(k, x - step0, y, &bx0, &by0); opencv_distortion_delta(k, x + step0, y, &fx0, &fy0);
			opencv_distortion_delta(k, x, y - step1, &bx1, &by1); opencv_distortion_delta(k, x, y + step1, &fx1, &fy1);
		}
		// J is column-major: J[c][r]
		const float J00 = 1.0f + (fx0 - bx0) / (2.0f * step0), J10 = (fx1 - bx1) / (2.0f * step1);
		const float J01 = (fy0 - by0) / (2.0f * step0), J11 = 1.0f + (fy1 - by1) / (2.0f * step1);
		const float det = J00 * J11 - J01 * J10;
		const float I00 = J11 / det, I01 = -J01 / det, I10 = -J10 / det, I11 = J00 / det; // adjoint / determinant
		const float rx = x + dx - x0, ry = y + dy - y0;
		float sx = 0.0f, sy = 0.0f;
		sx += I00 * rx; sy += I01 * rx;
		sx += I10 * ry; sy += I11 * ry;
		x -= sx; y -= sy;
		if (sx * sx + sy * sy < 1e-10f) break;
	}
	*u = x; *v = y;
}
__device__ __forceinline__ float2 square2disk_shirley(float a, float b) { // random_val.cuh:112-128
	float phi, r;
	if (a * a > b * b) { r = a; phi = (HN_PI_F / 4.0f) * (b / a); }
	else { r = b; phi = (HN_PI_F / 2.0f) - (HN_PI_F / 4.0f) * (a / b); }
	float sp, cp;
	sincosf(phi, &sp, &cp);
	return make_float2(r * cp, r * sp);
}
// ---- foveation (common_device.cuh:117-246): one axis; c = al bl cl am bm ar br cr switch_left switch_right inv_switch_left inv_switch_right
__host__ __device__ __forceinline__ float fov_warp(const float* c, float x) {
	x = fminf(fmaxf(x, 0.0f), 1.0f);
	if (x < c[8]) return c[0] * x * x + c[1] * x + c[2];
	if (x > c[9]) return c[5] * x * x + c[6] * x + c[7];
	return c[3] * x + c[4];
}
__host__ __device__ __forceinline__ float fov_unwarp(const float* c, float y) {
	y = fminf(fmaxf(y, 0.0f), 1.0f);
	if (y < c[10]) return (sqrtf(-4 * c[0] * c[2] + 4 * c[0] * y + c[1] * c[1]) - c[1]) / (2 * c[0]);
	if (y > c[11]) return (sqrtf(-4 * c[5] * c[7] + 4 * c[5] * y + c[6] * c[6]) - c[6]) / (2 * c[5]);
	return (y - c[4]) / c[3];
}
// Buffer2DView::at / at_lerp (common.h:268-305)
__device__ __forceinline__ uint8_t mask_at(const uint8_t* __restrict__ data, int w, int h, float u, float v) {
	const int px = min(max((int)((float)w * u), 0), w - 1), py = min(max((int)((float)h * v), 0), h - 1);
	return __ldg(data + px + py * w);
}
__device__ __forceinline__ float2 grid_at_lerp(const float2* __restrict__ data, int w, int h, float u, float v) {
	const float fx = (float)w * u, fy = (float)h * v;
	const int px = (int)fx, py = (int)fy;
	const float wx = fx - (float)px, wy = fy - (float)py;
	const int x0 = min(max(px, 0), w - 1), x1 = min(max(px + 1, 0), w - 1), y0 = min(max(py, 0), h - 1), y1 = min(max(py + 1, 0), h - 1);
	const float2 a = __ldg(data + x0 + y0 * w), b = __ldg(data + x1 + y0 * w), c = __ldg(data + x0 + y1 * w), d = __ldg(data + x1 + y1 * w);
	const float w00 = (1 - wx) * (1 - wy), w10 = wx * (1 - wy), w01 = (1 - wx) * wy, w11 = wx * wy;
	return make_float2(w00 * a.x + w10 * b.x + w01 * c.x + w11 * d.x, w00 * a.y + w10 * b.y + w01 * c.y + w11 * d.y);
}

// the lens models of uv_to_ray (common_device.cuh:439-463): camera-space direction of (u, v); false = invalid ray (f-theta outside its
// image circle). mode = ELensMode (0 Perspective, 1 OpenCV, 2 FTheta, 3 LatLong, 4 OpenCVFisheye, 5 Equirectangular). Shared by the
// render-time camera (camera_ray) and the per-image cameras of the training set (hn_train.cu).
__device__ __forceinline__ bool lens_direction(int mode, const float* __restrict__ lp, float u, float v, float sc0, float sc1, int w, int h, float f0, float f1, V3& d) {
	if (mode == 2) {
		const float xpix = (u - sc0) * lp[5], ypix = (v - sc1) * lp[6];
		const float norm = sqrtf(xpix * xpix + ypix * ypix);
		const float alpha = lp[0] + norm * (lp[1] + norm * (lp[2] + norm * (lp[3] + norm * lp[4])));
		float sa, ca;
		sincosf(alpha, &sa, &ca);
		if (ca <= FLT_MIN || norm == 0.f) return false;
		sa *= 1.f / norm;
		d = mk(sa * xpix, sa * ypix, ca);
		if (d.x == 0.0f && d.y == 0.0f && d.z == 0.0f) return false;
	} else if (mode == 3) {
		const float theta = (v - 0.5f) * HN_PI_F, phi = (u - 0.5f) * HN_PI_F * 2.0f;
		float sp, cp, st, ct;
		sincosf(theta, &st, &ct); sincosf(phi, &sp, &cp);
		d = mk(sp * ct, st, cp * ct);
	} else if (mode == 5) {
		const float ct = (v - 0.5f) * 2.0f, st = sqrtf(fmaxf(1.0f - ct * ct, 0.0f)), phi = (u - 0.5f) * HN_PI_F * 2.0f;
		float sp, cp;
		sincosf(phi, &sp, &cp);
		d = mk(sp * st, ct, cp * st);
	} else {
		d = mk((u - sc0) * (float)w / f0, (v - sc1) * (float)h / f1, 1.0f);
		if (mode == 1) iterative_undistortion<false>(lp, &d.x, &d.y);
		else if (mode == 4) iterative_undistortion<true>(lp, &d.x, &d.y);
	}
	return true;
}

// uv_to_ray (common_device.cuh:416-483). c0..c2 = camera rotation columns, pos = camera[3]. `use_lens` is false on the mesh pass
// (its ray init takes foveation and the hidden-area mask but neither Lens nor distortion grid, geometry :520-536).
// Returns false for an invalid ray (f-theta outside the image circle, or a pixel the mask hides). The direction is NOT normalised.
__device__ __forceinline__ bool camera_ray(const FrameArgs& a, bool use_lens, float u_in, float v_in, V3 c0, V3 c1, V3 c2, V3 pos, V3& origin, V3& dir) {
	V3 d;
	const int mode = use_lens ? a.lens_mode : 0;
	float u = u_in, v = v_in; // warped_uv; the depth-of-field pixel hash below keeps the un-warped uv (:470)
	if (a.foveated) { u = fov_warp(a.fov[0], u_in); v = fov_warp(a.fov[1], v_in); }
	if (a.hidden_mask && !mask_at(a.hidden_mask, a.hidden_w, a.hidden_h, u, v)) return false;
	if (!lens_direction(mode, a.lens_params, u, v, a.screen_center[0], a.screen_center[1], a.w, a.h, a.focal[0], a.focal[1], d)) return false;
	if (use_lens && a.distortion) { const float2 g = grid_at_lerp(a.distortion, a.dist_w, a.dist_h, u, v); d.x += g.x; d.y += g.y; }
	const V3 head = mk(a.parallax_shift[0], a.parallax_shift[1], 0.0f);
	d = d - head * a.parallax_shift[2];
	V3 r = mk(0.0f, 0.0f, 0.0f);
	r.x += c0.x * d.x; r.y += c0.y * d.x; r.z += c0.z * d.x;
	r.x += c1.x * d.y; r.y += c1.y * d.y; r.z += c1.z * d.y;
	r.x += c2.x * d.z; r.y += c2.y * d.z; r.z += c2.z * d.z;
	d = r;
	V3 o = mk(0.0f, 0.0f, 0.0f);
	o.x += c0.x * head.x; o.y += c0.y * head.x; o.z += c0.z * head.x;
	o.x += c1.x * head.y; o.y += c1.y * head.y; o.z += c1.z * head.y;
	o.x += c2.x * head.z; o.y += c2.y * head.z; o.z += c2.z * head.z;
	o = o + pos;
	if (a.aperture_size != 0.0f) {
		const V3 lookat = o + d * a.focus_z;
		const int px = (int)(u_in * (float)a.w), py = (int)(v_in * (float)a.h);
		const float2 rnd = hn_ld_random_val_2d(a.spp, (uint32_t)px * 19349663u + (uint32_t)py * 96925573u);
		const float2 disk = square2disk_shirley(rnd.x * 2.0f - 1.0f, rnd.y * 2.0f - 1.0f);
		const float bx = a.aperture_size * disk.x, by = a.aperture_size * disk.y;
		V3 sh = mk(0.0f, 0.0f, 0.0f);
		sh.x += c0.x * bx; sh.y += c0.y * bx; sh.z += c0.z * bx;
		sh.x += c1.x * by; sh.y += c1.y * by; sh.z += c1.z * by;
		o = o + sh;
		d = (lookat - o) / a.focus_z;
	}
	origin = o + d * a.near_distance;
	dir = d;
	return true;
}

// pos_to_uv (common_device.cuh:505-537) with a perspective lens, as ERenderMode::Distortion calls it (init_rays :1513-1515);
// inverse(mat3) = adjoint / determinant in tcnn's operation order (vec.h:771-829,861-864)
__device__ __forceinline__ float det2(float a, float b, float c, float d) { return a * d - b * c; }
__device__ __forceinline__ float2 pos_to_uv(const FrameArgs& a, V3 p, V3 c0, V3 c1, V3 c2, V3 cpos) {
	const V3 head = mk(a.parallax_shift[0], a.parallax_shift[1], 0.0f);
	V3 o = mk(0.0f, 0.0f, 0.0f);
	o.x += c0.x * head.x; o.y += c0.y * head.x; o.z += c0.z * head.x;
	o.x += c1.x * head.y; o.y += c1.y * head.y; o.z += c1.z * head.y;
	o.x += c2.x * head.z; o.y += c2.y * head.z; o.z += c2.z * head.z;
	o = o + cpos;
	const V3 d = p - o;
	// m(c, r): column c, row r
	const float m00_ = c0.x, m01_ = c0.y, m02_ = c0.z, m10_ = c1.x, m11_ = c1.y, m12_ = c1.z, m20_ = c2.x, m21_ = c2.y, m22_ = c2.z;
	const float det = m00_ * (m11_ * m22_ - m21_ * m12_) + -m10_ * (m01_ * m22_ - m21_ * m02_) + m20_ * (m01_ * m12_ - m11_ * m02_);
	const float a00 = det2(m11_, m21_, m12_, m22_), a01 = det2(m01_, m21_, m02_, m22_), a02 = det2(m01_, m11_, m02_, m12_);
	const float a10 = det2(m10_, m20_, m12_, m22_), a11 = det2(m00_, m20_, m02_, m22_), a12 = det2(m00_, m10_, m02_, m12_);
	const float a20 = det2(m10_, m20_, m11_, m21_), a21 = det2(m00_, m20_, m01_, m21_), a22 = det2(m00_, m10_, m01_, m11_);
	const V3 i0 = mk(a00, -a01, a02) / det, i1 = mk(-a10, a11, -a12) / det, i2 = mk(a20, -a21, a22) / det;
	V3 r = mk(0.0f, 0.0f, 0.0f);
	r.x += i0.x * d.x; r.y += i0.y * d.x; r.z += i0.z * d.x;
	r.x += i1.x * d.y; r.y += i1.y * d.y; r.z += i1.z * d.y;
	r.x += i2.x * d.z; r.y += i2.y * d.z; r.z += i2.z * d.z;
	r = r / r.z;
	r = r + head * a.parallax_shift[2];
	float u = r.x * a.focal[0] / (float)a.w + a.screen_center[0], v = r.y * a.focal[1] / (float)a.h + a.screen_center[1];
	if (a.foveated) { u = fov_unwarp(a.fov[0], u); v = fov_unwarp(a.fov[1], v); }
	return make_float2(u, v);
}
// hsv_to_rgb / to_rgb (common_device.cuh:713-739)
__device__ __forceinline__ V3 dir_to_rgb(float dx, float dy) {
	float h = atan2f(dy, dx) / (2.0f * HN_PI_F) + 0.5f;
	const float v = sqrtf(dx * dx + dy * dy); // s = 1
	h = fmodf(h, 1.0f) * 6.0f;
	const int i = (int)h;
	const float f = h - (float)i;
	const float p = v * (1.0f - 1.0f), q = v * (1.0f - 1.0f * f), t = v * (1.0f - 1.0f * (1.0f - f));
	switch (i) {
		case 0: return mk(v, t, p);
		case 1: return mk(q, v, p);
		case 2: return mk(p, v, t);
		case 3: return mk(p, q, v);
		case 4: return mk(t, p, v);
		default: return mk(v, p, q);
	}
}

// ---- a-16: environment map + light-direction sampling (random_val.cuh:62-72,130-185, envmap.cuh:24-50) ----
__device__ __forceinline__ float2 dir_to_spherical_unorm(V3 d) {
	const float cos_theta = fminf(fmaxf(d.z, -1.0f), 1.0f);
	const float theta = acosf(cos_theta);
	const float phi = atan2f(d.y, d.x);
	return make_float2(theta / HN_PI_F, (phi / (2.0f * HN_PI_F) + 0.5f));
}
__device__ __forceinline__ V3 cosine_hemisphere(float2 u) {
	const float r = sqrtf(u.x);
	const float phi = 2.0f * HN_PI_F * u.y;
	V3 p;
	p.x = r * cosf(phi);
	p.y = r * sinf(phi);
	p.z = sqrtf(fmaxf(0.0f, 1.0f - p.x * p.x - p.y * p.y));
	return p;
}
// compute_local_frame (random_val.cuh:167-185) applied to a local vector: columns X, Y, Z = n
__device__ __forceinline__ V3 local_frame_apply(V3 n, V3 l) {
	const float sz = (n.z >= 0) ? 1.0f : -1.0f;
	const float a = 1 / (sz + n.z);
	const float ya = n.y * a;
	const float b = n.x * ya;
	const float c = n.x * sz;
	const V3 X = mk(c * n.x * a - 1, sz * b, c), Y = mk(b, n.y * ya - sz, n.y);
	V3 r = mk(0.0f, 0.0f, 0.0f);
	r.x += X.x * l.x; r.y += X.y * l.x; r.z += X.z * l.x;
	r.x += Y.x * l.y; r.y += Y.y * l.y; r.z += Y.z * l.y;
	r.x += n.x * l.z; r.y += n.y * l.z; r.z += n.z * l.z;
	return r;
}
// light direction k of K of pixel idx (OUR definition, oracle env_light_dir)
__device__ __forceinline__ V3 env_light_dir(uint32_t idx, uint32_t spp, uint32_t K, uint32_t k, V3 n) {
	const float2 u = hn_ld_random_val_2d(k + K * spp, idx * 786433u + 0x9E3779B9u);
	return normalize3(local_frame_apply(n, cosine_hemisphere(u)));
}
// ---- a-16 irradiance probes: the pieces the reference ships (src/testbed_nerf.cu:1546-1557, random_val.cuh:130-160,338-350) ----
__device__ __forceinline__ V3 cylindrical_to_dir(float px, float py) {
	const float cos_theta = -2.0f * px + 1.0f;
	const float phi = 2.0f * HN_PI_F * (py - 0.5f);
	const float sin_theta = sqrtf(fmaxf(1.0f - cos_theta * cos_theta, 0.0f));
	float sin_phi, cos_phi;
	sincosf(phi, &sin_phi, &cos_phi);
	return mk(sin_theta * cos_phi, sin_theta * sin_phi, cos_theta);
}
__device__ __forceinline__ float halton(uint32_t base, uint32_t idx) {
	float f = 1, result = 0;
	while (idx > 0) {
		f /= (float)base;
		result += f * (float)(idx % base);
		idx /= base;
	}
	return result;
}
__device__ __forceinline__ V3 spherical_cap_uniform(float2 u, float cap_height) {
	const float z = (float)((double)(1.0f - cap_height) * (1.0 - (double)u.x) + (double)u.x); // the reference mixes a double literal in (random_val.cuh:147)
	const float r = sqrtf(fmaxf(0.0f, 1.0f - z * z));
	const float phi = 2.0f * HN_PI_F * u.y;
	return mk(r * cosf(phi), r * sinf(phi), z);
}
// Probe texture (OUR definition; the reference only declares m_envmap_tex / computeEnvmap*): texel (i, j) holds the radiance of probe
// ray idx = i + n_theta * j; lookup by direction = the inverse of cylindrical_to_dir, nearest texel, phi wraps, theta clamps.
__device__ __forceinline__ float4 read_probe(const float4* __restrict__ tex, int n_theta, int n_phi, V3 dir) {
	const float px = (1.0f - fminf(fmaxf(dir.z, -1.0f), 1.0f)) * 0.5f;
	const float py = atan2f(dir.y, dir.x) / (2.0f * HN_PI_F) + 0.5f;
	int i = (int)floorf(px * (float)n_theta), j = (int)floorf(py * (float)n_phi);
	i = max(0, min(i, n_theta - 1));
	j = ((j % n_phi) + n_phi) % n_phi;
	return __ldg(tex + i + (size_t)n_theta * j);
}
// Light direction k of K of pixel idx towards a spherical cap (axis, cap_height) and its estimator weight for a Lambertian
// surface with normal n: cos / (pi pdf) with pdf = 1 / (2 pi cap_height) (warp_square_to_spherical_cap_uniform_invpdf)
__device__ __forceinline__ V3 cap_light_dir(uint32_t idx, uint32_t spp, uint32_t K, uint32_t k, V3 axis, float cap_height, V3 n, float& weight) {
	const float2 u = hn_ld_random_val_2d(k + K * spp, idx * 786433u + 0x9E3779B9u);
	const V3 w = normalize3(local_frame_apply(axis, spherical_cap_uniform(u, cap_height)));
	weight = 2.0f * cap_height * fmaxf(0.0f, dot3(n, w));
	return w;
}

__device__ __forceinline__ float4 read_envmap(const float4* __restrict__ env, int w, int h, V3 dir) {
	const float2 cyl = dir_to_spherical_unorm(mk(dir.z, -dir.x, dir.y));
	const float fx = cyl.y * (float)(w - 1), fy = cyl.x * (float)(h - 1);
	const int tx = (int)fx, ty = (int)fy;
	const float wx = fx - (float)tx, wy = fy - (float)ty;
	int x0 = tx, x1 = tx + 1;
	if (x0 < 0) x0 += w; else if (x0 >= w) x0 -= w;
	if (x1 < 0) x1 += w; else if (x1 >= w) x1 -= w;
	const int y0 = max(min(ty, h - 1), 0), y1 = max(min(ty + 1, h - 1), 0);
	const float4 a = __ldg(env + x0 + (size_t)y0 * w), b = __ldg(env + x1 + (size_t)y0 * w);
	const float4 c = __ldg(env + x0 + (size_t)y1 * w), d = __ldg(env + x1 + (size_t)y1 * w);
	const float w00 = (1 - wx) * (1 - wy), w10 = wx * (1 - wy), w01 = (1 - wx) * wy, w11 = wx * wy;
	return make_float4(w00 * a.x + w10 * b.x + w01 * c.x + w11 * d.x, w00 * a.y + w10 * b.y + w01 * c.y + w11 * d.y,
		w00 * a.z + w10 * b.z + w01 * c.z + w11 * d.z, w00 * a.w + w10 * b.w + w01 * c.w + w11 * d.w);
}

} // namespace hn
