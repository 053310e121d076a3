// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h). NeRF half of the path: Sobol jitter, ray generation,
// occupancy DDA, hash grid, SH, MLPs, compositing, final blend. Build with -ffp-contract=off (no FMA contraction),
// every function cites the reference file:line it restates (paths relative to /root/reference).
#include "orc_common.h"
#include "orc_api.h"
#include <thread>
#include <atomic>

namespace orc {

// ---------------------------------------------------------------- random_val.cuh:207-370
// Sobol direction numbers, dimensions 0 and 1 (the only ones the path consumes): dim 0 is the identity
// (bit-reversal) sequence, dim 1 follows v_j = v_{j-1} ^ (v_{j-1} >> 1) (Joe & Kuo primitive polynomial x+1).
static uint32_t sobol_dir(uint32_t dim, uint32_t bit) {
	if (dim == 0) return 0x80000000u >> bit;
	uint32_t v = 0x80000000u;
	for (uint32_t i = 0; i < bit; ++i) v ^= v >> 1;
	return v;
}
static uint32_t sobol(uint32_t index, uint32_t dim) {
	uint32_t X = 0;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		if ((index >> bit) & 1) X ^= sobol_dir(dim, bit);
	}
	return X;
}
static uint32_t hash_combine(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
static uint32_t reverse_bits(uint32_t x) {
	x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
	x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
	x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
	x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
	return (x >> 16) | (x << 16);
}
static uint32_t laine_karras(uint32_t x, uint32_t seed) {
	x += seed;
	x ^= x * 0x6c50b47cu;
	x ^= x * 0xb82f1e52u;
	x ^= x * 0xc7afe638u;
	x ^= x * 0x8d22f6e6u;
	return x;
}
static uint32_t owen_scramble(uint32_t x, uint32_t seed) { return reverse_bits(laine_karras(reverse_bits(x), seed)); }
static const float TWO_M32 = (float)(1.0 / 4294967296.0);
float ld_random_val(uint32_t index, uint32_t seed, uint32_t dim = 0) {
	index = owen_scramble(index, seed);
	return (float)owen_scramble(sobol(index, dim), hash_combine(seed, dim)) * TWO_M32;
}
vec2 ld_random_val_2d(uint32_t index, uint32_t seed) {
	index = owen_scramble(index, seed);
	uint32_t x = owen_scramble(sobol(index, 0), hash_combine(seed, 0));
	uint32_t y = owen_scramble(sobol(index, 1), hash_combine(seed, 1));
	return {(float)x * TWO_M32, (float)y * TWO_M32};
}
static float fractf(float x) { return x - std::floor(x); }
vec2 ld_random_pixel_offset(uint32_t spp) {
	vec2 a = ld_random_val_2d(0, 0xdeadbeefu), b = ld_random_val_2d(spp, 0xdeadbeefu);
	return {fractf(0.5f - a.x + b.x), fractf(0.5f - a.y + b.y)};
}

// ---------------------------------------------------------------- tcnn vec.h:1060-1187 (quaternion slerp of the camera)
struct quat { float w, x, y, z; };
static quat quat_from(const mat3& m) {
	quat q;
	float tr = m.c[0].x + m.c[1].y + m.c[2].z;
	if (tr > 0.0f) {
		float S = std::sqrt(tr + 1.0f) * 2.0f;
		q.w = 0.25f * S;
		q.x = (m.c[1].z - m.c[2].y) / S;
		q.y = (m.c[2].x - m.c[0].z) / S;
		q.z = (m.c[0].y - m.c[1].x) / S;
	} else if (m.c[0].x > m.c[1].y && m.c[0].x > m.c[2].z) {
		float S = std::sqrt(1.0f + m.c[0].x - m.c[1].y - m.c[2].z) * 2.0f;
		q.w = (m.c[1].z - m.c[2].y) / S;
		q.x = 0.25f * S;
		q.y = (m.c[1].x + m.c[0].y) / S;
		q.z = (m.c[2].x + m.c[0].z) / S;
	} else if (m.c[1].y > m.c[2].z) {
		float S = std::sqrt(1.0f + m.c[1].y - m.c[0].x - m.c[2].z) * 2.0f;
		q.w = (m.c[2].x - m.c[0].z) / S;
		q.x = (m.c[1].x + m.c[0].y) / S;
		q.y = 0.25f * S;
		q.z = (m.c[2].y + m.c[1].z) / S;
	} else {
		float S = std::sqrt(1.0f + m.c[2].z - m.c[0].x - m.c[1].y) * 2.0f;
		q.w = (m.c[0].y - m.c[1].x) / S;
		q.x = (m.c[2].x + m.c[0].z) / S;
		q.y = (m.c[2].y + m.c[1].z) / S;
		q.z = 0.25f * S;
	}
	return q;
}
static float qdot(quat a, quat b) { return (a.w * b.w + a.x * b.x) + (a.y * b.y + a.z * b.z); }
static quat qnormalize(quat a) {
	float len = std::sqrt(qdot(a, a));
	if (len <= 0.0f) return {1.0f, 0.0f, 0.0f, 0.0f};
	return {a.w / len, a.x / len, a.y / len, a.z / len};
}
static quat qscale(quat a, float s) { return {a.w * s, a.x * s, a.y * s, a.z * s}; }
static quat qadd(quat a, quat b) { return {a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z}; }
static quat qslerp(quat x, quat y, float t) {
	quat z = y;
	float cos_theta = qdot(x, y);
	if (cos_theta < 0.0f) {
		z = {-y.w, -y.x, -y.y, -y.z};
		cos_theta = -cos_theta;
	}
	if (cos_theta > 1.0f - FLT_EPSILON) {
		return qadd(qscale(x, 1.0f - t), qscale(z, t));
	}
	float angle = std::acos(cos_theta);
	quat r = qadd(qscale(x, std::sin((1.0f - t) * angle)), qscale(z, std::sin(t * angle)));
	float s = std::sin(angle);
	return {r.w / s, r.x / s, r.y / s, r.z / s};
}
static mat3 quat_to_mat3(quat q) {
	float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z;
	float qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
	float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
	mat3 m;
	m.c[0] = {1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy)};
	m.c[1] = {2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx)};
	m.c[2] = {2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy)};
	return m;
}
// common_device.cuh:651-659
static cam_t camera_slerp(const cam_t& a, const cam_t& b, float t) {
	mat3 ma = {{a.c[0], a.c[1], a.c[2]}}, mb = {{b.c[0], b.c[1], b.c[2]}};
	mat3 rot = quat_to_mat3(qnormalize(qslerp(qnormalize(quat_from(ma)), qnormalize(quat_from(mb)), t)));
	cam_t r;
	r.c[0] = rot.c[0]; r.c[1] = rot.c[1]; r.c[2] = rot.c[2];
	r.c[3] = mix(a.c[3], b.c[3], t);
	return r;
}

static constexpr float PI_F = 3.14159265358979323846f;
// common_device.cuh:249-321: OpenCV / OpenCV-fisheye distortion deltas and their Newton inversion
static void opencv_distortion_delta(const float* k, float u, float v, float* du, float* dv) {
	float u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2;
	float radial = k[0] * r2 + k[1] * r2 * r2;
	*du = u * radial + 2.0f * k[2] * uv + k[3] * (r2 + 2.0f * u2);
	*dv = v * radial + 2.0f * k[3] * uv + k[2] * (r2 + 2.0f * v2);
}
static void fisheye_distortion_delta(const float* k, float u, float v, float* du, float* dv) {
	float r = std::sqrt(u * u + v * v);
	if (r > (float)std::numeric_limits<double>::epsilon()) {
		float theta = std::atan(r), theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
		float thetad = theta * (1.0f + k[0] * theta2 + k[1] * theta4 + k[2] * theta6 + k[3] * theta8);
		*du = u * thetad / r - u;
		*dv = v * thetad / r - v;
	} else { *du = 0.0f; *dv = 0.0f; }
}
typedef void (*distortion_fn)(const float*, float, float, float*, float*);
static void iterative_undistortion(const float* k, float* u, float* v, distortion_fn fn) {
	const float x0 = *u, y0 = *v;
	float x = x0, y = y0;
	for (uint32_t i = 0; i < 100; ++i) {
		float step0 = std::max(std::numeric_limits<float>::epsilon(), std::fabs(1e-6f * x));
		float step1 = std::max(std::numeric_limits<float>::epsilon(), std::fabs(1e-6f * y));
		float dx, dy, bx0, by0, fx0, fy0, bx1, by1, fx1, fy1;
		fn(k, x, y, &dx, &dy); fn(k, x - step0, y, &bx0, &by0); fn(k, x + step0, y, &fx0, &fy0); fn(k, x, y - step1, &bx1, &by1); fn(k, x, y + step1, &fx1, &fy1);
		float J00 = 1.0f + (fx0 - bx0) / (2.0f * step0), J10 = (fx1 - bx1) / (2.0f * step1);
		float J01 = (fy0 - by0) / (2.0f * step0), J11 = 1.0f + (fy1 - by1) / (2.0f * step1);
		float det = J00 * J11 - J01 * J10; // tcnn inverse = adjoint / determinant (vec.h:803-808,861-863)
		float I00 = J11 / det, I01 = -J01 / det, I10 = -J10 / det, I11 = J00 / det;
		float rx = x + dx - x0, ry = y + dy - y0;
		float sx = 0.0f, sy = 0.0f;
		sx += I00 * rx; sy += I01 * rx;
		sx += I10 * ry; sy += I11 * ry;
		x -= sx; y -= sy;
		if (sx * sx + sy * sy < 1e-10f) break;
	}
	*u = x; *v = y;
}
// random_val.cuh:112-128
static vec2 square2disk_shirley(vec2 sq) {
	float phi, r, a = sq.x, b = sq.y;
	if (a * a > b * b) { r = a; phi = (PI_F / 4.0f) * (b / a); }
	else { r = b; phi = (PI_F / 2.0f) - (PI_F / 4.0f) * (a / b); }
	return {r * std::cos(phi), r * std::sin(phi)};
}

#include "orc_fov.inc"
static vec2 pos_to_uv(vec3 pos, int w, int h, vec2 focal, const cam_t& cam, vec2 sc, const float parallax[3], const Fov& fov, int lens_mode, const float* lp) {
	vec3 head_pos = {parallax[0], parallax[1], 0.0f};
	mat3 m = {{cam.c[0], cam.c[1], cam.c[2]}};
	vec3 origin = mul(m, head_pos) + cam.c[3];
	vec3 dir = pos - origin;
	dir = mul(inverse3(m), dir);
	dir = dir / dir.z;
	dir = dir + head_pos * parallax[2];
	float du = 0.0f, dv = 0.0f;
	if (lens_mode == 1) opencv_distortion_delta(lp, dir.x, dir.y, &du, &dv);
	else if (lens_mode == 4) fisheye_distortion_delta(lp, dir.x, dir.y, &du, &dv);
	dir.x += du; dir.y += dv;
	vec2 uv = {dir.x * focal.x / (float)w + sc.x, dir.y * focal.y / (float)h + sc.y};
	return fov.unwarp(uv);
}

// common_device.cuh:416-483 uv_to_ray - direction not normalised; valid = false for an f-theta pixel outside the image circle or a
// pixel the hidden-area mask hides. lens_mode follows ELensMode (common.h:223-230). The mesh pass (use_lens = false) takes
// foveation and the mask but neither Lens nor distortion grid (src/testbed_geometry_training.cu:520-536).
Ray uv_to_ray(const orc_render_params& p, uint32_t spp, vec2 uv_in, const cam_t& cam, bool use_lens, bool* valid) {
	const int w = p.width, h = p.height;
	const vec2 focal = {p.focal[0], p.focal[1]}, sc = {p.screen_center[0], p.screen_center[1]};
	const int mode = use_lens ? p.lens_mode : 0;
	const float* lp = p.lens_params;
	if (valid) *valid = true;
	const vec2 uv = Fov(p.foveation).warp(uv_in); // warped_uv; the depth-of-field pixel hash below keeps the un-warped uv (:470)
	if (p.hidden_area_mask && !mask_at(p.hidden_area_mask, p.hidden_w, p.hidden_h, uv)) { if (valid) *valid = false; return {{0, 0, 0}, {0, 0, 0}}; }
	vec3 dir;
	if (mode == 2) { // f_theta_undistortion :360-374
		float xpix = (uv.x - sc.x) * lp[5], ypix = (uv.y - sc.y) * lp[6];
		float norm = std::sqrt(xpix * xpix + ypix * ypix);
		float alpha = lp[0] + norm * (lp[1] + norm * (lp[2] + norm * (lp[3] + norm * lp[4])));
		float sa = std::sin(alpha), ca = std::cos(alpha);
		if (ca <= std::numeric_limits<float>::min() || norm == 0.f) { if (valid) *valid = false; return {{0, 0, 0}, {0, 0, 0}}; }
		sa *= 1.f / norm;
		dir = {sa * xpix, sa * ypix, ca};
		if (dir.x == 0.0f && dir.y == 0.0f && dir.z == 0.0f) { if (valid) *valid = false; return {{0, 0, 0}, {0, 0, 0}}; }
	} else if (mode == 3) { // latlong_to_dir :376-383
		float theta = (uv.y - 0.5f) * PI_F, phi = (uv.x - 0.5f) * PI_F * 2.0f;
		float st = std::sin(theta), ct = std::cos(theta), sp = std::sin(phi), cp = std::cos(phi);
		dir = {sp * ct, st, cp * ct};
	} else if (mode == 5) { // equirectangular_to_dir :385-392
		float ct = (uv.y - 0.5f) * 2.0f, st = std::sqrt(std::max(1.0f - ct * ct, 0.0f)), phi = (uv.x - 0.5f) * PI_F * 2.0f;
		float sp = std::sin(phi), cp = std::cos(phi);
		dir = {sp * st, ct, cp * st};
	} else {
		dir = {(uv.x - sc.x) * (float)w / focal.x, (uv.y - sc.y) * (float)h / focal.y, 1.0f};
		if (mode == 1) iterative_undistortion(lp, &dir.x, &dir.y, opencv_distortion_delta);
		else if (mode == 4) iterative_undistortion(lp, &dir.x, &dir.y, fisheye_distortion_delta);
	}
	if (use_lens && p.distortion) { vec2 dd = grid_at_lerp(p.distortion, p.distortion_w, p.distortion_h, uv); dir.x += dd.x; dir.y += dd.y; }
	vec3 head_pos = {p.parallax_shift[0], p.parallax_shift[1], 0.0f};
	dir = dir - head_pos * p.parallax_shift[2];
	mat3 m = {{cam.c[0], cam.c[1], cam.c[2]}};
	dir = mul(m, dir);
	vec3 origin = mul(m, head_pos) + cam.c[3];
	if (p.aperture_size != 0.0f) {
		vec3 lookat = origin + dir * p.focus_z;
		int px = (int)(uv_in.x * (float)w), py = (int)(uv_in.y * (float)h);
		vec2 rnd = ld_random_val_2d(spp, (uint32_t)px * 19349663u + (uint32_t)py * 96925573u);
		vec2 disk = square2disk_shirley({rnd.x * 2.0f - 1.0f, rnd.y * 2.0f - 1.0f});
		float bx = p.aperture_size * disk.x, by = p.aperture_size * disk.y;
		vec3 sh = {0.0f, 0.0f, 0.0f};
		sh.x += cam.c[0].x * bx; sh.y += cam.c[0].y * bx; sh.z += cam.c[0].z * bx;
		sh.x += cam.c[1].x * by; sh.y += cam.c[1].y * by; sh.z += cam.c[1].z * by;
		origin = origin + sh;
		dir = (lookat - origin) / p.focus_z;
	}
	origin = origin + dir * p.near_distance;
	return {origin, dir};
}

// testbed_nerf.cu:1428-1476: uv + per-pixel rolling-shutter camera
static void pixel_uv_cam(const orc_render_params& p, uint32_t x, uint32_t y, vec2& uv, cam_t& cam) {
	uint32_t idx = x + (uint32_t)p.width * y;
	vec2 off = ld_random_pixel_offset(p.snap_to_pixel_centers ? 0 : p.spp);
	uv = {((float)x + off.x) / (float)p.width, ((float)y + off.y) / (float)p.height};
	float mb = ld_random_val(p.spp, idx * 72239731u);
	const float* rs = p.rolling_shutter;
	float pixel_t = rs[0] + rs[1] * uv.x + rs[2] * uv.y + rs[3] * mb;
	cam = camera_slerp(cam_from(p.cam0), cam_from(p.cam1), pixel_t);
}
Ray pixel_ray(const orc_render_params& p, uint32_t x, uint32_t y, bool* valid) {
	vec2 uv; cam_t cam;
	pixel_uv_cam(p, x, y, uv, cam);
	return uv_to_ray(p, p.spp, uv, cam, true, valid);
}

// ---------------------------------------------------------------- nerf_device.cuh:316-534 (occupancy DDA)
static uint32_t expand_bits(uint32_t v) {
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}
uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) { return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2); }
static uint32_t morton3D_invert(uint32_t x) {
	x = x & 0x49249249u;
	x = (x | (x >> 2)) & 0xc30c30c3u;
	x = (x | (x >> 4)) & 0x0f00f00fu;
	x = (x | (x >> 8)) & 0xff0000ffu;
	x = (x | (x >> 16)) & 0x0000ffffu;
	return x;
}
static uint32_t cascaded_grid_idx_at(vec3 pos, uint32_t mip) {
	float mip_scale = std::scalbnf(1.0f, -(int)mip);
	pos = pos - 0.5f;
	pos = pos * mip_scale;
	pos = pos + 0.5f;
	int ix = (int)(pos.x * (float)NERF_GRIDSIZE), iy = (int)(pos.y * (float)NERF_GRIDSIZE), iz = (int)(pos.z * (float)NERF_GRIDSIZE);
	if (ix < 0 || ix >= (int)NERF_GRIDSIZE || iy < 0 || iy >= (int)NERF_GRIDSIZE || iz < 0 || iz >= (int)NERF_GRIDSIZE) return 0xFFFFFFFFu;
	return morton3D(ix, iy, iz);
}
static bool occupied_at(vec3 pos, const uint8_t* bits, uint32_t mip) {
	uint32_t idx = cascaded_grid_idx_at(pos, mip);
	if (idx == 0xFFFFFFFFu) return false;
	return bits[idx / 8 + (NERF_GRID_N_CELLS * mip) / 8] & (1 << (idx % 8));
}
static float signf(float a) { return std::copysign(1.0f, a); }
static float distance_to_next_voxel(vec3 pos, vec3 dir, vec3 idir, float res) {
	vec3 p = res * (pos - 0.5f);
	float tx = (std::floor(p.x + 0.5f + 0.5f * signf(dir.x)) - p.x) * idir.x;
	float ty = (std::floor(p.y + 0.5f + 0.5f * signf(dir.y)) - p.y) * idir.y;
	float tz = (std::floor(p.z + 0.5f + 0.5f * signf(dir.z)) - p.z) * idir.z;
	float t = std::min(std::min(tx, ty), tz);
	return std::fmax(t / res, 0.0f);
}
static float to_stepping_space(float t, float cone_angle) {
	if (cone_angle <= 1e-5f) return t / MIN_CONE_STEPSIZE;
	float log1p_c = std::log(1.0f + cone_angle);
	float a = (std::log(MIN_CONE_STEPSIZE) - std::log(log1p_c)) / log1p_c;
	float b = (std::log(MAX_CONE_STEPSIZE) - std::log(log1p_c)) / log1p_c;
	float at = std::exp(a * log1p_c), bt = std::exp(b * log1p_c);
	if (t <= at) return (t - at) / MIN_CONE_STEPSIZE + a;
	if (t <= bt) return std::log(t) / log1p_c;
	return (t - bt) / MAX_CONE_STEPSIZE + b;
}
static float from_stepping_space(float n, float cone_angle) {
	if (cone_angle <= 1e-5f) return n * MIN_CONE_STEPSIZE;
	float log1p_c = std::log(1.0f + cone_angle);
	float a = (std::log(MIN_CONE_STEPSIZE) - std::log(log1p_c)) / log1p_c;
	float b = (std::log(MAX_CONE_STEPSIZE) - std::log(log1p_c)) / log1p_c;
	float at = std::exp(a * log1p_c), bt = std::exp(b * log1p_c);
	if (n <= a) return (n - a) * MIN_CONE_STEPSIZE + at;
	if (n <= b) return std::exp(n * log1p_c);
	return (n - b) * MAX_CONE_STEPSIZE + bt;
}
static float advance_n_steps(float t, float cone_angle, float n) { return from_stepping_space(to_stepping_space(t, cone_angle) + n, cone_angle); }
float calc_dt(float t, float cone_angle) { return advance_n_steps(t, cone_angle, 1.0f) - t; }
static float advance_to_next_voxel(float t, float cone_angle, vec3 pos, vec3 dir, vec3 idir, uint32_t mip) {
	float res = std::scalbnf((float)NERF_GRIDSIZE, -(int)mip);
	float t_target = t + distance_to_next_voxel(pos, dir, idir, res);
	t = to_stepping_space(t, cone_angle);
	t_target = to_stepping_space(t_target, cone_angle);
	return from_stepping_space(t + std::ceil(std::fmax(t_target - t, 0.5f)), cone_angle);
}
static uint32_t mip_from_pos(vec3 pos, uint32_t max_cascade = NERF_CASCADES - 1) {
	int exponent;
	float maxval = std::max(std::max(std::fabs(pos.x - 0.5f), std::fabs(pos.y - 0.5f)), std::fabs(pos.z - 0.5f));
	std::frexp(maxval, &exponent);
	return (uint32_t)std::min(std::max(exponent + 1, 0), (int)max_cascade);
}
// nerf_device.cuh:461-495 (iter_cap == 0) and :497-534 (iter_cap == 200, loop runs for i in [1, cap))
float advance_to_occupied(float t, float cone_angle, const Ray& ray, vec3 idir, const uint8_t* bits, uint32_t min_mip, uint32_t max_mip,
	const BoundingBox& aabb, int iter_cap, const mat3* to_local) {
	for (int i = 1; iter_cap <= 0 || i < iter_cap; ++i) {
		vec3 pos = ray.at(t);
		if (t >= MAX_DEPTH || !aabb.contains(to_local ? mul(*to_local, pos) : pos)) return MAX_DEPTH; // aabb_to_local * pos (nerf_device.cuh:470)
		uint32_t mip = std::min(std::max(mip_from_pos(pos), min_mip), max_mip);
		if (!bits || occupied_at(pos, bits, mip)) return t;
		while (mip < max_mip && !occupied_at(pos, bits, mip + 1)) ++mip;
		t = advance_to_next_voxel(t, cone_angle, pos, ray.d, idir, mip);
	}
	return MAX_DEPTH;
}

// m_render_aabb_to_local of the frame (nullptr = axis-aligned crop box)
static const mat3* crop_matrix(const orc_render_params& p, mat3& store) {
	if (!p.crop_rotated) return nullptr;
	const float* R = p.render_aabb_to_local;
	store = {{{R[0], R[1], R[2]}, {R[3], R[4], R[5]}, {R[6], R[7], R[8]}}};
	return &store;
}

// ---------------------------------------------------------------- testbed_nerf.cu:284-331, 2863-2880
float bitfield_from_grid(const float* grid, uint32_t n_cascades, uint8_t* out) {
	// mean of max(v,0) over cascade 0 (reduce_sum; summation order: the oracle uses double, the threshold only
	// matters when it drops below 0.01 — tests use grids whose mean is well above or feed 0/1 grids).
	double sum = 0.0;
	for (uint32_t i = 0; i < NERF_GRID_N_CELLS; ++i) sum += std::max(grid[i], 0.0f);
	float mean = (float)(sum / NERF_GRID_N_CELLS);
	float thresh = std::min(NERF_MIN_OPTICAL_THICKNESS, mean);
	const uint32_t n_bytes_per_mip = NERF_GRID_N_CELLS / 8;
	std::memset(out, 0, (size_t)n_bytes_per_mip * NERF_CASCADES);
	for (uint32_t i = 0; i < n_bytes_per_mip * n_cascades; ++i) {
		uint8_t b = 0;
		for (int j = 0; j < 8; ++j) b |= grid[(size_t)i * 8 + j] > thresh ? (uint8_t)(1 << j) : 0;
		out[i] = b;
	}
	for (uint32_t level = 1; level < NERF_CASCADES; ++level) {
		const uint8_t* prev = out + (size_t)n_bytes_per_mip * (level - 1);
		uint8_t* next = out + (size_t)n_bytes_per_mip * level;
		for (uint32_t i = 0; i < n_bytes_per_mip / 8; ++i) {
			uint8_t b = 0;
			for (int j = 0; j < 8; ++j) b |= prev[i * 8 + j] > 0 ? (uint8_t)(1 << j) : 0;
			uint32_t x = morton3D_invert(i >> 0) + NERF_GRIDSIZE / 8;
			uint32_t y = morton3D_invert(i >> 1) + NERF_GRIDSIZE / 8;
			uint32_t z = morton3D_invert(i >> 2) + NERF_GRIDSIZE / 8;
			next[morton3D(x, y, z)] |= b;
		}
	}
	return thresh;
}

// ---------------------------------------------------------------- tcnn grid.h:48-212, 668-721; common_device.h:651-718, 842-854
struct GridLayout {
	uint32_t L, F;
	std::vector<uint32_t> offsets, res;
	std::vector<float> scales;
};
GridLayout grid_layout(const orc_net_config& c) {
	GridLayout g;
	g.L = c.n_levels; g.F = c.n_features_per_level;
	g.offsets.resize(g.L + 1); g.res.resize(g.L); g.scales.resize(g.L);
	float log2_scale = std::log2(c.per_level_scale);
	uint32_t offset = 0;
	for (uint32_t i = 0; i < g.L; ++i) {
		float scale = std::exp2f(i * log2_scale) * c.base_resolution - 1.0f;
		uint32_t resolution = (uint32_t)std::ceil(scale) + 1;
		uint32_t max_params = UINT32_MAX / 2;
		uint32_t n = std::pow((float)resolution, 3) > (float)max_params ? max_params : resolution * resolution * resolution;
		n = (n + 7) / 8 * 8;
		n = std::min(n, 1u << c.log2_hashmap_size);
		g.offsets[i] = offset;
		offset += n;
		g.scales[i] = scale;
		g.res[i] = resolution;
		if (c.has_device_scales) { // what kernel_grid itself computes (offsets above stay host arithmetic, grid.h:692-721)
			g.scales[i] = c.device_scales[i];
			g.res[i] = (uint32_t)std::ceil(c.device_scales[i]) + 1;
		}
	}
	g.offsets[g.L] = offset;
	return g;
}
static uint32_t grid_index(uint32_t hashmap_size, uint32_t res, uint32_t x, uint32_t y, uint32_t z) {
	uint32_t p[3] = {x, y, z};
	uint32_t stride = 1, index = 0;
	for (uint32_t dim = 0; dim < 3 && stride <= hashmap_size; ++dim) {
		index += p[dim] * stride;
		stride *= res;
	}
	if (hashmap_size < stride) index = (x * 1u) ^ (y * 2654435761u) ^ (z * 805459861u);
	return index % hashmap_size;
}
// One sample: out[L*F] fp16. Weight is rounded to fp16, accumulation is an fp16 fma per corner (grid.h:162).
static void encode_one(const GridLayout& g, const half_t* table, const float* pos_in, half_t* out) {
	for (uint32_t level = 0; level < g.L; ++level) {
		const half_t* grid = table + (size_t)g.offsets[level] * g.F;
		uint32_t hashmap_size = g.offsets[level + 1] - g.offsets[level];
		float scale = g.scales[level];
		uint32_t res = g.res[level];
		float pos[3];
		uint32_t pg[3];
		for (int d = 0; d < 3; ++d) {
			float v = std::fmaf(scale, pos_in[d], 0.5f);
			float fl = std::floor(v);
			pg[d] = (uint32_t)(int)fl;
			pos[d] = v - fl;
		}
		half_t acc[8] = {};
		for (uint32_t idx = 0; idx < 8; ++idx) {
			float weight = 1;
			uint32_t pl[3];
			for (int d = 0; d < 3; ++d) {
				if ((idx & (1u << d)) == 0) { weight *= 1 - pos[d]; pl[d] = pg[d]; }
				else { weight *= pos[d]; pl[d] = pg[d] + 1; }
			}
			uint32_t index = grid_index(hashmap_size, res, pl[0], pl[1], pl[2]) * g.F;
			half_t w = f2h(weight);
			for (uint32_t f = 0; f < g.F; ++f) acc[f] = hfma(w, grid[index + f], acc[f]);
		}
		for (uint32_t f = 0; f < g.F; ++f) out[level * g.F + f] = acc[f];
	}
}

// ---------------------------------------------------------------- tcnn common_device.h:340-366 (degree 4 real SH)
static void sh4(float x, float y, float z, half_t* out) {
	float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
	out[0] = f2h(0.28209479177387814f);
	out[1] = f2h(-0.48860251190291987f * y);
	out[2] = f2h(0.48860251190291987f * z);
	out[3] = f2h(-0.48860251190291987f * x);
	out[4] = f2h(1.0925484305920792f * xy);
	out[5] = f2h(-1.0925484305920792f * yz);
	out[6] = f2h(0.94617469575755997f * z2 - 0.31539156525251999f);
	out[7] = f2h(-1.0925484305920792f * xz);
	out[8] = f2h(0.54627421529603959f * x2 - 0.54627421529603959f * y2);
	out[9] = f2h(0.59004358992664352f * y * (-3.0f * x2 + y2));
	out[10] = f2h(2.8906114426405538f * xy * z);
	out[11] = f2h(0.45704579946446572f * y * (1.0f - 5.0f * z2));
	out[12] = f2h(0.3731763325901154f * z * (5.0f * z2 - 3.0f));
	out[13] = f2h(0.45704579946446572f * x * (1.0f - 5.0f * z2));
	out[14] = f2h(1.4453057213202769f * z * (x2 - y2));
	out[15] = f2h(0.59004358992664352f * x * (-x2 + 3.0f * y2));
}

// ---------------------------------------------------------------- tcnn fully_fused_mlp.cu:47-129,315-476,635-678
// No biases; ReLU on hidden layers; weights row-major [out][in]; activations stored as fp16 between layers.
// The reference accumulates in fp16 (wmma half accumulator); the oracle accumulates in fp32.
struct MlpShape { uint32_t in, hidden_layers, out_padded; };
static size_t mlp_n_params(const MlpShape& s) { return (size_t)64 * s.in + (size_t)(s.hidden_layers - 1) * 64 * 64 + (size_t)s.out_padded * 64; }
static void mlp_forward(const MlpShape& s, const half_t* w, const half_t* in, half_t* out) {
	half_t a[64], b[64];
	for (uint32_t o = 0; o < 64; ++o) {
		float acc = 0.0f;
		for (uint32_t i = 0; i < s.in; ++i) acc += h2f(w[o * s.in + i]) * h2f(in[i]);
		a[o] = f2h(std::max(acc, 0.0f));
	}
	w += 64 * s.in;
	for (uint32_t l = 0; l + 1 < s.hidden_layers; ++l) {
		for (uint32_t o = 0; o < 64; ++o) {
			float acc = 0.0f;
			for (uint32_t i = 0; i < 64; ++i) acc += h2f(w[o * 64 + i]) * h2f(a[i]);
			b[o] = f2h(std::max(acc, 0.0f));
		}
		std::memcpy(a, b, sizeof(a));
		w += 64 * 64;
	}
	for (uint32_t o = 0; o < s.out_padded; ++o) {
		float acc = 0.0f;
		for (uint32_t i = 0; i < 64; ++i) acc += h2f(w[o * 64 + i]) * h2f(a[i]);
		out[o] = f2h(acc);
	}
}

struct Net {
	orc_net_config cfg;
	GridLayout grid;
	MlpShape density, rgb;
	const half_t *w_density, *w_rgb, *table;
};
Net make_net(const orc_net_config& c, const uint16_t* params) {
	Net n;
	n.cfg = c;
	n.grid = grid_layout(c);
	n.density = {c.n_levels * c.n_features_per_level, c.density_hidden_layers, 16};
	n.rgb = {32, c.rgb_hidden_layers, 16};
	// nerf_network.h:356-371: [density MLP | rgb MLP | pos encoding | dir encoding (0 params)]
	n.w_density = (const half_t*)params;
	n.w_rgb = n.w_density + mlp_n_params(n.density);
	n.table = n.w_rgb + mlp_n_params(n.rgb);
	return n;
}
// nerf_network.h:105-139: coords = {warped pos[3], warped dt, warped dir[3]} -> raw (r,g,b,sigma) fp16
void network_one(const Net& n, const float* coord, half_t out[4]) {
	half_t enc[32], rgb_in[32], rgb_out[16];
	encode_one(n.grid, n.table, coord, enc);
	mlp_forward(n.density, n.w_density, enc, rgb_in); // 16 density outputs -> rgb input rows 0..15
	sh4(coord[4] * 2.f - 1.f, coord[5] * 2.f - 1.f, coord[6] * 2.f - 1.f, rgb_in + 16);
	mlp_forward(n.rgb, n.w_rgb, rgb_in, rgb_out);
	out[0] = rgb_out[0]; out[1] = rgb_out[1]; out[2] = rgb_out[2];
	out[3] = rgb_in[0]; // extract_density: density MLP channel 0 -> output row 3
}

// density-only evaluation (NerfNetwork::density, nerf_network.h:291-318): encoding + density MLP, channel 0
half_t density_one(const Net& n, const float* warped_pos) {
	half_t enc[32], d_out[16];
	encode_one(n.grid, n.table, warped_pos, enc);
	mlp_forward(n.density, n.w_density, enc, d_out);
	return d_out[0];
}

// NerfNetwork::input_gradient(dim 3) (src/testbed_nerf.cu:2106-2107): gradient of the raw density output with respect to
// the warped position, through the density MLP (fully_fused_mlp backward: transposed weights, ReLU masks of the forward
// activations) and the grid encoding (grid.h:170-212: per feature, scale * sum over the corner pairs along the dimension of
// w_other * (v_right - v_left)). fp32 here; the reference runs the MLP part in fp16. Returns the raw density (fp16-rounded).
float density_input_gradient(const Net& n, const float* wp, float grad[3]) {
	half_t enc[32];
	encode_one(n.grid, n.table, wp, enc);
	const MlpShape& s = n.density;
	const half_t* w = n.w_density;
	std::vector<std::vector<float>> act(s.hidden_layers, std::vector<float>(64));
	for (uint32_t o = 0; o < 64; ++o) {
		float acc = 0.0f;
		for (uint32_t i = 0; i < s.in; ++i) acc += h2f(w[o * s.in + i]) * h2f(enc[i]);
		act[0][o] = h2f(f2h(std::max(acc, 0.0f)));
	}
	const half_t* wl = w + 64 * s.in;
	for (uint32_t l = 1; l < s.hidden_layers; ++l) {
		for (uint32_t o = 0; o < 64; ++o) {
			float acc = 0.0f;
			for (uint32_t i = 0; i < 64; ++i) acc += h2f(wl[o * 64 + i]) * act[l - 1][i];
			act[l][o] = h2f(f2h(std::max(acc, 0.0f)));
		}
		wl += 64 * 64;
	}
	float raw = 0.0f;
	std::vector<float> g(64), gp(64);
	for (uint32_t i = 0; i < 64; ++i) { float wv = h2f(wl[i]); raw += wv * act[s.hidden_layers - 1][i]; g[i] = act[s.hidden_layers - 1][i] > 0.0f ? wv : 0.0f; }
	raw = h2f(f2h(raw));
	for (uint32_t l = s.hidden_layers - 1; l >= 1; --l) {
		const half_t* wm = w + 64 * s.in + (size_t)(l - 1) * 64 * 64;
		for (uint32_t i = 0; i < 64; ++i) {
			float acc = 0.0f;
			for (uint32_t o = 0; o < 64; ++o) acc += h2f(wm[o * 64 + i]) * g[o];
			gp[i] = act[l - 1][i] > 0.0f ? acc : 0.0f;
		}
		g = gp;
	}
	float genc[32] = {};
	for (uint32_t o = 0; o < 64; ++o) for (uint32_t i = 0; i < s.in; ++i) genc[i] += h2f(w[o * s.in + i]) * g[o];
	grad[0] = grad[1] = grad[2] = 0.0f;
	const GridLayout& gl = n.grid;
	for (uint32_t level = 0; level < gl.L; ++level) {
		const half_t* grid = n.table + (size_t)gl.offsets[level] * gl.F;
		uint32_t hashmap_size = gl.offsets[level + 1] - gl.offsets[level];
		float scale = gl.scales[level];
		uint32_t res = gl.res[level];
		float pos[3]; uint32_t pg[3];
		for (int d = 0; d < 3; ++d) { float v = std::fmaf(scale, wp[d], 0.5f); float fl = std::floor(v); pg[d] = (uint32_t)(int)fl; pos[d] = v - fl; }
		for (uint32_t f = 0; f < gl.F; ++f) {
			float ge = genc[level * gl.F + f] * scale;
			for (int gd = 0; gd < 3; ++gd) {
				float acc = 0.0f;
				for (uint32_t idx = 0; idx < 4; ++idx) {
					float weight = 1.0f;
					uint32_t pl[3];
					for (int nd = 0; nd < 2; ++nd) {
						int dim = nd >= gd ? nd + 1 : nd;
						if ((idx & (1u << nd)) == 0) { weight *= 1 - pos[dim]; pl[dim] = pg[dim]; }
						else { weight *= pos[dim]; pl[dim] = pg[dim] + 1; }
					}
					pl[gd] = pg[gd];
					float left = h2f(grid[grid_index(hashmap_size, res, pl[0], pl[1], pl[2]) * gl.F + f]);
					pl[gd] = pg[gd] + 1;
					float right = h2f(grid[grid_index(hashmap_size, res, pl[0], pl[1], pl[2]) * gl.F + f]);
					acc += weight * (right - left);
				}
				grad[gd] += ge * acc;
			}
		}
	}
	return raw;
}

// ---------------------------------------------------------------- nerf_device.cuh:203-314
static float logistic(float x) { return 1.0f / (1.0f + std::exp(-x)); }
static float network_to_rgb(float v, int act) {
	switch (act) {
		case 0: return v;
		case 1: return v > 0.0f ? v : 0.0f;
		case 2: return logistic(v);
		case 3: return std::exp(clampf(v, -10.0f, 10.0f));
	}
	return 0.0f;
}
static float network_to_density(float v, int act) {
	switch (act) {
		case 0: return v;
		case 1: return v > 0.0f ? v : 0.0f;
		case 2: return logistic(v);
		case 3: return std::exp(v);
	}
	return 0.0f;
}
// nerf_device.cuh:245-254
static float network_to_density_derivative(float v, int act) {
	switch (act) {
		case 0: return 1.0f;
		case 1: return v > 0.0f ? 1.0f : 0.0f;
		case 2: { float d = logistic(v); return d * (1 - d); }
		case 3: return std::exp(std::min(std::max(v, -15.0f), 15.0f));
	}
	return 0.0f;
}
static float warp_dt(float dt) {
	float max_stepsize = MIN_CONE_STEPSIZE * (1 << (NERF_CASCADES - 1));
	return (dt - MIN_CONE_STEPSIZE) / (max_stepsize - MIN_CONE_STEPSIZE);
}
static float unwarp_dt(float dt) {
	float max_stepsize = MIN_CONE_STEPSIZE * (1 << (NERF_CASCADES - 1));
	return dt * (max_stepsize - MIN_CONE_STEPSIZE) + MIN_CONE_STEPSIZE;
}
float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : std::pow((s + 0.055f) / 1.055f, 2.4f); }

// ---------------------------------------------------------------- a-16: random_val.cuh:62-72,130-185, envmap.cuh:24-50
vec2 dir_to_spherical_unorm(vec3 d) {
	float cos_theta = std::fmin(std::fmax(d.z, -1.0f), 1.0f);
	float theta = std::acos(cos_theta);
	float phi = std::atan2(d.y, d.x);
	return {theta / PI_F, (phi / (2.0f * PI_F) + 0.5f)};
}
vec3 cosine_hemisphere(vec2 u) {
	float r = std::sqrt(u.x);
	float phi = 2.0f * PI_F * u.y;
	vec3 p;
	p.x = r * std::cos(phi);
	p.y = r * std::sin(phi);
	p.z = std::sqrt(std::fmax(0.0f, 1.0f - p.x * p.x - p.y * p.y));
	return p;
}
vec3 spherical_cap_uniform(vec2 u, float cap_height) {
	float z = (float)((double)(1.0f - cap_height) * (1.0 - (double)u.x) + (double)u.x); // the reference mixes a double literal in (random_val.cuh:147)
	float r = std::sqrt(std::fmax(0.0f, 1.0f - z * z));
	float phi = 2.0f * PI_F * u.y;
	return {r * std::cos(phi), r * std::sin(phi), z};
}
mat3 compute_local_frame(vec3 lz) {
	float x = lz.x, y = lz.y, z = lz.z;
	float sz = (z >= 0) ? 1.0f : -1.0f;
	float a = 1 / (sz + z);
	float ya = y * a;
	float b = x * ya;
	float c = x * sz;
	mat3 f;
	f.c[0] = {c * x * a - 1, sz * b, c};
	f.c[1] = {b, y * ya - sz, y};
	f.c[2] = lz;
	return f;
}
// envmap.cuh:24-50: lat-long lookup, bilinear, wrap in x, clamp in y
vec4 read_envmap(const float* env, int w, int h, vec3 dir) {
	vec2 cyl = dir_to_spherical_unorm({dir.z, -dir.x, dir.y});
	float fx = cyl.y * (float)(w - 1), fy = cyl.x * (float)(h - 1);
	int tx = (int)fx, ty = (int)fy;
	float wx = fx - (float)tx, wy = fy - (float)ty;
	auto rd = [&](int px, int py) -> vec4 {
		if (px < 0) px += w; else if (px >= w) px -= w;
		py = std::max(std::min(py, h - 1), 0);
		const float* t = env + 4 * ((size_t)px + (size_t)py * w);
		return {t[0], t[1], t[2], t[3]};
	};
	vec4 a = rd(tx, ty), b = rd(tx + 1, ty), c = rd(tx, ty + 1), d = rd(tx + 1, ty + 1);
	float w00 = (1 - wx) * (1 - wy), w10 = wx * (1 - wy), w01 = (1 - wx) * wy, w11 = wx * wy;
	return {w00 * a.x + w10 * b.x + w01 * c.x + w11 * d.x, w00 * a.y + w10 * b.y + w01 * c.y + w11 * d.y,
		w00 * a.z + w10 * b.z + w01 * c.z + w11 * d.z, w00 * a.w + w10 * b.w + w01 * c.w + w11 * d.w};
}
// OUR definition (the reference's shading kernels are absent, SURVEY a-16): light direction k of pixel idx
vec3 env_light_dir(uint32_t idx, uint32_t spp, uint32_t K, uint32_t k, vec3 n) {
	vec2 u = ld_random_val_2d(k + K * spp, idx * 786433u + 0x9E3779B9u);
	return normalize(mul(compute_local_frame(n), cosine_hemisphere(u)));
}

// ---------------------------------------------------------------- per-pixel ray set-up: testbed_nerf.cu:1428-1544 + 333-381
struct RayState { vec3 o, d; float t; bool alive; };
RayState init_ray(const orc_render_params& p, const uint8_t* bits, uint32_t x, uint32_t y, float* depth_px) {
	RayState r = {};
	uint32_t idx = x + (uint32_t)p.width * y;
	bool valid = true;
	Ray ray = pixel_ray(p, x, y, &valid);
	if (depth_px && *depth_px < 0.01f) *depth_px = MAX_DEPTH; // :1490-1493 (keeps a mesh depth already present)
	if (!valid) { r.o = ray.o; r.d = {0.0f, 0.0f, 0.0f}; r.alive = false; return r; } // :1494-1498
	ray.d = normalize(ray.d);
	BoundingBox render_aabb = bb_from(p.render_aabb);
	mat3 crop_store; const mat3* crop = crop_matrix(p, crop_store);
	float t = std::fmax((crop ? render_aabb.ray_intersect(mul(*crop, ray.o), mul(*crop, ray.d)) : render_aabb.ray_intersect(ray.o, ray.d)).x, 0.0f) + 1e-6f; // :1530
	r.o = ray.o; r.d = ray.d;
	if (!render_aabb.contains(crop ? mul(*crop, ray.at(t)) : ray.at(t))) { r.alive = false; return r; }
	// advance_pos_nerf :333-362
	vec3 idir = vec3{1.0f, 1.0f, 1.0f} / ray.d;
	float cone = p.cone_angle_constant;
	t = advance_n_steps(t, cone, ld_random_val(p.spp, idx * 786433u));
	t = advance_to_occupied(t, cone, ray, idir, bits, 0, p.max_mip, render_aabb, 0, crop);
	if (t >= MAX_DEPTH) { r.alive = false; return r; }
	r.t = t; r.alive = true;
	return r;
}

template <typename F>
static void parallel_rows(int n_rows, int n_threads, F&& f) {
	if (n_threads <= 1) { for (int y = 0; y < n_rows; ++y) f(y); return; }
	std::atomic<int> next{0};
	std::vector<std::thread> th;
	for (int i = 0; i < n_threads; ++i) th.emplace_back([&] { for (int y; (y = next.fetch_add(1)) < n_rows;) f(y); });
	for (auto& t : th) t.join();
}

// testbed_nerf.cu:2022-2144 (trace) + 430-477 (inputs) + 528-735 (composite) + 1361-1401 / geometry :1826-1871 (shade).
// The pass/compaction chunking of the reference does not change a ray's result, so the oracle walks each ray to its end.
uint64_t render_nerf(const orc_render_params& p, const Net& net, const uint8_t* bits, float* frame, float* depth, uint32_t* n_steps_out, bool depth_test) {
	BoundingBox render_aabb = bb_from(p.render_aabb), train_aabb = bb_from(p.train_aabb);
	cam_t cam1 = cam_from(p.cam1);
	vec3 cam_fwd = cam1.c[2], cam_pos = cam1.c[3];
	mat3 crop_store; const mat3* crop = crop_matrix(p, crop_store);
	std::atomic<uint64_t> total{0};
	parallel_rows(p.height, p.n_threads, [&](int y) {
		uint64_t row_total = 0;
		for (int x = 0; x < p.width; ++x) {
			uint32_t idx = x + p.width * y;
			if (n_steps_out) n_steps_out[idx] = 0;
			if (p.pixel_mask && !p.pixel_mask[idx]) continue;
			if (p.render_mode == 8 || p.render_mode == 10) { // the 2-D modes never trace (render_nerf :2397-2460): every valid pixel is shaded at once
				vec2 uv; cam_t cam;
				pixel_uv_cam(p, x, y, uv, cam);
				orc_render_params q = p;
				if (p.render_mode == 10) q.aperture_size = 0.0f; // plane_z < 0 (:1465-1467)
				bool valid = true;
				Ray ray = uv_to_ray(q, p.spp, uv, cam, true, &valid);
				if (depth[idx] < 0.01f) depth[idx] = MAX_DEPTH;
				if (!valid) continue;
				float* fb = frame + 4 * (size_t)idx;
				if (p.render_mode == 10) { // Slice (:1500-1510, 2431-2459): the network at the point where the pixel's ray meets the plane z = focus_z
					float nlen = length(ray.d);
					vec3 dir = (1.0f / nlen) * ray.d;
					float t = p.focus_z * nlen;
					depth[idx] = p.focus_z;
					vec3 wp = train_aabb.relative_pos(ray.o + dir * t);
					vec3 wd = (dir + 1.0f) * 0.5f;
					float coord[7] = {wp.x, wp.y, wp.z, warp_dt(MIN_CONE_STEPSIZE), wd.x, wd.y, wd.z};
					half_t out[4];
					network_one(net, coord, out);
					row_total += 1;
					if (n_steps_out) n_steps_out[idx] = 1;
					// compute_nerf_rgba(depth = 0.01, :408-421) then shade_kernel_nerf (:1361-1401; Slice leaves the depth buffer alone)
					float alpha = clampf(1.f - std::exp(-network_to_density(h2f(out[3]), p.density_activation) * 0.01f), 0.0f, 1.0f);
					vec4 tmp = {network_to_rgb(h2f(out[0]), p.rgb_activation) * alpha, network_to_rgb(h2f(out[1]), p.rgb_activation) * alpha, network_to_rgb(h2f(out[2]), p.rgb_activation) * alpha, alpha};
					if (!p.linear_colors) { tmp.x = srgb_to_linear(tmp.x); tmp.y = srgb_to_linear(tmp.y); tmp.z = srgb_to_linear(tmp.z); }
					float ia = 1.0f - tmp.w;
					fb[0] = tmp.x + fb[0] * ia; fb[1] = tmp.y + fb[1] * ia; fb[2] = tmp.z + fb[2] * ia; fb[3] = tmp.w + fb[3] * ia;
				} else { // Distortion (:1513-1521): where the lens + distortion grid moved the pixel, as a hue
					vec2 after = pos_to_uv(ray.at(1.0f), p.width, p.height, {p.focal[0], p.focal[1]}, cam, {p.screen_center[0], p.screen_center[1]}, p.parallax_shift, Fov(p.foveation), 0, p.lens_params);
					vec3 c = dir_to_rgb({(after.x - uv.x) * 64.0f, (after.y - uv.y) * 64.0f});
					fb[0] = c.x; fb[1] = c.y; fb[2] = c.z; fb[3] = 1.0f;
					depth[idx] = 1.0f;
				}
				continue;
			}
			RayState r = init_ray(p, bits, x, y, &depth[idx]);
			// :1526-1528 envmap -> frame buffer for every valid ray; in hybrid frames only where the mesh pass left nothing
			if (p.envmap && (r.d.x != 0.0f || r.d.y != 0.0f || r.d.z != 0.0f) && !(depth_test && depth[idx] < MAX_DEPTH)) { // invalid rays return before this (:1494-1498)
				vec4 e = read_envmap(p.envmap, p.env_w, p.env_h, r.d);
				float* fb = frame + 4 * (size_t)idx;
				fb[0] = e.x; fb[1] = e.y; fb[2] = e.z; fb[3] = e.w;
			}
			if (!r.alive) continue;
			Ray ray = {r.o, r.d};
			vec3 idir = vec3{1.0f, 1.0f, 1.0f} / r.d;
			float cone = p.cone_angle_constant;
			float t = r.t;
			vec4 rgba = {0, 0, 0, 0};
			float local_depth = 0.0f, max_weight = 0.0f;
			uint32_t n = 0;
			float clip_depth = (p.depth_merge_clip && depth_test) ? depth[idx] : MAX_DEPTH;
			while (true) {
				t = advance_to_occupied(t, cone, ray, idir, bits, 0, p.max_mip, render_aabb, p.dda_iter_cap, crop);
				if (t >= MAX_DEPTH) break;
				float dt = calc_dt(t, cone);
				vec3 wp = train_aabb.relative_pos(ray.at(t));
				if (p.depth_merge_clip && depth_test) {
					vec3 up = train_aabb.min + wp * train_aabb.diag();
					if (dot(cam_fwd, up - cam_pos) > clip_depth) break;
				}
				vec3 wd = (r.d + 1.0f) * 0.5f;
				float coord[7] = {wp.x, wp.y, wp.z, warp_dt(dt), wd.x, wd.y, wd.z};
				half_t out[4];
				network_one(net, coord, out);
				++n;
				vec3 pos = train_aabb.min + wp * train_aabb.diag();
				float T = 1.f - rgba.w;
				float dtu = unwarp_dt(coord[3]);
				float alpha = 1.f - std::exp(-network_to_density(h2f(out[3]), p.density_activation) * dtu);
				float weight = alpha * T;
				vec3 rgb = {network_to_rgb(h2f(out[0]), p.rgb_activation), network_to_rgb(h2f(out[1]), p.rgb_activation), network_to_rgb(h2f(out[2]), p.rgb_activation)};
				// :686-702 visualisation modes replace the sample colour
				if (p.render_mode == 6) rgb = (pos - 0.5f) / 2.0f + 0.5f;
				else if (p.render_mode == 12) rgb = wp;
				else if (p.render_mode == 7) { float z = dot(cam_fwd, pos - r.o) * p.depth_scale; rgb = {z, z, z}; }
				else if (p.render_mode == 0) rgb = {alpha, alpha, alpha};
				else if (p.render_mode == 5) { // :688-693 normal = -d sigma/d raw * (gradient of the raw density w.r.t. the network input), normalised; fp16 like the product's sample record
					float gr[3];
					float raw = density_input_gradient(net, coord, gr);
					float der = network_to_density_derivative(raw, p.density_activation);
					vec3 nrm = normalize(vec3{-der * gr[0], -der * gr[1], -der * gr[2]});
					rgb = {h2f(f2h(nrm.x)), h2f(f2h(nrm.y)), h2f(f2h(nrm.z))};
				}
				rgba.x += rgb.x * weight; rgba.y += rgb.y * weight; rgba.z += rgb.z * weight; rgba.w += weight;
				if (weight > max_weight) { max_weight = weight; local_depth = dot(cam_fwd, pos - cam_pos); }
				if (rgba.w > (1.0f - p.min_transmittance)) {
					float a = rgba.w;
					rgba.x /= a; rgba.y /= a; rgba.z /= a; rgba.w /= a;
					break;
				}
				t += dt;
			}
			row_total += n;
			if (n_steps_out) n_steps_out[idx] = n;
			if (!(rgba.w > 0.001f)) continue;                    // compact_kernel_nerf :1419
			if (depth_test && !p.depth_merge_clip && local_depth > depth[idx]) continue; // geometry :1843-1846
			vec4 tmp = rgba;
			if (p.render_mode == 9) { float col = (float)n / 128.0f; tmp = {col, col, col, 1.0f}; } // shade_kernel_nerf :1381-1383
			if (p.render_mode == 5) { vec3 nn = normalize(vec3{rgba.x, rgba.y, rgba.z}); tmp.x = (0.5f * nn.x + 0.5f) * rgba.w; tmp.y = (0.5f * nn.y + 0.5f) * rgba.w; tmp.z = (0.5f * nn.z + 0.5f) * rgba.w; } // :1379-1381
			if (!p.linear_colors && p.render_mode >= 1 && p.render_mode <= 4) { tmp.x = srgb_to_linear(tmp.x); tmp.y = srgb_to_linear(tmp.y); tmp.z = srgb_to_linear(tmp.z); } // :1391-1394
			float* fb = frame + 4 * (size_t)idx;
			float ia = 1.0f - tmp.w;
			fb[0] = tmp.x + fb[0] * ia; fb[1] = tmp.y + fb[1] * ia; fb[2] = tmp.z + fb[2] * ia; fb[3] = tmp.w + fb[3] * ia;
			if (tmp.w > 0.2f) depth[idx] = local_depth;
		}
		total += row_total;
	});
	return total;
}

// OUR definition (a-16): transmittance of the volume along a light ray from a mesh hit. Same AABB entry, DDA and step
// sizes as the camera rays (no start jitter); density-only network evaluation; stops once T < min_transmittance.
float nerf_transmittance(const orc_render_params& p, const Net& net, const uint8_t* bits, vec3 o, vec3 d, uint32_t* n_out) {
	BoundingBox render_aabb = bb_from(p.render_aabb), train_aabb = bb_from(p.train_aabb);
	if (n_out) *n_out = 0;
	mat3 crop_store; const mat3* crop = crop_matrix(p, crop_store);
	float t = std::fmax((crop ? render_aabb.ray_intersect(mul(*crop, o), mul(*crop, d)) : render_aabb.ray_intersect(o, d)).x, 0.0f) + 1e-6f;
	Ray ray = {o, d};
	if (!render_aabb.contains(crop ? mul(*crop, ray.at(t)) : ray.at(t))) return 1.0f;
	vec3 idir = vec3{1.0f, 1.0f, 1.0f} / d;
	float cone = p.cone_angle_constant;
	float T = 1.0f;
	uint32_t n = 0;
	while (true) {
		t = advance_to_occupied(t, cone, ray, idir, bits, 0, p.max_mip, render_aabb, p.dda_iter_cap, crop);
		if (t >= MAX_DEPTH) break;
		float dt = calc_dt(t, cone);
		vec3 wp = train_aabb.relative_pos(ray.at(t));
		float coord[3] = {wp.x, wp.y, wp.z};
		float sigma = network_to_density(h2f(density_one(net, coord)), p.density_activation);
		T *= std::exp(-sigma * unwarp_dt(warp_dt(dt)));
		++n;
		if (T < p.min_transmittance) break;
		t += dt;
	}
	if (n_out) *n_out = n;
	return T;
}

#include "orc_probe.inc"
#include "orc_train.inc"
#include "orc_grid.inc"
#include "orc_errmap.inc"

Net* net_new(const orc_net_config& c, const uint16_t* params) { return new Net(make_net(c, params)); }
void net_delete(Net* n) { delete n; }

} // namespace orc

using namespace orc;

extern "C" {

float orc_ld_random_val(uint32_t index, uint32_t seed) { return ld_random_val(index, seed); }
void orc_ld_random_pixel_offset(uint32_t spp, float out[2]) { vec2 v = ld_random_pixel_offset(spp); out[0] = v.x; out[1] = v.y; }
void orc_bbox_ray_intersect(const float bb[6], const float o[3], const float d[3], float out[2]) {
	vec2 r = bb_from(bb).ray_intersect({o[0], o[1], o[2]}, {d[0], d[1], d[2]});
	out[0] = r.x; out[1] = r.y;
}
uint32_t orc_morton3D(uint32_t x, uint32_t y, uint32_t z) { return morton3D(x, y, z); }
float orc_advance_to_occupied(float t, float cone_angle, const float o[3], const float d[3], const uint8_t* bitfield,
	uint32_t min_mip, uint32_t max_mip, const float aabb[6], int32_t iter_cap) {
	Ray ray = {{o[0], o[1], o[2]}, {d[0], d[1], d[2]}};
	vec3 idir = vec3{1.0f, 1.0f, 1.0f} / ray.d;
	return advance_to_occupied(t, cone_angle, ray, idir, bitfield, min_mip, max_mip, bb_from(aabb), iter_cap, nullptr);
}
float orc_calc_dt(float t, float cone_angle) { return calc_dt(t, cone_angle); }
float orc_srgb_to_linear(float v) { return srgb_to_linear(v); }

void orc_foveation_coeffs(const float f[6], float out[24]) {
	Fov fv(f);
	const FovWarp* w[2] = {&fv.x, &fv.y};
	for (int k = 0; k < 2; ++k) {
		float* o = out + 12 * k;
		o[0] = w[k]->al; o[1] = w[k]->bl; o[2] = w[k]->cl; o[3] = w[k]->am; o[4] = w[k]->bm; o[5] = w[k]->ar; o[6] = w[k]->br; o[7] = w[k]->cr;
		o[8] = w[k]->switch_left; o[9] = w[k]->switch_right; o[10] = w[k]->inv_switch_left; o[11] = w[k]->inv_switch_right;
	}
}
void orc_foveation_warp(const float f[6], const float uv[2], float warped[2], float unwarped[2], float* density) {
	Fov fv(f);
	vec2 a = fv.warp({uv[0], uv[1]}), b = fv.unwarp({uv[0], uv[1]});
	warped[0] = a.x; warped[1] = a.y; unwarped[0] = b.x; unwarped[1] = b.y; *density = fv.density({uv[0], uv[1]});
}
void orc_pos_to_uv(const orc_render_params* p, const float pos[3], float out[2]) {
	vec2 r = pos_to_uv({pos[0], pos[1], pos[2]}, p->width, p->height, {p->focal[0], p->focal[1]}, cam_from(p->cam0), {p->screen_center[0], p->screen_center[1]},
		p->parallax_shift, Fov(p->foveation), p->lens_mode, p->lens_params);
	out[0] = r.x; out[1] = r.y;
}
void orc_to_rgb(const float dir[2], float out[3]) { vec3 c = dir_to_rgb({dir[0], dir[1]}); out[0] = c.x; out[1] = c.y; out[2] = c.z; }
int32_t orc_uv_to_ray(const orc_render_params* p, uint32_t spp, const float uv[2], int32_t use_lens, float out_o[3], float out_d[3]) {
	bool valid = true;
	Ray r = uv_to_ray(*p, spp, {uv[0], uv[1]}, cam_from(p->cam0), use_lens != 0, &valid);
	out_o[0] = r.o.x; out_o[1] = r.o.y; out_o[2] = r.o.z; out_d[0] = r.d.x; out_d[1] = r.d.y; out_d[2] = r.d.z;
	return valid ? 1 : 0;
}
// src/render_buffer.cu:264-321 tonemap curve, :323-343 exposure / colour space, :529-560 tonemap_kernel (linear accumulation)
static void tonemap_curve(float* c, int curve) {
	if (curve == 0) return;
	for (int i = 0; i < 3; ++i) c[i] = std::max(c[i], 0.0f);
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
		float Y = 0.2126f * c[0] + 0.7152f * c[1] + 0.0722f * c[2];
		float sc = 1.f / (Y + 1.0f);
		for (int i = 0; i < 3; ++i) c[i] *= sc;
		return;
	}
	for (int i = 0; i < 3; ++i) c[i] = (c[i] * c[i] * k0 + k1 * c[i] + k2) / (k3 * (c[i] * c[i]) + k4 * c[i] + k5);
}
static float linear_to_srgb_f(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * std::pow(l, 0.41666f) - 0.055f; }
void orc_tonemap(const float* accum, uint32_t n, int32_t to_srgb, float exposure, const float bg[4], int32_t curve, int32_t clamp_output,
	int32_t unmultiply_alpha, float* out) {
	const float scale = std::exp2(exposure);
	for (uint32_t i = 0; i < n; ++i) {
		float c[4] = {accum[4 * i], accum[4 * i + 1], accum[4 * i + 2], accum[4 * i + 3]};
		float w = (1.f - c[3]) * bg[3];
		for (int k = 0; k < 3; ++k) c[k] += srgb_to_linear(bg[k]) * w;
		c[3] += w;
		for (int k = 0; k < 3; ++k) c[k] *= scale;
		tonemap_curve(c, curve);
		if (to_srgb) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb_f(c[k]);
		if (unmultiply_alpha && c[3] > 0.0f) for (int k = 0; k < 3; ++k) c[k] /= c[3];
		if (clamp_output) for (int k = 0; k < 4; ++k) c[k] = std::min(std::max(c[k], 0.f), 1.f);
		for (int k = 0; k < 4; ++k) out[4 * i + k] = c[k];
	}
}
void orc_ld_random_val_2d(uint32_t index, uint32_t seed, float out[2]) { vec2 v = ld_random_val_2d(index, seed); out[0] = v.x; out[1] = v.y; }
void orc_dir_to_spherical_unorm(const float d[3], float out[2]) { vec2 v = dir_to_spherical_unorm({d[0], d[1], d[2]}); out[0] = v.x; out[1] = v.y; }
void orc_cosine_hemisphere(const float u[2], float out[3]) { vec3 v = cosine_hemisphere({u[0], u[1]}); out[0] = v.x; out[1] = v.y; out[2] = v.z; }
void orc_spherical_cap(const float u[2], float cap_height, float out[3]) { vec3 v = spherical_cap_uniform({u[0], u[1]}, cap_height); out[0] = v.x; out[1] = v.y; out[2] = v.z; }
void orc_compute_local_frame(const float z[3], float out[9]) {
	mat3 f = compute_local_frame({z[0], z[1], z[2]});
	for (int c = 0; c < 3; ++c) { out[3 * c] = f.c[c].x; out[3 * c + 1] = f.c[c].y; out[3 * c + 2] = f.c[c].z; }
}
void orc_read_envmap(const float* envmap, int32_t w, int32_t h, const float dir[3], float out[4]) {
	vec4 v = read_envmap(envmap, w, h, {dir[0], dir[1], dir[2]});
	out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = v.w;
}
void orc_env_light_dirs(uint32_t idx, uint32_t spp, uint32_t K, const float n[3], float* out) {
	for (uint32_t k = 0; k < K; ++k) { vec3 v = env_light_dir(idx, spp, K, k, {n[0], n[1], n[2]}); out[3 * k] = v.x; out[3 * k + 1] = v.y; out[3 * k + 2] = v.z; }
}
float orc_nerf_transmittance(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield,
	const float o[3], const float d[3], uint32_t* n_samples_out) {
	Net net = make_net(*c, params);
	return nerf_transmittance(*p, net, bitfield, {o[0], o[1], o[2]}, {d[0], d[1], d[2]}, n_samples_out);
}

// loss_type 0 L2, 1 Huber: out = loss[3], gradient[3] (nerf_device.cuh:74-112,640-655)
void orc_loss_and_gradient(const float target[3], const float pred[3], int32_t loss_type, float out[6]) {
	for (int k = 0; k < 3; ++k) out[k] = loss_grad(loss_type, target[k], pred[k], out[3 + k]);
}
// Normals mode: raw density and its gradient w.r.t. the warped position (returns raw)
float orc_density_input_gradient(const orc_net_config* c, const uint16_t* params, const float wp[3], float grad[3]) {
	Net net = make_net(*c, params);
	return density_input_gradient(net, wp, grad);
}
// all-fp32 raw density of the same weights (finite-difference witness for the gradient above; tests only)
float orc_density_raw_f32(const orc_net_config* c, const uint16_t* params, const float wp[3]) {
	Net net = make_net(*c, params);
	std::vector<float> acts(mlp_act_size(net.density));
	encode_f32(net.grid, net.table, wp, acts.data(), nullptr, nullptr);
	mlp_forward_f32(net.density, net.w_density, acts.data());
	return acts[net.density.in + 64 * net.density.hidden_layers];
}
float orc_train_gradients(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield, uint32_t n_rays,
	const float* origins, const float* dirs, const float* t_start, const float* targets, const float* backgrounds, int32_t loss_type, float loss_scale,
	uint32_t max_samples, float* grads, uint32_t* n_steps_out, uint32_t* n_samples_total) {
	Net net = make_net(*c, params);
	return train_gradients(*p, net, bitfield, n_rays, origins, dirs, t_start, targets, backgrounds, loss_type, loss_scale, max_samples, grads, n_steps_out, n_samples_total);
}

float orc_train_gradients_depth(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield, uint32_t n_rays,
	const float* origins, const float* dirs, const float* t_start, const float* targets, const float* backgrounds, const float* target_depth, float depth_lambda,
	int32_t depth_loss_type, int32_t loss_type, float loss_scale, uint32_t max_samples, float* grads, uint32_t* n_steps_out, uint32_t* n_samples_total) {
	Net net = make_net(*c, params);
	return train_gradients(*p, net, bitfield, n_rays, origins, dirs, t_start, targets, backgrounds, loss_type, loss_scale, max_samples, grads, n_steps_out, n_samples_total,
		target_depth, depth_lambda, depth_loss_type);
}

void orc_mark_untrained_grid(float* grid, uint32_t n_cascades, uint32_t n_images, const float* cams, int32_t w, int32_t h, const float focal[2], const float principal[2]) {
	mark_untrained_grid(grid, n_cascades, n_images, cams, w, h, focal, principal);
}
void orc_mark_untrained_grid_per_image(float* grid, uint32_t n_cascades, uint32_t n_images, const float* cams, int32_t w, int32_t h, const float* focal, const float* principal,
	const int32_t* lens_mode, const float* lens_params) {
	ImageCams pc = {focal, principal, lens_mode, lens_params};
	const float f0[2] = {1.0f, 1.0f}, p0[2] = {0.5f, 0.5f};
	mark_untrained_grid(grid, n_cascades, n_images, cams, w, h, f0, p0, &pc);
}
void orc_update_density_grid(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, float* grid, uint32_t n_cascades, uint32_t ema_step,
	uint32_t n_uniform, uint32_t n_nonuniform, float decay, uint32_t seed, uint32_t rng_calls) {
	Net net = make_net(*c, params);
	update_density_grid(*p, net, grid, n_cascades, ema_step, n_uniform, n_nonuniform, decay, seed, rng_calls);
}
uint32_t orc_pcg32(uint64_t seed, uint64_t advance, uint32_t n, uint32_t* out) {
	Pcg32 r(seed); r.advance(advance);
	for (uint32_t i = 0; i < n; ++i) out[i] = r.next_uint();
	return n;
}

float orc_bitfield_from_grid(const float* grid, uint32_t n_cascades, uint8_t* out) { return bitfield_from_grid(grid, n_cascades, out); }

void orc_init_rays(const orc_render_params* p, const uint8_t* bitfield, float* origins, float* dirs, float* t, uint8_t* alive) {
	parallel_rows(p->height, p->n_threads, [&](int y) {
		for (int x = 0; x < p->width; ++x) {
			uint32_t idx = x + p->width * y;
			RayState r = init_ray(*p, bitfield, x, y, nullptr);
			origins[3 * idx] = r.o.x; origins[3 * idx + 1] = r.o.y; origins[3 * idx + 2] = r.o.z;
			dirs[3 * idx] = r.d.x; dirs[3 * idx + 1] = r.d.y; dirs[3 * idx + 2] = r.d.z;
			t[idx] = r.alive ? r.t : MAX_DEPTH;
			alive[idx] = r.alive;
		}
	});
}

void orc_march_all(const orc_render_params* p, const uint8_t* bitfield, uint32_t n_rays, const float* origins, const float* dirs,
	const float* t0, const uint8_t* alive, uint32_t* counts, float* samples, uint32_t max_per_ray) {
	BoundingBox render_aabb = bb_from(p->render_aabb);
	for (uint32_t i = 0; i < n_rays; ++i) {
		counts[i] = 0;
		if (!alive[i]) continue;
		Ray ray = {{origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]}, {dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]}};
		vec3 idir = vec3{1.0f, 1.0f, 1.0f} / ray.d;
		float t = t0[i];
		uint32_t n = 0;
		while (true) {
			mat3 crop_store; const mat3* crop = crop_matrix(*p, crop_store);
			t = advance_to_occupied(t, p->cone_angle_constant, ray, idir, bitfield, 0, p->max_mip, render_aabb, p->dda_iter_cap, crop);
			if (t >= MAX_DEPTH) break;
			float dt = calc_dt(t, p->cone_angle_constant);
			if (samples && n < max_per_ray) {
				vec3 pos = ray.at(t);
				float* s = samples + ((size_t)i * max_per_ray + n) * 4;
				s[0] = pos.x; s[1] = pos.y; s[2] = pos.z; s[3] = dt;
			}
			++n;
			t += dt;
		}
		counts[i] = n;
	}
}

size_t orc_n_params(const orc_net_config* c) {
	GridLayout g = grid_layout(*c);
	MlpShape d = {c->n_levels * c->n_features_per_level, c->density_hidden_layers, 16}, r = {32, c->rgb_hidden_layers, 16};
	return mlp_n_params(d) + mlp_n_params(r) + (size_t)g.offsets[g.L] * g.F;
}
size_t orc_grid_param_offset(const orc_net_config* c) {
	MlpShape d = {c->n_levels * c->n_features_per_level, c->density_hidden_layers, 16}, r = {32, c->rgb_hidden_layers, 16};
	return mlp_n_params(d) + mlp_n_params(r);
}
void orc_grid_layout(const orc_net_config* c, uint32_t* offsets, float* scales, uint32_t* resolutions) {
	GridLayout g = grid_layout(*c);
	for (uint32_t i = 0; i <= g.L; ++i) offsets[i] = g.offsets[i];
	for (uint32_t i = 0; i < g.L; ++i) { scales[i] = g.scales[i]; resolutions[i] = g.res[i]; }
}
void orc_encode(const orc_net_config* c, const uint16_t* params, const float* positions, uint32_t n, uint16_t* out) {
	Net net = make_net(*c, params);
	for (uint32_t i = 0; i < n; ++i) encode_one(net.grid, net.table, positions + 3 * (size_t)i, (half_t*)out + 32 * (size_t)i);
}
void orc_sh4(const float* d, uint32_t n, uint16_t* out) {
	for (uint32_t i = 0; i < n; ++i) sh4(d[3 * i] * 2.f - 1.f, d[3 * i + 1] * 2.f - 1.f, d[3 * i + 2] * 2.f - 1.f, (half_t*)out + 16 * (size_t)i);
}
void orc_network(const orc_net_config* c, const uint16_t* params, const float* coords, uint32_t n, uint16_t* out) {
	Net net = make_net(*c, params);
	for (uint32_t i = 0; i < n; ++i) network_one(net, coords + 7 * (size_t)i, (half_t*)out + 4 * (size_t)i);
}

uint64_t orc_render_nerf(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield,
	float* frame, float* depth, uint32_t* n_steps_out, int32_t depth_test) {
	Net net = make_net(*c, params);
	return render_nerf(*p, net, bitfield, frame, depth, n_steps_out, depth_test != 0);
}

// ---- a-16 irradiance probes (orc_probe.inc)
void orc_cylindrical_to_dir(const float p[2], float out[3]) { vec3 d = cylindrical_to_dir({p[0], p[1]}); out[0] = d.x; out[1] = d.y; out[2] = d.z; }
float orc_halton(uint32_t base, uint64_t idx) { return halton(base, (size_t)idx); }
uint32_t orc_probe_rays(int32_t mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float box[6], const float origin[3], float* o_out, float* d_out, uint32_t* texel_out) {
	const uint32_t n = n_theta * n_phi * (mode == 2 ? n_origin * n_origin : 1u);
	BoundingBox b = bb_from(box);
	for (uint32_t r = 0; r < n; ++r) {
		vec3 o, d; uint32_t t = 0;
		probe_ray(mode, n_theta, n_phi, n_origin, b, {origin[0], origin[1], origin[2]}, r, o, d, t);
		o_out[3 * r] = o.x; o_out[3 * r + 1] = o.y; o_out[3 * r + 2] = o.z; d_out[3 * r] = d.x; d_out[3 * r + 1] = d.y; d_out[3 * r + 2] = d.z; texel_out[r] = t;
	}
	return n;
}
uint64_t orc_render_probe(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield, int32_t mode,
	uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float origin[3], float* tex) {
	Net net = make_net(*c, params);
	return render_probe(*p, net, bitfield, mode, n_theta, n_phi, n_origin, {origin[0], origin[1], origin[2]}, tex);
}
void orc_read_probe(const float* tex, uint32_t n_theta, uint32_t n_phi, const float* dirs, uint32_t n, float* out) {
	for (uint32_t i = 0; i < n; ++i) { vec4 v = read_probe(tex, n_theta, n_phi, {dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]}); out[4 * i] = v.x; out[4 * i + 1] = v.y; out[4 * i + 2] = v.z; out[4 * i + 3] = v.w; }
}

} // extern "C"
