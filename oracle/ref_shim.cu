// ORACLE — TEST INFRASTRUCTURE ONLY. Thin C wrapper that compiles the reference's OWN NGP_HOST_DEVICE functions for
// the host (nvcc, no GPU needed to run them) straight from /root/reference, so the CPU restatement in orc_*.cpp can be
// pinned against them bit for bit. No reference source is copied: this file only #includes the headers where they lie.
// Output goes to oracle/_ref/ (git-ignored). Built by `make -C oracle ref`.
#include <cstdio>
#include <neural-graphics-primitives/common.h>
#include <neural-graphics-primitives/common_device.cuh>
#include <neural-graphics-primitives/bounding_box.cuh>
#include <neural-graphics-primitives/random_val.cuh>
#include <neural-graphics-primitives/nerf_device.cuh>
#include <neural-graphics-primitives/triangle.cuh>

using namespace ngp;

extern "C" {

float ref_ld_random_val(uint32_t index, uint32_t seed) { return ld_random_val(index, seed); }

void ref_ld_random_pixel_offset(uint32_t spp, float out[2]) {
	vec2 v = ld_random_pixel_offset(spp);
	out[0] = v.x; out[1] = v.y;
}

void ref_bbox_ray_intersect(const float bb[6], const float o[3], const float d[3], float out[2]) {
	BoundingBox b(vec3(bb[0], bb[1], bb[2]), vec3(bb[3], bb[4], bb[5]));
	vec2 r = b.ray_intersect(vec3(o[0], o[1], o[2]), vec3(d[0], d[1], d[2]));
	out[0] = r.x; out[1] = r.y;
}

float ref_triangle_ray_intersect(const float tri[9], const float o[3], const float d[3]) {
	Triangle t{vec3(tri[0], tri[1], tri[2]), vec3(tri[3], tri[4], tri[5]), vec3(tri[6], tri[7], tri[8])};
	return t.ray_intersect(vec3(o[0], o[1], o[2]), vec3(d[0], d[1], d[2]));
}

uint32_t ref_morton3D(uint32_t x, uint32_t y, uint32_t z) { return tcnn::morton3D(x, y, z); }

float ref_advance_to_occupied(float t, float cone_angle, const float o[3], const float d[3], const uint8_t* bitfield,
	uint32_t min_mip, uint32_t max_mip, const float aabb[6], int32_t iter_cap) {
	Ray ray{vec3(o[0], o[1], o[2]), vec3(d[0], d[1], d[2])};
	vec3 idir = vec3(1.0f) / ray.d;
	BoundingBox b(vec3(aabb[0], aabb[1], aabb[2]), vec3(aabb[3], aabb[4], aabb[5]));
	if (iter_cap > 0) {
		return if_unoccupied_advance_to_next_occupied_voxel_geometry(t, cone_angle, ray, idir, bitfield, min_mip, max_mip, b);
	}
	return if_unoccupied_advance_to_next_occupied_voxel(t, cone_angle, ray, idir, bitfield, min_mip, max_mip, b);
}

float ref_calc_dt(float t, float cone_angle) { return calc_dt(t, cone_angle); }
float ref_advance_n_steps(float t, float cone_angle, float n) { return advance_n_steps(t, cone_angle, n); }
float ref_srgb_to_linear(float v) { return srgb_to_linear(v); }
float ref_warp_dt(float dt) { return warp_dt(dt); }
float ref_unwarp_dt(float dt) { return unwarp_dt(dt); }

// init_rays_with_payload_kernel_nerf (src/testbed_nerf.cu:1428-1544) up to the normalised ray: the kernel itself is
// __global__, but every function it calls for the pinhole path is host-callable.
void ref_pixel_ray(uint32_t spp, uint32_t x, uint32_t y, int32_t w, int32_t h, const float focal[2], const float cam0[12], const float cam1[12],
	const float rolling_shutter[4], const float screen_center[2], int32_t snap_to_pixel_centers, float near_distance, float out_o[3], float out_d[3]) {
	ivec2 resolution(w, h);
	uint32_t idx = x + resolution.x * y;
	mat4x3 c0, c1;
	for (int i = 0; i < 4; ++i) { c0[i] = vec3(cam0[3*i], cam0[3*i+1], cam0[3*i+2]); c1[i] = vec3(cam1[3*i], cam1[3*i+1], cam1[3*i+2]); }
	vec4 rs(rolling_shutter[0], rolling_shutter[1], rolling_shutter[2], rolling_shutter[3]);
	vec2 pixel_offset = ld_random_pixel_offset(snap_to_pixel_centers ? 0 : spp);
	vec2 uv = vec2{(float)x + pixel_offset.x, (float)y + pixel_offset.y} / vec2(resolution);
	mat4x3 camera = get_xform_given_rolling_shutter({c0, c1}, rs, uv, ld_random_val(spp, idx * 72239731));
	Ray ray = uv_to_ray(spp, uv, resolution, vec2(focal[0], focal[1]), camera, vec2(screen_center[0], screen_center[1]), vec3(0.0f), near_distance, 1.0f, 0.0f);
	ray.d = normalize(ray.d);
	out_o[0] = ray.o.x; out_o[1] = ray.o.y; out_o[2] = ray.o.z;
	out_d[0] = ray.d.x; out_d[1] = ray.d.y; out_d[2] = ray.d.z;
}

// advance_pos_nerf (src/testbed_nerf.cu:333-362) body, host-callable pieces only.
float ref_start_t(uint32_t spp, uint32_t idx, const float o[3], const float d[3], const float aabb[6], float cone_angle,
	const uint8_t* bitfield, uint32_t max_mip) {
	Ray ray{vec3(o[0], o[1], o[2]), vec3(d[0], d[1], d[2])};
	BoundingBox b(vec3(aabb[0], aabb[1], aabb[2]), vec3(aabb[3], aabb[4], aabb[5]));
	float t = fmaxf(b.ray_intersect(ray.o, ray.d).x, 0.0f) + 1e-6f;
	if (!b.contains(ray(t))) return MAX_DEPTH();
	vec3 idir = vec3(1.0f) / ray.d;
	t = advance_n_steps(t, cone_angle, ld_random_val(spp, idx * 786433));
	return if_unoccupied_advance_to_next_occupied_voxel(t, cone_angle, ray, idir, bitfield, 0, max_mip, b);
}

// uv_to_ray (common_device.cuh:416-483) with a Lens, parallax shift and depth of field; foveation / masks / distortion grid default.
int32_t ref_uv_to_ray_lens(uint32_t spp, const float uv[2], int32_t w, int32_t h, const float focal[2], const float cam[12], const float screen_center[2],
	const float parallax[3], float near_distance, float focus_z, float aperture_size, int32_t lens_mode, const float lens_params[7], float out_o[3], float out_d[3]) {
	mat4x3 c;
	for (int i = 0; i < 4; ++i) c[i] = vec3(cam[3*i], cam[3*i+1], cam[3*i+2]);
	Lens lens;
	lens.mode = (ELensMode)lens_mode;
	for (int i = 0; i < 7; ++i) lens.params[i] = lens_params[i];
	Ray ray = uv_to_ray(spp, vec2(uv[0], uv[1]), ivec2(w, h), vec2(focal[0], focal[1]), c, vec2(screen_center[0], screen_center[1]),
		vec3(parallax[0], parallax[1], parallax[2]), near_distance, focus_z, aperture_size, Foveation{}, Buffer2DView<const uint8_t>{}, lens);
	out_o[0] = ray.o.x; out_o[1] = ray.o.y; out_o[2] = ray.o.z;
	out_d[0] = ray.d.x; out_d[1] = ray.d.y; out_d[2] = ray.d.z;
	return ray.is_valid() ? 1 : 0;
}

// Triangle::distance_sq / closest_point (triangle.cuh:66-147), NGP_HOST_DEVICE
float ref_triangle_distance_sq(const float tri[9], const float p[3]) {
	Triangle t{vec3(tri[0], tri[1], tri[2]), vec3(tri[3], tri[4], tri[5]), vec3(tri[6], tri[7], tri[8])};
	return t.distance_sq(vec3(p[0], p[1], p[2]));
}
void ref_triangle_closest_point(const float tri[9], const float p[3], float out[3]) {
	Triangle t{vec3(tri[0], tri[1], tri[2]), vec3(tri[3], tri[4], tri[5]), vec3(tri[6], tri[7], tri[8])};
	vec3 c = t.closest_point(vec3(p[0], p[1], p[2]));
	out[0] = c.x; out[1] = c.y; out[2] = c.z;
}

// raystab sign of the mesh SDF (triangle_bvh.cu:299-317,561-590): stab directions and the per-element random offset
void ref_fibonacci_dir32(uint32_t i, const float offset[2], float out[3]) {
	vec3 d = fibonacci_dir<32>(i, vec2(offset[0], offset[1]));
	out[0] = d.x; out[1] = d.y; out[2] = d.z;
}
void ref_default_rng_offset(uint64_t advance, float out[2]) {
	default_rng_t rng;
	rng.advance(advance);
	vec2 v = random_val_2d(rng);
	out[0] = v.x; out[1] = v.y;
}

// a-16 sampling helpers (random_val.cuh:62-72,130-185), all NGP_HOST_DEVICE
void ref_ld_random_val_2d(uint32_t index, uint32_t seed, float out[2]) { vec2 v = ld_random_val_2d(index, seed); out[0] = v.x; out[1] = v.y; }
void ref_dir_to_spherical_unorm(const float d[3], float out[2]) { vec2 v = dir_to_spherical_unorm(vec3(d[0], d[1], d[2])); out[0] = v.x; out[1] = v.y; }
void ref_cosine_hemisphere(const float u[2], float out[3]) { vec3 v = cosine_hemisphere(vec2(u[0], u[1])); out[0] = v.x; out[1] = v.y; out[2] = v.z; }
void ref_spherical_cap(const float u[2], float cap_height, float out[3]) {
	vec3 v = warp_square_to_spherical_cap_uniform(vec2(u[0], u[1]), cap_height); out[0] = v.x; out[1] = v.y; out[2] = v.z;
}
void ref_compute_local_frame(const float z[3], float out[9]) {
	mat3 f = compute_local_frame(vec3(z[0], z[1], z[2]));
	for (int c = 0; c < 3; ++c) { out[3 * c] = f[c].x; out[3 * c + 1] = f[c].y; out[3 * c + 2] = f[c].z; }
}

// a-21 loss functions (nerf_device.cuh:74-112,640-655): loss_type 0 = L2, 1 = Huber
void ref_loss_and_gradient(const float target[3], const float pred[3], int32_t loss_type, float out[6]) {
	LossAndGradient lg = loss_and_gradient(vec3(target[0], target[1], target[2]), vec3(pred[0], pred[1], pred[2]), (loss_type == 1 ? ELossType::Huber : loss_type == 2 ? ELossType::RelativeL2 : loss_type == 3 ? ELossType::L1 : loss_type == 4 ? ELossType::Mape : loss_type == 5 ? ELossType::Smape : loss_type == 6 ? ELossType::LogL1 : ELossType::L2));
	out[0] = lg.loss.x; out[1] = lg.loss.y; out[2] = lg.loss.z; out[3] = lg.gradient.x; out[4] = lg.gradient.y; out[5] = lg.gradient.z;
}

// pcg32 (tcnn/dependencies/pcg32): seed, skip-ahead, raw 32-bit outputs
uint32_t ref_pcg32(uint64_t seed, uint64_t advance, uint32_t n, uint32_t* out) {
	tcnn::pcg32 r{seed};
	r.advance((int64_t)advance);
	for (uint32_t i = 0; i < n; ++i) out[i] = r.next_uint();
	return n;
}

// ---- training error map: binary_search (common.h:242-264), sample_cdf_2d / image_idx (nerf_device.cuh:538-568, 617-639)
uint32_t ref_binary_search(float val, const float* data, uint32_t length) { return binary_search(val, data, length); }
void ref_sample_cdf_2d(float sample[2], uint32_t img, int32_t res_x, int32_t res_y, const float* cdf_x_cond_y, const float* cdf_y, float* pdf) {
	*pdf = 1.0f;
	vec2 r = sample_cdf_2d(vec2(sample[0], sample[1]), img, ivec2(res_x, res_y), cdf_x_cond_y, cdf_y, pdf);
	sample[0] = r.x; sample[1] = r.y;
}
uint32_t ref_image_idx(uint32_t base_idx, uint32_t n_rays, uint32_t n_images, const float* cdf, float* pdf) { return image_idx(base_idx, n_rays, 0, n_images, cdf, pdf); }

// ---- foveation (common_device.cuh:117-246): constructor (20-step bisection) + warp / unwarp / density
static Foveation make_fov(const float f[6]) { return Foveation(vec2(f[0], f[1]), vec2(f[2], f[3]), vec2(f[4], f[5])); }
void ref_foveation_coeffs(const float f[6], float out[24]) {
	Foveation fv = make_fov(f);
	const FoveationPiecewiseQuadratic* w[2] = {&fv.warp_x, &fv.warp_y};
	for (int k = 0; k < 2; ++k) {
		float* o = out + 12 * k;
		o[0] = w[k]->al; o[1] = w[k]->bl; o[2] = w[k]->cl; o[3] = w[k]->am; o[4] = w[k]->bm; o[5] = w[k]->ar; o[6] = w[k]->br; o[7] = w[k]->cr;
		o[8] = w[k]->switch_left; o[9] = w[k]->switch_right; o[10] = w[k]->inv_switch_left; o[11] = w[k]->inv_switch_right;
	}
}
void ref_foveation_warp(const float f[6], const float uv[2], float warped[2], float unwarped[2], float* density) {
	Foveation fv = make_fov(f);
	vec2 a = fv.warp(vec2(uv[0], uv[1])), b = fv.unwarp(vec2(uv[0], uv[1]));
	warped[0] = a.x; warped[1] = a.y; unwarped[0] = b.x; unwarped[1] = b.y; *density = fv.density(vec2(uv[0], uv[1]));
}
// uv_to_ray (common_device.cuh:416-483) with everything: foveation, hidden-area mask, lens, distortion grid
int32_t ref_uv_to_ray_full(uint32_t spp, const float uv[2], int32_t w, int32_t h, const float focal[2], const float cam[12], const float screen_center[2],
	const float parallax[3], float near_distance, float focus_z, float aperture_size, const float fov[6], const uint8_t* mask, int32_t mask_w, int32_t mask_h,
	int32_t lens_mode, const float lens_params[7], const float* distortion, int32_t dist_w, int32_t dist_h, float out_o[3], float out_d[3]) {
	mat4x3 c;
	for (int i = 0; i < 4; ++i) c[i] = vec3(cam[3*i], cam[3*i+1], cam[3*i+2]);
	Lens lens;
	lens.mode = (ELensMode)lens_mode;
	for (int i = 0; i < 7; ++i) lens.params[i] = lens_params[i];
	Buffer2DView<const uint8_t> m; m.data = mask; m.resolution = ivec2(mask_w, mask_h);
	Buffer2DView<const vec2> d; d.data = (const vec2*)distortion; d.resolution = ivec2(dist_w, dist_h);
	Ray ray = uv_to_ray(spp, vec2(uv[0], uv[1]), ivec2(w, h), vec2(focal[0], focal[1]), c, vec2(screen_center[0], screen_center[1]),
		vec3(parallax[0], parallax[1], parallax[2]), near_distance, focus_z, aperture_size, make_fov(fov), m, lens, d);
	out_o[0] = ray.o.x; out_o[1] = ray.o.y; out_o[2] = ray.o.z;
	out_d[0] = ray.d.x; out_d[1] = ray.d.y; out_d[2] = ray.d.z;
	return ray.is_valid() ? 1 : 0;
}
// pos_to_uv (common_device.cuh:505-537) and the colour of ERenderMode::Distortion (to_rgb, :737-739)
void ref_pos_to_uv(const float pos[3], int32_t w, int32_t h, const float focal[2], const float cam[12], const float screen_center[2], const float parallax[3],
	const float fov[6], int32_t lens_mode, const float lens_params[7], float out[2]) {
	mat4x3 c;
	for (int i = 0; i < 4; ++i) c[i] = vec3(cam[3*i], cam[3*i+1], cam[3*i+2]);
	Lens lens;
	lens.mode = (ELensMode)lens_mode;
	for (int i = 0; i < 7; ++i) lens.params[i] = lens_params[i];
	vec2 r = pos_to_uv(vec3(pos[0], pos[1], pos[2]), ivec2(w, h), vec2(focal[0], focal[1]), c, vec2(screen_center[0], screen_center[1]),
		vec3(parallax[0], parallax[1], parallax[2]), make_fov(fov), lens);
	out[0] = r.x; out[1] = r.y;
}
void ref_to_rgb(const float dir[2], float out[3]) { vec3 c = to_rgb(vec2(dir[0], dir[1])); out[0] = c.x; out[1] = c.y; out[2] = c.z; }

}
