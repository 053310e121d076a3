// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h header). C interface used by tests/ and bench.py via ctypes.
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

// Network shape (configs/nerf/base.json, configs/geometry/base.json). MLP width is 64, outputs 16-padded.
typedef struct {
	uint32_t n_levels;
	uint32_t n_features_per_level;
	uint32_t log2_hashmap_size;
	uint32_t base_resolution;
	float per_level_scale;
	uint32_t density_hidden_layers; // "network.n_hidden_layers"
	uint32_t rgb_hidden_layers;     // "rgb_network.n_hidden_layers"
	// The reference evaluates grid_scale() on the GPU under --use_fast_math (MUFU.EX2 + one FFMA, see csrc/hn_config.cpp
	// level_scale_kernel); a CPU cannot reproduce MUFU.EX2 bit for bit, so the per-level scales the device computed can be
	// handed in. 0 = derive them on the host with exp2f (exact for power-of-two per_level_scale, e.g. nerf/base.json).
	uint32_t has_device_scales;
	float device_scales[16];
} orc_net_config;

typedef struct {
	int32_t width, height;
	uint32_t spp; // sample index (render_buffer.spp)
	float cam0[12], cam1[12];
	float rolling_shutter[4];
	float screen_center[2];
	float focal[2]; // pixels
	int32_t snap_to_pixel_centers;
	float near_distance;
	// NeRF
	float render_aabb[6];
	float train_aabb[6];
	uint32_t max_mip;
	float cone_angle_constant;
	float min_transmittance;
	int32_t rgb_activation;     // 0 None 1 ReLU 2 Logistic 3 Exponential (ENerfActivation order)
	int32_t density_activation;
	int32_t linear_colors;      // nerf.training.linear_colors
	int32_t dda_iter_cap;       // 0 = unlimited (nerf_device.cuh:461), 200 = the _geometry variant (:497)
	// Mesh
	float mesh_aabb[6];
	float sun_dir[3];
	float up_dir[3];
	float floor_y;
	float brdf[15]; // metallic subsurface specular roughness sheen clearcoat clearcoat_gloss basecolor[3] ambientcolor[3] pad[2]
	int32_t compat_nearest_aabb_only; // 1 = reference top-level rule (geometry_bvh.cu:166-200)
	int32_t depth_merge_clip;         // 0 = reference per-pixel depth test, 1 = stop marching at the mesh depth
	int32_t n_threads;                // host threads for the frame loops (cpu baseline)
	// a-16 environment map (envmap.cuh:24-50): lat-long RGBA32F, row-major [env_h][env_w][4]; NULL = none.
	// env_samples = K cosine-weighted light directions per mesh hit (0 = envmap is background only).
	int32_t env_w, env_h, env_samples, env_pad;
	const float* envmap;
	// ERenderMode of the NeRF pass (common.h:58-72; 0 AO, 1 Shade, 6 Positions, 7 Depth, 9 Cost, 12 EncodingVis) and 1 / dataset scale
	int32_t render_mode; float depth_scale;
	// camera model of uv_to_ray (common_device.cuh:416-483): ELensMode + params (NeRF ray init only), depth of field, parallax shift
	int32_t lens_mode; float lens_params[7];
	float aperture_size, focus_z;
	float parallax_shift[3];
	// m_render_aabb_to_local (column-major mat3): crop box rotation, applied when crop_rotated != 0 (src/testbed_nerf.cu:1530-1532)
	int32_t crop_rotated; float render_aabb_to_local[9];
	// optional: uint8[width*height], only pixels with a non-zero entry are rendered by orc_render_frame / orc_render_nerf /
	// orc_render_mesh (the others keep the cleared value). Lets a test check a stratified subset of a full-size frame.
	int32_t mask_pad;
	const uint8_t* pixel_mask;
	// training: mean of max(density grid, 0) over cascade 0 (m_nerf.density_grid_mean); below NERF_MIN_OPTICAL_THICKNESS (0.01) the
	// loss adds the unscaled L1 density regulariser (src/testbed_nerf.cu:1113,1161-1164). 0 here = "not given" = regulariser off.
	float density_grid_mean; int32_t pad2;
	// a-16 probe as light of the mesh pass (OUR definition): env_source 1 = probe texture [probe_n_phi][probe_n_theta][4] instead of the
	// environment map (no NeRF transmittance along the light rays); env_sampling 0 = cosine hemisphere, 1 = uniform spherical cap around
	// the direction to the bounding sphere of the NeRF box (train_aabb)
	int32_t env_source, env_sampling, probe_n_theta, probe_n_phi;
	const float* probe;
	// VR / 2-D pieces of uv_to_ray (common_device.cuh:117-246,416-483): foveation = center_pixel_steepness xy, center_inverse_piecewise_y xy,
	// center_radius xy (all zero = none); hidden-area mask uint8 [hidden_h][hidden_w] (NULL = none, 0 = hidden); distortion grid of vec2
	// [distortion_h][distortion_w][2] (NULL = none, NeRF rays only). render_mode 8 = Distortion, 10 = Slice (plane at camera depth focus_z).
	float foveation[6];
	int32_t hidden_w, hidden_h;
	const uint8_t* hidden_area_mask;
	int32_t distortion_w, distortion_h;
	const float* distortion;
} orc_render_params;

// ---- scalar pieces (pinned against oracle/_ref) ----
float orc_ld_random_val(uint32_t index, uint32_t seed);
void orc_ld_random_pixel_offset(uint32_t spp, float out[2]);
void orc_bbox_ray_intersect(const float bb[6], const float o[3], const float d[3], float out[2]);
float orc_triangle_ray_intersect(const float tri[9], const float o[3], const float d[3]);
uint32_t orc_morton3D(uint32_t x, uint32_t y, uint32_t z);
float orc_advance_to_occupied(float t, float cone_angle, const float o[3], const float d[3], const uint8_t* bitfield,
	uint32_t min_mip, uint32_t max_mip, const float aabb[6], int32_t iter_cap);
float orc_calc_dt(float t, float cone_angle);
float orc_srgb_to_linear(float v);

// ---- a-16 sampling helpers (random_val.cuh:62-72,130-185; pinned against oracle/_ref) and the envmap lookup ----
// uv_to_ray with the lens / depth-of-field / parallax fields of p (direction not normalised); returns 0 for an invalid ray
int32_t orc_uv_to_ray(const orc_render_params* p, uint32_t spp, const float uv[2], int32_t use_lens, float out_o[3], float out_d[3]);
// foveation pieces and the 2-D helpers (orc_fov.inc): coefficients [2][12] (al bl cl am bm ar br cr switch_left switch_right inv_switch_left inv_switch_right)
void orc_foveation_coeffs(const float f[6], float out[24]);
void orc_foveation_warp(const float f[6], const float uv[2], float warped[2], float unwarped[2], float* density);
void orc_pos_to_uv(const orc_render_params* p, const float pos[3], float out[2]);
void orc_to_rgb(const float dir[2], float out[3]);
// tonemap_kernel (src/render_buffer.cu:529-560) on a linear accumulation buffer; curve = ETonemapCurve (0 Identity, 1 ACES, 2 Hable, 3 Reinhard)
float orc_density_input_gradient(const orc_net_config* c, const uint16_t* params, const float wp[3], float grad[3]);
float orc_density_raw_f32(const orc_net_config* c, const uint16_t* params, const float wp[3]);
void orc_tonemap(const float* accum, uint32_t n, int32_t to_srgb, float exposure, const float bg[4], int32_t curve, int32_t clamp_output,
	int32_t unmultiply_alpha, float* out);
void orc_ld_random_val_2d(uint32_t index, uint32_t seed, float out[2]);
void orc_dir_to_spherical_unorm(const float d[3], float out[2]);
void orc_cosine_hemisphere(const float u[2], float out[3]);
void orc_spherical_cap(const float u[2], float cap_height, float out[3]);
void orc_compute_local_frame(const float z[3], float out[9]); // 3 columns: X, Y, Z
void orc_read_envmap(const float* envmap, int32_t w, int32_t h, const float dir[3], float out[4]);
// K light directions of pixel `idx` at sample index spp around the unit normal n: out [K][3]
void orc_env_light_dirs(uint32_t idx, uint32_t spp, uint32_t K, const float n[3], float* out);
// transmittance of the NeRF along a ray (density-only network evaluation), as the env-light pass uses it
float orc_nerf_transmittance(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield,
	const float o[3], const float d[3], uint32_t* n_samples_out);

// ---- occupancy (a-20) ----
// grid: float[128^3 * n_cascades] Morton order. out: uint8[128^3/8 * 8]. Returns the threshold used.
float orc_bitfield_from_grid(const float* grid, uint32_t n_cascades, uint8_t* out);

// ---- ray generation (a-2..a-6) ----
// per pixel: origin[3], dir[3], t (after AABB entry, start jitter and advance to first occupied voxel), alive
void orc_init_rays(const orc_render_params* p, const uint8_t* bitfield, float* origins, float* dirs, float* t, uint8_t* alive);

// ---- marching without termination (a-7,a-8): all samples a ray emits until it leaves the box ----
// returns per-ray counts; if positions != NULL writes up to max_per_ray samples as {x,y,z,dt} (un-warped).
void orc_march_all(const orc_render_params* p, const uint8_t* bitfield, uint32_t n_rays, const float* origins, const float* dirs,
	const float* t0, const uint8_t* alive, uint32_t* counts, float* samples, uint32_t max_per_ray);

// ---- network (a-10..a-13) ----
size_t orc_n_params(const orc_net_config* c);            // number of fp16 parameters in the blob
size_t orc_grid_param_offset(const orc_net_config* c);   // offset of the hash table in the blob (elements)
void orc_grid_layout(const orc_net_config* c, uint32_t* offsets /*L+1*/, float* scales /*L*/, uint32_t* resolutions /*L*/);
void orc_encode(const orc_net_config* c, const uint16_t* params, const float* positions /*[n][3] in [0,1]*/, uint32_t n, uint16_t* out /*[n][32]*/);
void orc_sh4(const float* dirs01 /*[n][3] warped dirs in [0,1]*/, uint32_t n, uint16_t* out /*[n][16]*/);
// coords: [n][7] = warped pos, warped dt, warped dir. out: [n][4] fp16 raw network output (r,g,b,sigma) before activations.
void orc_network(const orc_net_config* c, const uint16_t* params, const float* coords, uint32_t n, uint16_t* out);

// ---- full NeRF pass (3.2) into an existing frame/depth buffer ----
// n_steps_out (optional, per pixel): number of samples composited.
uint64_t orc_render_nerf(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield,
	float* frame /*[h][w][4]*/, float* depth /*[h][w]*/, uint32_t* n_steps_out, int32_t depth_test);

// ---- a-16 irradiance probes (orc_probe.inc): probe ray set-up as the reference ships it (mode 0 centre, 1 looking in from `origin`,
// 2 Halton-jittered centres), NerfTracer::trace_mesh per ray, probe texture [n_phi][n_theta][4] (premultiplied linear RGBA)
void orc_cylindrical_to_dir(const float p[2], float out[3]);
float orc_halton(uint32_t base, uint64_t idx);
uint32_t orc_probe_rays(int32_t mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float box[6], const float origin[3], float* o_out, float* d_out, uint32_t* texel_out);
uint64_t orc_render_probe(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield, int32_t mode,
	uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float origin[3], float* tex);
void orc_read_probe(const float* tex, uint32_t n_theta, uint32_t n_phi, const float* dirs, uint32_t n, float* out);

// ---- training batch (a-21): loss + parameter gradients, fp32 forward / analytic fp32 backward (orc_train.inc) ----
// rays: origin, unit dir, start t (first sample search starts there); targets/backgrounds [n][3] in the training colour space.
// loss_type 0 = L2, 1 = Huber(0.1)/5. grads: fp32 [orc_n_params], scaled by loss_scale. Returns the mean loss over the rays.
void orc_loss_and_gradient(const float target[3], const float pred[3], int32_t loss_type, float out[6]);
float orc_train_gradients(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield, uint32_t n_rays,
	const float* origins, const float* dirs, const float* t_start, const float* targets, const float* backgrounds, int32_t loss_type, float loss_scale,
	uint32_t max_samples, float* grads, uint32_t* n_steps_out, uint32_t* n_samples_total);

// + depth supervision (src/testbed_nerf.cu:1063-1067,1149-1154): target_depth [n] = distance along the ray the ray should reach (<= 0: none),
// weighted by depth_lambda, compared with the composited ray depth through depth_loss_type (the reference's default is L1 = 3)
float orc_train_gradients_depth(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, const uint8_t* bitfield, uint32_t n_rays,
	const float* origins, const float* dirs, const float* t_start, const float* targets, const float* backgrounds, const float* target_depth, float depth_lambda,
	int32_t depth_loss_type, int32_t loss_type, float loss_scale, uint32_t max_samples, float* grads, uint32_t* n_steps_out, uint32_t* n_samples_total);

// ---- occupancy-grid maintenance from the live network (orc_grid.inc) ----
void orc_mark_untrained_grid(float* grid, uint32_t n_cascades, uint32_t n_images, const float* cams, int32_t w, int32_t h, const float focal[2], const float principal[2]);
// per-image focal [n][2] / principal [n][2] / ELensMode [n] / lens params [n][7] (TrainingImageMetadata); lens arrays may be NULL
void orc_mark_untrained_grid_per_image(float* grid, uint32_t n_cascades, uint32_t n_images, const float* cams, int32_t w, int32_t h, const float* focal, const float* principal,
	const int32_t* lens_mode, const float* lens_params);
void orc_update_density_grid(const orc_render_params* p, const orc_net_config* c, const uint16_t* params, float* grid, uint32_t n_cascades, uint32_t ema_step,
	uint32_t n_uniform, uint32_t n_nonuniform, float decay, uint32_t seed, uint32_t rng_calls);
uint32_t orc_pcg32(uint64_t seed, uint64_t advance, uint32_t n, uint32_t* out);

// ---- training error map (orc_errmap.inc) ----
uint32_t orc_binary_search(float val, const float* data, uint32_t length);
uint32_t orc_image_idx(uint32_t base_idx, uint32_t n_rays, uint32_t n_images, const float* cdf, float* pdf);
void orc_sample_cdf_2d(float sample[2], uint32_t img, int32_t res_x, int32_t res_y, const float* cdf_x_cond_y, const float* cdf_y, float* pdf);
void orc_error_map_cdf(uint32_t n_images, uint32_t height, uint32_t width, const float* data, float* cdf_x_cond_y, float* cdf_y, float* cdf_img, float* pmf_img);
void orc_sample_training_pixels(uint32_t n_rays, uint32_t ray_offset, uint32_t n_rays_total, uint32_t seed, uint32_t step, uint32_t n_images, int32_t res_x, int32_t res_y,
	const float* cdf_x_cond_y, const float* cdf_y, const float* cdf_img, uint32_t* img_out, float* uv_out, float* pdf_out);
void orc_error_map_deposit(uint32_t n, const uint32_t* img, const float* uv, const float* val, int32_t res_x, int32_t res_y, int32_t img_w, int32_t img_h, float* map);

// ---- meshes (a-17..a-19) ----
typedef struct orc_scene orc_scene;
orc_scene* orc_scene_create(void);
void orc_scene_destroy(orc_scene*);
// triangles: [n][9] already in scene space (normalisation is the loader's job).
void orc_scene_add_mesh(orc_scene*, const float* triangles, uint32_t n_tris, const float center[3]);
void orc_scene_build(orc_scene*);
void orc_scene_aabb(const orc_scene*, float out[6]);
// nearest hit per ray; brute = 1 tests every triangle (no BVH). t = MAX_DIST(100) and tri=-1 on miss.
// triangle.cuh:66-147 restated (pinned on the reference's host functions) and the brute-force scene distance (mode 0 unsigned, 1 watertight-signed, 2 raystab-signed)
void orc_triangle_distance_sq(const float tri[9], const float p[3], float* out);
void orc_triangle_closest_point(const float tri[9], const float p[3], float out[3]);
void orc_fibonacci_dir32(uint32_t i, const float offset[2], float out[3]);
void orc_default_rng_offset(uint64_t advance, float out[2]);
void orc_scene_distance(const orc_scene*, uint32_t n, const float* positions, int32_t mode, float max_distance, int32_t n_threads, float* out, int32_t* tri_out);
void orc_scene_trace(const orc_scene*, const orc_render_params* p, uint32_t n, const float* origins, const float* dirs,
	int32_t* mesh_idx, int32_t* tri_idx, float* t, int32_t brute);
// mesh pass (3.3): writes frame/depth.
void orc_render_mesh(const orc_scene*, const orc_render_params* p, float* frame, float* depth);

// ---- whole frame (3.1): clear, mesh pass (if scene), NeRF pass (if params) ----
uint64_t orc_render_frame(const orc_scene* scene, const orc_render_params* p, const orc_net_config* c, const uint16_t* params,
	const uint8_t* bitfield, float* frame, float* depth, uint32_t* n_steps_out);

#ifdef __cplusplus
}
#endif
