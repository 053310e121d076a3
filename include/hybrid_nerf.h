/* hybrid_nerf.h — C ABI of the B200-native hybrid NeRF + mesh render path.
 *
 * Drop-in boundary for the render path of ngp::Testbed (reference: include/neural-graphics-primitives/testbed.h).
 * Plain pointers and sizes only; no C++/torch types. Every call returns 0 on success, non-zero on failure;
 * hn_last_error() gives the message (the reference throws std::runtime_error instead, src/main.cu:209-223).
 * The library has no CPU fallback: every entry point that computes needs a CUDA device (sm_100a).
 *
 * Conventions (identical to the reference):
 *  - camera: float[12] = mat4x3, 4 columns x vec3 {right, down, forward, position} in NeRF space
 *  - frame buffer: float4 per pixel, row-major idx = x + W*y, premultiplied linear RGBA (CudaRenderBufferView,
 *    render_buffer.h:162-171); depth buffer: float per pixel, view-space z, 16384 = nothing
 *  - parameter blob: fp16 [density MLP | rgb MLP | hash grid] (nerf_network.h:356-371), MLP weights row-major [out][in]
 */
#ifndef HYBRID_NERF_H
#define HYBRID_NERF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hn_context hn_context;

/* ---- lifetime / errors ---- */
int hn_create(hn_context** out, int cuda_device);            /* replaces Testbed::Testbed(ETestbedMode), testbed.h:69 */
void hn_destroy(hn_context* ctx);
const char* hn_last_error(const hn_context* ctx);             /* replaces exceptions across the boundary */
const char* hn_version(void);

/* ---- network configuration (configs/nerf/*.json, configs/geometry/*.json incl. "parent" merge) ---- */
int hn_load_network_config(hn_context* ctx, const char* json_path);        /* Testbed::reload_network_from_file, testbed.h:607 */
int hn_load_network_config_string(hn_context* ctx, const char* json_text); /* Testbed::reload_network_from_json, testbed.h:608 */
int hn_set_aabb_scale(hn_context* ctx, int aabb_scale);                    /* nerf.training.dataset.aabb_scale (drives per_level_scale, max_cascade, cone angle) */

typedef struct {
	uint32_t n_levels, n_features_per_level, log2_hashmap_size, base_resolution;
	float per_level_scale;
	uint32_t density_hidden_layers, rgb_hidden_layers;
	uint64_t n_params;            /* fp16 elements in the blob */
	uint64_t grid_param_offset;   /* element offset of the hash table */
	uint32_t level_offsets[17];   /* GridOffsetTable, tcnn grid.h:692-721 */
	float level_scales[16];
	uint32_t level_resolutions[16];
} hn_network_info;
int hn_get_network_info(const hn_context* ctx, hn_network_info* out);

/* dtype: 0 = fp16, 1 = fp32 (converted to fp16 like Trainer's inference copy). Host pointer. */
int hn_set_params(hn_context* ctx, const void* blob, size_t n_elements, int dtype);   /* Trainer::set_params / deserialize, tcnn trainer.h:275-351 */

/* Snapshots (.ingp = zlib-wrapped MessagePack, .msgpack = plain): network config + fp16/fp32 parameters + fp16 density grid +
 * aabb / camera / sun / exposure, the reference's own file format (Testbed::load_snapshot / save_snapshot,
 * src/testbed.cu:5219-5488, testbed.h:780-782; Trainer::serialize, tcnn trainer.h:275-322). */
int hn_load_snapshot(hn_context* ctx, const char* ingp_or_msgpack_path);
int hn_save_snapshot(hn_context* ctx, const char* ingp_or_msgpack_path);

/* ---- occupancy (a-20) ---- */
/* grid: float[128^3 * n_cascades] in Morton order (host). Builds the 8-mip bitfield on the device. */
int hn_set_density_grid(hn_context* ctx, const float* grid, uint32_t n_cascades);     /* update_density_grid_mean_and_bitfield, src/testbed_nerf.cu:2863-2880 */
int hn_set_occupancy_bitfield(hn_context* ctx, const uint8_t* bits, size_t n_bytes);  /* 128^3/8*8 bytes */
int hn_get_occupancy_bitfield(hn_context* ctx, uint8_t* out, size_t n_bytes);

/* ---- scene (data/geometry/*.json: {"geometry":[{"center","path","type":"Mesh"|"Nerf"}]}) ---- */
int hn_load_scene(hn_context* ctx, const char* scene_json_path);                      /* Testbed::load_scene, src/testbed_geometry_training.cu:3101-3210 */
/* triangles: float[n][9] host. normalize != 0 applies load_mesh's unit-cube normalisation + centre offset (:2805-2823). */
int hn_add_mesh(hn_context* ctx, const float* triangles, size_t n_triangles, const float center[3], int normalize);
int hn_load_mesh_file(hn_context* ctx, const char* obj_or_stl_path, const float center[3]); /* Testbed::load_mesh, :2786-2866 */
int hn_build_scene(hn_context* ctx);                                                  /* GeometryBvh::build_mesh + m_aabb = root inflated by 4 (:3180-3189) */
int hn_clear_scene(hn_context* ctx);
int hn_get_scene_aabb(const hn_context* ctx, float out[6]);                           /* root box of all meshes (before inflation) */

/* ---- environment map (a-16) ---- */
/* rgba: float[height][width][4] HOST, lat-long (read_envmap, envmap.cuh:24-50; Testbed::m_envmap, testbed.h:1227-1241).
 * Background of camera rays (src/testbed_nerf.cu:1526-1528) and, with option "env_light_samples" = K > 0, light source of the
 * mesh pass: K cosine-weighted light rays per mesh hit, each a shadow ray through the BVH and a transmittance march through
 * the NeRF (the reference declares computeEnvmap / shade_from_nerf, testbed.h:394-407,709-743, without defining them).
 * rgba == NULL removes the map. */
int hn_set_envmap(hn_context* ctx, const float* rgba, int width, int height);

/* ---- VR pieces of uv_to_ray (common_device.cuh:117-246,416-483) and the 2-D render modes ----
 * option "foveation"[6] = Foveation(center_pixel_steepness xy, center_inverse_piecewise_y xy, center_radius xy) of the view
 * (Testbed::m_views[i].foveation, src/testbed.cu:3150-3164); all zero = none. Camera rays of the NeRF pass and of the mesh pass are
 * generated at foveation.warp(uv); the frame that comes out is the WARPED one, hn_unwarp_foveated resamples it.
 * hn_set_hidden_area_mask: render_buffer.hidden_area_mask (uint8 [height][width], 0 = no ray for that pixel); NULL removes it.
 * hn_set_distortion_map: m_distortion.inference_view() with m_nerf.render_with_lens_distortion (vec2 [height][width], added to the
 * camera-space direction of NeRF rays, bilinear); NULL removes it.
 * render_mode 8 (Distortion, src/testbed_nerf.cu:1513-1521) and 10 (Slice, :1500-1510,2431-2459; plane at camera depth "focus_z" =
 * m_slice_plane_z + m_scale) are 2-D modes of the NeRF pass: every valid pixel is shaded at once, nothing is traced. */
int hn_set_hidden_area_mask(hn_context* ctx, const uint8_t* mask, int width, int height);
int hn_set_distortion_map(hn_context* ctx, const float* grid, int width, int height);
/* Testbed::blit_texture + its fragment shader (src/testbed.cu:2600-2640,2664-2720) as a CUDA kernel: output pixel centre uv ->
 * foveation.unwarp(uv) -> bilinear fetch of the rendered (warped) frame / depth (device pointers; depth may be NULL). The GL y flip
 * of the shader is a window-system convention and is left out; the GL path itself is unavailable on this box. */
int hn_unwarp_foveated(hn_context* ctx, const float foveation[6], int width_in, int height_in, const float* frame_in_dev, const float* depth_in_dev,
	int width_out, int height_out, float* frame_out_dev, float* depth_out_dev, void* cuda_stream);
/* FoveationPiecewiseQuadratic's constructor (common_device.cuh:119-182) as the library runs it on the host: coefficients [2][12] =
 * al bl cl am bm ar br cr switch_left switch_right inv_switch_left inv_switch_right per axis */
int hn_op_foveation(hn_context* ctx, const float foveation[6], float coeffs[24]);

/* a-16 irradiance probes. What the reference ships of this path: the probe ray set-up NerfTracer::init_rays_from_center /
 * init_rays_from_center_outward / init_rays_from_multiple_center (src/testbed_nerf.cu:1924-2019, kernels :1559-1773,
 * cylindrical_to_dir_nerf :1546-1557, halton<2/3/5> random_val.cuh:338-350) and NerfTracer::trace_mesh (:2146-2262: t = 0, no start
 * jitter, the 200-iteration DDA). What it only declares (Testbed::computeEnvmap{,Multiple,MultipleMain,Grid}, NerfTracer::shade_from_nerf,
 * m_envmap_tex / gridSize: testbed.h:394-407,709-743,949-950) is defined HERE ("parity unpinned" for that part): the probe texture
 * [n_phi][n_theta] of premultiplied linear RGBA (texel (i, j) = ray idx i + n_theta * j; mode 2 averages n_origin^2 jittered centres),
 * its nearest-texel lookup by direction, and the mesh shading from it (options env_light_source = 1, env_light_sampling = 0 cosine /
 * 1 uniform spherical cap towards the NeRF, warp_square_to_spherical_cap_uniform random_val.cuh:144-165).
 * mode 0: rays from the centre of the NeRF box; 1: from `origin` looking in (frame around normalize(origin)); 2: Halton centres.
 * The texture stays on the device as the light of the mesh pass; rgba_host (optional) receives a copy. */
int hn_render_probe(hn_context* ctx, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float origin[3], float* rgba_host);
int hn_set_probe(hn_context* ctx, const float* rgba /* [n_phi][n_theta][4] or NULL */, uint32_t n_theta, uint32_t n_phi);
/* the probe rays alone (origins, unit directions, alive = the DDA found an occupied cell), n = n_theta*n_phi (* n_origin^2 in mode 2) */
int hn_op_probe_rays(hn_context* ctx, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float origin[3],
	float* origins, float* dirs, uint8_t* alive);
/* a-16 operator entry points (host in/out): envmap lookup for n unit directions -> out[n][4];
 * the K light directions of pixel `pixel_index` at sample index spp around unit normal n -> out[K][3];
 * NeRF transmittance along n rays (origin, unit dir) -> T[n], samples taken [n] (may be NULL). */
int hn_op_read_envmap(hn_context* ctx, const float* dirs, uint32_t n, float* out);
int hn_op_env_light_dirs(hn_context* ctx, uint32_t pixel_index, uint32_t spp_index, uint32_t K, const float normal[3], float* out);
int hn_op_transmittance(hn_context* ctx, uint32_t n_rays, const float* origins, const float* dirs, float* T, uint32_t* n_samples);

/* ---- training step (a-21; Testbed::train -> train_nerf -> train_nerf_step, src/testbed_nerf.cu:2929-3075,3183-3430;
 *      tcnn Trainer::training_step, trainer.h:163-190; optimizer / loss from the network config JSON) ---- */
/* In-memory dataset: n images of w x h, float RGBA (linear, straight alpha; NerfDataset pixels, src/nerf_loader.cu),
 * one camera mat4x3 each (float[12], same convention as the render camera), shared pinhole intrinsics
 * (focal length in pixels, principal point in [0,1]^2). Host pointers; copied to the device. */
int hn_train_set_dataset(hn_context* ctx, const float* rgba, uint32_t n_images, int width, int height, const float* cameras,
	const float focal_px[2], const float principal[2]);
/* Testbed::load_training_data(path) (testbed.h:600, src/testbed_nerf.cu:2674-2716 -> load_nerf, src/nerf_loader.cu:273-770):
 * a transforms.json (or the folder holding one) with 8-bit PNG or baseline-JPEG frames: camera matrices through nerf_matrix_to_ngp (scale 0.33,
 * offset 0.5), aabb_scale; focal length from x_fov | fl_x | camera_angle_x (and y), principal point cx / w, cy / h and the lens
 * (read_lens, :175-241: k1 k2 p1 p2 / is_fisheye k1..k4, ftheta_p0..4, latlong, equirectangular) at file level AND per frame, the
 * frame's values over the file's (:678-696). One resolution for all frames; JPEG / EXR, depth or ray files and rolling-shutter
 * frame pairs are errors / ignored. Call it before hn_set_params when aabb_scale != 1: a new aabb_scale re-derives the network layout. */
int hn_load_training_data(hn_context* ctx, const char* transforms_json_or_dir);
/* The loader's image decoder on its own (stbi_load(path, &w, &h, &comp, 4) in src/nerf_loader.cu:575-590): 8-bit PNG (grey, grey + alpha,
 * RGB, RGBA, non-interlaced) or baseline / extended-sequential JPEG (grey, YCbCr, Adobe RGB; any sampling up to 4:1, restart intervals)
 * -> RGBA8, row-major. Needs no context and no GPU. rgba may be NULL to query the size; on failure returns 1 and writes the reason to
 * `error`. EXR and progressive JPEG are not read. */
int hn_decode_image_file(const char* path, int* width, int* height, uint8_t* rgba, size_t capacity_bytes, char* error, size_t error_capacity);
/* TrainingImageMetadata of every image after hn_train_set_dataset (which gives all images one pinhole): focal_px [n][2], principal [n][2]
 * (may be NULL = keep), lens_mode [n] = ELensMode 0..5 (may be NULL), lens_params [n][7] (may be NULL). Training rays go through the
 * image's own lens (generate_training_samples_nerf, src/testbed_nerf.cu:806-808); mark_untrained_density_grid projects with it (:111-139). */
int hn_train_set_image_metadata(hn_context* ctx, uint32_t n_images, const float* focal_px, const float* principal, const int32_t* lens_mode, const float* lens_params);
/* (Re)creates the trainer state from the current parameters: fp32 master copy, zeroed Adam moments, EMA, step counter
 * (Trainer ctor + Testbed::reset_network, src/testbed.cu:3933-4200). seed drives ray / pixel / background sampling. */
int hn_train_init(hn_context* ctx, uint32_t seed);
/* One batch: sample rays from the dataset, march, forward, loss (loss_scale 128, common.h:232), backward ->
 * fp32 gradient buffer [n_params] (scaled by loss_scale, summed over the batch). n_rays = 0: adaptive rays per batch
 * towards the target batch size of 2^18 samples (src/testbed_nerf.cu:2943-2944). rank/world shard the ray stream. */
int hn_train_accumulate(hn_context* ctx, uint32_t n_rays, float* loss_out);
/* Device pointer of the gradient buffer (for the data-parallel all-reduce between hn_train_accumulate and hn_train_apply). */
int hn_train_gradient_buffer(hn_context* ctx, void** dev_ptr, size_t* n_floats);
/* Optimizer step on the gradient buffer: Adam (+ ExponentialDecay, + EMA) as configured; refreshes the fp16 inference copy.
 * grad_scale multiplies the gradients first (1/world after a sum all-reduce). */
int hn_train_apply(hn_context* ctx, float grad_scale);
/* Put the training / set-up calls on the caller's stream (the one its collectives use): accumulate -> all-reduce -> apply then needs
 * no host synchronisation in between. NULL = back to the context's own stream. */
int hn_set_stream(hn_context* ctx, void* cuda_stream);
/* n_steps x (accumulate + apply) on one GPU; loss_out = mean loss of the last step (Testbed::train + m_loss_scalar). */
int hn_train_step(hn_context* ctx, uint32_t n_steps, float* loss_out);
/* Occupancy-grid maintenance from the live network (Testbed::update_density_grid_nerf, src/testbed_nerf.cu:2772-2861; called by
 * training_prep_nerf :3433-3446, which hn_train_accumulate does on the reference's schedule unless option "train_auto_grid" = 0):
 * n_uniform samples over all cells + n_nonuniform among occupied cells, density network, max-splat, grid = max(grid * decay, new),
 * then mean -> threshold -> bitfield + mips (update_density_grid_mean_and_bitfield :2863-2880). */
int hn_update_density_grid(hn_context* ctx, float decay, uint32_t n_uniform, uint32_t n_nonuniform);
/* mark_untrained_density_grid (:74-145): cells no training camera sees become -1 (never occupied), the rest 0. */
int hn_mark_untrained_density_grid(hn_context* ctx);
int hn_get_density_grid(hn_context* ctx, float* out_host, size_t n_floats);     /* float[128^3 * (max_cascade+1)], Morton order */
typedef struct {
	uint32_t step, rays_per_batch, n_rays_last, n_samples_last;
	float loss_last, learning_rate;
	float ms_sample, ms_forward, ms_loss, ms_backward, ms_optimizer; /* filled when option "profile" = 1 */
} hn_train_stats;
int hn_get_train_stats(const hn_context* ctx, hn_train_stats* out);
/* operator entry point for the parity test: explicit rays (origin, unit dir, start t), targets and backgrounds [n][3] in the
 * training colour space -> mean loss, fp32 gradients [n_params] (host, scaled by loss_scale), samples used per ray.
 * loss_type (ELossType, nerf_device.cuh:74-142,640-655; also option "train_loss_type" and "loss"."otype" of the config):
 * 0 L2, 1 Huber(0.1)/5, 2 RelativeL2, 3 L1, 4 Mape, 5 Smape, 6 LogL1. */
int hn_op_train_gradients(hn_context* ctx, uint32_t n_rays, const float* origins, const float* dirs, const float* t_start,
	const float* targets, const float* backgrounds, int loss_type, float loss_scale, float* loss_out, float* grads_host, uint32_t* n_steps);
/* depth supervision (nerf.h:98,124; src/testbed_nerf.cu:1063-1067,1149-1154): hn_train_set_depth gives every training image a float depth image
 * [n][height][width] (metadata[img].depth: z-depth in scene units, <= 0 = no measurement; NULL removes them); with option
 * "depth_supervision_lambda" > 0 the loss kernel compares the composited ray depth with |un-normalised ray| x depth through
 * "depth_loss_type" (ELossType, default 3 = L1) and adds lambda x that gradient to dL/dsigma. The operator variant takes the per-ray targets. */
int hn_train_set_depth(hn_context* ctx, const float* depth, uint32_t n_images, int width, int height);
int hn_op_train_gradients_depth(hn_context* ctx, uint32_t n_rays, const float* origins, const float* dirs, const float* t_start, const float* targets, const float* backgrounds,
	const float* target_depth, float depth_lambda, int depth_loss_type, int loss_type, float loss_scale, float* loss_out, float* grads_host, uint32_t* n_steps);

/* ---- error map (nerf.h:104-118; src/testbed_nerf.cu:1080-1106,1776-1832,2986-3075; nerf_device.cuh:536-639) ----
 * hn_train_accumulate deposits every ray's mean loss into a coarse per-image map and rebuilds the sampling CDFs from it every
 * n_steps_between_error_map_updates steps (128, then x1.5 each time). Options "sample_focal_plane_proportional_to_error" and
 * "sample_image_proportional_to_error" (both 0 by default, as in the reference) let the CDFs pick the pixel / the image of a ray.
 * hn_get_error_map: m_nerf.training.error_map.{resolution,data}; res[2] = (x, y), data [n_images][res_y][res_x] (data may be NULL). */
int hn_get_error_map(hn_context* ctx, int res[2], uint32_t* n_images, float* data, size_t n_floats);
/* operators for the parity tests (host buffers): construct_cdf_2d + construct_cdf_1d + the host image CDF on `data`
 * [n_images][height][width] -> cdf_x_cond_y (same shape), cdf_y [n_images][height], cdf_img / pmf_img [n_images] (outputs may be NULL) */
int hn_op_error_map_cdf(hn_context* ctx, uint32_t n_images, uint32_t height, uint32_t width, const float* data, float* cdf_x_cond_y, float* cdf_y,
	float* cdf_img, float* pmf_img);
/* image_idx + nerf_random_image_pos_training as the training ray generator runs them for global ray ray_offset + i of step `step`
 * (pcg32(seed + step << 32) advanced 16 per ray; CDFs may be NULL = uniform): image, un-snapped uv [n][2], pdf = img_pdf * uv_pdf */
int hn_op_sample_training_pixels(hn_context* ctx, uint32_t n_rays, uint32_t ray_offset, uint32_t n_rays_total, uint32_t seed, uint32_t step, uint32_t n_images,
	int cdf_res_x, int cdf_res_y, const float* cdf_x_cond_y, const float* cdf_y, const float* cdf_img, uint32_t* img, float* uv, float* pdf);
/* the bilinear deposit of compute_loss_kernel_train_nerf (:1080-1106) of n values at (img, uv) into a zeroed map [n_images][res_y][res_x] */
int hn_op_error_map_deposit(hn_context* ctx, uint32_t n, const uint32_t* img, const float* uv, const float* val, uint32_t n_images, int res_x, int res_y,
	int img_w, int img_h, float* map);

/* ---- render state (public Testbed members; names follow src/python_api.cu:482-567) ---- */
int hn_set_option(hn_context* ctx, const char* name, const float* values, int n_values);
int hn_get_option(const hn_context* ctx, const char* name, float* values, int n_values);

/* ---- rendering ---- */
/* Testbed::render_frame(stream, cam0, cam1, prev, screen_center, relative_focal_length, rolling_shutter, ...) testbed.h:561-575
 * = clear + mesh pass + NeRF pass (render_frame_main, src/testbed.cu:4722-4894). frame_dev/depth_dev are DEVICE pointers
 * (W*H*4 floats / W*H floats). Asynchronous on `cuda_stream` (a cudaStream_t, NULL = default stream) except for the
 * marching loop's alive-ray read-backs. spp_index = CudaRenderBufferView::spp. */
int hn_render_frame(hn_context* ctx, const float cam0[12], const float cam1[12], const float screen_center[2],
	const float relative_focal_length[2], const float rolling_shutter[4], uint32_t spp_index,
	int width, int height, float* frame_dev, float* depth_dev, void* cuda_stream);

/* pyngp.Testbed.render(width, height, spp, linear) -> float32[h][w][4] (src/python_api.cu:124-202, 417-426):
 * spp passes accumulated (render_buffer.cu:228), optional sRGB tonemap (linear == 0), result copied to HOST memory. */
int hn_render_to_host(hn_context* ctx, int width, int height, int spp, int linear, float* rgba_host, float* depth_host);
int hn_set_camera(hn_context* ctx, const float cam[12]);                              /* set_nerf_camera_matrix, python_api.cu:443 */

/* multi-GPU: this process renders only the screen tiles t with t % world == rank (tile_w x tile_h pixels). */
int hn_set_tile_shard(hn_context* ctx, int rank, int world, int tile_w, int tile_h);
/* pack this rank's tiles of (frame, depth) into a contiguous buffer [n_tiles][tile_h*tile_w][5 floats] and back. */
int hn_tile_count(const hn_context* ctx, int width, int height, int rank, int world, uint32_t* n_tiles);
int hn_pack_tiles(hn_context* ctx, int width, int height, int rank, int world, const float* frame_dev, const float* depth_dev, float* packed_dev, void* cuda_stream);
int hn_unpack_tiles(hn_context* ctx, int width, int height, int rank, int world, const float* packed_dev, float* frame_dev, float* depth_dev, void* cuda_stream);

/* Frame gather over peer memory (NVLink / NVSwitch) without a collective, one process per GPU. The reference moves whole frames
 * between its devices with cudaMemcpyPeerAsync (Testbed::sync_device / use_device, src/testbed.cu:5523-5610); here the root rank
 * allocates frame + depth through hn_shared_alloc and publishes the 64-byte CUDA IPC handles, every other rank maps them once with
 * hn_shared_open and per frame copies the tiles it owns straight into them (hn_scatter_tiles: plain peer stores, one launch), then
 * all ranks pass one barrier. dev_ptr from hn_shared_alloc is an ordinary device pointer on the root (render into it directly). */
int hn_shared_alloc(hn_context* ctx, size_t bytes, void** dev_ptr, unsigned char handle[64]);
int hn_shared_open(hn_context* ctx, const unsigned char handle[64], void** dev_ptr);
int hn_shared_release(hn_context* ctx, void* dev_ptr, int owner);
int hn_scatter_tiles(hn_context* ctx, int width, int height, int rank, int world, const float* frame_dev, const float* depth_dev,
	float* frame_remote, float* depth_remote, void* cuda_stream);
/* The same scatter with HOST memory as its target - the sharded counterpart of render_to_cpu's device -> host copy
 * (src/python_api.cu:124-202). host_ptr: `bytes` of host memory every rank's process has mapped (one POSIX shared-memory segment);
 * hn_host_register page-locks it and maps it into this process's device address space, the pointer returned is a valid frame_remote /
 * depth_remote (depth_remote may be NULL: colour only) of hn_scatter_tiles. Every rank then writes the tiles it owns straight into host
 * memory over its own PCIe link, instead of the root rank reading the whole gathered frame back through one link. The stores are
 * complete on the host once the scattering stream has been synchronised. */
int hn_host_register(hn_context* ctx, void* host_ptr, size_t bytes, void** dev_ptr);
int hn_host_unregister(hn_context* ctx, void* host_ptr);

typedef struct {
	uint64_t n_samples;        /* network samples evaluated in the last frame (ERenderMode::Cost sum) */
	uint64_t n_samples_composited;
	uint32_t n_rays_alive;     /* rays that entered the volume */
	uint32_t n_passes;         /* march/eval/composite rounds */
	uint32_t n_kernel_launches;
	float ms_mesh, ms_raygen, ms_march, ms_eval, ms_composite, ms_total; /* filled when option "profile" = 1 */
	/* a-16 environment light pass */
	float ms_envlight;         /* whole pass (generation + shadow rays + volume march + reduce) */
	uint32_t n_env_hits;       /* mesh hits that gathered light */
	uint64_t n_light_rays;     /* light rays traced through the BVH */
	uint64_t n_light_rays_volume; /* of those, rays that entered the NeRF volume */
	uint64_t n_light_samples;  /* density-network samples evaluated for them */
	uint32_t n_bvh_stack_overflows; /* BVH children dropped because the traversal stack was full, since hn_build_scene: must be 0
	                                   (the reference prints "WARNING TOO BIG", triangle_bvh.cuh:139-144) */
	uint32_t pad_;
} hn_frame_stats;
int hn_get_frame_stats(const hn_context* ctx, hn_frame_stats* out);

/* ---- operator-level entry points (the reference's kernels taken one at a time; used by the parity tests) ---- */
/* a-4..a-6: per pixel origin[3], dir[3], t, alive after init_rays + advance_pos (host outputs). */
int hn_op_init_rays(hn_context* ctx, const float cam0[12], const float cam1[12], const float screen_center[2], const float relative_focal_length[2],
	const float rolling_shutter[4], uint32_t spp_index, int width, int height, float* origins, float* dirs, float* t, uint8_t* alive);
/* a-7/a-8: march every ray to the end of the box without termination; counts[n], samples [n][max_per_ray][4] = x,y,z,dt (host). */
int hn_op_march_all(hn_context* ctx, uint32_t n_rays, const float* origins, const float* dirs, const float* t0, const uint8_t* alive,
	uint32_t* counts, float* samples, uint32_t max_per_ray);
/* a-10: hash-grid encoding of n positions in [0,1]^3 -> fp16 [n][32] (host). */
int hn_op_encode(hn_context* ctx, const float* positions, uint32_t n, uint16_t* out);
/* hn_op_encode without any device allocation. No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Same arguments and the
 * same (bit-identical) result as hn_op_encode; the device staging is a static __device__ array of the library (4.75 MiB per GPU,
 * held until the process exits) instead of the context's growable buffers. n > HN_NOALLOC_MAX_POSITIONS is rejected (returns 1,
 * hn_last_error says why). The array is shared by every call and context on a device and keeps the last call's contents: calls
 * that overlap on the GPU race - serialise them. Calls no cudaMalloc / cudaFree / allocating library routine. (csrc/hn_noalloc.cu) */
#define HN_NOALLOC_MAX_POSITIONS 65536u
int hn_op_encode_noalloc(hn_context* ctx, const float* positions, uint32_t n, uint16_t* out);
/* a-10..a-13: samples come in ray chunks of 32 sharing one direction. positions [n][4] = un-warped x,y,z,dt (n % 32 == 0),
 * dirs [n/32][3] unit vectors; out fp16 [n][4] raw (r,g,b,sigma). */
int hn_op_network(hn_context* ctx, const float* positions, const float* dirs, uint32_t n, uint16_t* out);
/* a-17/a-18: nearest hit per ray: mesh index, ORIGINAL triangle index, t (100 = miss) (host in/out). */
/* Distance from each position to the nearest triangle of the scene, through the BVH (TriangleBvh::signed_distance_gpu,
 * src/triangle_bvh.cu:195-303,330-376; include/neural-graphics-primitives/triangle.cuh:66-147). mode 0 = unsigned_distance,
 * mode 1 = signed_distance_watertight (sign of dot(average normal around the closest point, p - closest point)), mode 2 =
 * signed_distance_raystab (32 Fibonacci stab rays with the kernel's pcg32 offset; inside when none escapes); path-escape
 * (OptiX only in the reference) is not built. max_distance bounds the search (the reference's MAX_DIST); tri_idx (optional) receives the
 * original index of the nearest triangle inside its mesh, -1 when nothing lies within max_distance. Host pointers. */
int hn_op_mesh_distance(hn_context* ctx, uint32_t n, const float* positions, int mode, float max_distance, float* distances, int32_t* tri_idx);
int hn_op_trace(hn_context* ctx, uint32_t n_rays, const float* origins, const float* dirs, int32_t* mesh_idx, int32_t* tri_idx, float* t);

#ifdef __cplusplus
}
#endif
#endif
