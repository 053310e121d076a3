// Host-side state behind the C ABI (the render-relevant subset of ngp::Testbed, testbed.h:67-1345).
#pragma once
#include "hn_internal.h"
#include "../../include/hybrid_nerf.h"
#include <stdexcept>
#include <string>
#include <vector>
#include <array>

namespace hn {

struct CudaError : std::runtime_error { using std::runtime_error::runtime_error; };
#define HN_CUDA(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) throw hn::CudaError(std::string(#expr) + " failed: " + cudaGetErrorString(e__)); } while (0)

template <typename T>
struct DevBuf {
	T* p = nullptr;
	size_t n = 0;
	void ensure(size_t count) {
		if (count <= n) return;
		release();
		HN_CUDA(cudaMalloc((void**)&p, count * sizeof(T)));
		n = count;
	}
	void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
	~DevBuf() { release(); }
	DevBuf() = default;
	DevBuf(const DevBuf&) = delete;
	DevBuf& operator=(const DevBuf&) = delete;
};

struct Tri { float a[3], b[3], c[3]; };

struct HostMesh {
	std::vector<Tri> tris;      // original order, scene space
	float center[3] = {0, 0, 0};
	float bb[6];                // exact box
	float mesh_scale = 1.f;
	// built BVH
	std::vector<Node4> nodes;   // local node indices, root = 0
	std::vector<uint32_t> order; // leaf order -> original triangle index
};

struct Options {
	// defaults follow testbed.h:852-1091 and nerf.h:150-186
	int snap_to_pixel_centers = 0;
	float near_distance = 0.0f;
	int lens_mode = 0;                          // m_nerf.render_lens.mode when render_with_lens_distortion (ELensMode, common.h:223-230)
	float lens_params[7] = {0, 0, 0, 0, 0, 0, 0};
	float aperture_size = 0.0f;                 // m_aperture_size
	float focus_z = 1.0f;                       // m_slice_plane_z + m_scale
	float parallax_shift[3] = {0, 0, 0};        // m_parallax_shift
	float foveation[6] = {0, 0, 0, 0, 0, 0};    // Foveation(center_pixel_steepness xy, center_inverse_piecewise_y xy, center_radius xy); all zero = Foveation{} (none)
	float render_aabb[6] = {0, 0, 0, 1, 1, 1};
	float render_aabb_to_local[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; // m_render_aabb_to_local, column-major (crop box rotation)
	float train_aabb[6] = {0, 0, 0, 1, 1, 1};   // m_geometry.nerfBoundingBox / m_aabb
	int max_cascade = 0;
	float cone_angle_constant = 0.0f;           // 0 for aabb_scale <= 1 (src/testbed_nerf.cu:2736), else 1/256
	float min_transmittance = 0.01f;
	int rgb_activation = 2;                     // Logistic
	int density_activation = 3;                 // Exponential
	int linear_colors = 0;
	int render_mode = 1;                        // ERenderMode::Shade (common.h:58-72)
	float depth_scale = 1.0f / 0.33f;           // 1 / nerf.training.dataset.scale (NERF_SCALE)
	int dda_iter_cap = 0;
	float sun_dir[3] = {0.57735026919f, 0.57735026919f, 0.57735026919f};
	float up_dir[3] = {0, 1, 0};
	int floor_enable = 0;
	float brdf[13] = {0.f, 0.f, 1.f, 0.5f, 0.f, 0.f, 0.f, 0.8f, 0.8f, 0.8f, 0.f, 0.f, 0.f};
	int compat_nearest_aabb_only = 0;
	int depth_merge_clip = 0;
	int render_nerf = 1, render_mesh = 1;
	float exposure = 0.f;
	int tonemap_curve = 0;                      // ETonemapCurve: 0 Identity, 1 ACES, 2 Hable, 3 Reinhard (render_buffer.cu:264-321)
	int clamp_output_color = 1, unmultiply_alpha = 0; // render_buffer.cu:552-558
	float background_color[4] = {0, 0, 0, 1};
	float fov_axis = 1;
	float relative_focal_length[2] = {1.057161212f, 1.057161212f}; // fov 50.625 deg on axis 1 (src/testbed.cu:541-568)
	float screen_center[2] = {0.5f, 0.5f};
	float zoom = 1.f;
	int profile = 0;
	int fuse_composite = 0;                     // 1 = composite inside the tcgen05 network kernel (eval_impl 1 only; measured slower, kept as an option)
	int composite_impl = 1;                     // 1 = one warp per ray + compact kernel, 0 = 32 rays per warp with in-kernel compaction
	int eval_gather = 2;                        // tcgen05 kernel, base layout: 1 / 2 = mask-only gather with / without paired fetches on hashed levels, 0 = generic gather
	int eval_texsplit = 0;                      // TEX / LSU split of the fine hash-grid levels in the tcgen05 kernel (see launch_eval_tc)
	int eval_groups = 4;                        // tile groups per CTA of the tcgen05 kernel (4 or 3)
	int eval_ray_sh = 1;                        // 1 = SH coefficients of a ray are computed once per frame (ray_sh_kernel), 0 = by every sample
	int eval_tex_gather = 1;                    // 1 = tcgen05 kernel gathers the fine hash-grid levels through the TEX pipe (F = 4), 0 = LSU only
	int eval_impl = 1;                          // 1 = tcgen05/TMEM network kernel (hn_network_tc.cu), 0 = warp-MMA kernel (hn_network.cu)
	int snapshot_optimizer_state = 0;           // hn_save_snapshot: include the Trainer's optimizer state (save_snapshot(path, include_optimizer_state))
	int env_light_source = 0;                   // light of the mesh pass: 0 = environment map (hn_set_envmap), 1 = the NeRF's irradiance probe (hn_render_probe)
	int env_light_sampling = 0;                 // 0 = cosine hemisphere around the normal, 1 = uniform spherical cap around the direction to the NeRF (probe source)
	int env_light_samples = 16;                 // K light rays per mesh hit when an envmap is set (0 = background only)
	int env_light_batch = 32 << 20;             // light rays in flight per batch (bounds the workspace: ~200 bytes per ray)
	int bvh_sah = 1;                            // BVH build: 1 = binned surface-area heuristic, 0 = the reference's median split (A/B); read by hn_build_scene
	int overlap_raygen = 1;                     // 1: hybrid frames set their NeRF rays up on a second stream while the mesh pass runs (A/B)
	int march_simple = 1;                       // 1: compile-time specialised raygen / march kernels when cone angle 0, one cascade, axis-aligned crop (A/B)
	int tight_t_far = 1;                        // 1: rays are cut off at the last flagged coarse occupancy cell (occ_t_far_tight), 0: at the box of the set cells (A/B)
	int mesh_aabb_user = 0;
	float mesh_aabb[6] = {0, 0, 0, 1, 1, 1};
};

// a-21 trainer state (tcnn Trainer + optimizer stack, trainer.h:37-360, optimizers/adam.h, ema.h, exponential_decay.h)
struct TrainCfg {
	int loss_type = 1;                      // 0 L2, 1 Huber (configs/nerf/base.json:2-4)
	float lr = 1e-2f, beta1 = 0.9f, beta2 = 0.99f, epsilon = 1e-15f, l2_reg = 1e-6f;
	bool ema = true; float ema_decay = 0.95f;
	bool exp_decay = true; uint32_t decay_start = 20000, decay_interval = 10000; float decay_base = 0.33f;
	float loss_scale = 128.0f;              // LOSS_SCALE(), common.h:232
	uint32_t target_batch = 1u << 18;       // samples per batch (src/testbed_nerf.cu:2943)
	int random_bg = 1;                      // nerf.training.random_bg_color
	int mlp_impl = 1;                       // MLP forward / backward: 1 = own tcgen05 kernels (hn_train_tc.cu), 0 = cuBLAS GEMMs + elementwise kernels (round-1 path, A/B)
	float depth_supervision_lambda = 0.0f;  // nerf.training.depth_supervision_lambda (nerf.h:124): weight of the depth term, needs hn_train_set_depth
	int depth_loss_type = 3;                // nerf.training.depth_loss_type (nerf.h:98): L1
	int sample_focal_plane_proportional_to_error = 0; // nerf.training.sample_focal_plane_proportional_to_error (nerf.h:112): pixels from the error-map CDF
	int sample_image_proportional_to_error = 0;       // nerf.training.sample_image_proportional_to_error (nerf.h:113): images from the error-map CDF
	int auto_grid = 1;                      // refresh the occupancy grid from the network on the reference's schedule (training_prep_nerf)
	float grid_decay = 0.95f;               // nerf.training.density_grid_decay (nerf.h:120)
};
struct TrainState {
	bool ready = false, has_dataset = false;
	int srgb_space = 0;                     // 1: images came from 8-bit sRGB files -> target blended in sRGB (EColorSpace::SRGB), 0: linear float images
	uint32_t seed = 1337, step = 0, ema_step = 0, rays_per_batch = 4096, n_rays_last = 0, n_samples_last = 0;
	float loss_last = 0.f, lr_last = 0.f;
	float ms[5] = {0, 0, 0, 0, 0};
	// dataset
	DevBuf<float4> images; DevBuf<float> cams;
	DevBuf<float> depth; bool has_depth = false; DevBuf<float> ray_target_depth; // depth supervision: [n][h][w] depth images, per-ray target distance
	DevBuf<ImageMeta> meta; std::vector<ImageMeta> meta_host; // per-image focal length, principal point, lens
	uint32_t n_images = 0; int img_w = 0, img_h = 0; float focal[2] = {0, 0}, principal[2] = {0.5f, 0.5f};
	// parameters
	DevBuf<float> master, m1, m2, ema, grads;
	DevBuf<uint32_t> param_steps;
	DevBuf<__half> raw;                     // fp16 copy of the raw (non-EMA) weights the training forward pass uses
	// batch workspace
	DevBuf<float> ray_o, ray_d, ray_t, ray_target, ray_bg;
	DevBuf<uint32_t> ray_numsteps, ray_base, counters;
	DevBuf<float4> coords;                  // warped position, dt per sample
	DevBuf<uint32_t> sample_ray;
	DevBuf<float4> march_staging;           // [n_rays][1024] per-ray staging rows of the single-walk sample generation
	DevBuf<__half> act[12];                 // activations: enc, density hidden..., density out, rgb in, rgb hidden..., out
	DevBuf<__half> dact[12];                // gradients: [0] dL/d(out), [1] dL/d(density out), [2] dL/d(enc), [3] dL/d(sigma) fp32, [4+l] colour hidden l, [8+l] density hidden l (cuBLAS path: [1],[2] ping-pong)
	DevBuf<float> loss_acc;
	// error map (nerf.h:104-118): accumulated per-image loss, its CDFs, and where / with which pdf each ray of the batch was drawn
	DevBuf<uint32_t> ray_img; DevBuf<float> ray_uv, ray_pdf;
	DevBuf<float> error_map, cdf_x_cond_y, cdf_y, cdf_img;
	int err_res[2] = {0, 0}, cdf_res[2] = {0, 0};
	bool cdf_valid = false;
	uint32_t n_steps_since_error_map_update = 0, n_steps_between_error_map_updates = 128;
	std::vector<float> pmf_img;             // error_map.pmf_img_cpu
	// occupancy-grid maintenance (update_density_grid_nerf)
	DevBuf<float> grid_tmp; DevBuf<float4> grid_pos; DevBuf<uint32_t> grid_idx; DevBuf<uint2> grid_out;
	uint32_t grid_seed = 1337, grid_rng_calls = 0, grid_ema_step = 0;
	void* cublas = nullptr;
};

} // namespace hn

struct hn_context {
	int device = 0;
	std::string error;
	cudaStream_t stream = nullptr; // stream of the non-frame calls: the context's own, or the caller's (hn_set_stream)
	cudaStream_t own_stream = nullptr;

	// network
	hn::NetCfg net;
	bool has_net = false;
	int aabb_scale = 1;
	std::string config_blob; // merged network config (MessagePack), kept so hn_set_aabb_scale can re-derive per_level_scale
	hn::GridLevels levels;
	hn::DevBuf<__half> params;
	hn::ParamTextures param_tex;                // texture views of `params` (TEX-pipe gather); rebuilt when the buffer moves
	std::vector<__half> params_host;        // fp16 inference copy on the host (snapshot writer)
	std::vector<float> density_grid_host;   // last density grid given to hn_set_density_grid (snapshot writer)
	bool has_params = false;

	// occupancy
	hn::DevBuf<uint8_t> bitfield;
	bool has_bitfield = false;
	hn::DevBuf<float4> probe_tex, probe_rays;   // a-16 probe: radiance texture [n_phi][n_theta] and the per-ray colours it was reduced from
	hn::DevBuf<float> probe_depth;
	int probe_n_theta = 0, probe_n_phi = 0;
	hn::DevBuf<uint4> ray_sh;               // [ray][2]: fp16 SH coefficients of the ray direction (colour MLP input), per frame
	hn::DevBuf<float> density_grid;         // device copy of the float grid [128^3 x grid_cascades] (Morton order), kept for the live updates
	uint32_t grid_cascades = 0;
	float density_grid_mean = 1.0f;         // m_nerf.density_grid_mean (mean of max(v, 0) over cascade 0); 1 = unknown (bitfield set directly)
	hn::DevBuf<uint32_t> occ_coarse;        // [CASCADES][1024] dilated 32^3 occupancy flags per mip (occ_t_far_tight), refreshed with the bitfield
	hn::DevBuf<float> ray_tfar;
	int occ_cells[hn::CASCADES * 6] = {};   // per-mip bounds of the set cells (cell coordinates), refreshed with the bitfield

	// scene
	std::vector<hn::HostMesh> meshes;
	bool scene_built = false;
	float scene_bb[6] = {0, 0, 0, 1, 1, 1};
	hn::DevBuf<hn::Node4> d_nodes;
	hn::DevBuf<float4> d_tris;
	hn::DevBuf<hn::MeshDev> d_meshes;

	// environment map (a-16)
	hn::DevBuf<uint8_t> hidden_mask; int hidden_w = 0, hidden_h = 0;   // render_buffer.hidden_area_mask (0 = hidden), empty = none
	hn::DevBuf<float2> distortion; int dist_w = 0, dist_h = 0;         // m_distortion.inference_view() (m_nerf.render_with_lens_distortion), empty = none
	hn::DevBuf<float> slice_dirs; hn::DevBuf<uint32_t> slice_pixels;   // ERenderMode::Slice workspace
	hn::DevBuf<float4> envmap;
	int env_w = 0, env_h = 0;
	hn::DevBuf<float4> hit_pos_pix, hit_nrm, hit_albedo;
	hn::DevBuf<uint32_t> hit_count;

	hn::TrainCfg train_cfg;
	hn::TrainState train;

	hn::Options opt;
	float camera[12] = {1, 0, 0, 0, -1, 0, 0, 0, -1, 0.5f, 0.5f, 2.0f}; // reset_camera default (src/testbed.cu:541-568)

	// tile shard
	int rank = 0, world = 1, tile_w = 64, tile_h = 32;

	// per-frame workspace (persistent, grows to the largest frame seen; no per-frame allocation)
	hn::DevBuf<float4> ray_o_t, ray_d_mw, ray_rgba;
	hn::DevBuf<float> ray_depth;
	hn::DevBuf<uint32_t> ray_pixel, ray_nsteps, alive_a, alive_b, sample_counts, counters;
	hn::DevBuf<float4> samples;
	hn::DevBuf<uint2> net_out;
	hn::DevBuf<uint32_t> survive;
	uint32_t* pinned_counters = nullptr; // ring of 4 x CNT_N uint32 (pass-loop read-backs)
	cudaEvent_t pass_events[4] = {nullptr, nullptr, nullptr, nullptr};
	cudaStream_t side_stream = nullptr;     // ray set-up of a hybrid frame runs here beside the mesh pass (option overlap_raygen)
	cudaEvent_t fork_event = nullptr, join_event = nullptr;
	// render_to_host buffers
	hn::DevBuf<float4> frame, accum, tonemapped;
	hn::DevBuf<float> depth;
	// scratch for operator entry points
	hn::DevBuf<uint8_t> scratch[8];

	hn_frame_stats stats = {};
	hn::DevBuf<uint32_t> d_stack_overflow; // see SceneDev::stack_overflow; zeroed when the scene is built, cumulative afterwards

	hn::SceneDev scene_dev() const {
		hn::SceneDev s;
		s.nodes = d_nodes.p; s.tris = d_tris.p; s.meshes = d_meshes.p; s.n_meshes = (uint32_t)meshes.size();
		s.stack_overflow = d_stack_overflow.p;
		return s;
	}
};

namespace hn {
// hn_config.cpp
struct Json;
void load_network_config_text(hn_context* ctx, const std::string& text, const std::string& base_dir);
void apply_network_config(hn_context* ctx, const Json& cfg);
void reapply_network_config(hn_context* ctx);
// hn_snapshot.cpp
Json msgpack_parse(const std::string& bytes);
std::string msgpack_write(const Json& j);
std::string zlib_inflate(const std::string& in);
std::string zlib_store(const std::string& in);
std::string gzip_inflate(const std::string& in);
std::string gzip_store(const std::string& in);
std::string snapshot_unwrap(const std::string& in); // gzip / zlib / plain, detected by magic like zstr::istream
void load_snapshot_file(hn_context* ctx, const std::string& path);
void save_snapshot_file(hn_context* ctx, const std::string& path);
// hn_api.cpp (shared by the snapshot loader)
void upload_params(hn_context* ctx, const void* blob, size_t n_elements, int dtype);
void upload_density_grid(hn_context* ctx, const float* grid, uint32_t n_cascades);
void finalize_network(hn_context* ctx);
// hn_train.cu
void train_set_dataset(hn_context* ctx, const float* rgba, uint32_t n, int w, int h, const float* cams, const float focal[2], const float principal[2]);
void train_init(hn_context* ctx, uint32_t seed);
void train_accumulate(hn_context* ctx, uint32_t n_rays, float* loss_out);
void train_apply(hn_context* ctx, float grad_scale);
void train_op_gradients(hn_context* ctx, uint32_t n_rays, const float* o, const float* d, const float* t, const float* targets, const float* bgs,
	int loss_type, float loss_scale, float* loss_out, float* grads_host, uint32_t* n_steps, const float* target_depth = nullptr, float depth_lambda = 0.0f, int depth_loss_type = 3);
void train_set_depth(hn_context* ctx, const float* depth, uint32_t n, int w, int h);
void train_release(hn_context* ctx);
void train_export_state(hn_context* ctx, std::vector<float>& m1, std::vector<float>& m2, std::vector<uint32_t>& steps);
void train_import_state(hn_context* ctx, uint32_t step, const float* m1, const float* m2, const uint32_t* steps, size_t n);
void decode_jpeg(const std::string& path, const std::string& file_bytes, int& w, int& h, std::vector<uint8_t>& rgba); // hn_jpeg.cpp
void decode_image_file(const std::string& path, int& w, int& h, std::vector<uint8_t>& rgba);                        // hn_dataset.cpp: PNG or JPEG by content
void train_update_density_grid(hn_context* ctx, float decay, uint32_t n_uniform, uint32_t n_nonuniform);
void train_set_image_metadata(hn_context* ctx, uint32_t n, const float* focal, const float* principal, const int32_t* lens_mode, const float* lens_params);
void train_op_error_map_cdf(hn_context* ctx, uint32_t n_images, uint32_t height, uint32_t width, const float* data, float* cdf_x_cond_y, float* cdf_y, float* cdf_img, float* pmf_img);
void train_op_sample_pixels(hn_context* ctx, uint32_t n_rays, uint32_t ray_offset, uint32_t n_rays_total, uint32_t seed, uint32_t step, uint32_t n_images, int cdf_res_x, int cdf_res_y,
	const float* cdf_x_cond_y, const float* cdf_y, const float* cdf_img, uint32_t* img, float* uv, float* pdf);
void train_op_error_map_deposit(hn_context* ctx, uint32_t n, const uint32_t* img, const float* uv, const float* val, uint32_t n_images, int res_x, int res_y, int img_w, int img_h, float* map);
void train_get_error_map(hn_context* ctx, int res[2], uint32_t* n_images, float* data, size_t n_floats);
void train_mark_untrained(hn_context* ctx);
void refresh_bitfield_from_device_grid(hn_context* ctx); // hn_api.cpp: mean -> threshold -> bitfield + mips -> occupancy bounds
// hn_dataset.cpp
void load_training_data_file(hn_context* ctx, const std::string& path);
// hn_scene.cpp
void add_mesh(hn_context* ctx, const float* tris, size_t n, const float center[3], bool normalize);
void load_mesh_file(hn_context* ctx, const std::string& path, const float center[3]);
void load_scene_file(hn_context* ctx, const std::string& path);
void build_scene(hn_context* ctx);
std::string read_text_file(const std::string& path);
} // namespace hn
