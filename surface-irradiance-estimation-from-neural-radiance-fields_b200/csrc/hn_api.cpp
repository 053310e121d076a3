// C ABI (include/hybrid_nerf.h): host orchestration of the hybrid frame. One context per process/GPU; all per-frame
// scratch comes from a persistent workspace owned by the context (the reference re-allocates from tcnn's arena per frame).
#include "hn_context.h"
#include "hn_device_math.cuh"
#include "hn_shade.cuh"
#include <cmath>
#include <cstring>
#include <map>
#include <chrono>

using namespace hn;

#define HN_TRY(ctx) try {
#define HN_CATCH(ctx) return 0; } catch (const std::exception& e) { if (ctx) (ctx)->error = e.what(); return 1; } catch (...) { if (ctx) (ctx)->error = "unknown error"; return 1; }

static void need(bool cond, const char* msg) { if (!cond) throw std::runtime_error(msg); }

// Texture views (16-byte and 8-byte texels, point sampled, element type) of the parameter buffer for the TEX-pipe gather of the
// tcgen05 kernel. Rebuilt lazily when the buffer has moved; nullptr = gather through the LSU pipe only.
static void release_param_textures(hn_context* ctx) {
	if (ctx->param_tex.tex16) cudaDestroyTextureObject(ctx->param_tex.tex16);
	if (ctx->param_tex.tex8) cudaDestroyTextureObject(ctx->param_tex.tex8);
	ctx->param_tex = hn::ParamTextures{};
}
static const hn::ParamTextures* param_textures(hn_context* ctx) {
	if (!ctx->opt.eval_tex_gather || !ctx->params.p || ctx->net.F != 4) return nullptr;
	if (ctx->param_tex.ptr == ctx->params.p) return &ctx->param_tex;
	release_param_textures(ctx);
	{ // a table beyond the 1D linear texture limit (e.g. log2_hashmap_size 24, F = 4) falls back to the LSU gather instead of failing
		int max_texels = 0;
		HN_CUDA(cudaDeviceGetAttribute(&max_texels, cudaDevAttrMaxTexture1DLinearWidth, ctx->device));
		if ((ctx->net.n_params * sizeof(__half)) / 8 > (size_t)max_texels) { ctx->param_tex.ptr = nullptr; return nullptr; }
	}
	cudaResourceDesc rd = {};
	rd.resType = cudaResourceTypeLinear;
	rd.res.linear.devPtr = (void*)ctx->params.p;
	rd.res.linear.sizeInBytes = (ctx->net.n_params * sizeof(__half)) & ~(size_t)15;
	cudaTextureDesc td = {};
	td.readMode = cudaReadModeElementType; td.filterMode = cudaFilterModePoint; td.addressMode[0] = cudaAddressModeClamp;
	rd.res.linear.desc = cudaCreateChannelDesc(32, 32, 32, 32, cudaChannelFormatKindUnsigned);
	HN_CUDA(cudaCreateTextureObject(&ctx->param_tex.tex16, &rd, &td, nullptr));
	rd.res.linear.desc = cudaCreateChannelDesc(32, 32, 0, 0, cudaChannelFormatKindUnsigned);
	HN_CUDA(cudaCreateTextureObject(&ctx->param_tex.tex8, &rd, &td, nullptr));
	ctx->param_tex.ptr = ctx->params.p;
	return &ctx->param_tex;
}

extern "C" {

const char* hn_version(void) { return "hybrid-nerf-b200 0.1 (sm_100a)"; }

int hn_create(hn_context** out, int cuda_device) {
	if (!out) return 1;
	*out = nullptr;
	hn_context* ctx = new hn_context();
	*out = ctx; // returned even on failure so that hn_last_error can explain
	HN_TRY(ctx)
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) throw std::runtime_error(std::string("no CUDA device: this library has no CPU fallback (") + cudaGetErrorString(e) + ")");
	need(cuda_device >= 0 && cuda_device < n, "invalid CUDA device ordinal");
	ctx->device = cuda_device;
	HN_CUDA(cudaSetDevice(cuda_device));
	cudaDeviceProp prop;
	HN_CUDA(cudaGetDeviceProperties(&prop, cuda_device));
	if (prop.major != 10) throw std::runtime_error(std::string("device '") + prop.name + "' is not sm_100: kernels are built for sm_100a only");
	HN_CUDA(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
	ctx->stream = ctx->own_stream;
	HN_CUDA(cudaMallocHost((void**)&ctx->pinned_counters, CNT_N * sizeof(uint32_t) * 4));
	for (auto& e : ctx->pass_events) HN_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
	HN_CUDA(cudaStreamCreateWithFlags(&ctx->side_stream, cudaStreamNonBlocking));
	HN_CUDA(cudaEventCreateWithFlags(&ctx->fork_event, cudaEventDisableTiming));
	HN_CUDA(cudaEventCreateWithFlags(&ctx->join_event, cudaEventDisableTiming));
	ctx->counters.ensure(CNT_N);
	ctx->bitfield.ensure(BITFIELD_BYTES);
	HN_CATCH(ctx)
}

void hn_destroy(hn_context* ctx) {
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	cudaDeviceSynchronize();
	hn::train_release(ctx);
	release_param_textures(ctx);
	if (ctx->pinned_counters) cudaFreeHost(ctx->pinned_counters);
	for (auto& e : ctx->pass_events) if (e) cudaEventDestroy(e);
	if (ctx->fork_event) cudaEventDestroy(ctx->fork_event);
	if (ctx->join_event) cudaEventDestroy(ctx->join_event);
	if (ctx->side_stream) cudaStreamDestroy(ctx->side_stream);
	if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
	delete ctx;
}

const char* hn_last_error(const hn_context* ctx) { return ctx ? ctx->error.c_str() : "null context"; }

// ---------------------------------------------------------------- network
int hn_load_network_config(hn_context* ctx, const char* json_path) {
	HN_TRY(ctx)
	std::string path = json_path;
	size_t p = path.find_last_of('/');
	load_network_config_text(ctx, read_text_file(path), p == std::string::npos ? "." : path.substr(0, p));
	HN_CATCH(ctx)
}
int hn_load_network_config_string(hn_context* ctx, const char* json_text) {
	HN_TRY(ctx)
	load_network_config_text(ctx, json_text, ".");
	HN_CATCH(ctx)
}
int hn_set_aabb_scale(hn_context* ctx, int aabb_scale) {
	HN_TRY(ctx)
	need(aabb_scale >= 1 && aabb_scale <= 128 && (aabb_scale & (aabb_scale - 1)) == 0, "aabb_scale must be a power of two in [1,128]");
	ctx->aabb_scale = aabb_scale;
	if (ctx->has_net) reapply_network_config(ctx);
	// the render box follows the training box: [0.5 - s/2, 0.5 + s/2]^3 (src/testbed_nerf.cu:2722-2727)
	for (int d = 0; d < 3; ++d) { ctx->opt.render_aabb[d] = 0.5f - 0.5f * aabb_scale; ctx->opt.render_aabb[3 + d] = 0.5f + 0.5f * aabb_scale; }
	HN_CATCH(ctx)
}
int hn_get_network_info(const hn_context* ctx, hn_network_info* out) {
	if (!ctx || !out || !ctx->has_net) return 1;
	const NetCfg& n = ctx->net;
	std::memset(out, 0, sizeof(*out));
	out->n_levels = n.L; out->n_features_per_level = n.F; out->log2_hashmap_size = n.log2_T; out->base_resolution = n.base_res;
	out->per_level_scale = n.per_level_scale; out->density_hidden_layers = n.density_hidden; out->rgb_hidden_layers = n.rgb_hidden;
	out->n_params = n.n_params; out->grid_param_offset = n.grid_offset;
	for (uint32_t i = 0; i <= n.L; ++i) out->level_offsets[i] = n.offsets[i];
	for (uint32_t i = 0; i < n.L; ++i) { out->level_scales[i] = n.scales[i]; out->level_resolutions[i] = n.res[i]; }
	return 0;
}
int hn_set_params(hn_context* ctx, const void* blob, size_t n_elements, int dtype) {
	HN_TRY(ctx)
	upload_params(ctx, blob, n_elements, dtype);
	HN_CATCH(ctx)
}

// ---------------------------------------------------------------- occupancy
static void refresh_occ_bounds(hn_context* ctx) {
	DevBuf<int> d_b;
	d_b.ensure(CASCADES * 6);
	launch_occ_bounds(ctx->bitfield.p, d_b.p, ctx->stream);
	ctx->occ_coarse.ensure(CASCADES * 1024);
	launch_occ_coarse(ctx->bitfield.p, ctx->occ_coarse.p, ctx->stream);
	HN_CUDA(cudaMemcpyAsync(ctx->occ_cells, d_b.p, sizeof(ctx->occ_cells), cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CUDA(cudaGetLastError());
}
int hn_set_density_grid(hn_context* ctx, const float* grid, uint32_t n_cascades) {
	HN_TRY(ctx)
	upload_density_grid(ctx, grid, n_cascades);
	HN_CATCH(ctx)
}
} // extern "C"
namespace hn {
void upload_params(hn_context* ctx, const void* blob, size_t n_elements, int dtype) {
	need(ctx->has_net, "load a network config first");
	need(n_elements == ctx->net.n_params, "parameter blob has the wrong number of elements");
	need(dtype == 0 || dtype == 1, "dtype must be 0 (fp16) or 1 (fp32)");
	// host copy first (hn_save_snapshot writes it back out); Trainer keeps an fp16 inference copy the same way
	ctx->params_host.resize(n_elements);
	if (dtype == 0) std::memcpy(ctx->params_host.data(), blob, n_elements * sizeof(__half));
	else { const float* f = (const float*)blob; for (size_t i = 0; i < n_elements; ++i) ctx->params_host[i] = __float2half_rn(f[i]); }
	HN_CUDA(cudaSetDevice(ctx->device));
	ctx->params.ensure(n_elements);
	HN_CUDA(cudaMemcpy(ctx->params.p, ctx->params_host.data(), n_elements * sizeof(__half), cudaMemcpyHostToDevice));
	ctx->has_params = true;
	ctx->train.ready = false; // a new blob restarts the trainer state (Trainer::set_params)
}
void upload_density_grid(hn_context* ctx, const float* grid, uint32_t n_cascades) {
	need(n_cascades >= 1 && n_cascades <= CASCADES, "n_cascades must be in [1,8]");
	ctx->density_grid_host.assign(grid, grid + (size_t)GRID_CELLS * n_cascades);
	HN_CUDA(cudaSetDevice(ctx->device));
	ctx->density_grid.ensure((size_t)GRID_CELLS * n_cascades);
	ctx->grid_cascades = n_cascades;
	HN_CUDA(cudaMemcpyAsync(ctx->density_grid.p, grid, (size_t)GRID_CELLS * n_cascades * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	refresh_bitfield_from_device_grid(ctx);
}
void refresh_bitfield_from_device_grid(hn_context* ctx) {
	const uint32_t n_cascades = ctx->grid_cascades;
	float* d_grid_p = ctx->density_grid.p;
	struct { float* p; } d_grid = {d_grid_p};
	DevBuf<double> d_sum;
	d_sum.ensure(1);
	launch_grid_mean(d_grid.p, (float*)d_sum.p, ctx->stream);
	double sum = 0.0;
	HN_CUDA(cudaMemcpyAsync(&sum, d_sum.p, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	const float mean = (float)(sum / GRID_CELLS);
	ctx->density_grid_mean = mean;
	const float thresh = std::min(0.01f, mean); // NERF_MIN_OPTICAL_THICKNESS, src/testbed_nerf.cu:297
	launch_grid_to_bitfield(d_grid.p, n_cascades, thresh, ctx->bitfield.p, ctx->stream);
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CUDA(cudaGetLastError());
	refresh_occ_bounds(ctx);
	ctx->has_bitfield = true;
}
} // namespace hn
extern "C" {
int hn_load_snapshot(hn_context* ctx, const char* path) { HN_TRY(ctx) load_snapshot_file(ctx, path); HN_CATCH(ctx) }
int hn_save_snapshot(hn_context* ctx, const char* path) { HN_TRY(ctx) save_snapshot_file(ctx, path); HN_CATCH(ctx) }
int hn_set_occupancy_bitfield(hn_context* ctx, const uint8_t* bits, size_t n_bytes) {
	HN_TRY(ctx)
	need(n_bytes == BITFIELD_BYTES, "bitfield must be 128^3/8*8 bytes");
	HN_CUDA(cudaSetDevice(ctx->device));
	HN_CUDA(cudaMemcpy(ctx->bitfield.p, bits, n_bytes, cudaMemcpyHostToDevice));
	ctx->density_grid_host.clear();
	ctx->density_grid_mean = 1.0f;
	refresh_occ_bounds(ctx);
	ctx->has_bitfield = true;
	HN_CATCH(ctx)
}
int hn_get_occupancy_bitfield(hn_context* ctx, uint8_t* out, size_t n_bytes) {
	HN_TRY(ctx)
	need(n_bytes == BITFIELD_BYTES && ctx->has_bitfield, "no bitfield / wrong size");
	HN_CUDA(cudaSetDevice(ctx->device));
	HN_CUDA(cudaMemcpy(out, ctx->bitfield.p, n_bytes, cudaMemcpyDeviceToHost));
	HN_CATCH(ctx)
}

// ---------------------------------------------------------------- scene
int hn_load_scene(hn_context* ctx, const char* path) { HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); load_scene_file(ctx, path); HN_CATCH(ctx) }
int hn_add_mesh(hn_context* ctx, const float* tris, size_t n, const float center[3], int normalize) { HN_TRY(ctx) add_mesh(ctx, tris, n, center, normalize != 0); HN_CATCH(ctx) }
int hn_load_mesh_file(hn_context* ctx, const char* path, const float center[3]) { HN_TRY(ctx) load_mesh_file(ctx, path, center); HN_CATCH(ctx) }
int hn_build_scene(hn_context* ctx) { HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); build_scene(ctx); HN_CATCH(ctx) }
int hn_clear_scene(hn_context* ctx) { HN_TRY(ctx) ctx->meshes.clear(); ctx->scene_built = false; HN_CATCH(ctx) }
int hn_get_scene_aabb(const hn_context* ctx, float out[6]) { if (!ctx || !ctx->scene_built) return 1; std::memcpy(out, ctx->scene_bb, 24); return 0; }

// ---------------------------------------------------------------- options
struct OptRef { void* p; int n; bool is_int; };
static std::map<std::string, OptRef> option_table(hn_context* c) {
	Options& o = c->opt;
	return {
		{"snap_to_pixel_centers", {&o.snap_to_pixel_centers, 1, true}}, {"render_near_distance", {&o.near_distance, 1, false}},
		{"lens_mode", {&o.lens_mode, 1, true}}, {"lens_params", {o.lens_params, 7, false}}, {"aperture_size", {&o.aperture_size, 1, false}},
		{"focus_z", {&o.focus_z, 1, false}}, {"parallax_shift", {o.parallax_shift, 3, false}}, {"foveation", {o.foveation, 6, false}},
		{"render_aabb", {o.render_aabb, 6, false}}, {"render_aabb_to_local", {o.render_aabb_to_local, 9, false}}, {"nerf_aabb", {o.train_aabb, 6, false}}, {"max_cascade", {&o.max_cascade, 1, true}},
		{"cone_angle_constant", {&o.cone_angle_constant, 1, false}}, {"render_min_transmittance", {&o.min_transmittance, 1, false}},
		{"rgb_activation", {&o.rgb_activation, 1, true}}, {"density_activation", {&o.density_activation, 1, true}},
		{"linear_colors", {&o.linear_colors, 1, true}}, {"render_mode", {&o.render_mode, 1, true}}, {"tonemap_curve", {&o.tonemap_curve, 1, true}},
		{"clamp_output_color", {&o.clamp_output_color, 1, true}}, {"unmultiply_alpha", {&o.unmultiply_alpha, 1, true}}, {"eval_tex_gather", {&o.eval_tex_gather, 1, true}}, {"eval_gather", {&o.eval_gather, 1, true}}, {"eval_groups", {&o.eval_groups, 1, true}}, {"eval_texsplit", {&o.eval_texsplit, 1, true}}, {"eval_ray_sh", {&o.eval_ray_sh, 1, true}}, {"depth_scale", {&o.depth_scale, 1, false}}, {"dda_iter_cap", {&o.dda_iter_cap, 1, true}},
		{"sun_dir", {o.sun_dir, 3, false}}, {"up_dir", {o.up_dir, 3, false}}, {"floor_enable", {&o.floor_enable, 1, true}},
		{"brdf", {o.brdf, 13, false}}, {"compat_nearest_aabb_only", {&o.compat_nearest_aabb_only, 1, true}},
		{"depth_merge_clip", {&o.depth_merge_clip, 1, true}}, {"render_nerf", {&o.render_nerf, 1, true}}, {"render_mesh", {&o.render_mesh, 1, true}},
		{"exposure", {&o.exposure, 1, false}}, {"background_color", {o.background_color, 4, false}},
		{"relative_focal_length", {o.relative_focal_length, 2, false}}, {"screen_center", {o.screen_center, 2, false}}, {"zoom", {&o.zoom, 1, false}},
		{"fov_axis", {&o.fov_axis, 1, false}}, {"profile", {&o.profile, 1, true}}, {"eval_impl", {&o.eval_impl, 1, true}}, {"fuse_composite", {&o.fuse_composite, 1, true}}, {"composite_impl", {&o.composite_impl, 1, true}}, {"mesh_aabb", {o.mesh_aabb, 6, false}},
		{"train_auto_grid", {&c->train_cfg.auto_grid, 1, true}}, {"train_mlp_impl", {&c->train_cfg.mlp_impl, 1, true}}, {"train_random_bg", {&c->train_cfg.random_bg, 1, true}},
		{"train_grid_decay", {&c->train_cfg.grid_decay, 1, false}}, {"train_loss_type", {&c->train_cfg.loss_type, 1, true}},
		{"depth_supervision_lambda", {&c->train_cfg.depth_supervision_lambda, 1, false}}, {"depth_loss_type", {&c->train_cfg.depth_loss_type, 1, true}},
		{"sample_focal_plane_proportional_to_error", {&c->train_cfg.sample_focal_plane_proportional_to_error, 1, true}},
		{"sample_image_proportional_to_error", {&c->train_cfg.sample_image_proportional_to_error, 1, true}},
		{"snapshot_optimizer_state", {&o.snapshot_optimizer_state, 1, true}},
		{"env_light_samples", {&o.env_light_samples, 1, true}}, {"env_light_source", {&o.env_light_source, 1, true}}, {"env_light_sampling", {&o.env_light_sampling, 1, true}}, {"env_light_batch", {&o.env_light_batch, 1, true}}, {"tight_t_far", {&o.tight_t_far, 1, true}}, {"march_simple", {&o.march_simple, 1, true}}, {"overlap_raygen", {&o.overlap_raygen, 1, true}}, {"bvh_sah", {&o.bvh_sah, 1, true}},
	};
}
int hn_set_option(hn_context* ctx, const char* name, const float* values, int n_values) {
	HN_TRY(ctx)
	auto t = option_table(ctx);
	auto it = t.find(name);
	if (it == t.end()) throw std::runtime_error(std::string("unknown option '") + name + "'");
	need(n_values == it->second.n, "wrong number of values for option");
	for (int i = 0; i < n_values; ++i) {
		if (it->second.is_int) ((int*)it->second.p)[i] = (int)values[i];
		else ((float*)it->second.p)[i] = values[i];
	}
	if (std::string(name) == "mesh_aabb") ctx->opt.mesh_aabb_user = 1;
	HN_CATCH(ctx)
}
int hn_get_option(const hn_context* cctx, const char* name, float* values, int n_values) {
	hn_context* ctx = const_cast<hn_context*>(cctx);
	HN_TRY(ctx)
	auto t = option_table(ctx);
	auto it = t.find(name);
	if (it == t.end()) throw std::runtime_error(std::string("unknown option '") + name + "'");
	need(n_values == it->second.n, "wrong number of values for option");
	for (int i = 0; i < n_values; ++i) values[i] = it->second.is_int ? (float)((int*)it->second.p)[i] : ((float*)it->second.p)[i];
	HN_CATCH(ctx)
}
int hn_set_camera(hn_context* ctx, const float cam[12]) { if (!ctx) return 1; std::memcpy(ctx->camera, cam, 48); return 0; }
int hn_set_tile_shard(hn_context* ctx, int rank, int world, int tile_w, int tile_h) {
	HN_TRY(ctx)
	need(world >= 1 && rank >= 0 && rank < world && tile_w > 0 && tile_h > 0, "bad tile shard");
	ctx->rank = rank; ctx->world = world; ctx->tile_w = tile_w; ctx->tile_h = tile_h;
	HN_CATCH(ctx)
}
int hn_get_frame_stats(const hn_context* ctx, hn_frame_stats* out) {
	if (!ctx || !out) return 1;
	*out = ctx->stats;
	out->n_bvh_stack_overflows = 0;
	if (ctx->d_stack_overflow.p) { // cumulative since hn_build_scene; read on demand (not on the frame path)
		cudaSetDevice(ctx->device);
		cudaMemcpy(&out->n_bvh_stack_overflows, ctx->d_stack_overflow.p, sizeof(uint32_t), cudaMemcpyDeviceToHost);
	}
	return 0;
}

} // extern "C"

// ---------------------------------------------------------------- frame
namespace {

// Sobol dimension 1 (v_j = v_{j-1} ^ (v_{j-1} >> 1)) for ld_random_pixel_offset (random_val.cuh:360-366); per-frame, host side.
uint32_t sobol_dim1(uint32_t index) {
	uint32_t X = 0, v = 0x80000000u;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		if ((index >> bit) & 1u) X ^= v;
		v ^= v >> 1;
	}
	return X;
}
void ld_random_val_2d(uint32_t index, uint32_t seed, float out[2]) {
	index = hn_owen(index, seed);
	out[0] = (float)hn_owen(hn_reverse_bits(index), hn_hash_combine(seed, 0u)) * 2.3283064365386963e-10f;
	out[1] = (float)hn_owen(sobol_dim1(index), hn_hash_combine(seed, 1u)) * 2.3283064365386963e-10f;
}
void ld_random_pixel_offset(uint32_t spp, float out[2]) {
	float a[2], b[2];
	ld_random_val_2d(0, 0xdeadbeefu, a);
	ld_random_val_2d(spp, 0xdeadbeefu, b);
	for (int i = 0; i < 2; ++i) { float o = 0.5f - a[i] + b[i]; out[i] = o - std::floor(o); }
}

bool foveation_coeffs(const float f[6], float out[24]);
FrameArgs make_args(hn_context* ctx, const float cam0[12], const float cam1[12], const float screen_center[2], const float rel_focal[2],
	const float rolling_shutter[4], uint32_t spp, int w, int h) {
	const Options& o = ctx->opt;
	FrameArgs a;
	std::memset(&a, 0, sizeof(a));
	a.w = w; a.h = h; a.spp = spp;
	std::memcpy(a.cam0, cam0, 48); std::memcpy(a.cam1, cam1, 48);
	std::memcpy(a.rolling_shutter, rolling_shutter, 16);
	const int res_axis = ((int)o.fov_axis == 0) ? w : h;
	for (int i = 0; i < 2; ++i) {
		a.focal[i] = rel_focal[i] * (float)res_axis * o.zoom;             // calc_focal_length, src/testbed.cu:4474-4476
		a.screen_center[i] = (0.5f - screen_center[i]) * o.zoom + 0.5f;    // render_screen_center, :4478-4481
	}
	ld_random_pixel_offset(o.snap_to_pixel_centers ? 0 : spp, a.pixel_offset);
	a.near_distance = o.near_distance;
	a.lens_mode = o.lens_mode;
	for (int i = 0; i < 7; ++i) a.lens_params[i] = o.lens_params[i];
	a.aperture_size = o.aperture_size; a.focus_z = o.focus_z;
	for (int i = 0; i < 3; ++i) a.parallax_shift[i] = o.parallax_shift[i];
	a.foveated = foveation_coeffs(o.foveation, &a.fov[0][0]) ? 1 : 0;
	a.hidden_mask = ctx->hidden_w > 0 ? ctx->hidden_mask.p : nullptr; a.hidden_w = ctx->hidden_w; a.hidden_h = ctx->hidden_h;
	a.distortion = ctx->dist_w > 0 ? ctx->distortion.p : nullptr; a.dist_w = ctx->dist_w; a.dist_h = ctx->dist_h;
	std::memcpy(a.render_aabb, o.render_aabb, 24); std::memcpy(a.train_aabb, o.train_aabb, 24);
	{
		static const float ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
		std::memcpy(a.aabb_to_local, o.render_aabb_to_local, 36);
		a.crop_rotated = std::memcmp(o.render_aabb_to_local, ident, 36) != 0;
	}
	// world-space union of the set cells of mips 0..max_cascade, each inflated by half a cell of its mip
	for (int k = 0; k < 3; ++k) { a.occ_aabb[k] = 1e30f; a.occ_aabb[3 + k] = -1e30f; }
	for (int m = 0; m <= o.max_cascade && m < (int)CASCADES; ++m) {
		const int* c = ctx->occ_cells + m * 6;
		if (c[3] < c[0]) continue;
		const float scale = std::ldexp(1.0f, m), pad = scale / 256.0f;
		for (int k = 0; k < 3; ++k) {
			a.occ_aabb[k] = std::min(a.occ_aabb[k], ((float)c[k] / 128.0f - 0.5f) * scale + 0.5f - pad);
			a.occ_aabb[3 + k] = std::max(a.occ_aabb[3 + k], ((float)(c[3 + k] + 1) / 128.0f - 0.5f) * scale + 0.5f + pad);
		}
	}
	a.occ_coarse = ctx->opt.tight_t_far ? ctx->occ_coarse.p : nullptr; a.simple_walk = ctx->opt.march_simple;
	a.max_mip = (uint32_t)o.max_cascade; a.cone_angle = o.cone_angle_constant; a.min_transmittance = o.min_transmittance;
	a.rgb_act = o.rgb_activation; a.density_act = o.density_activation; a.linear_colors = o.linear_colors; a.dda_iter_cap = o.dda_iter_cap;
	a.depth_test = 0; a.depth_clip = o.depth_merge_clip;
	a.render_mode = o.render_mode; a.depth_scale = o.depth_scale;
	a.eval_gather = o.eval_gather; a.eval_groups = o.eval_groups; a.eval_texsplit = o.eval_texsplit;
	a.rank = ctx->rank; a.world = ctx->world; a.tile_w = ctx->tile_w; a.tile_h = ctx->tile_h;
	std::memcpy(a.mesh_aabb, o.mesh_aabb, 24);
	std::memcpy(a.sun, o.sun_dir, 12); std::memcpy(a.up, o.up_dir, 12);
	a.floor_y = o.floor_enable ? o.mesh_aabb[1] + 0.001f : -10000.f; // get_floor_y, testbed.h:1019
	std::memcpy(a.brdf, o.brdf, sizeof(a.brdf));
	a.compat_nearest_aabb_only = o.compat_nearest_aabb_only;
	a.envmap = ctx->envmap.p; a.env_w = ctx->env_w; a.env_h = ctx->env_h; // nullptr = no map
	a.env_samples = ctx->envmap.p ? std::max(o.env_light_samples, 0) : 0;
	a.probe = nullptr; a.probe_n_theta = a.probe_n_phi = 0; a.env_source = 0; a.env_sampling = o.env_light_sampling;
	if (o.env_light_source == 1 && ctx->probe_tex.p && ctx->probe_n_theta > 0) { // the NeRF's probe lights the meshes
		a.probe = ctx->probe_tex.p; a.probe_n_theta = ctx->probe_n_theta; a.probe_n_phi = ctx->probe_n_phi; a.env_source = 1;
		a.env_samples = std::max(o.env_light_samples, 0);
	}
	return a;
}

// FoveationPiecewiseQuadratic's constructor (common_device.cuh:119-182) on the host: 12 coefficients of one axis
// (al bl cl am bm ar br cr switch_left switch_right inv_switch_left inv_switch_right)
void foveation_axis(float center_pixel_steepness, float center_inverse_piecewise_y, float center_radius, float c[12]) {
	const float ident[12] = {0, 0, 0, 1, 0, 0, 0, 0, 0, 1, 0, 1};
	std::memcpy(c, ident, sizeof(ident));
	const float center_inverse_radius = center_radius * center_pixel_steepness;
	float left_inv = center_inverse_piecewise_y - center_inverse_radius, right_inv = center_inverse_piecewise_y + center_inverse_radius;
	if (left_inv < 0) left_inv = 0.0f;
	if (right_inv > 1) right_inv = 1.0f;
	const float am = center_pixel_steepness;
	const float d = (right_inv - left_inv) / center_pixel_steepness / 2;
	float bm = 0.0f, m_min = 0.0f, m_max = 1.0f;
	for (uint32_t i = 0; i < 20; i++) {
		const float m = (m_min + m_max) / 2.0f;
		const float l = m - d, r = m + d;
		bm = -((am - 1) * l * l) / (r * r - 2 * r + l * l + 1);
		const float l_actual = (left_inv - bm) / am, r_actual = (right_inv - bm) / am;
		const float m_actual = (l_actual + r_actual) / 2;
		if (m_actual > m) m_min = m; else m_max = m;
	}
	const float l = (left_inv - bm) / am, r = (right_inv - bm) / am;
	if ((l == 0.0f && r == 1.0f) || (am == 1.0f)) return; // fully linear
	c[8] = l; c[9] = r;
	c[3] = am;
	c[0] = (am - 1) / (r * r - 2 * r + l * l + 1);
	c[1] = (am * (r * r - 2 * r + 1) + am * l * l + (2 - 2 * am) * l) / (r * r - 2 * r + l * l + 1);
	c[2] = 0;
	c[4] = -((am - 1) * l * l) / (r * r - 2 * r + l * l + 1);
	c[5] = -(am - 1) / (r * r - 2 * r + l * l + 1);
	c[6] = (am * (r * r + 1) - 2 * r + am * l * l) / (r * r - 2 * r + l * l + 1);
	c[7] = -(am * r * r - r * r + (am - 1) * l * l) / (r * r - 2 * r + l * l + 1);
	c[10] = c[3] * c[8] + c[4];
	c[11] = c[3] * c[9] + c[4];
}
bool foveation_coeffs(const float f[6], float out[24]) {
	const bool on = f[0] != 0.0f || f[1] != 0.0f;
	if (on) { foveation_axis(f[0], f[2], f[4], out); foveation_axis(f[1], f[3], f[5], out + 12); }
	else { foveation_axis(1.0f, 0.5f, 0.5f, out); foveation_axis(1.0f, 0.5f, 0.5f, out + 12); }
	return on;
}

RaySoA ray_soa(hn_context* ctx) {
	RaySoA r;
	r.o_t = ctx->ray_o_t.p; r.d_mw = ctx->ray_d_mw.p; r.rgba = ctx->ray_rgba.p; r.depth = ctx->ray_depth.p;
	r.pixel = ctx->ray_pixel.p; r.n_steps = ctx->ray_nsteps.p; r.t_far = ctx->ray_tfar.p;
	r.sh = nullptr; // set by the camera-ray path after ray_sh_kernel has run
	return r;
}

void ensure_rays(hn_context* ctx, size_t n) {
	ctx->ray_o_t.ensure(n); ctx->ray_d_mw.ensure(n); ctx->ray_rgba.ensure(n); ctx->ray_depth.ensure(n);
	ctx->ray_pixel.ensure(n); ctx->ray_nsteps.ensure(n); ctx->ray_tfar.ensure(n); ctx->alive_a.ensure(n); ctx->alive_b.ensure(n); ctx->sample_counts.ensure(n);
}

template <typename T> T* scratch(hn_context* ctx, int slot, size_t count) { ctx->scratch[slot].ensure(count * sizeof(T) + 16); return (T*)ctx->scratch[slot].p; }

struct Timer {
	cudaEvent_t a = nullptr, b = nullptr; bool on; cudaStream_t s;
	Timer(bool enabled, cudaStream_t st) : on(enabled), s(st) { if (on) { cudaEventCreate(&a); cudaEventCreate(&b); } }
	~Timer() { if (on) { cudaEventDestroy(a); cudaEventDestroy(b); } }
	void start() { if (on) cudaEventRecord(a, s); }
	void stop(float& acc) { if (on) { cudaEventRecord(b, s); cudaEventSynchronize(b); float ms = 0; cudaEventElapsedTime(&ms, a, b); acc += ms; } }
};

// The march / network / composite rounds over a live-ray list whose initial count sits in counters[CNT_ALIVE_A].
// light == false: camera rays (colour compositing into the frame); light == true: a-16 light rays (density-only network
// evaluation, transmittance only). Returns the number of rays that entered the loop.
// No host round trip per pass (the reference blocks on the alive counter every pass, src/testbed_nerf.cu:2071-2072): every
// kernel reads its ray count from device memory, the host only needs an upper bound for grid and buffer sizes (the
// live-ray count never grows) and the moment it reaches zero. The count entering pass j is copied to a pinned ring slot
// after pass j-1 and looked at LAG passes later, so the stream always holds LAG passes of queued work; at most LAG empty
// passes run at the end.
uint32_t run_passes(hn_context* ctx, const FrameArgs& a, RaySoA rays, float4* frame, float* depth, bool light, cudaStream_t s, Timer& tm) {
	hn_frame_stats& st = ctx->stats;
	uint32_t* cnt = ctx->counters.p;
	uint32_t* lists[2] = {ctx->alive_a.p, ctx->alive_b.p};
	int cur = 0;
	const int LAG = ctx->opt.profile ? 0 : 1;
	constexpr int RING = 4;
	HN_CUDA(cudaMemcpyAsync(ctx->pinned_counters, cnt, CNT_N * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	uint32_t n_alive = ctx->pinned_counters[CNT_ALIVE_A]; // exact for the first pass, an upper bound afterwards
	const uint32_t n_first = n_alive;
	need(n_alive <= ctx->alive_a.n, "alive counter corrupt");
	uint32_t passes = 0;
	float ms_dummy = 0.f;
	for (int k = 0; n_alive > 0; ++k) {
		const int j = k - LAG; // newest pass whose entry count must be known before pass k is queued
		if (j >= 1) {
			HN_CUDA(cudaEventSynchronize(ctx->pass_events[j % RING]));
			const uint32_t n = ctx->pinned_counters[(j % RING) * CNT_N + ((j & 1) ? CNT_ALIVE_B : CNT_ALIVE_A)];
			if (n == 0) { passes = (uint32_t)j; break; }
			need(n <= n_alive, "alive counter corrupt");
			n_alive = n;
		}
		passes = (uint32_t)k + 1;
		ctx->samples.ensure((size_t)n_alive * CHUNK);
		ctx->net_out.ensure((size_t)n_alive * CHUNK);
		uint32_t* n_alive_ptr = cnt + (cur == 0 ? CNT_ALIVE_A : CNT_ALIVE_B);
		uint32_t* n_next_ptr = cnt + (cur == 0 ? CNT_ALIVE_B : CNT_ALIVE_A);
		HN_CUDA(cudaMemsetAsync(n_next_ptr, 0, sizeof(uint32_t), s));
		tm.start();
		launch_march(a, ctx->bitfield.p, rays, lists[cur], n_alive_ptr, ctx->samples.p, ctx->sample_counts.p, cnt, depth, n_alive, s);
		tm.stop(light ? ms_dummy : st.ms_march);
		tm.start();
		const bool fused = !light && ctx->opt.eval_impl == 1 && ctx->opt.fuse_composite;
		if (fused) ctx->survive.ensure(n_alive);
		launch_eval(ctx->net, ctx->levels, ctx->params.p, a, ctx->samples.p, ctx->sample_counts.p, lists[cur], rays, nullptr, n_alive_ptr, n_alive, ctx->net_out.p,
			ctx->opt.eval_impl, fused ? frame : nullptr, depth, ctx->survive.p, cnt, s, light, param_textures(ctx));
		if (!light && a.render_mode == 5) // Normals: the density gradient replaces the colour of every sample
			launch_normals(ctx->net, ctx->levels, ctx->params.p, a, ctx->samples.p, ctx->sample_counts.p, n_alive_ptr, n_alive, ctx->net_out.p, s);
		tm.stop(light ? ms_dummy : st.ms_eval);
		tm.start();
		const bool per_ray = !light && !fused && ctx->opt.composite_impl == 1;
		if (light) {
			ctx->survive.ensure(n_alive);
			launch_transmittance(a, rays, lists[cur], n_alive_ptr, ctx->samples.p, ctx->sample_counts.p, ctx->net_out.p, ctx->survive.p, cnt, n_alive, s);
			st.n_kernel_launches += 1;
		} else if (per_ray) {
			ctx->survive.ensure(n_alive);
			launch_composite_rays(make_comp_args(a, rays, lists[cur], ctx->samples.p, frame, depth, ctx->survive.p, cnt), n_alive_ptr, ctx->sample_counts.p, ctx->net_out.p, n_alive, s);
			st.n_kernel_launches += 1;
		}
		if (light || fused || per_ray) launch_compact(lists[cur], n_alive_ptr, ctx->survive.p, lists[cur ^ 1], n_next_ptr, n_alive, s);
		else launch_composite(a, rays, lists[cur], n_alive_ptr, lists[cur ^ 1], cnt, ctx->samples.p, ctx->sample_counts.p, ctx->net_out.p, frame, depth, n_alive, s);
		tm.stop(light ? ms_dummy : st.ms_composite);
		st.n_kernel_launches += 3;
		cur ^= 1;
		HN_CUDA(cudaMemcpyAsync(ctx->pinned_counters + ((k + 1) % RING) * CNT_N, cnt, CNT_N * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
		HN_CUDA(cudaEventRecord(ctx->pass_events[(k + 1) % RING], s));
	}
	if (!light) st.n_passes = passes;
	HN_CUDA(cudaMemcpyAsync(ctx->pinned_counters, cnt, CNT_N * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	return n_first;
}

EnvHits env_hits(hn_context* ctx) {
	EnvHits h;
	h.pos_pix = ctx->hit_pos_pix.p; h.nrm = ctx->hit_nrm.p; h.albedo = ctx->hit_albedo.p; h.count = ctx->hit_count.p;
	return h;
}

// a-16: environment light on the mesh hits the mesh pass listed. Batches of hits so that at most env_light_batch light
// rays are in flight: generate (direction, shadow ray, envmap radiance) -> enter the volume -> transmittance passes -> reduce.
void env_light_pass(hn_context* ctx, const FrameArgs& a_in, float4* frame, float* depth, cudaStream_t s, Timer& tm) {
	hn_frame_stats& st = ctx->stats;
	FrameArgs a = a_in;
	a.depth_test = 0; a.depth_clip = 0; // light rays are not clipped by the camera's mesh depth
	uint32_t n_hits = 0;
	HN_CUDA(cudaMemcpyAsync(&n_hits, ctx->hit_count.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	st.n_env_hits = n_hits;
	if (!n_hits) return;
	const uint32_t K = (uint32_t)a.env_samples;
	// probe light: the texture already holds what the NeRF emits towards the scene, no transmittance march along the light rays
	const bool volume = a.env_source != 1 && ctx->opt.render_nerf && ctx->has_net && ctx->has_params && ctx->has_bitfield;
	const uint32_t per_batch = std::max<uint32_t>(1u, (uint32_t)std::max(ctx->opt.env_light_batch, 1) / K);
	uint32_t* cnt = ctx->counters.p;
	for (uint32_t first = 0; first < n_hits; first += per_batch) {
		const uint32_t n = std::min(per_batch, n_hits - first);
		const size_t n_rays = (size_t)n * K;
		ensure_rays(ctx, n_rays);
		RaySoA rays = ray_soa(ctx);
		launch_envlight_gen(a, ctx->scene_dev(), env_hits(ctx), first, n, rays, s);
		st.n_kernel_launches += 1;
		st.n_light_rays += n_rays;
		if (volume) {
			HN_CUDA(cudaMemsetAsync(cnt, 0, CNT_N * sizeof(uint32_t), s));
			launch_secondary_start(a, ctx->bitfield.p, rays, (uint32_t)n_rays, ctx->alive_a.p, cnt, s);
			st.n_kernel_launches += 1;
			st.n_light_rays_volume += run_passes(ctx, a, rays, frame, depth, true, s, tm);
			st.n_light_samples += (uint64_t)ctx->pinned_counters[CNT_SAMPLES_LO] | ((uint64_t)ctx->pinned_counters[CNT_SAMPLES_HI] << 32);
		}
		launch_envlight_reduce(a, env_hits(ctx), first, n, rays, frame, s);
		st.n_kernel_launches += 1;
	}
}

// ERenderMode::Distortion / Slice of the NeRF pass. Slice: one network sample per pixel, in batches of 1 Mi pixels (a chunk of the
// network kernel carries one direction, so each pixel takes a chunk of its own with one valid sample).
void render_2d(hn_context* ctx, const FrameArgs& a_in, float4* frame, float* depth, cudaStream_t s) {
	hn_frame_stats& st = ctx->stats;
	if (a_in.render_mode == 8) { launch_distortion(a_in, frame, depth, s); st.n_kernel_launches += 1; return; }
	FrameArgs a = a_in;
	a.aperture_size = 0.0f; // plane_z < 0 switches depth of field off (:1465-1467)
	a.render_mode = 1;
	const uint32_t n_threads = owned_pixel_threads(a), batch = 1u << 20;
	const uint32_t cap = std::min(n_threads, batch);
	ctx->samples.ensure((size_t)cap * CHUNK); ctx->net_out.ensure((size_t)cap * CHUNK); ctx->sample_counts.ensure(cap);
	ctx->slice_dirs.ensure((size_t)cap * 3); ctx->slice_pixels.ensure(cap);
	RaySoA none = {};
	for (uint32_t first = 0; first < n_threads; first += batch) {
		const uint32_t n = std::min(batch, n_threads - first);
		launch_slice_raygen(a, ctx->opt.focus_z, first, n, ctx->samples.p, ctx->sample_counts.p, ctx->slice_dirs.p, ctx->slice_pixels.p, depth, s);
		launch_eval(ctx->net, ctx->levels, ctx->params.p, a, ctx->samples.p, ctx->sample_counts.p, nullptr, none, ctx->slice_dirs.p, nullptr, n, ctx->net_out.p,
			ctx->opt.eval_impl, nullptr, nullptr, nullptr, nullptr, s, false, param_textures(ctx));
		launch_slice_shade(n, ctx->slice_pixels.p, ctx->net_out.p, a.rgb_act, a.density_act, a.linear_colors, frame, s);
		st.n_kernel_launches += 3; st.n_samples += n;
	}
}

void render_frame_impl(hn_context* ctx, const FrameArgs& args_in, float4* frame, float* depth, cudaStream_t s) {
	FrameArgs a = args_in;
	const size_t n_px = (size_t)a.w * a.h;
	hn_frame_stats& st = ctx->stats;
	st = hn_frame_stats{};
	Timer tm(ctx->opt.profile != 0, s), total(ctx->opt.profile != 0, s);
	total.start();
	// CudaRenderBufferView::clear (render_buffer.cu:603-607)
	HN_CUDA(cudaMemsetAsync(frame, 0, n_px * sizeof(float4), s));
	HN_CUDA(cudaMemsetAsync(depth, 0, n_px * sizeof(float), s));
	const bool do_mesh = ctx->opt.render_mesh && !ctx->meshes.empty();
	const bool do_nerf = ctx->opt.render_nerf && ctx->has_net;
	const bool nerf_2d = ctx->opt.render_mode == 8 || ctx->opt.render_mode == 10;
	const bool env_light = do_mesh && (a.envmap || a.env_source == 1) && a.env_samples > 0;
	if (do_mesh) need(ctx->scene_built, "scene has meshes but hn_build_scene was not called");
	if (do_nerf) {
		need(ctx->has_params, "network parameters not set");
		need(ctx->has_bitfield, "occupancy grid not set");
		if (!nerf_2d) {
			const int m = ctx->opt.render_mode;
			need(m == 0 || (m >= 1 && m <= 7) || m == 9 || m == 12, "render_mode must be one of ERenderMode: 0-10 or 12 (EncodingVis)");
			need(m != 5 || ctx->net.density_hidden <= 2, "render_mode Normals: at most two hidden layers in the density network");
			need((m >= 1 && m <= 4) || (ctx->opt.composite_impl == 1 && !ctx->opt.fuse_composite) || (ctx->opt.eval_impl == 1 && ctx->opt.fuse_composite), "visualisation render modes need composite_impl 1");
		}
	}
	// init rays + advance to the first occupied cell (+ SH of the direction) for every pixel; depth_or_null: see raygen_kernel
	RaySoA rays = {};
	auto set_up_rays = [&](const FrameArgs& fa, float* depth_or_null, cudaStream_t rs) {
		uint32_t* cnt = ctx->counters.p;
		HN_CUDA(cudaMemsetAsync(cnt, 0, CNT_N * sizeof(uint32_t), rs));
		launch_raygen(fa, ctx->bitfield.p, rays, ctx->alive_a.p, cnt, depth_or_null, frame, rs);
		st.n_kernel_launches += 1;
		if (ctx->opt.eval_impl == 1 && ctx->opt.eval_ray_sh) { // SH input of the colour MLP, once per ray instead of once per sample
			launch_ray_sh(ctx->alive_a.p, cnt + CNT_ALIVE_A, (uint32_t)n_px, rays.d_mw, ctx->ray_sh.p, rs);
			st.n_kernel_launches += 1;
		}
	};
	if (do_nerf && !nerf_2d) { // buffers first: growing one is a cudaMalloc, which would serialise the two streams below
		ensure_rays(ctx, n_px);
		rays = ray_soa(ctx);
		if (ctx->opt.eval_impl == 1 && ctx->opt.eval_ray_sh) { ctx->ray_sh.ensure(2 * n_px); rays.sh = ctx->ray_sh.p; }
	}
	// Hybrid frame: the ray set-up needs nothing from the mesh pass (the mesh depth only clips the march), so it runs beside it on a second
	// stream - both kernels are latency-bound walks (BVH / occupancy grid) that leave issue slots free. Not with an environment map (its
	// background write looks at the mesh depth) and not when per-kernel times are being taken.
	const bool overlap = do_mesh && do_nerf && !nerf_2d && ctx->opt.overlap_raygen && !ctx->opt.profile && !env_light && !a.envmap;
	if (do_mesh) {
		EnvHits hits = {nullptr, nullptr, nullptr, nullptr};
		if (env_light) {
			ctx->hit_pos_pix.ensure(n_px); ctx->hit_nrm.ensure(n_px); ctx->hit_albedo.ensure(n_px); ctx->hit_count.ensure(1);
			HN_CUDA(cudaMemsetAsync(ctx->hit_count.p, 0, sizeof(uint32_t), s));
			hits = env_hits(ctx);
		}
		if (overlap) {
			FrameArgs fa = a; fa.depth_test = 1;
			HN_CUDA(cudaEventRecord(ctx->fork_event, s));
			HN_CUDA(cudaStreamWaitEvent(ctx->side_stream, ctx->fork_event, 0));
			set_up_rays(fa, nullptr, ctx->side_stream);
			HN_CUDA(cudaEventRecord(ctx->join_event, ctx->side_stream));
		}
		tm.start();
		launch_mesh_pass(a, ctx->scene_dev(), frame, depth, hits, s);
		st.n_kernel_launches += 1;
		tm.stop(st.ms_mesh);
		if (overlap) HN_CUDA(cudaStreamWaitEvent(s, ctx->join_event, 0));
		if (env_light) {
			Timer te(ctx->opt.profile != 0, s);
			te.start();
			env_light_pass(ctx, a, frame, depth, s, tm);
			te.stop(st.ms_envlight);
		}
		a.depth_test = 1;
	}
	if (do_nerf) {
		if (nerf_2d) { // the 2-D modes never trace (render_nerf, src/testbed_nerf.cu:2397-2460)
			render_2d(ctx, a, frame, depth, s);
			total.stop(st.ms_total);
			HN_CUDA(cudaGetLastError());
			return;
		}
		if (!overlap) {
			rays = ray_soa(ctx); // the light pass may have grown (moved) the ray buffers
			if (ctx->opt.eval_impl == 1 && ctx->opt.eval_ray_sh) rays.sh = ctx->ray_sh.p;
			tm.start();
			set_up_rays(a, depth, s);
			tm.stop(st.ms_raygen);
		}
		st.n_rays_alive = run_passes(ctx, a, rays, frame, depth, false, s, tm);
		st.n_samples = (uint64_t)ctx->pinned_counters[CNT_SAMPLES_LO] | ((uint64_t)ctx->pinned_counters[CNT_SAMPLES_HI] << 32);
		st.n_samples_composited = (uint64_t)ctx->pinned_counters[CNT_COMPOSITED_LO] | ((uint64_t)ctx->pinned_counters[CNT_COMPOSITED_HI] << 32);
	}
	total.stop(st.ms_total);
	HN_CUDA(cudaGetLastError());
}

} // namespace

namespace hn {
FrameArgs train_frame_args(hn_context* ctx) {
	const float rs[4] = {0.f, 0.f, 0.f, 1.f};
	FrameArgs a = make_args(ctx, ctx->camera, ctx->camera, ctx->opt.screen_center, ctx->opt.relative_focal_length, rs, 0, 1, 1);
	// training clips and marches rays against the TRAINING box (m_aabb: generate_training_samples_nerf, src/testbed_nerf.cu:737-890),
	// never the render crop (m_render_aabb / m_render_aabb_to_local), so a crop set for viewing cannot change what is trained
	std::memcpy(a.render_aabb, a.train_aabb, sizeof(a.render_aabb));
	a.crop_rotated = 0;
	return a;
}
} // namespace hn

extern "C" {

// ---------------------------------------------------------------- training (a-21)
int hn_train_set_dataset(hn_context* ctx, const float* rgba, uint32_t n_images, int width, int height, const float* cameras,
	const float focal_px[2], const float principal[2]) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_set_dataset(ctx, rgba, n_images, width, height, cameras, focal_px, principal); HN_CATCH(ctx)
}
int hn_train_set_depth(hn_context* ctx, const float* depth, uint32_t n_images, int width, int height) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_set_depth(ctx, depth, n_images, width, height); HN_CATCH(ctx)
}
int hn_op_train_gradients_depth(hn_context* ctx, uint32_t n_rays, const float* origins, const float* dirs, const float* t_start, const float* targets, const float* backgrounds,
	const float* target_depth, float depth_lambda, int depth_loss_type, int loss_type, float loss_scale, float* loss_out, float* grads_host, uint32_t* n_steps) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device));
	need(target_depth != nullptr, "target_depth is required");
	train_op_gradients(ctx, n_rays, origins, dirs, t_start, targets, backgrounds, loss_type, loss_scale, loss_out, grads_host, n_steps, target_depth, depth_lambda, depth_loss_type);
	HN_CATCH(ctx)
}
int hn_train_set_image_metadata(hn_context* ctx, uint32_t n_images, const float* focal_px, const float* principal, const int32_t* lens_mode, const float* lens_params) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_set_image_metadata(ctx, n_images, focal_px, principal, lens_mode, lens_params); HN_CATCH(ctx)
}
int hn_decode_image_file(const char* path, int* width, int* height, uint8_t* rgba, size_t capacity_bytes, char* error, size_t error_capacity) {
	try {
		if (!path || !width || !height) throw std::runtime_error("bad arguments");
		std::vector<uint8_t> px;
		decode_image_file(path, *width, *height, px);
		if (rgba) {
			if (capacity_bytes < px.size()) throw std::runtime_error("pixel buffer too small");
			std::memcpy(rgba, px.data(), px.size());
		}
		return 0;
	} catch (const std::exception& e) {
		if (error && error_capacity) { std::strncpy(error, e.what(), error_capacity - 1); error[error_capacity - 1] = 0; }
		return 1;
	}
}
int hn_load_training_data(hn_context* ctx, const char* path) {
	HN_TRY(ctx) need(path != nullptr, "null path"); HN_CUDA(cudaSetDevice(ctx->device)); load_training_data_file(ctx, path); HN_CATCH(ctx)
}
int hn_train_init(hn_context* ctx, uint32_t seed) { HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_init(ctx, seed); HN_CATCH(ctx) }
int hn_train_accumulate(hn_context* ctx, uint32_t n_rays, float* loss_out) { HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_accumulate(ctx, n_rays, loss_out); HN_CATCH(ctx) }
int hn_train_gradient_buffer(hn_context* ctx, void** dev_ptr, size_t* n_floats) {
	HN_TRY(ctx)
	need(ctx->train.ready && dev_ptr && n_floats, "trainer not initialised (hn_train_init)");
	*dev_ptr = ctx->train.grads.p; *n_floats = ctx->net.n_params;
	HN_CATCH(ctx)
}
int hn_train_apply(hn_context* ctx, float grad_scale) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	train_apply(ctx, grad_scale);
	if (ctx->stream == ctx->own_stream) HN_CUDA(cudaStreamSynchronize(ctx->stream)); // on a caller's stream the step stays stream-ordered (no host wait)
	HN_CATCH(ctx)
}
// Run the training / set-up calls on the caller's stream (e.g. the one its NCCL collectives use), so that
// hn_train_accumulate -> all-reduce of hn_train_gradient_buffer -> hn_train_apply needs no host synchronisation in between.
// stream = NULL returns to the context's own stream.
int hn_set_stream(hn_context* ctx, void* cuda_stream) {
	HN_TRY(ctx)
	need(ctx->own_stream != nullptr, "no CUDA device");
	HN_CUDA(cudaSetDevice(ctx->device));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
	HN_CATCH(ctx)
}
int hn_train_step(hn_context* ctx, uint32_t n_steps, float* loss_out) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	for (uint32_t i = 0; i < n_steps; ++i) { train_accumulate(ctx, 0, loss_out); train_apply(ctx, 1.0f); }
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CATCH(ctx)
}
int hn_get_train_stats(const hn_context* ctx, hn_train_stats* out) {
	if (!ctx || !out) return 1;
	const TrainState& t = ctx->train;
	out->step = t.step; out->rays_per_batch = t.rays_per_batch; out->n_rays_last = t.n_rays_last; out->n_samples_last = t.n_samples_last;
	out->loss_last = t.loss_last; out->learning_rate = t.lr_last;
	out->ms_sample = t.ms[0]; out->ms_forward = t.ms[1]; out->ms_loss = t.ms[2]; out->ms_backward = t.ms[3]; out->ms_optimizer = t.ms[4];
	return 0;
}
int hn_update_density_grid(hn_context* ctx, float decay, uint32_t n_uniform, uint32_t n_nonuniform) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_update_density_grid(ctx, decay, n_uniform, n_nonuniform); HN_CATCH(ctx)
}
int hn_mark_untrained_density_grid(hn_context* ctx) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_mark_untrained(ctx); refresh_bitfield_from_device_grid(ctx); HN_CATCH(ctx)
}
int hn_get_density_grid(hn_context* ctx, float* out, size_t n_floats) {
	HN_TRY(ctx)
	need(ctx->density_grid.p && n_floats == (size_t)GRID_CELLS * ctx->grid_cascades, "no density grid / wrong size");
	HN_CUDA(cudaSetDevice(ctx->device));
	HN_CUDA(cudaMemcpy(out, ctx->density_grid.p, n_floats * sizeof(float), cudaMemcpyDeviceToHost));
	HN_CATCH(ctx)
}
int hn_op_train_gradients(hn_context* ctx, uint32_t n_rays, const float* origins, const float* dirs, const float* t_start,
	const float* targets, const float* backgrounds, int loss_type, float loss_scale, float* loss_out, float* grads_host, uint32_t* n_steps) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	train_op_gradients(ctx, n_rays, origins, dirs, t_start, targets, backgrounds, loss_type, loss_scale, loss_out, grads_host, n_steps);
	HN_CATCH(ctx)
}

int hn_op_error_map_cdf(hn_context* ctx, uint32_t n_images, uint32_t height, uint32_t width, const float* data, float* cdf_x_cond_y, float* cdf_y, float* cdf_img, float* pmf_img) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_op_error_map_cdf(ctx, n_images, height, width, data, cdf_x_cond_y, cdf_y, cdf_img, pmf_img); HN_CATCH(ctx)
}
int hn_op_sample_training_pixels(hn_context* ctx, uint32_t n_rays, uint32_t ray_offset, uint32_t n_rays_total, uint32_t seed, uint32_t step, uint32_t n_images,
	int cdf_res_x, int cdf_res_y, const float* cdf_x_cond_y, const float* cdf_y, const float* cdf_img, uint32_t* img, float* uv, float* pdf) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device));
	train_op_sample_pixels(ctx, n_rays, ray_offset, n_rays_total, seed, step, n_images, cdf_res_x, cdf_res_y, cdf_x_cond_y, cdf_y, cdf_img, img, uv, pdf);
	HN_CATCH(ctx)
}
int hn_op_error_map_deposit(hn_context* ctx, uint32_t n, const uint32_t* img, const float* uv, const float* val, uint32_t n_images, int res_x, int res_y, int img_w, int img_h, float* map) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_op_error_map_deposit(ctx, n, img, uv, val, n_images, res_x, res_y, img_w, img_h, map); HN_CATCH(ctx)
}
int hn_get_error_map(hn_context* ctx, int res[2], uint32_t* n_images, float* data, size_t n_floats) {
	HN_TRY(ctx) HN_CUDA(cudaSetDevice(ctx->device)); train_get_error_map(ctx, res, n_images, data, n_floats); HN_CATCH(ctx)
}

int hn_render_frame(hn_context* ctx, const float cam0[12], const float cam1[12], const float screen_center[2],
	const float relative_focal_length[2], const float rolling_shutter[4], uint32_t spp_index,
	int width, int height, float* frame_dev, float* depth_dev, void* cuda_stream) {
	HN_TRY(ctx)
	need(width > 0 && height > 0 && frame_dev && depth_dev, "bad frame arguments");
	HN_CUDA(cudaSetDevice(ctx->device));
	FrameArgs a = make_args(ctx, cam0, cam1, screen_center, relative_focal_length, rolling_shutter, spp_index, width, height);
	render_frame_impl(ctx, a, (float4*)frame_dev, depth_dev, (cudaStream_t)cuda_stream);
	HN_CATCH(ctx)
}

int hn_set_hidden_area_mask(hn_context* ctx, const uint8_t* mask, int width, int height) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	if (!mask || width <= 0 || height <= 0) { ctx->hidden_w = ctx->hidden_h = 0; return 0; }
	ctx->hidden_mask.ensure((size_t)width * height);
	HN_CUDA(cudaMemcpy(ctx->hidden_mask.p, mask, (size_t)width * height, cudaMemcpyHostToDevice));
	ctx->hidden_w = width; ctx->hidden_h = height;
	HN_CATCH(ctx)
}
int hn_set_distortion_map(hn_context* ctx, const float* grid, int width, int height) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	if (!grid || width <= 0 || height <= 0) { ctx->dist_w = ctx->dist_h = 0; return 0; }
	ctx->distortion.ensure((size_t)width * height);
	HN_CUDA(cudaMemcpy(ctx->distortion.p, grid, (size_t)width * height * sizeof(float2), cudaMemcpyHostToDevice));
	ctx->dist_w = width; ctx->dist_h = height;
	HN_CATCH(ctx)
}
int hn_unwarp_foveated(hn_context* ctx, const float foveation[6], int width_in, int height_in, const float* frame_in_dev, const float* depth_in_dev,
	int width_out, int height_out, float* frame_out_dev, float* depth_out_dev, void* cuda_stream) {
	HN_TRY(ctx)
	need(foveation && frame_in_dev && frame_out_dev && width_in > 0 && height_in > 0 && width_out > 0 && height_out > 0, "bad unwarp arguments");
	HN_CUDA(cudaSetDevice(ctx->device));
	float c[24];
	foveation_coeffs(foveation, c);
	launch_unwarp_foveated(c, width_in, height_in, (const float4*)frame_in_dev, depth_in_dev, width_out, height_out, (float4*)frame_out_dev, depth_out_dev, (cudaStream_t)cuda_stream);
	HN_CUDA(cudaGetLastError());
	HN_CATCH(ctx)
}
int hn_op_foveation(hn_context* ctx, const float foveation[6], float coeffs[24]) {
	HN_TRY(ctx)
	need(foveation && coeffs, "bad foveation arguments");
	foveation_coeffs(foveation, coeffs);
	HN_CATCH(ctx)
}

int hn_render_to_host(hn_context* ctx, int width, int height, int spp, int linear, float* rgba_host, float* depth_host) {
	HN_TRY(ctx)
	need(width > 0 && height > 0 && spp > 0 && rgba_host, "bad render arguments");
	HN_CUDA(cudaSetDevice(ctx->device));
	const size_t n = (size_t)width * height;
	ctx->frame.ensure(n); ctx->accum.ensure(n); ctx->tonemapped.ensure(n); ctx->depth.ensure(n);
	const float rs[4] = {0.f, 0.f, 0.f, 1.f}; // python_api.cu:177
	uint32_t launches = 0;
	hn_frame_stats acc = {};
	for (int i = 0; i < spp; ++i) {
		FrameArgs a = make_args(ctx, ctx->camera, ctx->camera, ctx->opt.screen_center, ctx->opt.relative_focal_length, rs, (uint32_t)i, width, height);
		render_frame_impl(ctx, a, ctx->frame.p, ctx->depth.p, ctx->stream);
		launch_accumulate(ctx->frame.p, ctx->accum.p, (uint32_t)n, (uint32_t)i, ctx->stream);
		launches += ctx->stats.n_kernel_launches + 1;
		acc.n_samples += ctx->stats.n_samples; acc.n_samples_composited += ctx->stats.n_samples_composited; acc.n_passes += ctx->stats.n_passes;
		acc.n_rays_alive = ctx->stats.n_rays_alive;
		acc.ms_mesh += ctx->stats.ms_mesh; acc.ms_raygen += ctx->stats.ms_raygen; acc.ms_march += ctx->stats.ms_march; acc.ms_eval += ctx->stats.ms_eval;
		acc.ms_composite += ctx->stats.ms_composite; acc.ms_total += ctx->stats.ms_total;
		acc.ms_envlight += ctx->stats.ms_envlight; acc.n_env_hits = ctx->stats.n_env_hits; acc.n_light_rays += ctx->stats.n_light_rays;
		acc.n_light_rays_volume += ctx->stats.n_light_rays_volume; acc.n_light_samples += ctx->stats.n_light_samples;
	}
	launch_tonemap(ctx->accum.p, ctx->tonemapped.p, (uint32_t)n, linear ? 0 : 1, ctx->opt.exposure, ctx->opt.background_color,
		ctx->opt.tonemap_curve, ctx->opt.clamp_output_color, ctx->opt.unmultiply_alpha, ctx->stream);
	launches += 1;
	HN_CUDA(cudaMemcpyAsync(rgba_host, ctx->tonemapped.p, n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
	if (depth_host) HN_CUDA(cudaMemcpyAsync(depth_host, ctx->depth.p, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	ctx->stats = acc;
	ctx->stats.n_kernel_launches = launches;
	HN_CATCH(ctx)
}

// ---------------------------------------------------------------- tiles
int hn_tile_count(const hn_context* ctx, int width, int height, int rank, int world, uint32_t* n_tiles) {
	if (!ctx || !n_tiles || world < 1) return 1;
	uint32_t tiles = ((width + ctx->tile_w - 1) / ctx->tile_w) * ((height + ctx->tile_h - 1) / ctx->tile_h);
	*n_tiles = (tiles + world - 1 - rank) / world;
	return 0;
}
int hn_pack_tiles(hn_context* ctx, int width, int height, int rank, int world, const float* frame_dev, const float* depth_dev, float* packed_dev, void* cuda_stream) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	launch_pack_tiles(width, height, rank, world, ctx->tile_w, ctx->tile_h, (const float4*)frame_dev, depth_dev, packed_dev, false, (cudaStream_t)cuda_stream);
	HN_CUDA(cudaGetLastError());
	HN_CATCH(ctx)
}
int hn_unpack_tiles(hn_context* ctx, int width, int height, int rank, int world, const float* packed_dev, float* frame_dev, float* depth_dev, void* cuda_stream) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	launch_pack_tiles(width, height, rank, world, ctx->tile_w, ctx->tile_h, (const float4*)frame_dev, depth_dev, (float*)packed_dev, true, (cudaStream_t)cuda_stream);
	HN_CUDA(cudaGetLastError());
	HN_CATCH(ctx)
}

// Peer-memory frame gather (one process per GPU): the root rank allocates frame + depth through the library and exports them with
// CUDA IPC, every other rank maps them once and then copies its tiles into them per frame (launch_scatter_tiles).
int hn_shared_alloc(hn_context* ctx, size_t bytes, void** dev_ptr, unsigned char handle[64]) {
	HN_TRY(ctx)
	need(dev_ptr && handle && bytes > 0, "hn_shared_alloc: bad arguments");
	static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
	HN_CUDA(cudaSetDevice(ctx->device));
	HN_CUDA(cudaMalloc(dev_ptr, bytes));
	HN_CUDA(cudaMemset(*dev_ptr, 0, bytes));
	cudaIpcMemHandle_t h;
	HN_CUDA(cudaIpcGetMemHandle(&h, *dev_ptr));
	std::memcpy(handle, &h, 64);
	HN_CATCH(ctx)
}
int hn_shared_open(hn_context* ctx, const unsigned char handle[64], void** dev_ptr) {
	HN_TRY(ctx)
	need(dev_ptr && handle, "hn_shared_open: bad arguments");
	HN_CUDA(cudaSetDevice(ctx->device));
	cudaIpcMemHandle_t h;
	std::memcpy(&h, handle, 64);
	HN_CUDA(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
	HN_CATCH(ctx)
}
int hn_shared_release(hn_context* ctx, void* dev_ptr, int owner) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	if (dev_ptr) { if (owner) HN_CUDA(cudaFree(dev_ptr)); else HN_CUDA(cudaIpcCloseMemHandle(dev_ptr)); }
	HN_CATCH(ctx)
}
int hn_scatter_tiles(hn_context* ctx, int width, int height, int rank, int world, const float* frame_dev, const float* depth_dev, float* frame_remote, float* depth_remote, void* cuda_stream) {
	HN_TRY(ctx)
	need(frame_dev && frame_remote && (depth_dev || !depth_remote) && world >= 1 && rank >= 0 && rank < world, "hn_scatter_tiles: bad arguments");
	HN_CUDA(cudaSetDevice(ctx->device));
	launch_scatter_tiles(width, height, rank, world, ctx->tile_w, ctx->tile_h, (const float4*)frame_dev, depth_dev, (float4*)frame_remote, depth_remote, (cudaStream_t)cuda_stream);
	HN_CUDA(cudaGetLastError());
	HN_CATCH(ctx)
}

// Host delivery of a sharded frame: the caller's host buffer (one POSIX shared-memory segment mapped by every rank's process) is page-locked
// and mapped into this process's device address space; the pointer returned is a valid frame_remote / depth_remote of hn_scatter_tiles, so
// every rank writes the tiles it owns straight into host memory over its own PCIe link (coalesced 16-byte stores, 1 KB per tile row)
// instead of the root rank reading the whole gathered frame back through one link.
int hn_host_register(hn_context* ctx, void* host_ptr, size_t bytes, void** dev_ptr) {
	HN_TRY(ctx)
	need(host_ptr && bytes > 0 && dev_ptr, "hn_host_register: bad arguments");
	HN_CUDA(cudaSetDevice(ctx->device));
	HN_CUDA(cudaHostRegister(host_ptr, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped));
	cudaError_t e = cudaHostGetDevicePointer(dev_ptr, host_ptr, 0);
	if (e != cudaSuccess) { cudaHostUnregister(host_ptr); HN_CUDA(e); }
	HN_CATCH(ctx)
}
int hn_host_unregister(hn_context* ctx, void* host_ptr) {
	HN_TRY(ctx)
	need(host_ptr != nullptr, "hn_host_unregister: bad arguments");
	HN_CUDA(cudaSetDevice(ctx->device));
	HN_CUDA(cudaDeviceSynchronize());
	HN_CUDA(cudaHostUnregister(host_ptr));
	HN_CATCH(ctx)
}

// ---------------------------------------------------------------- operator-level entry points

int hn_op_init_rays(hn_context* ctx, const float cam0[12], const float cam1[12], const float screen_center[2], const float relative_focal_length[2],
	const float rolling_shutter[4], uint32_t spp_index, int width, int height, float* origins, float* dirs, float* t, uint8_t* alive) {
	HN_TRY(ctx)
	need(ctx->has_bitfield, "occupancy grid not set");
	HN_CUDA(cudaSetDevice(ctx->device));
	FrameArgs a = make_args(ctx, cam0, cam1, screen_center, relative_focal_length, rolling_shutter, spp_index, width, height);
	a.world = 1;
	const size_t n = (size_t)width * height;
	float* d_o = scratch<float>(ctx, 0, n * 3); float* d_d = scratch<float>(ctx, 1, n * 3); float* d_t = scratch<float>(ctx, 2, n); uint8_t* d_a = scratch<uint8_t>(ctx, 3, n);
	launch_raygen_debug(a, ctx->bitfield.p, d_o, d_d, d_t, d_a, ctx->stream);
	HN_CUDA(cudaMemcpyAsync(origins, d_o, n * 12, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(dirs, d_d, n * 12, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(t, d_t, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(alive, d_a, n, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CATCH(ctx)
}

int hn_op_march_all(hn_context* ctx, uint32_t n, const float* origins, const float* dirs, const float* t0, const uint8_t* alive,
	uint32_t* counts, float* samples, uint32_t max_per_ray) {
	HN_TRY(ctx)
	need(ctx->has_bitfield, "occupancy grid not set");
	HN_CUDA(cudaSetDevice(ctx->device));
	const float rs[4] = {0, 0, 0, 1};
	FrameArgs a = make_args(ctx, ctx->camera, ctx->camera, ctx->opt.screen_center, ctx->opt.relative_focal_length, rs, 0, 1, 1);
	float* d_o = scratch<float>(ctx, 0, (size_t)n * 3); float* d_d = scratch<float>(ctx, 1, (size_t)n * 3); float* d_t = scratch<float>(ctx, 2, n);
	uint8_t* d_a = scratch<uint8_t>(ctx, 3, n); uint32_t* d_c = scratch<uint32_t>(ctx, 4, n);
	float* d_s = samples ? scratch<float>(ctx, 5, (size_t)n * max_per_ray * 4) : nullptr;
	HN_CUDA(cudaMemcpyAsync(d_o, origins, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(d_d, dirs, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(d_t, t0, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(d_a, alive, n, cudaMemcpyHostToDevice, ctx->stream));
	if (d_s) HN_CUDA(cudaMemsetAsync(d_s, 0, (size_t)n * max_per_ray * 16, ctx->stream));
	launch_march_all(a, ctx->bitfield.p, n, d_o, d_d, d_t, d_a, d_c, d_s, max_per_ray, ctx->stream);
	HN_CUDA(cudaMemcpyAsync(counts, d_c, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (d_s) HN_CUDA(cudaMemcpyAsync(samples, d_s, (size_t)n * max_per_ray * 16, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CATCH(ctx)
}

// a-16 irradiance probe: rays from the NeRF's centre (mode 0), from `origin` looking in (mode 1) or from n_origin^2 Halton-jittered
// centres (mode 2) on the n_theta x n_phi lattice of cylindrical_to_dir_nerf, traced through the NeRF like NerfTracer::trace_mesh
// (t = 0, no jitter, 200-iteration DDA) by the SAME pass loop as camera rays (march, tcgen05 network, composite), reduced into the
// probe texture. Kept on the device as the light of the mesh pass (option env_light_source = 1); optionally copied to the host.
int hn_render_probe(hn_context* ctx, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float origin[3], float* rgba_host) {
	HN_TRY(ctx)
	need(mode >= 0 && mode <= 2 && n_theta > 0 && n_phi > 0 && (mode != 2 || n_origin > 0), "hn_render_probe: bad arguments");
	need(mode != 1 || origin != nullptr, "hn_render_probe: mode 1 needs an origin");
	need(ctx->has_net && ctx->has_params && ctx->has_bitfield, "hn_render_probe: network, parameters and occupancy grid must be set");
	HN_CUDA(cudaSetDevice(ctx->device));
	const size_t n_rays = (size_t)n_theta * n_phi * (mode == 2 ? (size_t)n_origin * n_origin : 1u);
	need(n_rays < (1u << 30), "hn_render_probe: too many rays");
	cudaStream_t s = ctx->stream;
	const float rs[4] = {0.f, 0.f, 0.f, 1.f};
	FrameArgs a = make_args(ctx, ctx->camera, ctx->camera, ctx->opt.screen_center, ctx->opt.relative_focal_length, rs, 0, (int)n_rays, 1);
	a.depth_test = 0; a.depth_clip = 0; a.dda_iter_cap = 200; a.render_mode = 1; a.crop_rotated = 0;
	a.rank = 0; a.world = 1;
	ctx->probe_rays.ensure(n_rays); ctx->probe_depth.ensure(n_rays); ctx->probe_tex.ensure((size_t)n_theta * n_phi);
	HN_CUDA(cudaMemsetAsync(ctx->probe_rays.p, 0, n_rays * sizeof(float4), s));
	HN_CUDA(cudaMemsetAsync(ctx->probe_depth.p, 0, n_rays * sizeof(float), s));
	ensure_rays(ctx, n_rays);
	uint32_t* cnt = ctx->counters.p;
	HN_CUDA(cudaMemsetAsync(cnt, 0, CNT_N * sizeof(uint32_t), s));
	RaySoA rays = ray_soa(ctx);
	const float zero3[3] = {0.f, 0.f, 0.f};
	launch_probe_raygen(a, ctx->bitfield.p, mode, n_theta, n_phi, n_origin, origin ? origin : zero3, rays, ctx->alive_a.p, cnt, s);
	if (ctx->opt.eval_impl == 1 && ctx->opt.eval_ray_sh) {
		ctx->ray_sh.ensure(2 * n_rays);
		launch_ray_sh(ctx->alive_a.p, cnt + CNT_ALIVE_A, (uint32_t)n_rays, rays.d_mw, ctx->ray_sh.p, s);
		rays.sh = ctx->ray_sh.p;
	}
	hn_frame_stats keep = ctx->stats;
	Timer tm(false, s);
	run_passes(ctx, a, rays, ctx->probe_rays.p, ctx->probe_depth.p, false, s, tm);
	ctx->stats = keep;
	launch_probe_reduce(ctx->probe_rays.p, mode, n_theta, n_phi, n_origin, ctx->probe_tex.p, s);
	HN_CUDA(cudaGetLastError());
	ctx->probe_n_theta = (int)n_theta; ctx->probe_n_phi = (int)n_phi;
	if (rgba_host) HN_CUDA(cudaMemcpyAsync(rgba_host, ctx->probe_tex.p, (size_t)n_theta * n_phi * sizeof(float4), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	HN_CATCH(ctx)
}
// upload a probe texture computed elsewhere ([n_phi][n_theta][4]); nullptr removes it
int hn_set_probe(hn_context* ctx, const float* rgba, uint32_t n_theta, uint32_t n_phi) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	if (!rgba) { ctx->probe_tex.release(); ctx->probe_n_theta = ctx->probe_n_phi = 0; return 0; }
	need(n_theta > 0 && n_phi > 0, "hn_set_probe: empty texture");
	ctx->probe_tex.ensure((size_t)n_theta * n_phi);
	HN_CUDA(cudaMemcpy(ctx->probe_tex.p, rgba, (size_t)n_theta * n_phi * sizeof(float4), cudaMemcpyHostToDevice));
	ctx->probe_n_theta = (int)n_theta; ctx->probe_n_phi = (int)n_phi;
	HN_CATCH(ctx)
}
int hn_op_probe_rays(hn_context* ctx, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float origin[3], float* origins, float* dirs, uint8_t* alive) {
	HN_TRY(ctx)
	need(mode >= 0 && mode <= 2 && n_theta > 0 && n_phi > 0 && ctx->has_bitfield, "hn_op_probe_rays: bad arguments / no occupancy grid");
	HN_CUDA(cudaSetDevice(ctx->device));
	const size_t n_rays = (size_t)n_theta * n_phi * (mode == 2 ? (size_t)n_origin * n_origin : 1u);
	cudaStream_t s = ctx->stream;
	const float rs[4] = {0.f, 0.f, 0.f, 1.f};
	FrameArgs a = make_args(ctx, ctx->camera, ctx->camera, ctx->opt.screen_center, ctx->opt.relative_focal_length, rs, 0, (int)n_rays, 1);
	a.crop_rotated = 0;
	ensure_rays(ctx, n_rays);
	uint32_t* cnt = ctx->counters.p;
	HN_CUDA(cudaMemsetAsync(cnt, 0, CNT_N * sizeof(uint32_t), s));
	RaySoA rays = ray_soa(ctx);
	const float zero3[3] = {0.f, 0.f, 0.f};
	launch_probe_raygen(a, ctx->bitfield.p, mode, n_theta, n_phi, n_origin, origin ? origin : zero3, rays, ctx->alive_a.p, cnt, s);
	std::vector<float4> ot(n_rays), dm(n_rays);
	HN_CUDA(cudaMemcpyAsync(ot.data(), rays.o_t, n_rays * sizeof(float4), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaMemcpyAsync(dm.data(), rays.d_mw, n_rays * sizeof(float4), cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	for (size_t r = 0; r < n_rays; ++r) {
		origins[3 * r] = ot[r].x; origins[3 * r + 1] = ot[r].y; origins[3 * r + 2] = ot[r].z;
		dirs[3 * r] = dm[r].x; dirs[3 * r + 1] = dm[r].y; dirs[3 * r + 2] = dm[r].z;
		if (alive) alive[r] = ot[r].w < MAX_DEPTH_F ? 1 : 0;
	}
	HN_CATCH(ctx)
}

int hn_set_envmap(hn_context* ctx, const float* rgba, int width, int height) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	if (!rgba) { ctx->envmap.release(); ctx->env_w = ctx->env_h = 0; return 0; }
	need(width >= 2 && height >= 2, "envmap must be at least 2x2");
	ctx->envmap.release();
	ctx->envmap.ensure((size_t)width * height);
	HN_CUDA(cudaMemcpy(ctx->envmap.p, rgba, (size_t)width * height * sizeof(float4), cudaMemcpyHostToDevice));
	ctx->env_w = width; ctx->env_h = height;
	HN_CATCH(ctx)
}

int hn_op_read_envmap(hn_context* ctx, const float* dirs, uint32_t n, float* out) {
	HN_TRY(ctx)
	need(ctx->envmap.p != nullptr, "no environment map set");
	HN_CUDA(cudaSetDevice(ctx->device));
	const float rs[4] = {0, 0, 0, 1};
	FrameArgs a = make_args(ctx, ctx->camera, ctx->camera, ctx->opt.screen_center, ctx->opt.relative_focal_length, rs, 0, 1, 1);
	float* d_d = scratch<float>(ctx, 0, (size_t)n * 3); float4* d_o = scratch<float4>(ctx, 1, n);
	HN_CUDA(cudaMemcpyAsync(d_d, dirs, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
	launch_op_read_envmap(a, d_d, n, d_o, ctx->stream);
	HN_CUDA(cudaMemcpyAsync(out, d_o, (size_t)n * 16, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CATCH(ctx)
}

int hn_op_env_light_dirs(hn_context* ctx, uint32_t pixel_index, uint32_t spp_index, uint32_t K, const float normal[3], float* out) {
	HN_TRY(ctx)
	HN_CUDA(cudaSetDevice(ctx->device));
	float* d_o = scratch<float>(ctx, 0, (size_t)K * 3);
	launch_op_env_light_dirs(pixel_index, spp_index, K, normal, d_o, ctx->stream);
	HN_CUDA(cudaMemcpyAsync(out, d_o, (size_t)K * 12, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CATCH(ctx)
}

// the volume half of the env-light pass on caller-given rays: enter the volume, density-only passes, transmittance
int hn_op_transmittance(hn_context* ctx, uint32_t n, const float* origins, const float* dirs, float* T, uint32_t* n_samples) {
	HN_TRY(ctx)
	need(ctx->has_net && ctx->has_params && ctx->has_bitfield, "network, parameters and occupancy grid must be set");
	HN_CUDA(cudaSetDevice(ctx->device));
	if (!n) return 0;
	cudaStream_t s = ctx->stream;
	const float rs[4] = {0, 0, 0, 1};
	FrameArgs a = make_args(ctx, ctx->camera, ctx->camera, ctx->opt.screen_center, ctx->opt.relative_focal_length, rs, 0, 1, 1);
	a.depth_test = 0; a.depth_clip = 0;
	ensure_rays(ctx, n);
	RaySoA rays = ray_soa(ctx);
	std::vector<float4> h_o(n), h_d(n), h_c(n, make_float4(1.f, 1.f, 1.f, 1.f));
	for (uint32_t i = 0; i < n; ++i) {
		h_o[i] = make_float4(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2], 0.f);
		h_d[i] = make_float4(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2], 0.f);
	}
	HN_CUDA(cudaMemcpyAsync(rays.o_t, h_o.data(), (size_t)n * 16, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemcpyAsync(rays.d_mw, h_d.data(), (size_t)n * 16, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemcpyAsync(rays.rgba, h_c.data(), (size_t)n * 16, cudaMemcpyHostToDevice, s));
	HN_CUDA(cudaMemsetAsync(rays.n_steps, 0, (size_t)n * 4, s));
	HN_CUDA(cudaMemsetAsync(rays.pixel, 0, (size_t)n * 4, s));
	HN_CUDA(cudaMemsetAsync(ctx->counters.p, 0, CNT_N * sizeof(uint32_t), s));
	launch_secondary_start(a, ctx->bitfield.p, rays, n, ctx->alive_a.p, ctx->counters.p, s);
	Timer tm(false, s);
	ctx->stats = hn_frame_stats{};
	run_passes(ctx, a, rays, nullptr, nullptr, true, s, tm);
	HN_CUDA(cudaMemcpyAsync(h_c.data(), rays.rgba, (size_t)n * 16, cudaMemcpyDeviceToHost, s));
	if (n_samples) HN_CUDA(cudaMemcpyAsync(n_samples, rays.n_steps, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
	HN_CUDA(cudaStreamSynchronize(s));
	for (uint32_t i = 0; i < n; ++i) T[i] = h_c[i].w;
	HN_CATCH(ctx)
}

int hn_op_encode(hn_context* ctx, const float* positions, uint32_t n, uint16_t* out) {
	HN_TRY(ctx)
	need(ctx->has_params, "network parameters not set");
	HN_CUDA(cudaSetDevice(ctx->device));
	float* d_p = scratch<float>(ctx, 0, (size_t)n * 3); __half* d_o = scratch<__half>(ctx, 1, (size_t)n * 32);
	HN_CUDA(cudaMemcpyAsync(d_p, positions, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
	launch_encode(ctx->levels, ctx->net.F, ctx->params.p + ctx->net.grid_offset, d_p, n, d_o, ctx->stream);
	HN_CUDA(cudaMemcpyAsync(out, d_o, (size_t)n * 64, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CATCH(ctx)
}

int hn_op_network(hn_context* ctx, const float* positions, const float* dirs, uint32_t n, uint16_t* out) {
	HN_TRY(ctx)
	need(ctx->has_params, "network parameters not set");
	need(n % CHUNK == 0, "samples must come in ray chunks of 32");
	HN_CUDA(cudaSetDevice(ctx->device));
	const uint32_t chunks = n / CHUNK;
	float4* d_p = scratch<float4>(ctx, 0, n); float* d_d = scratch<float>(ctx, 1, (size_t)chunks * 3); uint2* d_o = scratch<uint2>(ctx, 2, n);
	HN_CUDA(cudaMemcpyAsync(d_p, positions, (size_t)n * 16, cudaMemcpyHostToDevice, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(d_d, dirs, (size_t)chunks * 12, cudaMemcpyHostToDevice, ctx->stream));
	const float rs[4] = {0, 0, 0, 1};
	FrameArgs a = make_args(ctx, ctx->camera, ctx->camera, ctx->opt.screen_center, ctx->opt.relative_focal_length, rs, 0, 1, 1);
	RaySoA none = {};
	launch_eval(ctx->net, ctx->levels, ctx->params.p, a, d_p, nullptr, nullptr, none, d_d, nullptr, chunks, d_o, ctx->opt.eval_impl, nullptr, nullptr, nullptr, nullptr, ctx->stream, false, param_textures(ctx));
	HN_CUDA(cudaMemcpyAsync(out, d_o, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CUDA(cudaGetLastError());
	HN_CATCH(ctx)
}

int hn_op_mesh_distance(hn_context* ctx, uint32_t n, const float* positions, int mode, float max_distance, float* distances, int32_t* tri_idx) {
	HN_TRY(ctx)
	need(ctx->scene_built && !ctx->meshes.empty(), "scene not built");
	need(mode >= 0 && mode <= 2, "mesh distance mode: 0 = unsigned, 1 = signed (watertight), 2 = signed (raystab); path-escape (OptiX only) is not built");
	need(max_distance > 0.0f, "max_distance must be positive");
	HN_CUDA(cudaSetDevice(ctx->device));
	float* d_p = scratch<float>(ctx, 0, (size_t)n * 3); float* d_d = scratch<float>(ctx, 1, n); int32_t* d_i = scratch<int32_t>(ctx, 2, n);
	HN_CUDA(cudaMemcpyAsync(d_p, positions, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
	launch_distance(ctx->scene_dev(), n, d_p, mode, max_distance, d_d, tri_idx ? d_i : nullptr, ctx->stream);
	HN_CUDA(cudaMemcpyAsync(distances, d_d, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (tri_idx) HN_CUDA(cudaMemcpyAsync(tri_idx, d_i, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CUDA(cudaGetLastError());
	HN_CATCH(ctx)
}
int hn_op_trace(hn_context* ctx, uint32_t n, const float* origins, const float* dirs, int32_t* mesh_idx, int32_t* tri_idx, float* t) {
	HN_TRY(ctx)
	need(ctx->scene_built && !ctx->meshes.empty(), "scene not built");
	HN_CUDA(cudaSetDevice(ctx->device));
	float* d_o = scratch<float>(ctx, 0, (size_t)n * 3); float* d_d = scratch<float>(ctx, 1, (size_t)n * 3);
	int32_t* d_m = scratch<int32_t>(ctx, 2, n); int32_t* d_i = scratch<int32_t>(ctx, 3, n); float* d_t = scratch<float>(ctx, 4, n);
	HN_CUDA(cudaMemcpyAsync(d_o, origins, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(d_d, dirs, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
	launch_trace(ctx->scene_dev(), ctx->opt.compat_nearest_aabb_only, n, d_o, d_d, d_m, d_i, d_t, ctx->stream);
	HN_CUDA(cudaMemcpyAsync(mesh_idx, d_m, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(tri_idx, d_i, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaMemcpyAsync(t, d_t, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	HN_CUDA(cudaStreamSynchronize(ctx->stream));
	HN_CUDA(cudaGetLastError());
	HN_CATCH(ctx)
}

} // extern "C"
