// Internal declarations shared by the host side (hn_*.cpp) and the CUDA translation units (hn_*.cu).
#pragma once
#include <cstdint>
#include <cstddef>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

namespace hn {

constexpr uint32_t GRIDSIZE = 128;                     // NERF_GRIDSIZE, nerf_device.cuh:23
constexpr uint32_t GRID_CELLS = GRIDSIZE * GRIDSIZE * GRIDSIZE;
constexpr uint32_t CASCADES = 8;                       // NERF_CASCADES, nerf_device.cuh:28
constexpr uint32_t BITFIELD_BYTES = GRID_CELLS / 8 * CASCADES;
constexpr float MAX_DEPTH_F = 16384.0f;                // MAX_DEPTH(), common_device.cuh:32
constexpr float MAX_DIST_F = 100.0f;                   // MAX_DIST in triangle_bvh.cu / geometry_bvh.cu
constexpr uint32_t CHUNK = 32;                         // samples per ray per pass (one warp = one ray chunk)
constexpr uint32_t MAX_LEVELS = 16;

// ---- network ----
struct NetCfg {
	uint32_t L = 0, F = 0, log2_T = 0, base_res = 0;
	float per_level_scale = 0.f;
	uint32_t density_hidden = 1, rgb_hidden = 2;
	uint32_t offsets[MAX_LEVELS + 1] = {};
	float scales[MAX_LEVELS] = {};
	uint32_t res[MAX_LEVELS] = {};
	size_t n_params = 0, grid_offset = 0, density_w_offset = 0, rgb_w_offset = 0;
};

// Level table handed to the gather kernel by value.
struct GridLevels {
	uint32_t offset[MAX_LEVELS];   // entry offset of the level
	uint32_t size[MAX_LEVELS];     // hashmap_size of the level
	uint32_t res[MAX_LEVELS];
	float scale[MAX_LEVELS];
	uint32_t hashed[MAX_LEVELS];   // 1 = coherent prime hash, 0 = dense index
	uint32_t mask[MAX_LEVELS];     // size - 1 when the size is a power of two (index reduction by AND), else 0
	uint32_t L;
};

// ---- per-frame arguments (by value to kernels) ----
struct FrameArgs {
	int w, h;
	uint32_t spp;
	float cam0[12], cam1[12];
	float rolling_shutter[4];
	float screen_center[2];
	float focal[2];
	float pixel_offset[2]; // ld_random_pixel_offset(snap ? 0 : spp), computed on the host
	float near_distance;
	int lens_mode;            // ELensMode of the NeRF ray init (0 = perspective)
	float lens_params[7];
	float aperture_size, focus_z;  // depth of field: focus_z = slice_plane_z + scale (src/testbed_nerf.cu:2342)
	float parallax_shift[3];
	// VR pieces of uv_to_ray (common_device.cuh:117-246,416-483): foveation coefficients per axis (al bl cl am bm ar br cr switch_left
	// switch_right inv_switch_left inv_switch_right; computed on the host by FoveationPiecewiseQuadratic's constructor), hidden-area mask
	// (0 = hidden) and the distortion grid of the NeRF rays
	int foveated; float fov[2][12];
	const uint8_t* hidden_mask; int hidden_w, hidden_h;
	const float2* distortion; int dist_w, dist_h;
	float render_aabb[6], train_aabb[6];
	float aabb_to_local[9]; int crop_rotated; // m_render_aabb_to_local (column-major); crop_rotated = 0 -> identity, skipped
	float occ_aabb[6];     // inflated bounds of every set occupancy cell of mips 0..max_mip (empty: lo > hi)
	const uint32_t* occ_coarse; // [CASCADES][32^3 bits] dilated coarse occupancy (occ_coarse_kernel), nullptr = slab bound only
	uint32_t max_mip;
	int simple_walk;       // option march_simple: allow the compile-time specialised walk kernels (A/B switch)
	float cone_angle;
	float min_transmittance;
	int rgb_act, density_act, linear_colors, dda_iter_cap;
	int depth_test, depth_clip;
	int render_mode;       // ERenderMode of the NeRF pass (1 = Shade)
	int eval_texsplit;
	int eval_gather, eval_groups; // options eval_gather / eval_groups (gather variant, tile groups per CTA of the tcgen05 kernel)
	float depth_scale;
	int rank, world, tile_w, tile_h;
	// mesh pass
	float mesh_aabb[6];
	float sun[3], up[3];
	float floor_y;
	float brdf[13]; // metallic subsurface specular roughness sheen clearcoat clearcoat_gloss basecolor[3] ambientcolor[3]
	int compat_nearest_aabb_only;
	// a-16 environment map (device, lat-long float4 [env_h][env_w]); env_samples = K light directions per mesh hit
	const float4* envmap;
	int env_w, env_h, env_samples;
	// a-16 probe: lat-long radiance texture of the NeRF (hn_render_probe) as light source of the mesh pass when env_source == 1;
	// env_sampling 0 = cosine hemisphere, 1 = uniform spherical cap around the direction to the NeRF's bounding sphere
	const float4* probe;
	int probe_n_theta, probe_n_phi, env_source, env_sampling;
};

// per-image camera of the training set (TrainingImageMetadata, nerf.h: focal_length, principal_point, lens)
struct ImageMeta {
	float focal[2];       // pixels
	float principal[2];   // fraction of the image
	int lens_mode;        // ELensMode
	float lens_params[7];
};

// mesh hits that gather environment light: filled by the mesh pass (one entry per shaded triangle hit, warp-aggregated append)
struct EnvHits {
	float4* pos_pix;   // hit position (already offset along the facing normal), pixel index bits in .w
	float4* nrm;       // viewer-facing unit normal
	float4* albedo;    // Lambertian albedo of the hit (floor checker or basecolor^2)
	uint32_t* count;   // device counter
};

// ---- ray state, SoA, indexed by ray id (assigned at ray generation) ----
struct RaySoA {
	float4* o_t;     // origin.xyz, t
	float4* d_mw;    // dir.xyz, max_weight
	float4* rgba;
	float* depth;
	uint32_t* pixel;
	uint32_t* n_steps;
	float* t_far;    // ray parameter beyond which no set occupancy cell can be met (occ_t_far_tight), written when the ray is started
	uint4* sh;       // [ray][2] fp16 SH coefficients of the direction, filled once per frame for the first live list (may be nullptr)
};

// ---- BVH (4-wide, 128 B nodes) ----
struct alignas(16) Node4 {
	float lo_x[4], lo_y[4], lo_z[4], hi_x[4], hi_y[4], hi_z[4];
	int32_t child[4];  // inner: node index (>0); leaf: ~first_triangle; unused slot: count == 0 and child == 0
	int32_t count[4];  // leaf: number of triangles (> 0); inner: -1; unused: 0
};
static_assert(sizeof(Node4) == 128, "Node4 must be one 128-byte line");

struct MeshDev {
	uint32_t node_root;   // index of the mesh's root node in the shared node array
	uint32_t tri_first;   // first triangle of the mesh in the shared (leaf-ordered) triangle array
	float bb[6];          // exact box of the mesh (geometry_bvh.cu:18-35)
};

struct SceneDev {
	const Node4* nodes;
	const float4* tris;   // 3 float4 per triangle: (a, orig_idx bits), (b, 0), (c, 0), leaf order
	const MeshDev* meshes;
	uint32_t n_meshes;
	uint32_t* stack_overflow; // device counter: children that did not fit the traversal stack (a dropped child = a possibly missed hit)
};

// ---- launchers (defined in the .cu files) ----
// exact TU (compiled with -fmad=false, no fast math): bit-exact against the oracle
void launch_grid_to_bitfield(const float* grid, uint32_t n_cascades, float thresh, uint8_t* bits, cudaStream_t s);
void launch_grid_mean(const float* grid, float* mean_out, cudaStream_t s);
void launch_occ_bounds(const uint8_t* bits, int* out, cudaStream_t s);
void launch_occ_coarse(const uint8_t* bits, uint32_t* coarse, cudaStream_t s);
// 2-D render modes and the foveated blit (hn_march.cu / hn_composite.cu)
void launch_distortion(const FrameArgs& a, float4* frame, float* depth_buffer, cudaStream_t s);
void launch_slice_raygen(const FrameArgs& a, float plane_z, uint32_t first, uint32_t n, float4* samples, uint32_t* sample_counts, float* dirs, uint32_t* pixels,
	float* depth_buffer, cudaStream_t s);
void launch_slice_shade(uint32_t n, const uint32_t* pixels, const uint2* net_out, int rgb_act, int density_act, int linear_colors, float4* frame, cudaStream_t s);
void launch_unwarp_foveated(const float coeffs[24], int w_in, int h_in, const float4* frame_in, const float* depth_in, int w_out, int h_out, float4* frame_out, float* depth_out, cudaStream_t s);
void launch_raygen(const FrameArgs& a, const uint8_t* bits, RaySoA rays, uint32_t* alive_list, uint32_t* counters,
	float* depth_buffer, float4* frame, cudaStream_t s);
void launch_raygen_debug(const FrameArgs& a, const uint8_t* bits, float* o, float* d, float* t, uint8_t* alive, cudaStream_t s);
void launch_march(const FrameArgs& a, const uint8_t* bits, RaySoA rays, const uint32_t* alive_list, const uint32_t* n_alive_ptr,
	float4* samples, uint32_t* sample_counts, uint32_t* counters, const float* depth_buffer, uint32_t max_rays, cudaStream_t s);
void launch_march_all(const FrameArgs& a, const uint8_t* bits, uint32_t n, const float* o, const float* d, const float* t0, const uint8_t* alive,
	uint32_t* counts, float* samples, uint32_t max_per_ray, cudaStream_t s);

void launch_train_march(const FrameArgs& a, const uint8_t* bits, uint32_t n_rays, const float* o, const float* d, const float* t, uint32_t max_samples,
	uint32_t* counter, uint32_t* numsteps, uint32_t* base, float4* coords, uint32_t* sample_ray, float4* staging /* [n_rays][1024] */, cudaStream_t s);

// network TU
// distance queries on the scene BVH (triangle_bvh.cu:195-303): mode 0 unsigned, 1 signed (watertight rule)
void launch_distance(const SceneDev& sc, uint32_t n, const float* pos, int mode, float max_distance, float* out, int32_t* tri_out, cudaStream_t s);
// ERenderMode::Normals: overwrite the (r,g,b) of the network output with the unit normal from the density gradient
void launch_normals(const NetCfg& cfg, const GridLevels& g, const __half* params, const FrameArgs& a, const float4* samples, const uint32_t* sample_counts,
	const uint32_t* n_chunks_ptr, uint32_t max_chunks, uint2* out, cudaStream_t s);
void launch_encode(const GridLevels& g, uint32_t F, const __half* table, const float* positions, uint32_t n, __half* out, cudaStream_t s);
// The parameter buffer seen as 16-byte and 8-byte texels: the tcgen05 kernel gathers the fine hash-grid levels through the TEX pipe.
struct ParamTextures { cudaTextureObject_t tex16 = 0, tex8 = 0; const void* ptr = nullptr; };
void launch_probe_raygen(const FrameArgs& a, const uint8_t* bits, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float origin[3], RaySoA rays,
	uint32_t* alive_list, uint32_t* counters, cudaStream_t s);                                                   // hn_march.cu
void launch_probe_reduce(const float4* ray_rgba, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, float4* tex, cudaStream_t s); // hn_march.cu
void launch_scatter_tiles(int w, int h, int rank, int world, int tw, int th, const float4* frame, const float* depth, float4* frame_remote, float* depth_remote, cudaStream_t s); // hn_composite.cu
void launch_ray_sh(const uint32_t* list, const uint32_t* n_ptr, uint32_t n_max, const float4* d_mw, uint4* out, cudaStream_t s); // hn_network_tc.cu
void launch_eval(const NetCfg& cfg, const GridLevels& g, const __half* params, const FrameArgs& a, const float4* samples,
	const uint32_t* sample_counts, const uint32_t* alive_list, RaySoA rays, const float* dirs_override, const uint32_t* n_chunks_ptr,
	uint32_t max_chunks, uint2* out, int impl, float4* fused_frame, float* fused_depth, uint32_t* survive, uint32_t* counters, cudaStream_t s,
	bool density_only = false, const ParamTextures* tex = nullptr);
// fused_frame != nullptr (impl 1 only): composite inside the network kernel, write survive[chunk]; then launch_compact builds the next list
void launch_compact(const uint32_t* alive_in, const uint32_t* n_alive_ptr, const uint32_t* survive, uint32_t* alive_out, uint32_t* n_alive_out, uint32_t max_rays, cudaStream_t s); // impl: 0 = warp MMA (mma.sync), 1 = tcgen05/TMEM

// composite TU
void launch_composite(const FrameArgs& a, RaySoA rays, const uint32_t* alive_in, const uint32_t* n_alive_ptr, uint32_t* alive_out,
	uint32_t* counters, const float4* samples, const uint32_t* sample_counts, const uint2* net_out,
	float4* frame, float* depth_buffer, uint32_t max_rays, cudaStream_t s);
void launch_accumulate(const float4* frame, float4* accum, uint32_t n, uint32_t spp_index, cudaStream_t s);
void launch_tonemap(const float4* accum, float4* out, uint32_t n, int to_srgb, float exposure, const float* background, int curve, int clamp_output, int unmultiply_alpha, cudaStream_t s);
void launch_pack_tiles(int w, int h, int rank, int world, int tw, int th, const float4* frame, const float* depth, float* packed, bool unpack, cudaStream_t s);

// mesh TU (exact flags as well)
void launch_mesh_pass(const FrameArgs& a, SceneDev scene, float4* frame, float* depth_buffer, EnvHits hits, cudaStream_t s);
// a-16: K light rays per hit [first, first+n): direction, mesh visibility, envmap radiance -> ray slots (hit-first)*K + k
void launch_envlight_gen(const FrameArgs& a, SceneDev scene, EnvHits hits, uint32_t first, uint32_t n, RaySoA rays, cudaStream_t s);
// frame[pixel].rgb += albedo / K * sum_k radiance_k * T_k
void launch_envlight_reduce(const FrameArgs& a, EnvHits hits, uint32_t first, uint32_t n, RaySoA rays, float4* frame, cudaStream_t s);
// light rays with radiance enter the volume: AABB entry + advance to the first occupied voxel; live ones are appended to the list
void launch_secondary_start(const FrameArgs& a, const uint8_t* bits, RaySoA rays, uint32_t n_rays, uint32_t* alive_list, uint32_t* counters, cudaStream_t s);
// one warp per ray chunk: T *= prod exp(-sigma dt), stop below min_transmittance; survive flags for launch_compact
void launch_transmittance(const FrameArgs& a, RaySoA rays, const uint32_t* alive_list, const uint32_t* n_alive_ptr, const float4* samples,
	const uint32_t* sample_counts, const uint2* net_out, uint32_t* survive, uint32_t* counters, uint32_t max_rays, cudaStream_t s);
void launch_trace(SceneDev scene, int compat, uint32_t n, const float* o, const float* d, int32_t* mesh, int32_t* tri, float* t, cudaStream_t s);

void launch_op_read_envmap(const FrameArgs& a, const float* dirs, uint32_t n, float4* out, cudaStream_t s);
void launch_op_env_light_dirs(uint32_t pix, uint32_t spp, uint32_t K, const float n[3], float* out, cudaStream_t s);

// counters layout (uint32 array on the device)
enum Counter : uint32_t { CNT_ALIVE_A = 0, CNT_ALIVE_B = 1, CNT_SAMPLES_LO = 2, CNT_SAMPLES_HI = 3, CNT_COMPOSITED_LO = 4, CNT_COMPOSITED_HI = 5, CNT_RAYS = 6, CNT_ERR = 7, CNT_N = 8 };

} // namespace hn
