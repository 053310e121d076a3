// (a) Ray generation + occupancy-grid DDA marching for sm_100a.
// IEEE-exact translation unit: built with -fmad=false and without fast math so that ray origins/directions, start t,
// per-ray sample counts and sample positions are bit-identical to the CPU oracle.
// Layout: SoA ray state (float4 loads/stores), warp-aggregated compaction of live rays (one atomic per warp).
#include "hn_device_math.cuh"

namespace hn {

// ---------------------------------------------------------------- camera (tcnn vec.h:1060-1187, common_device.cuh:651-659)
struct Quat { float w, x, y, z; };
struct M3 { V3 c[3]; };
struct Cam { V3 c[4]; };

__device__ __forceinline__ Cam ldcam(const float* f) {
	Cam m;
	#pragma unroll
	for (int i = 0; i < 4; ++i) m.c[i] = {f[3 * i], f[3 * i + 1], f[3 * i + 2]};
	return m;
}
__device__ __forceinline__ V3 mul3(const M3& m, V3 v) {
	V3 r = {0.0f, 0.0f, 0.0f};
	r.x += m.c[0].x * v.x; r.y += m.c[0].y * v.x; r.z += m.c[0].z * v.x;
	r.x += m.c[1].x * v.y; r.y += m.c[1].y * v.y; r.z += m.c[1].z * v.y;
	r.x += m.c[2].x * v.z; r.y += m.c[2].y * v.z; r.z += m.c[2].z * v.z;
	return r;
}
__device__ Quat quat_from(const M3& m) {
	Quat q;
	float tr = m.c[0].x + m.c[1].y + m.c[2].z;
	if (tr > 0.0f) {
		float S = sqrtf(tr + 1.0f) * 2.0f;
		q.w = 0.25f * S;
		q.x = (m.c[1].z - m.c[2].y) / S;
		q.y = (m.c[2].x - m.c[0].z) / S;
		q.z = (m.c[0].y - m.c[1].x) / S;
	} else if (m.c[0].x > m.c[1].y && m.c[0].x > m.c[2].z) {
		float S = sqrtf(1.0f + m.c[0].x - m.c[1].y - m.c[2].z) * 2.0f;
		q.w = (m.c[1].z - m.c[2].y) / S;
		q.x = 0.25f * S;
		q.y = (m.c[1].x + m.c[0].y) / S;
		q.z = (m.c[2].x + m.c[0].z) / S;
	} else if (m.c[1].y > m.c[2].z) {
		float S = sqrtf(1.0f + m.c[1].y - m.c[0].x - m.c[2].z) * 2.0f;
		q.w = (m.c[2].x - m.c[0].z) / S;
		q.x = (m.c[1].x + m.c[0].y) / S;
		q.y = 0.25f * S;
		q.z = (m.c[2].y + m.c[1].z) / S;
	} else {
		float S = sqrtf(1.0f + m.c[2].z - m.c[0].x - m.c[1].y) * 2.0f;
		q.w = (m.c[0].y - m.c[1].x) / S;
		q.x = (m.c[2].x + m.c[0].z) / S;
		q.y = (m.c[2].y + m.c[1].z) / S;
		q.z = 0.25f * S;
	}
	return q;
}
__device__ __forceinline__ float qdot(Quat a, Quat b) { return (a.w * b.w + a.x * b.x) + (a.y * b.y + a.z * b.z); }
__device__ __forceinline__ Quat qnormalize(Quat a) {
	float len = sqrtf(qdot(a, a));
	if (len <= 0.0f) return {1.0f, 0.0f, 0.0f, 0.0f};
	return {a.w / len, a.x / len, a.y / len, a.z / len};
}
__device__ __forceinline__ Quat qscale(Quat a, float s) { return {a.w * s, a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ Quat qadd(Quat a, Quat b) { return {a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ Quat qslerp(Quat x, Quat y, float t) {
	Quat z = y;
	float cos_theta = qdot(x, y);
	if (cos_theta < 0.0f) { z = {-y.w, -y.x, -y.y, -y.z}; cos_theta = -cos_theta; }
	if (cos_theta > 1.0f - FLT_EPSILON) return qadd(qscale(x, 1.0f - t), qscale(z, t));
	// cam0 != cam1: acosf/sinf are not bit-identical between libm and CUDA -> tolerance parity only in this branch
	float angle = acosf(cos_theta);
	Quat r = qadd(qscale(x, sinf((1.0f - t) * angle)), qscale(z, sinf(t * angle)));
	float s = sinf(angle);
	return {r.w / s, r.x / s, r.y / s, r.z / s};
}
__device__ __forceinline__ M3 quat_to_mat3(Quat q) {
	float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z;
	float qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
	float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
	M3 m;
	m.c[0] = {1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy)};
	m.c[1] = {2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx)};
	m.c[2] = {2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy)};
	return m;
}

// init_rays_with_payload_kernel_nerf (src/testbed_nerf.cu:1428-1476) + uv_to_ray pinhole (common_device.cuh:416-483)
// the un-normalised ray of a pixel together with its uv and its (rolling-shutter) camera: what the 2-D render modes need
__device__ bool pixel_ray_raw(const FrameArgs& a, uint32_t x, uint32_t y, V3& origin, V3& d, float& u, float& v, M3& rot, V3& pos) {
	uint32_t idx = x + (uint32_t)a.w * y;
	u = ((float)x + a.pixel_offset[0]) / (float)a.w;
	v = ((float)y + a.pixel_offset[1]) / (float)a.h;
	float mb = hn_ld_random_val(a.spp, idx * 72239731u);
	float pixel_t = a.rolling_shutter[0] + a.rolling_shutter[1] * u + a.rolling_shutter[2] * v + a.rolling_shutter[3] * mb;
	Cam c0 = ldcam(a.cam0), c1 = ldcam(a.cam1);
	M3 m0 = {{c0.c[0], c0.c[1], c0.c[2]}}, m1 = {{c1.c[0], c1.c[1], c1.c[2]}};
	rot = quat_to_mat3(qnormalize(qslerp(qnormalize(quat_from(m0)), qnormalize(quat_from(m1)), pixel_t)));
	pos = mix3(c0.c[3], c1.c[3], pixel_t);
	return camera_ray(a, true, u, v, rot.c[0], rot.c[1], rot.c[2], pos, origin, d);
}
__device__ bool pixel_ray(const FrameArgs& a, uint32_t x, uint32_t y, V3& origin, V3& dir) {
	V3 d, pos; M3 rot; float u, v;
	if (!pixel_ray_raw(a, x, y, origin, d, u, v, rot, pos)) { origin = mk(0.0f, 0.0f, 0.0f); dir = mk(0.0f, 0.0f, 0.0f); return false; } // Ray::invalid()
	dir = normalize3(d);
	return true;
}

// ---------------------------------------------------------------- occupancy DDA (nerf_device.cuh:316-534)
__device__ __forceinline__ uint32_t expand_bits(uint32_t v) {
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}
__device__ __forceinline__ uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) { return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2); }
__device__ __forceinline__ uint32_t morton3D_invert(uint32_t x) {
	x = x & 0x49249249u;
	x = (x | (x >> 2)) & 0xc30c30c3u;
	x = (x | (x >> 4)) & 0x0f00f00fu;
	x = (x | (x >> 8)) & 0xff0000ffu;
	x = (x | (x >> 16)) & 0x0000ffffu;
	return x;
}
__device__ __forceinline__ bool occupied_at(V3 pos, const uint8_t* __restrict__ bits, uint32_t mip) {
	float mip_scale = scalbnf(1.0f, -(int)mip);
	pos = pos - 0.5f;
	pos = pos * mip_scale;
	pos = pos + 0.5f;
	int ix = (int)(pos.x * (float)GRIDSIZE), iy = (int)(pos.y * (float)GRIDSIZE), iz = (int)(pos.z * (float)GRIDSIZE);
	if (ix < 0 || ix >= (int)GRIDSIZE || iy < 0 || iy >= (int)GRIDSIZE || iz < 0 || iz >= (int)GRIDSIZE) return false;
	uint32_t idx = morton3D(ix, iy, iz);
	return __ldg(bits + idx / 8 + (GRID_CELLS * mip) / 8) & (1 << (idx % 8));
}
// to_stepping_space / from_stepping_space / advance_n_steps / calc_dt: hn_device_math.cuh (shared with the training ray generator)
__device__ __forceinline__ float min_ba(float a, float b) { return b < a ? b : a; } // std::min
__device__ __forceinline__ float distance_to_next_voxel(V3 pos, V3 dir, V3 idir, float res) {
	V3 p = res * (pos - 0.5f);
	float tx = (floorf(p.x + 0.5f + 0.5f * copysignf(1.0f, dir.x)) - p.x) * idir.x;
	float ty = (floorf(p.y + 0.5f + 0.5f * copysignf(1.0f, dir.y)) - p.y) * idir.y;
	float tz = (floorf(p.z + 0.5f + 0.5f * copysignf(1.0f, dir.z)) - p.z) * idir.z;
	float t = min_ba(min_ba(tx, ty), tz);
	return fmaxf(t / res, 0.0f);
}
__device__ __forceinline__ float advance_to_next_voxel(float t, float cone, V3 pos, V3 dir, V3 idir, uint32_t mip) {
	float res = scalbnf((float)GRIDSIZE, -(int)mip);
	float t_target = t + distance_to_next_voxel(pos, dir, idir, res);
	t = to_stepping_space(t, cone);
	t_target = to_stepping_space(t_target, cone);
	return from_stepping_space(t + ceilf(fmaxf(t_target - t, 0.5f)), cone);
}
__device__ __forceinline__ uint32_t mip_from_pos(V3 pos) {
	int exponent;
	float maxval = fmaxf(fmaxf(fabsf(pos.x - 0.5f), fabsf(pos.y - 0.5f)), fabsf(pos.z - 0.5f));
	frexpf(maxval, &exponent);
	return (uint32_t)min(max(exponent + 1, 0), (int)CASCADES - 1);
}
// crop box test with the optional rotation render_aabb_to_local (src/testbed_nerf.cu:1530-1532, nerf_device.cuh:461-495):
// R = nullptr is the axis-aligned box (the common case costs one uniform predicate)
__device__ __forceinline__ V3 to_crop_local(const float* __restrict__ R, V3 v) {
	if (!R) return v;
	V3 r = mk(0.0f, 0.0f, 0.0f);
	r.x += R[0] * v.x; r.y += R[1] * v.x; r.z += R[2] * v.x;
	r.x += R[3] * v.y; r.y += R[4] * v.y; r.z += R[5] * v.y;
	r.x += R[6] * v.z; r.y += R[7] * v.z; r.z += R[8] * v.z;
	return r;
}
#define HN_CROP_R(a) ((a).crop_rotated ? (a).aabb_to_local : nullptr)
// nerf_device.cuh:461-495 (iter_cap == 0) / :497-534 (iter_cap == 200)
// t_far: parameter beyond which the ray has left the (inflated) bounds of every set occupancy cell. No sample can follow, so
// the reference's loop is bound to end in MAX_DEPTH there; returning it at once is exact and removes the cell-by-cell walk
// through empty space to the edge of the render box.
// SIMPLE: the common set-up (cone angle 0, one cascade, axis-aligned crop box, a bitfield) as compile-time constants - the stepping
// space is linear (no logf / expf paths), the mip is 0, the crop rotation and the null checks fold away: same arithmetic on the
// same values, about half the instructions per trip of the walk.
template <bool SIMPLE = false>
__device__ __forceinline__ float advance_to_occupied(float t, float cone_in, V3 o, V3 d, V3 idir, const uint8_t* __restrict__ bits_in, uint32_t min_mip_in, uint32_t max_mip_in,
	const Box& aabb, int iter_cap, float t_far, const float* __restrict__ R_in = nullptr) {
	const float cone = SIMPLE ? 0.0f : cone_in;
	const uint32_t min_mip = SIMPLE ? 0u : min_mip_in, max_mip = SIMPLE ? 0u : max_mip_in;
	const float* __restrict__ R = SIMPLE ? nullptr : R_in;
	const uint8_t* __restrict__ bits = bits_in;
	for (int i = 1; iter_cap <= 0 || i < iter_cap; ++i) {
		V3 pos = o + t * d;
		if (t >= MAX_DEPTH_F || t > t_far || !box_contains(aabb, to_crop_local(R, pos))) return MAX_DEPTH_F;
		uint32_t mip = SIMPLE ? 0u : min(max(mip_from_pos(pos), min_mip), max_mip);
		if ((!SIMPLE && !bits) || occupied_at(pos, bits, mip)) return t;
		while (mip < max_mip && !occupied_at(pos, bits, mip + 1)) ++mip;
		t = advance_to_next_voxel(t, cone, pos, d, idir, mip);
	}
	return MAX_DEPTH_F;
}

// Far end of the ray inside the occupancy bounds (conservative: robust slab test on the inflated box, plus a margin).
// A ray that misses the bounds gets -1: it can never meet a set cell.
__device__ __forceinline__ float occ_t_far(const FrameArgs& a, const uint8_t* bits, V3 o, V3 d) {
	if (!bits) return FLT_MAX;
	float tmin = -FLT_MAX, tmax = FLT_MAX;
	const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
	#pragma unroll
	for (int k = 0; k < 3; ++k) {
		const float lo = a.occ_aabb[k], hi = a.occ_aabb[3 + k];
		if (dd[k] == 0.0f) {
			if (oo[k] < lo || oo[k] > hi) return -1.0f;
		} else {
			const float t0 = (lo - oo[k]) / dd[k], t1 = (hi - oo[k]) / dd[k];
			tmin = fmaxf(tmin, fminf(t0, t1));
			tmax = fminf(tmax, fmaxf(t0, t1));
		}
	}
	if (!(tmin <= tmax)) return -1.0f;
	return tmax + 1e-3f;
}

// Tight far end: the slab test above bounds the ray by the BOX of the set cells; a ray that grazes that box, or has left the
// occupied region long before it leaves the box, still walks the empty cells one by one (up to ~150 skip trips through the
// corner of a sphere's box - the slowest warps of raygen_kernel and of every march pass). Each mip therefore also has a 32^3
// coarse grid (one coarse cell = 4x4x4 fine cells = one aligned 64-bit word of the Morton-ordered bitfield) whose flag is set when
// the cell OR one of its 26 neighbours holds a set fine cell (occ_coarse_kernel). The ray is walked BACKWARDS from its slab exit
// through that grid with a plain 3-D DDA: the first flagged cell met is the last one along the ray, and its far side bounds every
// t at which the fine walk can still find a set cell. The one-cell dilation makes the bound safe against the rounding of this
// DDA (a cell the ray clips within floating-point error has a flagged neighbour that the walk does visit at about the same t);
// the exact fine walk is unchanged, only where it is cut off, so counts and positions stay bit-identical.
__device__ float occ_t_far_tight(const FrameArgs& a, const uint8_t* bits, V3 o, V3 d) {
	const float coarse_far = occ_t_far(a, bits, o, d);
	if (!bits || !a.occ_coarse || coarse_far < 0.0f) return coarse_far;
	float best = -1.0f;
	const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
	for (uint32_t m = 0; m <= a.max_mip; ++m) {
		const float inv = scalbnf(1.0f, -(int)m); // mip-m grid coordinates u = (p - 0.5) 2^-m + 0.5 in [0,1]^3
		float q[3], e[3];
		float t0 = 0.0f, t1 = coarse_far;
		bool miss = false;
		#pragma unroll
		for (int k = 0; k < 3; ++k) {
			q[k] = (oo[k] - 0.5f) * inv + 0.5f; e[k] = dd[k] * inv;
			if (e[k] == 0.0f) { if (q[k] < 0.0f || q[k] > 1.0f) miss = true; }
			else { const float ta = (0.0f - q[k]) / e[k], tb = (1.0f - q[k]) / e[k]; t0 = fmaxf(t0, fminf(ta, tb)); t1 = fminf(t1, fmaxf(ta, tb)); }
		}
		if (miss || !(t0 <= t1)) continue;
		// reversed ray from the exit point, parameter s in [0, t1 - t0]; scalars per axis (no dynamically indexed arrays)
		int gx, gy, gz, sx, sy, sz;
		float mx, my, mz, dx, dy, dz;
		#define HN_AXIS(K, G, S, M, D) { \
			const float x = (q[K] + t1 * e[K]) * 32.0f, r = -e[K] * 32.0f; /* position and direction in coarse cells */ \
			G = min(max((int)floorf(x), 0), 31); \
			if (r > 0.0f) { S = 1; M = ((float)(G + 1) - x) / r; D = 1.0f / r; } \
			else if (r < 0.0f) { S = -1; M = ((float)G - x) / r; D = -1.0f / r; } \
			else { S = 0; M = FLT_MAX; D = FLT_MAX; } \
			M = fmaxf(M, 0.0f); }
		HN_AXIS(0, gx, sx, mx, dx) HN_AXIS(1, gy, sy, my, dy) HN_AXIS(2, gz, sz, mz, dz)
		#undef HN_AXIS
		const uint32_t* __restrict__ flags = a.occ_coarse + m * 1024u;
		const float s_end = t1 - t0;
		float s = 0.0f; // parameter at which the current cell was entered
		for (int it = 0; it < 100; ++it) {
			const uint32_t cell = (uint32_t)gx + 32u * (uint32_t)gy + 1024u * (uint32_t)gz;
			if (__ldg(flags + (cell >> 5)) & (1u << (cell & 31u))) { best = fmaxf(best, t1 - s); break; }
			if (mx <= my && mx <= mz) { s = mx; gx += sx; mx += dx; if (gx < 0 || gx > 31) break; }
			else if (my <= mz) { s = my; gy += sy; my += dy; if (gy < 0 || gy > 31) break; }
			else { s = mz; gz += sz; mz += dz; if (gz < 0 || gz > 31) break; }
			if (s > s_end) break;
		}
	}
	return best < 0.0f ? -1.0f : fminf(best + 1e-3f, coarse_far);
}

// per-pixel start state: init_rays (:1478-1543) + advance_pos_nerf (:333-362). Returns alive.
template <bool SIMPLE = false>
__device__ __forceinline__ bool start_ray(const FrameArgs& a, const uint8_t* __restrict__ bits, uint32_t x, uint32_t y, V3& o, V3& d, float& t_out, float& t_far_out) {
	t_out = MAX_DEPTH_F; t_far_out = -1.0f;
	if (!pixel_ray(a, x, y, o, d)) return false; // invalid ray (f-theta lens outside its image circle)
	uint32_t idx = x + (uint32_t)a.w * y;
	Box aabb = ldbox(a.render_aabb);
	const float* R = SIMPLE ? nullptr : HN_CROP_R(a);
	float t = fmaxf(box_ray_intersect(aabb, to_crop_local(R, o), to_crop_local(R, d)).x, 0.0f) + 1e-6f;
	t_out = MAX_DEPTH_F;
	if (!box_contains(aabb, to_crop_local(R, o + t * d))) return false;
	V3 idir = mk(1.0f, 1.0f, 1.0f) / d;
	t = advance_n_steps(t, SIMPLE ? 0.0f : a.cone_angle, hn_ld_random_val(a.spp, idx * 786433u));
	t_far_out = occ_t_far_tight(a, bits, o, d);
	t = advance_to_occupied<SIMPLE>(t, a.cone_angle, o, d, idir, bits, 0, a.max_mip, aabb, 0, t_far_out, R);
	if (t >= MAX_DEPTH_F) return false;
	t_out = t;
	return true;
}

__device__ __forceinline__ bool tile_owned(const FrameArgs& a, uint32_t x, uint32_t y) {
	if (a.world <= 1) return true;
	uint32_t tiles_x = (a.w + a.tile_w - 1) / a.tile_w;
	uint32_t tile = (y / a.tile_h) * tiles_x + (x / a.tile_w);
	return (int)(tile % (uint32_t)a.world) == a.rank;
}

// ---------------------------------------------------------------- kernels
// One thread per pixel. Live rays are appended to the SoA ray buffers through a warp-aggregated atomic (one atomicAdd
// per warp); the ray id is the slot. Order is non-deterministic, results are keyed by pixel so the image is not.
template <bool SIMPLE>
__global__ void __launch_bounds__(256) raygen_kernel(FrameArgs a, const uint8_t* __restrict__ bits, RaySoA rays, uint32_t* __restrict__ alive_list,
	uint32_t* __restrict__ counters, float* __restrict__ depth_buffer, float4* __restrict__ frame) {
	uint32_t x = 0, y = 0, idx = 0;
	bool alive = false;
	V3 o, d; float t = 0.0f, t_far = -1.0f;
	if (owned_pixel(a, blockIdx.x * blockDim.x + threadIdx.x, x, y)) {
		idx = x + (uint32_t)a.w * y;
		// src/testbed_nerf.cu:1490-1493: a cleared depth (0) becomes "nothing", a mesh depth survives
		// (depth_buffer == nullptr: the mesh pass, which writes every pixel's depth itself, is running beside this kernel)
		if (depth_buffer && depth_buffer[idx] < 0.01f) depth_buffer[idx] = MAX_DEPTH_F;
		alive = start_ray<SIMPLE>(a, bits, x, y, o, d, t, t_far);
		// :1526-1528: envmap -> frame buffer; in hybrid frames only where the mesh pass left nothing
		if (a.envmap && !(a.depth_test && depth_buffer[idx] < MAX_DEPTH_F)) if (d.x != 0.0f || d.y != 0.0f || d.z != 0.0f) frame[idx] = read_envmap(a.envmap, a.env_w, a.env_h, d);
	}
	uint32_t mask = __ballot_sync(0xffffffffu, alive);
	if (mask == 0) return;
	uint32_t lane = threadIdx.x & 31u;
	uint32_t base = 0;
	if (lane == (uint32_t)(__ffs(mask) - 1)) base = atomicAdd(&counters[CNT_ALIVE_A], __popc(mask));
	base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
	if (alive) {
		uint32_t r = base + __popc(mask & ((1u << lane) - 1u));
		rays.o_t[r] = make_float4(o.x, o.y, o.z, t);
		rays.d_mw[r] = make_float4(d.x, d.y, d.z, 0.0f);
		rays.rgba[r] = make_float4(0.f, 0.f, 0.f, 0.f);
		rays.depth[r] = 0.0f;
		rays.pixel[r] = idx;
		rays.n_steps[r] = 0;
		rays.t_far[r] = t_far;
		alive_list[r] = r;
	}
}

// ERenderMode::Distortion (init_rays_with_payload_kernel_nerf, src/testbed_nerf.cu:1513-1521): no tracing - every valid pixel shows, as a
// hue, where the lens model + distortion grid moved it: to_rgb((pos_to_uv(ray(1)) - uv) * 64), alpha 1, depth 1.
__global__ void __launch_bounds__(256) distortion_kernel(FrameArgs a, float4* __restrict__ frame, float* __restrict__ depth_buffer) {
	uint32_t x = 0, y = 0;
	if (!owned_pixel(a, blockIdx.x * blockDim.x + threadIdx.x, x, y)) return;
	const uint32_t idx = x + (uint32_t)a.w * y;
	if (depth_buffer[idx] < 0.01f) depth_buffer[idx] = MAX_DEPTH_F;
	V3 o, d, pos; M3 rot; float u, v;
	if (!pixel_ray_raw(a, x, y, o, d, u, v, rot, pos)) return;
	const float2 after = pos_to_uv(a, o + 1.0f * d, rot.c[0], rot.c[1], rot.c[2], pos);
	const V3 c = dir_to_rgb((after.x - u) * 64.0f, (after.y - v) * 64.0f);
	frame[idx] = make_float4(c.x, c.y, c.z, 1.0f);
	depth_buffer[idx] = 1.0f;
}
// ERenderMode::Slice (:1500-1510, 2431-2459): the point where the pixel's ray meets the plane at camera depth plane_z becomes ONE network
// sample (list slot p of this batch: chunk p holds it in lane 0), the depth buffer gets plane_z. The caller passes aperture_size = 0.
__global__ void __launch_bounds__(256) slice_raygen_kernel(FrameArgs a, float plane_z, uint32_t first, uint32_t n, float4* __restrict__ samples, uint32_t* __restrict__ sample_counts,
	float* __restrict__ dirs, uint32_t* __restrict__ pixels, float* __restrict__ depth_buffer) {
	const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= n) return;
	uint32_t x = 0, y = 0, count = 0, idx = 0xffffffffu;
	V3 dir = mk(0.f, 0.f, 1.f);
	if (owned_pixel(a, first + p, x, y)) {
		idx = x + (uint32_t)a.w * y;
		if (depth_buffer[idx] < 0.01f) depth_buffer[idx] = MAX_DEPTH_F;
		V3 o, d, pos; M3 rot; float u, v;
		if (pixel_ray_raw(a, x, y, o, d, u, v, rot, pos)) {
			const float len = length3(d);
			dir = (1.0f / len) * d;
			const float t = plane_z * len;
			const V3 q = o + dir * t;
			samples[(size_t)p * CHUNK] = make_float4(q.x, q.y, q.z, HN_STEPSIZE);
			depth_buffer[idx] = plane_z;
			count = 1;
		} else idx = 0xffffffffu;
	}
	sample_counts[p] = count;
	pixels[p] = idx;
	dirs[3 * p] = dir.x; dirs[3 * p + 1] = dir.y; dirs[3 * p + 2] = dir.z;
}

// a-16: light rays that carry radiance enter the volume like camera rays (AABB entry, advance to the first occupied voxel;
// no start jitter: oracle nerf_transmittance). Rays that never meet a set cell keep T = 1 and are not listed.
__global__ void __launch_bounds__(256) secondary_start_kernel(FrameArgs a, const uint8_t* __restrict__ bits, RaySoA rays, uint32_t n_rays,
	uint32_t* __restrict__ alive_list, uint32_t* __restrict__ counters) {
	const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
	bool alive = false;
	float t = 0.0f, t_far = -1.0f;
	if (r < n_rays) {
		const float4 L = rays.rgba[r];
		if (L.x != 0.0f || L.y != 0.0f || L.z != 0.0f) {
			const float4 ot = rays.o_t[r], dm = rays.d_mw[r];
			const V3 o = mk(ot.x, ot.y, ot.z), d = mk(dm.x, dm.y, dm.z);
			const Box aabb = ldbox(a.render_aabb);
			const float* R = HN_CROP_R(a);
			t = fmaxf(box_ray_intersect(aabb, to_crop_local(R, o), to_crop_local(R, d)).x, 0.0f) + 1e-6f;
			if (box_contains(aabb, to_crop_local(R, o + t * d))) {
				const V3 idir = mk(1.0f, 1.0f, 1.0f) / d;
				t_far = occ_t_far_tight(a, bits, o, d);
				t = advance_to_occupied(t, a.cone_angle, o, d, idir, bits, 0, a.max_mip, aabb, a.dda_iter_cap, t_far, R);
				alive = t < MAX_DEPTH_F;
			}
		}
	}
	const uint32_t mask = __ballot_sync(0xffffffffu, alive);
	if (!mask) return;
	const uint32_t lane = threadIdx.x & 31u;
	const int leader = __ffs(mask) - 1;
	uint32_t base = 0;
	if ((int)lane == leader) base = atomicAdd(&counters[CNT_ALIVE_A], __popc(mask));
	base = __shfl_sync(0xffffffffu, base, leader);
	if (alive) {
		rays.o_t[r].w = t;
		rays.t_far[r] = t_far;
		alive_list[base + __popc(mask & ((1u << lane) - 1u))] = r;
	}
}

// a-16 irradiance probes: ray set-up of init_rays_from_center{,_outward,_multiple_center}_with_payload_kernel_nerf
// (src/testbed_nerf.cu:1559-1773) - directions on the (cos theta, phi) lattice of cylindrical_to_dir_nerf, origin = centre of the NeRF
// box / a given point looking in / Halton<2,3,5>-jittered centres - followed by the start of NerfTracer::trace_mesh (:2146-2262):
// t = 0, no start jitter, rays alive when the 200-iteration DDA finds an occupied cell. Ray slot r keeps its own colour; the texel
// it belongs to is recomputed by probe_reduce_kernel.
__device__ __forceinline__ bool probe_ray(int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, V3 center, V3 origin_in, uint32_t r, V3& o, V3& d) {
	if (mode == 2) {
		const uint32_t wt = n_theta * n_origin;
		const uint32_t tm = r % wt, pm = r / wt;
		if (pm >= n_phi * n_origin) return false;
		const uint32_t theta_mul = tm / n_origin, theta_rem = tm % n_origin, phi_mul = pm / n_origin, phi_rem = pm % n_origin;
		const uint32_t hidx = theta_rem * n_origin + phi_rem;
		const float ox = (float)((double)halton(2, hidx) - 0.5), oy = (float)((double)halton(3, hidx) - 0.5), oz = (float)((double)halton(5, hidx) - 0.5);
		o = center + mk(ox, oy, oz);
		d = normalize3(cylindrical_to_dir((float)theta_mul / (float)n_theta, (float)phi_mul / (float)n_phi));
		return true;
	}
	if (r >= n_theta * n_phi) return false;
	const uint32_t theta_mul = r % n_theta, phi_mul = r / n_theta;
	const V3 local = cylindrical_to_dir((float)theta_mul / (float)n_theta, (float)phi_mul / (float)n_phi);
	if (mode == 0) { o = center; d = normalize3(local); }
	else { o = origin_in; const V3 w = normalize3(local_frame_apply(normalize3(origin_in), local)); d = mk(-w.x, -w.y, -w.z); }
	return true;
}
__global__ void __launch_bounds__(256) probe_raygen_kernel(FrameArgs a, const uint8_t* __restrict__ bits, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin,
	float ox, float oy, float oz, uint32_t n_rays, RaySoA rays, uint32_t* __restrict__ alive_list, uint32_t* __restrict__ counters) {
	const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
	bool alive = false;
	float t = 0.0f, t_far = -1.0f;
	V3 o = mk(0.f, 0.f, 0.f), d = mk(0.f, 0.f, 1.f);
	if (r < n_rays) {
		const Box train = ldbox(a.train_aabb);
		const V3 center = 0.5f * (train.hi + train.lo);
		if (probe_ray(mode, n_theta, n_phi, n_origin, center, mk(ox, oy, oz), r, o, d)) {
			const Box aabb = ldbox(a.render_aabb);
			const V3 idir = mk(1.0f, 1.0f, 1.0f) / d;
			t_far = occ_t_far_tight(a, bits, o, d);
			t = advance_to_occupied(0.0f, a.cone_angle, o, d, idir, bits, 0, a.max_mip, aabb, 200, t_far, nullptr);
			alive = t < MAX_DEPTH_F;
		}
		rays.o_t[r] = make_float4(o.x, o.y, o.z, alive ? t : MAX_DEPTH_F);
		rays.d_mw[r] = make_float4(d.x, d.y, d.z, 0.0f);
		rays.rgba[r] = make_float4(0.f, 0.f, 0.f, 0.f);
		rays.depth[r] = 0.0f;
		rays.pixel[r] = r;
		rays.n_steps[r] = 0;
		rays.t_far[r] = t_far;
	}
	const uint32_t mask = __ballot_sync(0xffffffffu, alive);
	if (!mask) return;
	const uint32_t lane = threadIdx.x & 31u;
	const int leader = __ffs(mask) - 1;
	uint32_t base = 0;
	if ((int)lane == leader) base = atomicAdd(&counters[CNT_ALIVE_A], __popc(mask));
	base = __shfl_sync(0xffffffffu, base, leader);
	if (alive) alive_list[base + __popc(mask & ((1u << lane) - 1u))] = r;
}
// probe texture [n_phi][n_theta]: mean over the rays of a texel (1 ray, or n_origin^2 jittered centres in mode 2), fixed order
__global__ void __launch_bounds__(256) probe_reduce_kernel(const float4* __restrict__ ray_rgba, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, float4* __restrict__ tex) {
	const uint32_t texel = blockIdx.x * blockDim.x + threadIdx.x;
	if (texel >= n_theta * n_phi) return;
	if (mode != 2) { tex[texel] = ray_rgba[texel]; return; }
	const uint32_t i = texel % n_theta, j = texel / n_theta, wt = n_theta * n_origin;
	float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
	for (uint32_t pr = 0; pr < n_origin; ++pr)
		for (uint32_t tr = 0; tr < n_origin; ++tr) {
			const float4 v = ray_rgba[(i * n_origin + tr) + (size_t)wt * (j * n_origin + pr)];
			acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
		}
	const float inv = 1.0f / (float)(n_origin * n_origin);
	tex[texel] = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
}

// One ray per lane, at most `limit` samples, in the flat warp-converged loop of march_kernel (each trip a lane either takes a
// sample or skips to the next voxel; the skip code runs only when some lane needs it): per-lane arithmetic and its order are
// those of advance_to_occupied + calc_dt, so counts and positions equal the sequential walk. WRITE = false only counts.
template <bool WRITE>
__device__ __forceinline__ uint32_t flat_march(const FrameArgs& a, const uint8_t* __restrict__ bits, bool valid, V3 o, V3 d, float t, uint32_t limit,
	float4* __restrict__ dst, uint32_t* __restrict__ dst_ray, uint32_t ray) {
	const Box aabb = ldbox(a.render_aabb), train = ldbox(a.train_aabb);
	const V3 tdiag = train.hi - train.lo;
	const V3 idir = mk(1.0f, 1.0f, 1.0f) / d;
	const float t_far = valid ? occ_t_far_tight(a, bits, o, d) : -1.f;
	bool active = valid && limit > 0;
	uint32_t count = 0;
	int it = 1;
	while (__any_sync(0xffffffffu, active)) {
		bool skip = false;
		uint32_t mip = 0;
		V3 pos = o;
		if (active) {
			pos = o + t * d;
			if (t >= MAX_DEPTH_F || t > t_far || !box_contains(aabb, to_crop_local(HN_CROP_R(a), pos))) {
				active = false;
			} else {
				mip = min(max(mip_from_pos(pos), 0u), a.max_mip);
				if (!bits || occupied_at(pos, bits, mip)) {
					const float dt = calc_dt(t, a.cone_angle);
					if (WRITE) {
						const V3 wp = (pos - train.lo) / tdiag;
						dst[count] = make_float4(wp.x, wp.y, wp.z, dt);
						if (dst_ray) dst_ray[count] = ray;
					}
					++count;
					t += dt;
					it = 1;
					if (count == limit) active = false;
				} else {
					skip = true;
				}
			}
		}
		if (__any_sync(0xffffffffu, skip)) {
			if (skip) {
				while (mip < a.max_mip && !occupied_at(pos, bits, mip + 1)) ++mip;
				t = advance_to_next_voxel(t, a.cone_angle, pos, d, idir, mip);
				++it;
				if (a.dda_iter_cap > 0 && it >= a.dda_iter_cap) active = false;
			}
		}
	}
	return count;
}

// a-21 generate_training_samples_nerf (src/testbed_nerf.cu:737-890): per training ray, walk the occupancy DDA (at most
// NERF_STEPS = 1024 samples), reserve that many slots (one atomic per WARP: warp prefix sum of the counts) and store warped
// position + dt. Rays whose samples do not fit below max_samples get none (:848-852). The reference walks every ray twice
// (count, then write); a batch has only a few thousand rays, so the kernel is bound by the latency of the longest serial
// walk: here the ONE walk writes into a per-ray staging row (n_rays x 1024 float4 in HBM, L2-resident for a batch) and the
// warp then copies each of its 32 rays' rows to the packed slots with coalesced accesses. Same samples, same packing rule.
__global__ void __launch_bounds__(64) train_march_kernel(FrameArgs a, const uint8_t* __restrict__ bits, uint32_t n_rays, const float* __restrict__ ray_o,
	const float* __restrict__ ray_d, const float* __restrict__ ray_t, uint32_t max_samples, uint32_t* __restrict__ counter,
	uint32_t* __restrict__ numsteps_out, uint32_t* __restrict__ base_out, float4* __restrict__ coords, uint32_t* __restrict__ sample_ray, float4* __restrict__ staging) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t lane = threadIdx.x & 31u;
	const bool valid = i < n_rays;
	V3 o = mk(0.f, 0.f, 0.f), d = mk(0.f, 0.f, 1.f);
	float t0 = MAX_DEPTH_F;
	if (valid) { o = ld3(ray_o + 3 * i); d = ld3(ray_d + 3 * i); t0 = ray_t[i]; }
	uint32_t n = flat_march<true>(a, bits, valid, o, d, t0, 1024u, staging + (size_t)(valid ? i : 0u) * 1024u, nullptr, 0u);
	// warp-aggregated slot reservation
	uint32_t incl = n;
	#pragma unroll
	for (int s = 1; s < 32; s <<= 1) { const uint32_t up = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= (uint32_t)s) incl += up; }
	const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
	uint32_t warp_base = 0;
	if (lane == 0 && total) warp_base = atomicAdd(counter, total);
	warp_base = __shfl_sync(0xffffffffu, warp_base, 0);
	const uint32_t base = warp_base + incl - n;
	if (base + n > max_samples) n = 0;
	if (valid) { numsteps_out[i] = n; base_out[i] = base; }
	__syncwarp();
	// staging rows -> packed slots, one ray of the warp at a time, all 32 lanes copying (coalesced both ways)
	const uint32_t ray0 = i - lane;
	for (uint32_t r = 0; r < 32u; ++r) {
		const uint32_t nr = __shfl_sync(0xffffffffu, n, r), br = __shfl_sync(0xffffffffu, base, r);
		const float4* src = staging + (size_t)(ray0 + r) * 1024u;
		for (uint32_t k = lane; k < nr; k += 32u) { coords[br + k] = src[k]; sample_ray[br + k] = ray0 + r; }
	}
}

__global__ void __launch_bounds__(256) raygen_debug_kernel(FrameArgs a, const uint8_t* __restrict__ bits, float* __restrict__ o_out, float* __restrict__ d_out,
	float* __restrict__ t_out, uint8_t* __restrict__ alive_out) {
	uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= (uint32_t)a.w * (uint32_t)a.h) return;
	V3 o, d; float t, t_far;
	bool alive = start_ray(a, bits, idx % (uint32_t)a.w, idx / (uint32_t)a.w, o, d, t, t_far);
	o_out[3 * idx] = o.x; o_out[3 * idx + 1] = o.y; o_out[3 * idx + 2] = o.z;
	d_out[3 * idx] = d.x; d_out[3 * idx + 1] = d.y; d_out[3 * idx + 2] = d.z;
	t_out[idx] = t;
	alive_out[idx] = alive ? 1 : 0;
}

// generate_next_nerf_network_inputs (src/testbed_nerf.cu:430-477): up to CHUNK samples per live ray per pass.
// Sample slot of list position p, step j = p*CHUNK + j (ray-major, so a warp of the network kernel owns one ray chunk).
// The reference nests a data-dependent empty-space loop inside the per-sample loop; with one ray per lane that serialises
// a warp on its slowest lane at every sample. Here both loops are flattened into ONE warp-converged loop: each trip a lane
// either emits a sample (cell occupied) or skips to the next voxel, and the skip code is entered only when some lane of
// the warp needs it (warp-uniform branch). Per-lane arithmetic and its order are unchanged -> bit-identical samples.
template <bool SIMPLE>
__global__ void __launch_bounds__(128) march_kernel(FrameArgs a, const uint8_t* __restrict__ bits, RaySoA rays, const uint32_t* __restrict__ alive_list,
	const uint32_t* __restrict__ n_alive_ptr, float4* __restrict__ samples, uint32_t* __restrict__ sample_counts, uint32_t* __restrict__ counters,
	const float* __restrict__ depth_buffer) {
	const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t n_alive = *n_alive_ptr;
	const bool valid = p < n_alive;
	bool active = valid;
	uint32_t count = 0, r = 0;
	int it = 1; // trip counter of the reference's inner loop (nerf_device.cuh:510: i < 200 in the _geometry variant)
	V3 o = {0.f, 0.f, 0.f}, d = {0.f, 0.f, 1.f}, idir = {0.f, 0.f, 1.f};
	float t = 0.f, t_far = -1.f, clip = MAX_DEPTH_F;
	const Box aabb = ldbox(a.render_aabb), train = ldbox(a.train_aabb);
	const V3 tdiag = train.hi - train.lo;
	const V3 cam_fwd = ld3(a.cam1 + 6), cam_pos = ld3(a.cam1 + 9);
	const bool clipping = SIMPLE ? false : (a.depth_clip && a.depth_test);
	const float cone = SIMPLE ? 0.0f : a.cone_angle;
	const uint32_t max_mip = SIMPLE ? 0u : a.max_mip;
	const float* __restrict__ R = SIMPLE ? nullptr : HN_CROP_R(a);
	if (valid) {
		r = alive_list[p];
		const float4 ot = rays.o_t[r], dm = rays.d_mw[r];
		o = mk(ot.x, ot.y, ot.z); d = mk(dm.x, dm.y, dm.z); t = ot.w;
		idir = mk(1.0f, 1.0f, 1.0f) / d;
		t_far = rays.t_far[r]; // occ_t_far_tight, computed once when the ray was started
		if (clipping) clip = depth_buffer[rays.pixel[r]];
	}
	float4* dst = samples + (size_t)p * CHUNK;
	while (__any_sync(0xffffffffu, active)) {
		bool skip = false;
		uint32_t mip = 0;
		V3 pos = o;
		if (active) {
			pos = o + t * d;
			if (t >= MAX_DEPTH_F || t > t_far || !box_contains(aabb, to_crop_local(R, pos))) {
				t = MAX_DEPTH_F; active = false;
			} else {
				mip = max_mip ? min(max(mip_from_pos(pos), 0u), max_mip) : 0u; // one cascade: the clamp makes it 0 whatever frexpf says
				if ((!SIMPLE && !bits) || occupied_at(pos, bits, mip)) {
					const float dt = calc_dt(t, cone);
					bool emit = true;
					if (clipping) {
						const V3 up = train.lo + ((pos - train.lo) / tdiag) * tdiag;
						if (dot3(cam_fwd, up - cam_pos) > clip) { emit = false; active = false; }
					}
					if (emit) {
						dst[count] = make_float4(pos.x, pos.y, pos.z, dt);
						++count;
						t += dt;
						it = 1;
						if (count == CHUNK) active = false;
					}
				} else {
					skip = true;
				}
			}
		}
		if (__any_sync(0xffffffffu, skip)) {
			if (skip) {
				while (mip < max_mip && !occupied_at(pos, bits, mip + 1)) ++mip;
				t = advance_to_next_voxel(t, cone, pos, d, idir, mip);
				++it;
				if (a.dda_iter_cap > 0 && it >= a.dda_iter_cap) { t = MAX_DEPTH_F; active = false; }
			}
		}
	}
	if (valid) {
		rays.o_t[r].w = t;
		sample_counts[p] = count;
	}
	// frame statistics: one 64-bit atomic per warp
	uint32_t total = count;
	#pragma unroll
	for (int s = 16; s > 0; s >>= 1) total += __shfl_xor_sync(0xffffffffu, total, s);
	if ((threadIdx.x & 31u) == 0 && total) atomicAdd((unsigned long long*)&counters[CNT_SAMPLES_LO], (unsigned long long)total);
}

__global__ void __launch_bounds__(128) march_all_kernel(FrameArgs a, const uint8_t* __restrict__ bits, uint32_t n, const float* __restrict__ o_in,
	const float* __restrict__ d_in, const float* __restrict__ t0, const uint8_t* __restrict__ alive, uint32_t* __restrict__ counts, float* __restrict__ samples, uint32_t max_per_ray) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t c = 0;
	if (alive[i]) {
		V3 o = ld3(o_in + 3 * i), d = ld3(d_in + 3 * i);
		V3 idir = mk(1.0f, 1.0f, 1.0f) / d;
		Box aabb = ldbox(a.render_aabb);
		float t = t0[i];
		const float t_far = occ_t_far_tight(a, bits, o, d);
		while (true) {
			t = advance_to_occupied(t, a.cone_angle, o, d, idir, bits, 0, a.max_mip, aabb, a.dda_iter_cap, t_far, HN_CROP_R(a));
			if (t >= MAX_DEPTH_F) break;
			float dt = calc_dt(t, a.cone_angle);
			if (samples && c < max_per_ray) {
				V3 pos = o + t * d;
				float* s = samples + ((size_t)i * max_per_ray + c) * 4;
				s[0] = pos.x; s[1] = pos.y; s[2] = pos.z; s[3] = dt;
			}
			++c;
			t += dt;
		}
	}
	counts[i] = c;
}

// ---------------------------------------------------------------- occupancy bitfield (src/testbed_nerf.cu:284-331)
__global__ void grid_to_bitfield_kernel(uint32_t n_elements, uint32_t n_nonzero, const float* __restrict__ grid, uint8_t* __restrict__ bits, float thresh) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	if (i >= n_nonzero) { bits[i] = 0; return; }
	const float4* g = (const float4*)(grid + (size_t)i * 8);
	float4 v0 = g[0], v1 = g[1];
	uint8_t b = 0;
	b |= v0.x > thresh ? 1 : 0; b |= v0.y > thresh ? 2 : 0; b |= v0.z > thresh ? 4 : 0; b |= v0.w > thresh ? 8 : 0;
	b |= v1.x > thresh ? 16 : 0; b |= v1.y > thresh ? 32 : 0; b |= v1.z > thresh ? 64 : 0; b |= v1.w > thresh ? 128 : 0;
	bits[i] = b;
}
__global__ void bitfield_max_pool_kernel(uint32_t n_elements, const uint8_t* __restrict__ prev, uint8_t* __restrict__ next) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	uint2 v = ((const uint2*)prev)[i];
	uint8_t b = 0;
	#pragma unroll
	for (int j = 0; j < 4; ++j) {
		b |= ((v.x >> (8 * j)) & 0xffu) ? (uint8_t)(1 << j) : 0;
		b |= ((v.y >> (8 * j)) & 0xffu) ? (uint8_t)(16 << j) : 0;
	}
	uint32_t x = morton3D_invert(i >> 0) + GRIDSIZE / 8;
	uint32_t y = morton3D_invert(i >> 1) + GRIDSIZE / 8;
	uint32_t z = morton3D_invert(i >> 2) + GRIDSIZE / 8;
	// each thread owns a distinct byte of the next level (Morton is a bijection), zeroed by the previous kernel
	next[morton3D(x, y, z)] |= b;
}
// mean of max(v, 0) over cascade 0, accumulated in double for an order-independent result (the threshold is
// min(0.01, mean): src/testbed_nerf.cu:297)
__global__ void grid_mean_kernel(const float* __restrict__ grid, double* __restrict__ sum) {
	double acc = 0.0;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < GRID_CELLS; i += gridDim.x * blockDim.x) acc += (double)fmaxf(grid[i], 0.0f);
	#pragma unroll
	for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
	if ((threadIdx.x & 31u) == 0) atomicAdd(sum, acc);
}

// per-mip bounds of the set cells in cell coordinates: out[mip] = {min x,y,z, max x,y,z} (min > max when the mip is empty)
__global__ void __launch_bounds__(1024) occ_bounds_kernel(const uint8_t* __restrict__ bits, int* __restrict__ out) {
	// one block = 1024 consecutive bytes of ONE mip (a mip has 262 144 bytes): warp reduction, block reduction in shared memory, six
	// atomics per block (was six atomics per non-empty byte, all onto the same 48 words: 1.05 ms per grid refresh)
	__shared__ int s_lo[32][3], s_hi[32][3];
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; // byte index over all mips
	const uint32_t per_mip = GRID_CELLS / 8;
	const uint32_t mip = (blockIdx.x * blockDim.x) / per_mip;
	int lo[3] = {1 << 30, 1 << 30, 1 << 30}, hi[3] = {-1, -1, -1};
	const uint32_t byte = i < BITFIELD_BYTES ? bits[i] : 0u;
	if (byte) {
		const uint32_t base = (i - mip * per_mip) * 8;
		for (uint32_t b = 0; b < 8; ++b) {
			if (!(byte & (1u << b))) continue;
			const uint32_t idx = base + b;
			const int c[3] = {(int)morton3D_invert(idx), (int)morton3D_invert(idx >> 1), (int)morton3D_invert(idx >> 2)};
			for (int k = 0; k < 3; ++k) { lo[k] = min(lo[k], c[k]); hi[k] = max(hi[k], c[k]); }
		}
	}
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	for (int k = 0; k < 3; ++k) { lo[k] = __reduce_min_sync(0xffffffffu, lo[k]); hi[k] = __reduce_max_sync(0xffffffffu, hi[k]); }
	if (lane == 0) for (int k = 0; k < 3; ++k) { s_lo[warp][k] = lo[k]; s_hi[warp][k] = hi[k]; }
	__syncthreads();
	if (warp == 0) {
		for (int k = 0; k < 3; ++k) {
			const int l = __reduce_min_sync(0xffffffffu, s_lo[lane][k]), h = __reduce_max_sync(0xffffffffu, s_hi[lane][k]);
			if (lane == 0 && h >= 0) { atomicMin(out + mip * 6 + k, l); atomicMax(out + mip * 6 + 3 + k, h); }
		}
	}
}

// coarse occupancy flags for occ_t_far_tight: [mip][32^3 bits], bit (cx + 32 cy + 1024 cz) = some fine cell is set in coarse cell
// (cx, cy, cz) or in one of its 26 neighbours. A coarse cell is the aligned 64-bit word morton3D(cx, cy, cz) of the mip's bitfield.
__global__ void __launch_bounds__(256) occ_coarse_kernel(const uint8_t* __restrict__ bits, uint32_t* __restrict__ coarse) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; // over CASCADES x 32^3
	const uint32_t mip = i >> 15, c = i & 32767u;
	const int cx = (int)(c & 31u), cy = (int)((c >> 5) & 31u), cz = (int)(c >> 10);
	const unsigned long long* __restrict__ words = reinterpret_cast<const unsigned long long*>(bits + (size_t)mip * (GRID_CELLS / 8));
	bool any = false;
	for (int dz = -1; dz <= 1; ++dz) for (int dy = -1; dy <= 1; ++dy) for (int dx = -1; dx <= 1; ++dx) {
		const int x = cx + dx, y = cy + dy, z = cz + dz;
		if (x < 0 || x > 31 || y < 0 || y > 31 || z < 0 || z > 31) continue;
		any = any || words[morton3D((uint32_t)x, (uint32_t)y, (uint32_t)z)] != 0ull;
	}
	const uint32_t word = __ballot_sync(0xffffffffu, any);
	if ((threadIdx.x & 31u) == 0) coarse[i >> 5] = word;
}

// ---------------------------------------------------------------- launchers
// the SIMPLE specialisation of the walk applies: linear stepping space, one cascade, axis-aligned crop box, a bitfield
static bool simple_walk(const FrameArgs& a, const uint8_t* bits) { return a.simple_walk && bits && a.cone_angle <= 1e-5f && a.max_mip == 0 && !a.crop_rotated; }
void launch_occ_coarse(const uint8_t* bits, uint32_t* coarse /* CASCADES x 1024 words, device */, cudaStream_t s) {
	occ_coarse_kernel<<<CASCADES * 32768 / 256, 256, 0, s>>>(bits, coarse);
}
void launch_occ_bounds(const uint8_t* bits, int* out /* CASCADES x 6, device */, cudaStream_t s) {
	int init[CASCADES * 6];
	for (uint32_t m = 0; m < CASCADES; ++m) for (int k = 0; k < 3; ++k) { init[m * 6 + k] = 1 << 30; init[m * 6 + 3 + k] = -1; }
	cudaMemcpyAsync(out, init, sizeof(init), cudaMemcpyHostToDevice, s);
	cudaStreamSynchronize(s); // `init` lives on this stack frame
	occ_bounds_kernel<<<(BITFIELD_BYTES + 1023) / 1024, 1024, 0, s>>>(bits, out);
}
void launch_grid_mean(const float* grid, float* mean_out, cudaStream_t s) {
	// mean_out is reinterpreted as a double accumulator (8 bytes, zeroed here)
	cudaMemsetAsync(mean_out, 0, sizeof(double), s);
	grid_mean_kernel<<<148 * 4, 256, 0, s>>>(grid, (double*)mean_out);
}
void launch_grid_to_bitfield(const float* grid, uint32_t n_cascades, float thresh, uint8_t* bits, cudaStream_t s) {
	const uint32_t per_mip = GRID_CELLS / 8;
	uint32_t n = per_mip * CASCADES;
	grid_to_bitfield_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, per_mip * n_cascades, grid, bits, thresh);
	for (uint32_t level = 1; level < CASCADES; ++level) {
		bitfield_max_pool_kernel<<<(per_mip / 8 + 255) / 256, 256, 0, s>>>(per_mip / 8, bits + (size_t)per_mip * (level - 1), bits + (size_t)per_mip * level);
	}
}
void launch_raygen(const FrameArgs& a, const uint8_t* bits, RaySoA rays, uint32_t* alive_list, uint32_t* counters, float* depth_buffer, float4* frame, cudaStream_t s) {
	uint32_t n = owned_pixel_threads(a);
	if (!n) return;
	if (simple_walk(a, bits)) raygen_kernel<true><<<(n + 255) / 256, 256, 0, s>>>(a, bits, rays, alive_list, counters, depth_buffer, frame);
	else raygen_kernel<false><<<(n + 255) / 256, 256, 0, s>>>(a, bits, rays, alive_list, counters, depth_buffer, frame);
}
void launch_distortion(const FrameArgs& a, float4* frame, float* depth_buffer, cudaStream_t s) {
	const uint32_t n = owned_pixel_threads(a);
	if (n) distortion_kernel<<<(n + 255) / 256, 256, 0, s>>>(a, frame, depth_buffer);
}
void launch_slice_raygen(const FrameArgs& a, float plane_z, uint32_t first, uint32_t n, float4* samples, uint32_t* sample_counts, float* dirs, uint32_t* pixels,
	float* depth_buffer, cudaStream_t s) {
	if (n) slice_raygen_kernel<<<(n + 255) / 256, 256, 0, s>>>(a, plane_z, first, n, samples, sample_counts, dirs, pixels, depth_buffer);
}
void launch_secondary_start(const FrameArgs& a, const uint8_t* bits, RaySoA rays, uint32_t n_rays, uint32_t* alive_list, uint32_t* counters, cudaStream_t s) {
	if (!n_rays) return;
	secondary_start_kernel<<<(n_rays + 255) / 256, 256, 0, s>>>(a, bits, rays, n_rays, alive_list, counters);
}
void launch_probe_raygen(const FrameArgs& a, const uint8_t* bits, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, const float origin[3], RaySoA rays,
	uint32_t* alive_list, uint32_t* counters, cudaStream_t s) {
	const uint32_t n_rays = n_theta * n_phi * (mode == 2 ? n_origin * n_origin : 1u);
	if (!n_rays) return;
	probe_raygen_kernel<<<(n_rays + 255) / 256, 256, 0, s>>>(a, bits, mode, n_theta, n_phi, n_origin, origin[0], origin[1], origin[2], n_rays, rays, alive_list, counters);
}
void launch_probe_reduce(const float4* ray_rgba, int mode, uint32_t n_theta, uint32_t n_phi, uint32_t n_origin, float4* tex, cudaStream_t s) {
	const uint32_t n = n_theta * n_phi;
	if (!n) return;
	probe_reduce_kernel<<<(n + 255) / 256, 256, 0, s>>>(ray_rgba, mode, n_theta, n_phi, n_origin, tex);
}
void launch_train_march(const FrameArgs& a, const uint8_t* bits, uint32_t n_rays, const float* o, const float* d, const float* t, uint32_t max_samples,
	uint32_t* counter, uint32_t* numsteps, uint32_t* base, float4* coords, uint32_t* sample_ray, float4* staging, cudaStream_t s) {
	if (!n_rays) return;
	train_march_kernel<<<(n_rays + 63) / 64, 64, 0, s>>>(a, bits, n_rays, o, d, t, max_samples, counter, numsteps, base, coords, sample_ray, staging);
}
void launch_raygen_debug(const FrameArgs& a, const uint8_t* bits, float* o, float* d, float* t, uint8_t* alive, cudaStream_t s) {
	uint32_t n = (uint32_t)a.w * (uint32_t)a.h;
	raygen_debug_kernel<<<(n + 255) / 256, 256, 0, s>>>(a, bits, o, d, t, alive);
}
void launch_march(const FrameArgs& a, const uint8_t* bits, RaySoA rays, const uint32_t* alive_list, const uint32_t* n_alive_ptr,
	float4* samples, uint32_t* sample_counts, uint32_t* counters, const float* depth_buffer, uint32_t max_rays, cudaStream_t s) {
	if (!max_rays) return;
	if (simple_walk(a, bits) && !(a.depth_clip && a.depth_test)) march_kernel<true><<<(max_rays + 127) / 128, 128, 0, s>>>(a, bits, rays, alive_list, n_alive_ptr, samples, sample_counts, counters, depth_buffer);
	else march_kernel<false><<<(max_rays + 127) / 128, 128, 0, s>>>(a, bits, rays, alive_list, n_alive_ptr, samples, sample_counts, counters, depth_buffer);
}
void launch_march_all(const FrameArgs& a, const uint8_t* bits, uint32_t n, const float* o, const float* d, const float* t0, const uint8_t* alive,
	uint32_t* counts, float* samples, uint32_t max_per_ray, cudaStream_t s) {
	if (!n) return;
	march_all_kernel<<<(n + 127) / 128, 128, 0, s>>>(a, bits, n, o, d, t0, alive, counts, samples, max_per_ray);
}

} // namespace hn
