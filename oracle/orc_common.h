// ORACLE — TEST INFRASTRUCTURE ONLY.
// CPU restatement of the reference's hybrid NeRF+mesh render path. Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may build, load or call anything under oracle/.
// The product (surface-irradiance-estimation-from-neural-radiance-fields_b200/) never links or calls it.
//
// Parity status: the reference ships no golden vectors for this path. The integer / IEEE
// scalar pieces of this restatement (Sobol, ray generation, AABB slab test, DDA stepping, triangle test,
// Morton bitfield) are pinned against the reference's own NGP_HOST_DEVICE functions compiled for the host
// (oracle/_ref, see oracle/ref_shim.cu + tests/golden/*.npz). The device-only kernels (hash grid, SH, MLP,
// compositing, shading) have no host-runnable reference: for those this oracle is "parity unpinned".
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cfloat>
#include <algorithm>
#include <vector>

namespace orc {

typedef _Float16 half_t;

static inline half_t f2h(float f) { return (half_t)f; }
static inline float h2f(half_t h) { return (float)h; }
static inline half_t bits2h(uint16_t b) { half_t h; std::memcpy(&h, &b, 2); return h; }
static inline uint16_t h2bits(half_t h) { uint16_t b; std::memcpy(&b, &h, 2); return b; }
// fp16 fused multiply-add with a single rounding (the product and sum are exact in double).
static inline half_t hfma(half_t a, half_t b, half_t c) { return (half_t)((double)a * (double)b + (double)c); }

struct vec2 { float x, y; };
struct vec3 {
	float x, y, z;
	float& operator[](int i) { return (&x)[i]; }
	float operator[](int i) const { return (&x)[i]; }
};
struct vec4 { float x, y, z, w; };

static inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
static inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
static inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
static inline vec3 operator+(vec3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
static inline vec3 operator-(vec3 a, float s) { return {a.x - s, a.y - s, a.z - s}; }
static inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
// tcnn vec.h:425 REDUCTION_OP(dot): result = 0; result += a[i]*b[i] in index order.
static inline float dot(vec3 a, vec3 b) { float r = 0.0f; r += a.x * b.x; r += a.y * b.y; r += a.z * b.z; return r; }
static inline float length2(vec3 a) { return dot(a, a); }
static inline float length(vec3 a) { return std::sqrt(length2(a)); }
// tcnn vec.h:466 normalize: v / len (three IEEE divides), (1,0,0) for zero length.
static inline vec3 normalize(vec3 v) { float l = length(v); if (l <= 0.0f) return {1.0f, 0.0f, 0.0f}; return v / l; }
static inline vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static inline float mixf(float a, float b, float c) { return a * (1.0f - c) + b * c; }
static inline vec3 mix(vec3 a, vec3 b, float c) { return a * (1.0f - c) + b * c; }
static inline vec3 mix(vec3 a, vec3 b, vec3 c) { return {mixf(a.x, b.x, c.x), mixf(a.y, b.y, c.y), mixf(a.z, b.z, c.z)}; }
static inline float clampf(float v, float lo, float hi) { return std::max(lo, std::min(v, hi)); }

// Column-major 3x3: m[col][row], matching tcnn::mat3 (vec.h:590: result[j] += m[i][j] * v[i]).
struct mat3 {
	vec3 c[3];
};
static inline vec3 mul(const mat3& m, vec3 v) {
	vec3 r = {0.0f, 0.0f, 0.0f};
	for (int i = 0; i < 3; ++i) {
		r.x += m.c[i].x * v[i];
		r.y += m.c[i].y * v[i];
		r.z += m.c[i].z * v[i];
	}
	return r;
}

// mat4x3 = 4 columns of vec3: {right, up/down, forward, position}
struct cam_t {
	vec3 c[4];
};
static inline cam_t cam_from(const float* f) {
	cam_t m;
	for (int i = 0; i < 4; ++i) m.c[i] = {f[3 * i], f[3 * i + 1], f[3 * i + 2]};
	return m;
}

struct Ray {
	vec3 o, d;
	vec3 at(float t) const { return o + t * d; }
};

struct BoundingBox {
	vec3 min, max;
	vec3 diag() const { return max - min; }
	vec3 center() const { return 0.5f * (max + min); } // bounding_box.cuh:99-101
	// bounding_box.cuh:313
	bool contains(vec3 p) const {
		return p.x >= min.x && p.x <= max.x && p.y >= min.y && p.y <= max.y && p.z >= min.z && p.z <= max.z;
	}
	// bounding_box.cuh:172-219. Explicit swaps, early-outs, FLT_MAX pair on miss.
	vec2 ray_intersect(vec3 pos, vec3 dir) const {
		float tmin = (min.x - pos.x) / dir.x;
		float tmax = (max.x - pos.x) / dir.x;
		if (tmin > tmax) std::swap(tmin, tmax);
		float tymin = (min.y - pos.y) / dir.y;
		float tymax = (max.y - pos.y) / dir.y;
		if (tymin > tymax) std::swap(tymin, tymax);
		if (tmin > tymax || tymin > tmax) return {FLT_MAX, FLT_MAX};
		if (tymin > tmin) tmin = tymin;
		if (tymax < tmax) tmax = tymax;
		float tzmin = (min.z - pos.z) / dir.z;
		float tzmax = (max.z - pos.z) / dir.z;
		if (tzmin > tzmax) std::swap(tzmin, tzmax);
		if (tmin > tzmax || tzmin > tmax) return {FLT_MAX, FLT_MAX};
		if (tzmin > tmin) tmin = tzmin;
		if (tzmax < tmax) tmax = tzmax;
		return {tmin, tmax};
	}
	vec3 relative_pos(vec3 p) const { return (p - min) / diag(); }
	void enlarge(vec3 p) {
		min = {std::min(min.x, p.x), std::min(min.y, p.y), std::min(min.z, p.z)};
		max = {std::max(max.x, p.x), std::max(max.y, p.y), std::max(max.z, p.z)};
	}
};
static inline BoundingBox bb_from(const float* f) { return {{f[0], f[1], f[2]}, {f[3], f[4], f[5]}}; }

static constexpr float MAX_DEPTH = 16384.0f;          // common_device.cuh:32
static constexpr float MAX_DIST = 100.0f;             // triangle_bvh.cuh (mint start)
static constexpr uint32_t NERF_GRIDSIZE = 128;        // nerf_device.cuh:23
static constexpr uint32_t NERF_GRID_N_CELLS = 128 * 128 * 128;
static constexpr uint32_t NERF_STEPS = 1024;          // nerf_device.cuh:27
static constexpr uint32_t NERF_CASCADES = 8;          // nerf_device.cuh:28
static constexpr float SQRT3 = 1.73205080757f;
static constexpr float STEPSIZE = SQRT3 / NERF_STEPS; // nerf_device.cuh:31
static constexpr float MIN_CONE_STEPSIZE = STEPSIZE;
static constexpr float MAX_CONE_STEPSIZE = STEPSIZE * (1 << (NERF_CASCADES - 1)) * NERF_STEPS / NERF_GRIDSIZE;
static constexpr float NERF_MIN_OPTICAL_THICKNESS = 0.01f;

} // namespace orc
