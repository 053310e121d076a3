// (e) Ray-mesh and shadow-ray traversal + mesh shading, fused into ONE kernel per frame for sm_100a:
// pixel -> primary ray -> 4-wide BVH traversal -> sun shadow ray through the same BVH -> Disney-style BRDF -> frame/depth.
// No intermediate ray buffers (the reference runs 6 kernels and 6 D2D copies over three RaysMeshSoa sets,
// src/testbed_geometry_training.cu:2202-2320). IEEE-exact translation unit (-fmad=false): hit (mesh, triangle, t) is
// bit-identical to the oracle's brute-force search. Nodes are 128-byte lines (4 child boxes SoA), triangles 3 x float4 in
// leaf order carrying their ORIGINAL index; the traversal stack lives in shared memory, one bank-conflict-free column
// per thread, the nearest-first child order comes from a register sorting network.
// NOTE on provenance: evaluate_shading() below follows evaluate_shading_geometry (src/testbed_geometry_training.cu:75-144) and
// tri_distance_sq() follows Triangle::distance_sq (include/neural-graphics-primitives/triangle.cuh:66-147) statement for
// statement, identifiers included: those ~60 lines of formula have to match the reference operation for operation for the frame
// (<= 1e-2) and distance (bit-exact vs brute force) parity tests, so they are transliterations, not re-derivations. Traversal, node
// and triangle layout, the fused pass and everything around them are this project's own.
#include "hn_device_math.cuh"

namespace hn {

constexpr int MESH_BLOCK = 128;
constexpr int STACK_DEPTH = 64; // the surface-area-heuristic tree is less balanced than the median one: deeper stacks (32 KB of shared memory per block)

struct Hit { float t; int mesh; int tri; uint32_t slot; }; // tri = ORIGINAL triangle index, slot = position in the leaf-ordered array

// triangle.cuh:44-58, same operation order
__device__ __forceinline__ float tri_intersect(V3 a, V3 b, V3 c, V3 ro, V3 rd) {
	V3 v1v0 = b - a, v2v0 = c - a, rov0 = ro - a;
	V3 n = cross3(v1v0, v2v0);
	V3 q = cross3(rov0, rd);
	float d = 1.0f / dot3(rd, n);
	float u = d * -dot3(q, v2v0);
	float v = d * dot3(q, v1v0);
	float t = d * -dot3(n, rov0);
	if (u < 0.0f || u > 1.0f || v < 0.0f || (u + v) > 1.0f || t < 0.0f) t = FLT_MAX;
	return t;
}

__device__ __forceinline__ void cswap(float& da, int& ia, float& db, int& ib) {
	if (da < db) { float t = da; da = db; db = t; int k = ia; ia = ib; ib = k; } // descending: nearest ends up last = popped first
}

// Nearest hit inside one mesh. Acceptance as triangle_bvh.cu:150-193 (mint starts at the caller's best, strict <) plus the
// shared tie rule: equal t inside the same mesh -> lower original triangle index.
// ANY = true: occlusion query - return at the first accepted triangle (same acceptance test, so "is there a hit" has the same answer as
// the nearest-hit search; which triangle is reported is arbitrary). Returns whether a triangle was accepted during this call.
template <bool ANY = false>
__device__ bool traverse_mesh(const SceneDev& sc, int mesh, V3 o, V3 d, Hit& best, int* stack /* shared, stride MESH_BLOCK */) {
	bool found = false;
	const MeshDev m = sc.meshes[mesh];
	const V3 inv = mk(1.0f, 1.0f, 1.0f) / d;
	int sp = 0;
	stack[0] = (int)m.node_root; sp = 1;
	while (sp > 0) {
		const int code = stack[(--sp) * MESH_BLOCK];
		if (code < 0) {
			const uint32_t v = (uint32_t)(-1 - code);
			const uint32_t first = m.tri_first + (v >> 4), cnt = v & 15u;
			for (uint32_t k = 0; k < cnt; ++k) {
				const float4 A = __ldg(sc.tris + (size_t)(first + k) * 3), B = __ldg(sc.tris + (size_t)(first + k) * 3 + 1), C = __ldg(sc.tris + (size_t)(first + k) * 3 + 2);
				const float t = tri_intersect(mk(A.x, A.y, A.z), mk(B.x, B.y, B.z), mk(C.x, C.y, C.z), o, d);
				const int orig = __float_as_int(A.w);
				if (t < best.t || (t == best.t && best.mesh == mesh && orig < best.tri)) {
					best.t = t; best.mesh = mesh; best.tri = orig; best.slot = first + k; found = true;
					if (ANY) return true;
				}
			}
			continue;
		}
		const float4* np = reinterpret_cast<const float4*>(sc.nodes + code);
		const float4 lox = __ldg(np), loy = __ldg(np + 1), loz = __ldg(np + 2), hix = __ldg(np + 3), hiy = __ldg(np + 4), hiz = __ldg(np + 5);
		const int4 child = __ldg(reinterpret_cast<const int4*>(np + 6)), count = __ldg(reinterpret_cast<const int4*>(np + 7));
		float dist[4]; int idx[4];
		const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, ly[4] = {loy.x, loy.y, loy.z, loy.w}, lz[4] = {loz.x, loz.y, loz.z, loz.w};
		const float hx[4] = {hix.x, hix.y, hix.z, hix.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
		const int ch[4] = {child.x, child.y, child.z, child.w}, cn[4] = {count.x, count.y, count.z, count.w};
		#pragma unroll
		for (int i = 0; i < 4; ++i) {
			// NaN-safe slab test on the (conservatively padded) child box; keep it when entry <= current best t
			float t0x = (lx[i] - o.x) * inv.x, t1x = (hx[i] - o.x) * inv.x;
			float t0y = (ly[i] - o.y) * inv.y, t1y = (hy[i] - o.y) * inv.y;
			float t0z = (lz[i] - o.z) * inv.z, t1z = (hz[i] - o.z) * inv.z;
			float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), 0.0f));
			float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), best.t));
			bool hit = cn[i] != 0 && tn <= tf;
			dist[i] = hit ? tn : -1.0f;
			idx[i] = ch[i];
		}
		// sorting network N=4 (5 compare-exchanges), descending by distance; misses (-1) sink to the end and are skipped
		cswap(dist[0], idx[0], dist[2], idx[2]);
		cswap(dist[1], idx[1], dist[3], idx[3]);
		cswap(dist[0], idx[0], dist[1], idx[1]);
		cswap(dist[2], idx[2], dist[3], idx[3]);
		cswap(dist[1], idx[1], dist[2], idx[2]);
		#pragma unroll
		for (int i = 0; i < 4; ++i) {
			if (dist[i] >= 0.0f) {
				if (sp < STACK_DEPTH) { stack[sp * MESH_BLOCK] = idx[i]; ++sp; }
				else if (sc.stack_overflow) atomicAdd(sc.stack_overflow, 1u); // never silent: surfaced as hn_frame_stats.n_bvh_stack_overflows (the reference prints "WARNING TOO BIG", triangle_bvh.cuh:139-144)
			}
		}
	}
	return found;
}
// a-16 shadow test of a light ray: is ANY triangle of any mesh hit (the light pass only asks whether the ray is blocked)
__device__ bool scene_occluded(const SceneDev& sc, V3 o, V3 d, int* stack) {
	Hit h; h.t = MAX_DIST_F; h.mesh = -1; h.tri = -1; h.slot = 0;
	for (uint32_t i = 0; i < sc.n_meshes; ++i) if (traverse_mesh<true>(sc, (int)i, o, d, h, stack)) return true;
	return false;
}

// geometry_bvh.cu:646-676 with the top level of :166-200. compat = pick the ONE mesh with the nearest box entry
// (reference rule, BoundingBox::ray_intersect on the exact mesh boxes); default = nearest hit over all meshes.
// `touched` = whether the reference kernel would have moved the ray at all.
__device__ void scene_nearest(const SceneDev& sc, int compat, V3 o, V3 d, Hit& h, bool& touched, int* stack) {
	h.t = MAX_DIST_F; h.mesh = -1; h.tri = -1; h.slot = 0;
	touched = false;
	if (compat) {
		float mint = MAX_DIST_F; int mesh = -1;
		for (uint32_t i = 0; i < sc.n_meshes; ++i) {
			Box b = ldbox(sc.meshes[i].bb);
			float2 tt = box_ray_intersect(b, o, d);
			if (tt.x < mint && tt.x > -FLT_MAX) { mint = tt.x; mesh = (int)i; }
		}
		if (mesh < 0) return;
		touched = true;
		traverse_mesh(sc, mesh, o, d, h, stack);
		h.mesh = mesh; // the reference reports the chosen mesh even when no triangle of it is hit
		return;
	}
	touched = true;
	for (uint32_t i = 0; i < sc.n_meshes; ++i) traverse_mesh(sc, (int)i, o, d, h, stack);
}
// scene_nearest for a ray of which only "is anything hit, and is that point inside the shading box" is asked (the sun shadow ray,
// write_shadow_ray_result_geometry): the same mesh choice, the traversal stops at the first accepted triangle. Every triangle lies inside
// the mesh box, which lies inside the shading box, so the answer is the nearest-hit search's.
__device__ void scene_any(const SceneDev& sc, int compat, V3 o, V3 d, Hit& h, bool& touched, int* stack) {
	h.t = MAX_DIST_F; h.mesh = -1; h.tri = -1; h.slot = 0;
	touched = false;
	if (compat) {
		float mint = MAX_DIST_F; int mesh = -1;
		for (uint32_t i = 0; i < sc.n_meshes; ++i) {
			Box b = ldbox(sc.meshes[i].bb);
			float2 tt = box_ray_intersect(b, o, d);
			if (tt.x < mint && tt.x > -FLT_MAX) { mint = tt.x; mesh = (int)i; }
		}
		if (mesh < 0) return;
		touched = true;
		traverse_mesh<true>(sc, mesh, o, d, h, stack);
		return;
	}
	touched = true;
	for (uint32_t i = 0; i < sc.n_meshes; ++i) if (traverse_mesh<true>(sc, (int)i, o, d, h, stack)) return;
}

__global__ void __launch_bounds__(MESH_BLOCK) trace_kernel(SceneDev sc, int compat, uint32_t n, const float* __restrict__ o_in, const float* __restrict__ d_in,
	int32_t* __restrict__ mesh_out, int32_t* __restrict__ tri_out, float* __restrict__ t_out) {
	__shared__ int stack_s[STACK_DEPTH * MESH_BLOCK];
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	Hit h; bool touched;
	scene_nearest(sc, compat, ld3(o_in + 3 * i), ld3(d_in + 3 * i), h, touched, stack_s + threadIdx.x);
	mesh_out[i] = h.mesh; tri_out[i] = h.tri; t_out[i] = h.t;
}

// ---------------------------------------------------------------- distance queries (triangle.cuh:39-41,66-147, triangle_bvh.cu:195-303)
__device__ __forceinline__ float sign1(float x) { return copysignf(1.0f, x); }
__device__ __forceinline__ float clamp01(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }
// Triangle::distance_sq (iq's formulation), same operation order as the reference
__device__ float tri_distance_sq(V3 a, V3 b, V3 c, V3 pos) {
	const V3 v21 = b - a, p1 = pos - a, v32 = c - b, p2 = pos - b, v13 = a - c, p3 = pos - c;
	const V3 nor = cross3(v21, v13);
	if (sign1(dot3(cross3(v21, nor), p1)) + sign1(dot3(cross3(v32, nor), p2)) + sign1(dot3(cross3(v13, nor), p3)) < 2.0f) {
		const V3 e1 = v21 * clamp01(dot3(v21, p1) / dot3(v21, v21)) - p1;
		const V3 e2 = v32 * clamp01(dot3(v32, p2) / dot3(v32, v32)) - p2;
		const V3 e3 = v13 * clamp01(dot3(v13, p3) / dot3(v13, v13)) - p3;
		return fminf(fminf(dot3(e1, e1), dot3(e2, e2)), dot3(e3, e3));
	}
	return dot3(nor, p1) * dot3(nor, p1) / dot3(nor, nor);
}
__device__ __forceinline__ V3 tri_normal(V3 a, V3 b, V3 c) { const V3 n = cross3(b - a, c - a); return n / sqrtf(dot3(n, n)); }
__device__ __forceinline__ V3 closest_on_segment(V3 a, V3 b, V3 c) {
	float t = dot3(c - a, b - a) / dot3(b - a, b - a);
	t = fmaxf(fminf(t, 1.0f), 0.0f);
	return a + t * (b - a);
}
__device__ V3 tri_closest_point(V3 a, V3 b, V3 c, V3 point) {
	const V3 n = tri_normal(a, b, c);
	point = point - dot3(n, point - a) * n;
	{ // point_in_triangle
		const V3 la = a - point, lb = b - point, lc = c - point;
		const V3 u = cross3(lb, lc), v = cross3(lc, la), w = cross3(la, lb);
		if (dot3(u, v) >= 0.0f && dot3(u, w) >= 0.0f) return point;
	}
	const V3 c1 = closest_on_segment(a, b, point), c2 = closest_on_segment(b, c, point), c3 = closest_on_segment(c, a, point);
	const V3 d1 = point - c1, d2 = point - c2, d3 = point - c3;
	const float m1 = dot3(d1, d1), m2 = dot3(d2, d2), m3 = dot3(d3, d3);
	const float mn = fminf(fminf(m1, m2), m3);
	if (mn == m1) return c1;
	if (mn == m2) return c2;
	return c3;
}
__device__ __forceinline__ float box_distance_sq(float lx, float ly, float lz, float hx, float hy, float hz, V3 p) {
	const float dx = fmaxf(fmaxf(lx - p.x, p.x - hx), 0.0f), dy = fmaxf(fmaxf(ly - p.y, p.y - hy), 0.0f), dz = fmaxf(fmaxf(lz - p.z, p.z - hz), 0.0f);
	return dx * dx + dy * dy + dz * dz;
}
// One stack walk over one mesh. NEAREST: closest triangle (`<=` as triangle_bvh.cu:211, so the last of equal candidates in
// walk order wins; the distance itself does not depend on that). !NEAREST: sum of the normals of every triangle within
// sqrt(1e-6) of `p` (avg_normal_around_point :246-283).
template <bool NEAREST>
__device__ void distance_walk(const SceneDev& sc, int mesh, V3 p, float& best_d2, uint32_t& best_slot, V3& nsum, float& nweight, int* stack) {
	const MeshDev m = sc.meshes[mesh];
	int sp = 0;
	stack[0] = (int)m.node_root; sp = 1;
	while (sp > 0) {
		const int code = stack[(--sp) * MESH_BLOCK];
		if (code < 0) {
			const uint32_t v = (uint32_t)(-1 - code);
			const uint32_t first = m.tri_first + (v >> 4), cnt = v & 15u;
			for (uint32_t k = 0; k < cnt; ++k) {
				const float4 A = __ldg(sc.tris + (size_t)(first + k) * 3), B = __ldg(sc.tris + (size_t)(first + k) * 3 + 1), C = __ldg(sc.tris + (size_t)(first + k) * 3 + 2);
				const V3 a = mk(A.x, A.y, A.z), b = mk(B.x, B.y, B.z), c = mk(C.x, C.y, C.z);
				const float d2 = tri_distance_sq(a, b, c, p);
				if (NEAREST) { if (d2 <= best_d2) { best_d2 = d2; best_slot = first + k; } }
				else if (d2 < 1e-6f) { nsum = nsum + tri_normal(a, b, c); nweight += 1.0f; }
			}
			continue;
		}
		const float4* np = reinterpret_cast<const float4*>(sc.nodes + code);
		const float4 lox = __ldg(np), loy = __ldg(np + 1), loz = __ldg(np + 2), hix = __ldg(np + 3), hiy = __ldg(np + 4), hiz = __ldg(np + 5);
		const int4 child = __ldg(reinterpret_cast<const int4*>(np + 6)), count = __ldg(reinterpret_cast<const int4*>(np + 7));
		const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, ly[4] = {loy.x, loy.y, loy.z, loy.w}, lz[4] = {loz.x, loz.y, loz.z, loz.w};
		const float hx[4] = {hix.x, hix.y, hix.z, hix.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
		const int ch[4] = {child.x, child.y, child.z, child.w}, cn[4] = {count.x, count.y, count.z, count.w};
		float dist[4]; int idx[4];
		#pragma unroll
		for (int i = 0; i < 4; ++i) {
			const float d2 = box_distance_sq(lx[i], ly[i], lz[i], hx[i], hy[i], hz[i], p); // padded boxes: a valid lower bound
			const bool keep = cn[i] != 0 && (NEAREST ? d2 <= best_d2 : d2 < 1e-6f);
			dist[i] = keep ? d2 : -1.0f;
			idx[i] = ch[i];
		}
		cswap(dist[0], idx[0], dist[2], idx[2]);
		cswap(dist[1], idx[1], dist[3], idx[3]);
		cswap(dist[0], idx[0], dist[1], idx[1]);
		cswap(dist[2], idx[2], dist[3], idx[3]);
		cswap(dist[1], idx[1], dist[2], idx[2]);
		#pragma unroll
		for (int i = 0; i < 4; ++i) {
			if (dist[i] >= 0.0f) {
				if (sp < STACK_DEPTH) { stack[sp * MESH_BLOCK] = idx[i]; ++sp; }
				else if (sc.stack_overflow) atomicAdd(sc.stack_overflow, 1u); // never silent: surfaced as hn_frame_stats.n_bvh_stack_overflows (the reference prints "WARNING TOO BIG", triangle_bvh.cuh:139-144)
			}
		}
	}
}
// mode 0: unsigned distance to the nearest triangle of the scene; mode 1: signed, watertight rule (:289-297: sign of
// dot(average normal around the closest point, p - closest point)); mode 2: signed, raystab rule (:299-317).
// max_distance bounds the search like MAX_DIST.
__global__ void __launch_bounds__(MESH_BLOCK) distance_kernel(SceneDev sc, uint32_t n, const float* __restrict__ pos, int mode, float max_distance,
	float* __restrict__ out, int32_t* __restrict__ tri_out) {
	__shared__ int stack_s[STACK_DEPTH * MESH_BLOCK];
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const V3 p = ld3(pos + 3 * i);
	float best = max_distance * max_distance;
	uint32_t slot = 0xFFFFFFFFu;
	int best_mesh = -1;
	V3 ns = mk(0.0f, 0.0f, 0.0f); float nw = 0.0f;
	for (uint32_t m = 0; m < sc.n_meshes; ++m) {
		uint32_t s = 0xFFFFFFFFu; float b = best;
		distance_walk<true>(sc, (int)m, p, b, s, ns, nw, stack_s + threadIdx.x);
		if (s != 0xFFFFFFFFu) { best = b; slot = s; best_mesh = (int)m; }
	}
	if (slot == 0xFFFFFFFFu) { out[i] = max_distance; if (tri_out) tri_out[i] = -1; return; }
	float d = sqrtf(best);
	const float4 A = __ldg(sc.tris + (size_t)slot * 3), B = __ldg(sc.tris + (size_t)slot * 3 + 1), C = __ldg(sc.tris + (size_t)slot * 3 + 2);
	if (tri_out) tri_out[i] = __float_as_int(A.w);
	if (mode == 1) {
		const V3 cp = tri_closest_point(mk(A.x, A.y, A.z), mk(B.x, B.y, B.z), mk(C.x, C.y, C.z), p);
		float dummy = 0.0f; uint32_t ds = 0;
		distance_walk<false>(sc, best_mesh, cp, dummy, ds, ns, nw, stack_s + threadIdx.x);
		const V3 avg = ns / nw;
		d = copysignf(d, dot3(avg, p - cp));
	} else if (mode == 2) { // signed_distance_raystab (:299-317): 32 Fibonacci stab rays, any escape -> outside
		// default_rng_t rng; rng.advance(i * 2); offset = random_val_2d(rng)   (triangle_bvh.cu:573-574, pcg32 default state)
		uint64_t state = 0x853c49e6748fea9bULL; const uint64_t inc = 0xda3e39cb94b95bdbULL;
		{
			uint64_t delta = (uint64_t)i * 2u, cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
			while (delta > 0) {
				if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
				cur_plus = (cur_mult + 1) * cur_plus;
				cur_mult *= cur_mult;
				delta /= 2;
			}
			state = acc_mult * state + acc_plus;
		}
		float off[2];
		#pragma unroll
		for (int k = 0; k < 2; ++k) {
			const uint64_t old = state;
			state = old * 0x5851f42d4c957f2dULL + inc;
			const uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
			off[k] = __uint_as_float((((xs >> rot) | (xs << ((~rot + 1u) & 31))) >> 9) | 0x3f800000u) - 1.0f;
		}
		bool inside = true;
		for (uint32_t k = 0; k < 32u && inside; ++k) {
			// fibonacci_dir<32> (random_val.cuh:45-54,84-101), epsilon = 1.33
			float px = ((float)k + 1.33f) / (32 - 1 + 2 * 1.33f) + off[0], py = (float)k / 1.6180339887498948482045868343656f + off[1];
			px -= floorf(px); py -= floorf(py);
			const float cos_theta = -2.0f * px + 1.0f, phi = 2.0f * HN_PI_F * (py - 0.5f);
			const float sin_theta = sqrtf(fmaxf(1.0f - cos_theta * cos_theta, 0.0f));
			float sp, cp;
			sincosf(phi, &sp, &cp);
			Hit h; bool touched;
			scene_nearest(sc, 0, p, mk(sin_theta * cp, sin_theta * sp, cos_theta), h, touched, stack_s + threadIdx.x);
			if (h.tri < 0) inside = false;
		}
		if (inside) d = -d;
	}
	out[i] = d;
}
void launch_distance(const SceneDev& sc, uint32_t n, const float* pos, int mode, float max_distance, float* out, int32_t* tri_out, cudaStream_t s) {
	if (!n) return;
	distance_kernel<<<(n + MESH_BLOCK - 1) / MESH_BLOCK, MESH_BLOCK, 0, s>>>(sc, n, pos, mode, max_distance, out, tri_out);
}

// ---------------------------------------------------------------- shading (testbed_geometry_training.cu:50-144)
__device__ __forceinline__ float sqr(float x) { return x * x; }
__device__ __forceinline__ float saturate1(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }
__device__ __forceinline__ float schlick(float u) { float m = saturate1(1.0f - u); return sqr(sqr(m)) * m; }
#define HN_PI 3.14159265358979323846f
__device__ __forceinline__ float G1f(float NdotH, float a) {
	if (a >= 1.0f) return 1.0f / HN_PI;
	float a2 = sqr(a);
	float t = 1.0f + (a2 - 1.0f) * NdotH * NdotH;
	return (a2 - 1.0f) / (HN_PI * logf(a2) * t);
}
__device__ __forceinline__ float G2f(float NdotH, float a) {
	float a2 = sqr(a);
	float t = 1.0f + (a2 - 1.0f) * NdotH * NdotH;
	return a2 / (HN_PI * t * t);
}
__device__ __forceinline__ float smithG(float NdotV, float alphaG) {
	float a = alphaG * alphaG, b = NdotV * NdotV;
	return 1.0f / (NdotV + sqrtf(a + b - a * b));
}
__device__ V3 evaluate_shading(V3 base_color, V3 ambient_color, V3 light_color, float metallic, float subsurface, float specular, float roughness,
	float specular_tint, float sheen, float sheen_tint, float clearcoat, float clearcoat_gloss, V3 L, V3 V, V3 N) {
	float NdotL = dot3(N, L), NdotV = dot3(N, V);
	V3 H = normalize3(L + V);
	float NdotH = dot3(N, H), LdotH = dot3(L, H);
	float FL = schlick(NdotL), FV = schlick(NdotV);
	V3 amb = ambient_color * mixf(0.2f, FV, metallic);
	amb = amb * base_color;
	if (NdotL < 0.f || NdotV < 0.f) return amb;
	float luminance = dot3(base_color, mk(0.3f, 0.6f, 0.1f));
	V3 one = mk(1.0f, 1.0f, 1.0f);
	V3 Ctint = base_color * (1.f / (luminance + 0.00001f));
	V3 Cspec0 = mix3(mix3(one, Ctint, specular_tint) * specular * 0.08f, base_color, metallic);
	V3 Csheen = mix3(one, Ctint, sheen_tint);
	float Fd90 = 0.5f + 2.0f * LdotH * LdotH * roughness;
	float Fd = mixf(1, Fd90, FL) * mixf(1.f, Fd90, FV);
	float Fss90 = LdotH * LdotH * roughness;
	float Fss = mixf(1.0f, Fss90, FL) * mixf(1.0f, Fss90, FV);
	float ss = 1.25f * (Fss * (1.f / (NdotL + NdotV) - 0.5f) + 0.5f);
	float a = fmaxf(0.001f, sqr(roughness));
	float Ds = G2f(NdotH, a);
	float FH = schlick(LdotH);
	V3 Fs = mix3(Cspec0, one, FH);
	float Gs = smithG(NdotL, a) * smithG(NdotV, a);
	V3 Fsheen = FH * sheen * Csheen;
	float Dr = G1f(NdotH, mixf(0.1f, 0.001f, clearcoat_gloss));
	float Fr = mixf(0.04f, 1.0f, FH);
	float Gr = smithG(NdotL, 0.25f) * smithG(NdotV, 0.25f);
	float CCs = 0.25f * clearcoat * Gr * Fr * Dr;
	V3 brdf = ((1.0f / HN_PI) * mixf(Fd, ss, subsurface) * base_color + Fsheen) * (1.0f - metallic) + Gs * Fs * Ds + mk(CCs, CCs, CCs);
	return (brdf * light_color) * NdotL + amb;
}

// Triangle::normal (triangle.cuh:40-42)
__device__ __forceinline__ V3 tri_normal(const SceneDev& sc, uint32_t slot) {
	const float4 A = __ldg(sc.tris + (size_t)slot * 3), B = __ldg(sc.tris + (size_t)slot * 3 + 1), C = __ldg(sc.tris + (size_t)slot * 3 + 2);
	V3 a = mk(A.x, A.y, A.z), b = mk(B.x, B.y, B.z), c = mk(C.x, C.y, C.z);
	return normalize3(cross3(b - a, c - a));
}

__device__ __forceinline__ bool mesh_tile_owned(const FrameArgs& a, uint32_t x, uint32_t y) {
	if (a.world <= 1) return true;
	uint32_t tiles_x = (a.w + a.tile_w - 1) / a.tile_w;
	uint32_t tile = (y / a.tile_h) * tiles_x + (x / a.tile_w);
	return (int)(tile % (uint32_t)a.world) == a.rank;
}

// render_geometry_mesh (src/testbed_geometry_training.cu:2202-2320), Shade mode, one thread per pixel.
// Per-pixel body; returns true when the pixel is a shaded triangle hit that gathers environment light (a-16).
__device__ bool mesh_pixel(const FrameArgs& a, const SceneDev& sc, float4* __restrict__ frame, float* __restrict__ depth_buffer, int* stack,
	uint32_t x, uint32_t y, V3& env_org, V3& env_nrm, V3& env_albedo) {
	const uint32_t idx = x + (uint32_t)a.w * y;
	const Box aabb = ldbox(a.mesh_aabb);
	// init_rays_with_payload_kernel_mesh_geometry (:488-579): pixel_to_ray with camera_matrix0 as is (no rolling shutter)
	const float u = ((float)x + a.pixel_offset[0]) / (float)a.w, v = ((float)y + a.pixel_offset[1]) / (float)a.h;
	const V3 c0 = ld3(a.cam0), c1 = ld3(a.cam0 + 3), c2 = ld3(a.cam0 + 6), cam_pos = ld3(a.cam0 + 9);
	V3 org, dir;
	const bool valid = camera_ray(a, false, u, v, c0, c1, c2, cam_pos, org, dir); // no Lens on this ray init; foveation, hidden-area mask, parallax shift and depth of field apply
	depth_buffer[idx] = MAX_DEPTH_F;
	if (!valid) return false; // hidden by the mask: the pixel stays cleared (the reference pushes a zero ray through its trace / shade kernels: undefined colour)
	dir = normalize3(dir);
	const float t0 = fmaxf(box_ray_intersect(aabb, org, dir).x, 0.0f);
	V3 pos = org + (t0 + 1e-6f) * dir;
	// primary hit (:2140-2155: every pixel is traced; the direction rides in the normal buffer)
	Hit h; bool touched;
	scene_nearest(sc, a.compat_nearest_aabb_only, pos, dir, h, touched, stack);
	V3 nrm = dir;
	if (touched) {
		pos = pos + h.t * dir;
		if (h.tri >= 0) nrm = tri_normal(sc, h.slot);
	}
	// sun shadow ray (:222-278)
	const V3 sun = normalize3(ld3(a.sun));
	float shadow;
	{
		const float sgn = -copysignf(1.0f, dot3(dir, nrm)); // faceforward(n, i, nref) = n * -sign(dot(i, nref))
		V3 view_pos = pos + normalize3(nrm * sgn) * 1e-3f;
		const float ts = fmaxf(box_ray_intersect(aabb, view_pos, sun).x + 1e-6f, 0.0f);
		view_pos = view_pos + ts * sun;
		const V3 sdir = box_contains(aabb, view_pos) ? sun : dir; // an origin outside the box keeps payload.dir (:239-244)
		Hit sh; bool st;
		scene_any(sc, a.compat_nearest_aabb_only, view_pos, sdir, sh, st, stack);
		const V3 end = st ? view_pos + sh.t * sdir : view_pos;
		shadow = box_contains(aabb, end) ? 0.0f : 1.0f; // write_shadow_ray_result_geometry
	}
	// shade_kernel_mesh_geometry (:280-355)
	if (!box_contains(aabb, pos)) return false;
	V3 normal = normalize3(nrm);
	bool floor_hit = false;
	if (pos.y < a.floor_y + 0.001f && dir.y < 0.f) { normal = mk(0.0f, 1.0f, 0.0f); floor_hit = true; }
	const V3 up = normalize3(ld3(a.up));
	const float skyam = -dot3(normal, up) * 0.5f + 0.5f;
	const V3 suncol = mk(255.f / 255.0f, 225.f / 255.0f, 195.f / 255.0f) * 4.f * shadow;
	const V3 skycol = mk(195.f / 255.0f, 215.f / 255.0f, 255.f / 255.0f) * 4.f * skyam;
	const float check_size = 8.f / (aabb.hi.x - aabb.lo.x);
	const float check = ((int(floorf(check_size * (pos.x - aabb.lo.x))) ^ int(floorf(check_size * (pos.z - aabb.lo.z)))) & 1) ? 0.8f : 0.2f;
	const V3 floorcol = mk(check * check * check, check * check, check);
	const float* B = a.brdf;
	const V3 basecolor = mk(B[7], B[8], B[9]), ambientcolor = mk(B[10], B[11], B[12]);
	const V3 color = evaluate_shading(floor_hit ? floorcol : basecolor * basecolor, ambientcolor * skycol, suncol,
		floor_hit ? 0.f : B[0], floor_hit ? 0.f : B[1], floor_hit ? 1.f : B[2], floor_hit ? 0.5f : B[3], 0.f, floor_hit ? 0.f : B[4], 0.f,
		floor_hit ? 0.f : B[5], B[6], sun, -normalize3(dir), normal);
	frame[idx] = make_float4(color.x, color.y, color.z, 1.0f);
	depth_buffer[idx] = dot3(c2, pos - cam_pos);
	if (!(touched && h.tri >= 0)) return false;
	env_nrm = normal * -copysignf(1.0f, dot3(dir, normal));
	env_org = pos + env_nrm * 1e-3f;
	env_albedo = floor_hit ? floorcol : basecolor * basecolor;
	return true;
}

// Thread -> pixel of the mesh pass: a warp covers an 8 x 4 pixel patch and a block of 128 threads a 16 x 8 one (camera rays of a patch stay
// together deeper into the tree than those of a 32 x 1 row segment). One rank: patches row-major over the frame. Tile-sharded: the same patches
// inside each of the rank's tiles when the tile size allows it, else the row-major order of owned_pixel. Results are keyed by pixel, so the
// order changes nothing but the speed.
__device__ __forceinline__ void patch_xy(uint32_t k, uint32_t patches_x, uint32_t& x, uint32_t& y) {
	const uint32_t patch = k / MESH_BLOCK, t = k % MESH_BLOCK, warp = t >> 5, lane = t & 31u;
	x = (patch % patches_x) * 16u + (warp & 1u) * 8u + (lane & 7u);
	y = (patch / patches_x) * 8u + (warp >> 1) * 4u + (lane >> 3);
}
__device__ __forceinline__ bool mesh_thread_pixel(const FrameArgs& a, uint32_t i, uint32_t& x, uint32_t& y) {
	const uint32_t w = (uint32_t)a.w, h = (uint32_t)a.h;
	if (a.world <= 1) {
		const uint32_t px = (w + 15u) / 16u, py = (h + 7u) / 8u;
		if (i >= px * py * MESH_BLOCK) return false;
		patch_xy(i, px, x, y);
		return x < w && y < h;
	}
	const uint32_t tw = (uint32_t)a.tile_w, th = (uint32_t)a.tile_h;
	if ((tw % 16u) || (th % 8u)) return owned_pixel(a, i, x, y);
	const uint32_t per_tile = tw * th, tiles_x = (w + tw - 1) / tw, tiles_y = (h + th - 1) / th;
	const uint32_t tile = (i / per_tile) * (uint32_t)a.world + (uint32_t)a.rank;
	if (tile >= tiles_x * tiles_y) return false;
	patch_xy(i % per_tile, tw / 16u, x, y);
	x += (tile % tiles_x) * tw; y += (tile / tiles_x) * th;
	return x < w && y < h;
}
__host__ inline uint32_t mesh_pass_threads(const FrameArgs& a) {
	if (a.world <= 1) return (uint32_t)((a.w + 15) / 16) * (uint32_t)((a.h + 7) / 8) * MESH_BLOCK;
	return owned_pixel_threads(a);
}

__global__ void __launch_bounds__(MESH_BLOCK) mesh_pass_kernel(FrameArgs a, SceneDev sc, float4* __restrict__ frame, float* __restrict__ depth_buffer, EnvHits hits) {
	__shared__ int stack_s[STACK_DEPTH * MESH_BLOCK];
	uint32_t x = 0, y = 0;
	bool lit = false;
	V3 org, nrm, alb;
	if (mesh_thread_pixel(a, blockIdx.x * blockDim.x + threadIdx.x, x, y)) lit = mesh_pixel(a, sc, frame, depth_buffer, stack_s + threadIdx.x, x, y, org, nrm, alb);
	if (!hits.count) return; // warp-uniform
	const uint32_t mask = __ballot_sync(0xffffffffu, lit);
	if (!mask) return;
	const uint32_t lane = threadIdx.x & 31u;
	const int leader = __ffs(mask) - 1;
	uint32_t base = 0;
	if ((int)lane == leader) base = atomicAdd(hits.count, __popc(mask));
	base = __shfl_sync(0xffffffffu, base, leader);
	if (lit) {
		const uint32_t slot = base + __popc(mask & ((1u << lane) - 1u));
		hits.pos_pix[slot] = make_float4(org.x, org.y, org.z, __uint_as_float(x + (uint32_t)a.w * y));
		hits.nrm[slot] = make_float4(nrm.x, nrm.y, nrm.z, 0.0f);
		hits.albedo[slot] = make_float4(alb.x, alb.y, alb.z, 0.0f);
	}
}

// a-16 (OUR definition, oracle render_mesh): light ray k of hit j -> ray slot (j - first) * K + k. The shadow test is an any-hit
// query over all meshes (same answer as the oracle's nearest-hit search); occluded rays carry zero radiance and never enter the volume.
__global__ void __launch_bounds__(MESH_BLOCK) envlight_gen_kernel(FrameArgs a, SceneDev sc, EnvHits hits, uint32_t first, uint32_t n, RaySoA rays) {
	__shared__ int stack_s[STACK_DEPTH * MESH_BLOCK];
	const uint32_t K = (uint32_t)a.env_samples;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n * K) return;
	const uint32_t j = first + i / K, k = i % K;
	const float4 pp = __ldg(hits.pos_pix + j), nn = __ldg(hits.nrm + j);
	const uint32_t pix = __float_as_uint(pp.w);
	const V3 org = mk(pp.x, pp.y, pp.z);
	const V3 nrm = mk(nn.x, nn.y, nn.z);
	V3 w;
	float weight = 1.0f; // estimator weight of this direction: 1 for cosine sampling of a Lambertian surface
	bool cap = false;
	if (a.env_source == 1 && a.env_sampling == 1) { // uniform spherical cap around the direction to the NeRF's bounding sphere
		const Box train = ldbox(a.train_aabb);
		const V3 c = 0.5f * (train.hi + train.lo), to_c = c - org;
		const float dist = length3(to_c), radius = 0.5f * length3(train.hi - train.lo);
		if (dist > radius) {
			const float cos_max = sqrtf(fmaxf(0.0f, 1.0f - (radius / dist) * (radius / dist)));
			w = cap_light_dir(pix, a.spp, K, k, to_c / dist, 1.0f - cos_max, nrm, weight);
			cap = true;
		}
	}
	if (!cap) w = env_light_dir(pix, a.spp, K, k, nrm);
	// zero-weight directions carry no radiance whatever the shadow test says: skip the traversal
	const bool blocked = weight > 0.0f ? scene_occluded(sc, org, w, stack_s + threadIdx.x) : true;
	float4 L = make_float4(0.f, 0.f, 0.f, 0.f);
	if (!blocked && weight > 0.0f) {
		L = a.env_source == 1 ? read_probe(a.probe, a.probe_n_theta, a.probe_n_phi, w) : read_envmap(a.envmap, a.env_w, a.env_h, w);
		L.x *= weight; L.y *= weight; L.z *= weight;
	}
	rays.o_t[i] = make_float4(org.x, org.y, org.z, 0.0f);
	rays.d_mw[i] = make_float4(w.x, w.y, w.z, 0.0f);
	rays.rgba[i] = make_float4(L.x, L.y, L.z, 1.0f); // .w = transmittance T of the volume along the light ray
	rays.pixel[i] = pix;
	rays.n_steps[i] = 0;
}

__global__ void __launch_bounds__(256) envlight_reduce_kernel(FrameArgs a, EnvHits hits, uint32_t first, uint32_t n, RaySoA rays, float4* __restrict__ frame) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t K = (uint32_t)a.env_samples;
	V3 sum = mk(0.f, 0.f, 0.f);
	for (uint32_t k = 0; k < K; ++k) {
		const float4 r = rays.rgba[(size_t)i * K + k];
		sum = sum + mk(r.x, r.y, r.z) * r.w;
	}
	const float4 alb = hits.albedo[first + i];
	const uint32_t pix = __float_as_uint(hits.pos_pix[first + i].w);
	const V3 add = (mk(alb.x, alb.y, alb.z) * sum) * (1.0f / (float)K);
	float4 f = frame[pix];
	f.x += add.x; f.y += add.y; f.z += add.z;
	frame[pix] = f;
}

void launch_mesh_pass(const FrameArgs& a, SceneDev scene, float4* frame, float* depth_buffer, EnvHits hits, cudaStream_t s) {
	uint32_t n = mesh_pass_threads(a);
	if (!n) return;
	mesh_pass_kernel<<<(n + MESH_BLOCK - 1) / MESH_BLOCK, MESH_BLOCK, 0, s>>>(a, scene, frame, depth_buffer, hits);
}
void launch_envlight_gen(const FrameArgs& a, SceneDev scene, EnvHits hits, uint32_t first, uint32_t n, RaySoA rays, cudaStream_t s) {
	if (!n) return;
	const uint32_t threads = n * (uint32_t)a.env_samples;
	envlight_gen_kernel<<<(threads + MESH_BLOCK - 1) / MESH_BLOCK, MESH_BLOCK, 0, s>>>(a, scene, hits, first, n, rays);
}
void launch_envlight_reduce(const FrameArgs& a, EnvHits hits, uint32_t first, uint32_t n, RaySoA rays, float4* frame, cudaStream_t s) {
	if (!n) return;
	envlight_reduce_kernel<<<(n + 255) / 256, 256, 0, s>>>(a, hits, first, n, rays, frame);
}
__global__ void op_read_envmap_kernel(FrameArgs a, const float* __restrict__ dirs, uint32_t n, float4* __restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = read_envmap(a.envmap, a.env_w, a.env_h, ld3(dirs + 3 * i));
}
__global__ void op_env_light_dirs_kernel(uint32_t pix, uint32_t spp, uint32_t K, float nx, float ny, float nz, float* __restrict__ out) {
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= K) return;
	const V3 w = env_light_dir(pix, spp, K, k, mk(nx, ny, nz));
	out[3 * k] = w.x; out[3 * k + 1] = w.y; out[3 * k + 2] = w.z;
}
void launch_op_read_envmap(const FrameArgs& a, const float* dirs, uint32_t n, float4* out, cudaStream_t s) {
	if (n) op_read_envmap_kernel<<<(n + 255) / 256, 256, 0, s>>>(a, dirs, n, out);
}
void launch_op_env_light_dirs(uint32_t pix, uint32_t spp, uint32_t K, const float n[3], float* out, cudaStream_t s) {
	if (K) op_env_light_dirs_kernel<<<(K + 63) / 64, 64, 0, s>>>(pix, spp, K, n[0], n[1], n[2], out);
}
void launch_trace(SceneDev scene, int compat, uint32_t n, const float* o, const float* d, int32_t* mesh, int32_t* tri, float* t, cudaStream_t s) {
	if (!n) return;
	trace_kernel<<<(n + MESH_BLOCK - 1) / MESH_BLOCK, MESH_BLOCK, 0, s>>>(scene, compat, n, o, d, mesh, tri, t);
}

} // namespace hn
