// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h). Mesh half of the path: triangle test, nearest-hit search
// (brute force = ground truth, plus a plain binary BVH so the 1 M-triangle CPU baseline finishes), top-level mesh
// selection, sun shadow rays, Disney-style shading, and the whole-frame driver.
#include "orc_common.h"
#include "orc_api.h"
#include <thread>
#include <atomic>
#include <numeric>

namespace orc {

// defined in orc_nerf.cpp
Ray pixel_ray(const orc_render_params& p, uint32_t x, uint32_t y, bool* valid);
vec2 ld_random_pixel_offset(uint32_t spp);
Ray uv_to_ray(const orc_render_params& p, uint32_t spp, vec2 uv, const cam_t& cam, bool use_lens, bool* valid);
struct Net;
Net* net_new(const orc_net_config& c, const uint16_t* params);
void net_delete(Net* n);
vec4 read_envmap(const float* env, int w, int h, vec3 dir);
vec3 env_light_dir(uint32_t idx, uint32_t spp, uint32_t K, uint32_t k, vec3 n);
vec3 cap_light_dir(uint32_t idx, uint32_t spp, uint32_t K, uint32_t k, vec3 axis, float cap_height, vec3 n, float& weight);
vec4 read_probe(const float* tex, uint32_t n_theta, uint32_t n_phi, vec3 dir);
float nerf_transmittance(const orc_render_params& p, const Net& net, const uint8_t* bits, vec3 o, vec3 d, uint32_t* n_out);

struct Triangle { vec3 a, b, c; };

// triangle.cuh:44-58
static float tri_ray_intersect(const Triangle& tr, vec3 ro, vec3 rd) {
	vec3 v1v0 = tr.b - tr.a;
	vec3 v2v0 = tr.c - tr.a;
	vec3 rov0 = ro - tr.a;
	vec3 n = cross(v1v0, v2v0);
	vec3 q = cross(rov0, rd);
	float d = 1.0f / dot(rd, n);
	float u = d * -dot(q, v2v0);
	float v = d * dot(q, v1v0);
	float t = d * -dot(n, rov0);
	if (u < 0.0f || u > 1.0f || v < 0.0f || (u + v) > 1.0f || t < 0.0f) t = FLT_MAX;
	return t;
}
static vec3 tri_normal(const Triangle& tr) { return normalize(cross(tr.b - tr.a, tr.c - tr.a)); }

struct BvhNode {
	BoundingBox bb;
	int left, right; // inner: child node indices; leaf: left = -(first+1), right = count
};

struct Mesh {
	std::vector<Triangle> tris; // original order
	vec3 center;
	BoundingBox bb;
	std::vector<BvhNode> nodes;
	std::vector<uint32_t> order; // BVH leaf order -> original triangle index
};

static BoundingBox tri_bounds(const Mesh& m, const uint32_t* begin, const uint32_t* end) {
	BoundingBox bb = {{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}};
	for (const uint32_t* it = begin; it != end; ++it) {
		bb.enlarge(m.tris[*it].a); bb.enlarge(m.tris[*it].b); bb.enlarge(m.tris[*it].c);
	}
	return bb;
}
static void pad(BoundingBox& bb) {
	// conservative padding so that the slab test can never reject a ray the triangle test accepts
	vec3 d = bb.diag();
	float e = 1e-5f * std::max(std::max(d.x, d.y), d.z) + 1e-6f;
	bb.min = bb.min - e; bb.max = bb.max + e;
}
static int build_rec(Mesh& m, uint32_t first, uint32_t count) {
	int idx = (int)m.nodes.size();
	m.nodes.emplace_back();
	BoundingBox bb = tri_bounds(m, m.order.data() + first, m.order.data() + first + count);
	BoundingBox padded = bb; pad(padded);
	m.nodes[idx].bb = padded;
	if (count <= 4) {
		m.nodes[idx].left = -(int)(first + 1);
		m.nodes[idx].right = (int)count;
		return idx;
	}
	vec3 d = bb.diag();
	int axis = d.x >= d.y && d.x >= d.z ? 0 : (d.y >= d.z ? 1 : 2);
	auto centroid = [&](uint32_t t) { return m.tris[t].a[axis] + m.tris[t].b[axis] + m.tris[t].c[axis]; };
	uint32_t mid = count / 2;
	std::nth_element(m.order.begin() + first, m.order.begin() + first + mid, m.order.begin() + first + count,
		[&](uint32_t x, uint32_t y) { float cx = centroid(x), cy = centroid(y); return cx < cy || (cx == cy && x < y); });
	int l = build_rec(m, first, mid);
	int r = build_rec(m, first + mid, count - mid);
	m.nodes[idx].left = l; m.nodes[idx].right = r;
	return idx;
}

// robust slab test (NaN-safe): returns entry distance or FLT_MAX
static float slab_entry(const BoundingBox& bb, vec3 o, vec3 inv, float tmax) {
	float t0 = 0.0f, t1 = tmax;
	for (int a = 0; a < 3; ++a) {
		float ta = (bb.min[a] - o[a]) * inv[a], tb = (bb.max[a] - o[a]) * inv[a];
		float lo = std::fmin(ta, tb), hi = std::fmax(ta, tb);
		t0 = std::fmax(t0, lo); t1 = std::fmin(t1, hi);
	}
	return t0 <= t1 ? t0 : FLT_MAX;
}

// Nearest hit inside one mesh. Tie rule shared with the product: smallest t, then smallest original triangle index.
// Same acceptance as triangle_bvh.cu:150-193: mint starts at MAX_DIST, strict t < mint.
static void mesh_nearest(const Mesh& m, vec3 o, vec3 d, bool brute, int& best, float& mint) {
	best = -1; mint = MAX_DIST;
	if (brute) {
		for (uint32_t i = 0; i < m.tris.size(); ++i) {
			float t = tri_ray_intersect(m.tris[i], o, d);
			if (t < mint) { mint = t; best = (int)i; }
		}
		return;
	}
	vec3 inv = vec3{1.0f, 1.0f, 1.0f} / d;
	int stack[64]; int sp = 0; stack[sp++] = 0;
	while (sp) {
		const BvhNode& n = m.nodes[stack[--sp]];
		if (slab_entry(n.bb, o, inv, mint) == FLT_MAX) continue;
		if (n.left < 0) {
			uint32_t first = (uint32_t)(-n.left - 1);
			for (uint32_t k = 0; k < (uint32_t)n.right; ++k) {
				uint32_t ti = m.order[first + k];
				float t = tri_ray_intersect(m.tris[ti], o, d);
				if (t < mint || (t == mint && best >= 0 && (int)ti < best)) { mint = t; best = (int)ti; }
			}
		} else {
			stack[sp++] = n.left; stack[sp++] = n.right;
		}
	}
}

} // namespace orc

using namespace orc;

struct orc_scene {
	std::vector<Mesh> meshes;
	BoundingBox root;
};

namespace orc {

// geometry_bvh.cu:646-676 + :166-200. Returns mesh index (or -1), triangle index (or -1) and t (MAX_DIST on a miss
// inside the chosen mesh). `touched` says whether the reference kernel would have modified the ray at all.
static void scene_nearest(const orc_scene& s, bool compat, vec3 o, vec3 d, bool brute, int& mesh, int& tri, float& t, bool& touched) {
	mesh = -1; tri = -1; t = MAX_DIST; touched = false;
	if (compat) {
		float mint = MAX_DIST;
		for (size_t i = 0; i < s.meshes.size(); ++i) {
			vec2 tt = s.meshes[i].bb.ray_intersect(o, d);
			if (tt.x < mint && tt.x > -FLT_MAX) { mint = tt.x; mesh = (int)i; }
		}
		if (mesh < 0) return;
		touched = true;
		mesh_nearest(s.meshes[mesh], o, d, brute, tri, t);
		return;
	}
	for (size_t i = 0; i < s.meshes.size(); ++i) {
		int ti; float tt;
		mesh_nearest(s.meshes[i], o, d, brute, ti, tt);
		if (ti >= 0 && tt < t) { t = tt; tri = ti; mesh = (int)i; }
	}
	touched = true; // the default mode always advances the ray (to the hit, or by MAX_DIST on a miss)
	if (tri < 0) mesh = -1;
}

// testbed_geometry_training.cu:50-144
static float sqr(float x) { return x * x; }
static float saturate(float x) { return std::min(std::max(x, 0.0f), 1.0f); }
static float SchlickFresnel(float u) { float m = saturate(1.0f - u); return sqr(sqr(m)) * m; }
static const float PI_F = 3.14159265358979323846f;
static float G1(float NdotH, float a) {
	if (a >= 1.0f) return 1.0f / PI_F;
	float a2 = sqr(a);
	float t = 1.0f + (a2 - 1.0f) * NdotH * NdotH;
	return (a2 - 1.0f) / (PI_F * std::log(a2) * t);
}
static float G2(float NdotH, float a) {
	float a2 = sqr(a);
	float t = 1.0f + (a2 - 1.0f) * NdotH * NdotH;
	return a2 / (PI_F * t * t);
}
static float SmithG_GGX(float NdotV, float alphaG) {
	float a = alphaG * alphaG, b = NdotV * NdotV;
	return 1.0f / (NdotV + std::sqrt(a + b - a * b));
}
static vec3 v3(float s) { return {s, s, s}; }
static vec3 evaluate_shading(vec3 base_color, vec3 ambient_color, vec3 light_color, float metallic, float subsurface, float specular,
	float roughness, float specular_tint, float sheen, float sheen_tint, float clearcoat, float clearcoat_gloss, vec3 L, vec3 V, vec3 N) {
	float NdotL = dot(N, L), NdotV = dot(N, V);
	vec3 H = normalize(L + V);
	float NdotH = dot(N, H), LdotH = dot(L, H);
	float FL = SchlickFresnel(NdotL), FV = SchlickFresnel(NdotV);
	vec3 amb = ambient_color * mixf(0.2f, FV, metallic);
	amb = amb * base_color;
	if (NdotL < 0.f || NdotV < 0.f) return amb;
	float luminance = dot(base_color, vec3{0.3f, 0.6f, 0.1f});
	vec3 Ctint = base_color * (1.f / (luminance + 0.00001f));
	vec3 Cspec0 = mix(mix(v3(1.0f), Ctint, specular_tint) * specular * 0.08f, base_color, metallic);
	vec3 Csheen = mix(v3(1.0f), Ctint, sheen_tint);
	float Fd90 = 0.5f + 2.0f * LdotH * LdotH * roughness;
	float Fd = mixf(1, Fd90, FL) * mixf(1.f, Fd90, FV);
	float Fss90 = LdotH * LdotH * roughness;
	float Fss = mixf(1.0f, Fss90, FL) * mixf(1.0f, Fss90, FV);
	float ss = 1.25f * (Fss * (1.f / (NdotL + NdotV) - 0.5f) + 0.5f);
	float a = std::max(0.001f, sqr(roughness));
	float Ds = G2(NdotH, a);
	float FH = SchlickFresnel(LdotH);
	vec3 Fs = mix(Cspec0, v3(1.0f), FH);
	float Gs = SmithG_GGX(NdotL, a) * SmithG_GGX(NdotV, a);
	vec3 Fsheen = FH * sheen * Csheen;
	float Dr = G1(NdotH, mixf(0.1f, 0.001f, clearcoat_gloss));
	float Fr = mixf(0.04f, 1.0f, FH);
	float Gr = SmithG_GGX(NdotL, 0.25f) * SmithG_GGX(NdotV, 0.25f);
	float CCs = 0.25f * clearcoat * Gr * Fr * Dr;
	vec3 brdf = ((1.0f / PI_F) * mixf(Fd, ss, subsurface) * base_color + Fsheen) * (1.0f - metallic) + Gs * Fs * Ds + v3(CCs);
	return (brdf * light_color) * NdotL + amb;
}

template <typename F>
static void parallel_rows(int n_rows, int n_threads, F&& f) {
	if (n_threads <= 1) { for (int y = 0; y < n_rows; ++y) f(y); return; }
	std::atomic<int> next{0};
	std::vector<std::thread> th;
	for (int i = 0; i < n_threads; ++i) th.emplace_back([&] { for (int y; (y = next.fetch_add(1)) < n_rows;) f(y); });
	for (auto& t : th) t.join();
}

// testbed_geometry_training.cu:2202-2320 (render_geometry_mesh), Shade mode with sun shadow rays.
// With an envmap and env_samples = K > 0 every shaded triangle hit also gathers environment light (a-16, OUR definition:
// the reference's integration/shading kernels are declared but absent): K cosine-weighted directions around the
// viewer-facing normal, each tested against the meshes (nearest-hit search = shadow ray) and attenuated by the
// transmittance of the NeRF; Lambertian estimator colour += albedo / K * sum_k L_env(w_k) * V_k * T_k.
static void render_mesh(const orc_scene& s, const orc_render_params& p, float* frame, float* depth, const Net* net = nullptr, const uint8_t* bits = nullptr) {
	BoundingBox aabb = bb_from(p.mesh_aabb);
	cam_t cam = cam_from(p.cam0);
	vec3 cam_pos = cam.c[3], cam_fwd = cam.c[2];
	vec3 sun = normalize(vec3{p.sun_dir[0], p.sun_dir[1], p.sun_dir[2]});
	vec3 up = normalize(vec3{p.up_dir[0], p.up_dir[1], p.up_dir[2]});
	const float* B = p.brdf;
	vec3 basecolor = {B[7], B[8], B[9]}, ambientcolor = {B[10], B[11], B[12]};
	bool compat = p.compat_nearest_aabb_only != 0;
	parallel_rows(p.height, p.n_threads, [&](int y) {
		for (int x = 0; x < p.width; ++x) {
			uint32_t idx = x + p.width * y;
			if (p.pixel_mask && !p.pixel_mask[idx]) continue;
			// :488-579 init_rays_with_payload_kernel_mesh_geometry (pixel_to_ray: no rolling shutter, camera_matrix as is)
			vec2 off = ld_random_pixel_offset(p.snap_to_pixel_centers ? 0 : p.spp);
			vec2 uv = {((float)x + off.x) / (float)p.width, ((float)y + off.y) / (float)p.height};
			bool valid = true;
			Ray ray = uv_to_ray(p, p.spp, uv, cam, false, &valid); // no Lens on this ray init (:520-536); foveation, hidden-area mask, parallax shift + depth of field apply
			depth[idx] = MAX_DEPTH;
			// a pixel the hidden-area mask hides: the reference marks the payload dead (:544-551) and still pushes its zero ray through the
			// trace / shade kernels (colour undefined); here such a pixel simply stays cleared
			if (!valid) continue;
			ray.d = normalize(ray.d);
			float t0 = std::max(aabb.ray_intersect(ray.o, ray.d).x, 0.0f);
			vec3 pos = ray.o + (t0 + 1e-6f) * ray.d;
			vec3 dir = ray.d;
			// :2140-2155 trace_mesh_bvh: every pixel is traced (alive or not), n_hit == n_pixels
			int mi, ti; float t; bool touched;
			scene_nearest(s, compat, pos, dir, false, mi, ti, t, touched);
			vec3 nrm = dir; // direction smuggled through the normal buffer
			if (touched) {
				pos = pos + t * dir;
				if (ti >= 0) nrm = tri_normal(s.meshes[mi].tris[ti]);
			}
			// :222-278 shadow ray
			float shadow;
			{
				float sgn = -std::copysign(1.0f, dot(dir, nrm)); // faceforward(n, i, nref) = n * -sign(dot(i, nref))
				vec3 view_pos = pos + normalize(nrm * sgn) * 1e-3f;
				float ts = std::fmax(aabb.ray_intersect(view_pos, sun).x + 1e-6f, 0.0f);
				view_pos = view_pos + ts * sun;
				// :239-244: a shadow origin outside the box keeps payload.dir (= the primary direction) and is still
				// traced by trace_mesh_bvh (:2149 copies payload.dir for every ray, alive or not).
				vec3 sdir = aabb.contains(view_pos) ? sun : dir;
				int smi, sti; float st; bool stouched;
				scene_nearest(s, compat, view_pos, sdir, false, smi, sti, st, stouched);
				vec3 end = stouched ? view_pos + st * sdir : view_pos;
				shadow = aabb.contains(end) ? 0.0f : 1.0f;
			}
			// :280-355 shade_kernel_mesh_geometry
			if (!aabb.contains(pos)) continue;
			vec3 normal = normalize(nrm);
			bool floor = false;
			if (pos.y < p.floor_y + 0.001f && dir.y < 0.f) { normal = {0.0f, 1.0f, 0.0f}; floor = true; }
			float skyam = -dot(normal, up) * 0.5f + 0.5f;
			vec3 suncol = vec3{255.f / 255.0f, 225.f / 255.0f, 195.f / 255.0f} * 4.f * shadow;
			vec3 skycol = vec3{195.f / 255.0f, 215.f / 255.0f, 255.f / 255.0f} * 4.f * skyam;
			float check_size = 8.f / aabb.diag().x;
			float check = ((int(std::floor(check_size * (pos.x - aabb.min.x))) ^ int(std::floor(check_size * (pos.z - aabb.min.z)))) & 1) ? 0.8f : 0.2f;
			vec3 floorcol = {check * check * check, check * check, check};
			vec3 color = evaluate_shading(floor ? floorcol : basecolor * basecolor, ambientcolor * skycol, suncol,
				floor ? 0.f : B[0], floor ? 0.f : B[1], floor ? 1.f : B[2], floor ? 0.5f : B[3], 0.f, floor ? 0.f : B[4], 0.f,
				floor ? 0.f : B[5], B[6], sun, -normalize(dir), normal);
			const bool probe_light = p.env_source == 1 && p.probe && p.probe_n_theta > 0;
			if ((p.envmap || probe_light) && p.env_samples > 0 && touched && ti >= 0) {
				const uint32_t K = (uint32_t)p.env_samples;
				vec3 nf = normal * -std::copysign(1.0f, dot(dir, normal));
				vec3 org = pos + nf * 1e-3f;
				vec3 albedo = floor ? floorcol : basecolor * basecolor;
				vec3 sum = {0.f, 0.f, 0.f};
				for (uint32_t k = 0; k < K; ++k) {
					vec3 w;
					float weight = 1.0f;
					bool cap = false;
					if (probe_light && p.env_sampling == 1) {
						BoundingBox train = bb_from(p.train_aabb);
						vec3 to_c = train.center() - org;
						float dist = length(to_c), radius = 0.5f * length(train.diag());
						if (dist > radius) {
							float cos_max = std::sqrt(std::fmax(0.0f, 1.0f - (radius / dist) * (radius / dist)));
							w = cap_light_dir(idx, p.spp, K, k, to_c / dist, 1.0f - cos_max, nf, weight);
							cap = true;
						}
					}
					if (!cap) w = env_light_dir(idx, p.spp, K, k, nf);
					int emi, eti; float et; bool etouched;
					scene_nearest(s, false, org, w, false, emi, eti, et, etouched);
					if (eti >= 0 || !(weight > 0.0f)) continue;
					vec4 L = probe_light ? read_probe(p.probe, (uint32_t)p.probe_n_theta, (uint32_t)p.probe_n_phi, w) : read_envmap(p.envmap, p.env_w, p.env_h, w);
					float T = (!probe_light && net && bits) ? nerf_transmittance(p, *net, bits, org, w, nullptr) : 1.0f;
					sum = sum + vec3{L.x, L.y, L.z} * (T * weight);
				}
				color = color + albedo * sum * (1.0f / (float)K);
			}
			float* fb = frame + 4 * (size_t)idx;
			fb[0] = color.x; fb[1] = color.y; fb[2] = color.z; fb[3] = 1.0f;
			depth[idx] = dot(cam_fwd, pos - cam_pos);
		}
	});
}

} // namespace orc

extern "C" {

float orc_triangle_ray_intersect(const float tri[9], const float o[3], const float d[3]) {
	Triangle t = {{tri[0], tri[1], tri[2]}, {tri[3], tri[4], tri[5]}, {tri[6], tri[7], tri[8]}};
	return tri_ray_intersect(t, {o[0], o[1], o[2]}, {d[0], d[1], d[2]});
}

orc_scene* orc_scene_create(void) { return new orc_scene(); }
void orc_scene_destroy(orc_scene* s) { delete s; }
void orc_scene_add_mesh(orc_scene* s, const float* tr, uint32_t n, const float center[3]) {
	s->meshes.emplace_back();
	Mesh& m = s->meshes.back();
	m.tris.resize(n);
	for (uint32_t i = 0; i < n; ++i) {
		const float* f = tr + 9 * (size_t)i;
		m.tris[i] = {{f[0], f[1], f[2]}, {f[3], f[4], f[5]}, {f[6], f[7], f[8]}};
	}
	m.center = {center[0], center[1], center[2]};
}
void orc_scene_build(orc_scene* s) {
	bool first = true;
	for (Mesh& m : s->meshes) {
		m.order.resize(m.tris.size());
		std::iota(m.order.begin(), m.order.end(), 0u);
		// geometry_bvh.cu:18-35: mesh box = exact min/max over its vertices
		m.bb = tri_bounds(m, m.order.data(), m.order.data() + m.order.size());
		m.nodes.clear();
		m.nodes.reserve(m.tris.size());
		build_rec(m, 0, (uint32_t)m.tris.size());
		if (first) { s->root = m.bb; first = false; }
		else { s->root.enlarge(m.bb.min); s->root.enlarge(m.bb.max); }
	}
}
void orc_scene_aabb(const orc_scene* s, float out[6]) {
	out[0] = s->root.min.x; out[1] = s->root.min.y; out[2] = s->root.min.z;
	out[3] = s->root.max.x; out[4] = s->root.max.y; out[5] = s->root.max.z;
}
// ---------------------------------------------------------------- triangle.cuh:39-41,66-147 (distance queries)
static float sign1(float x) { return std::copysign(1.0f, x); }
static float clamp01(float x) { return std::min(std::max(x, 0.0f), 1.0f); }
static float tri_distance_sq(const Triangle& t, vec3 pos) {
	vec3 v21 = t.b - t.a, p1 = pos - t.a, v32 = t.c - t.b, p2 = pos - t.b, v13 = t.a - t.c, p3 = pos - t.c;
	vec3 nor = cross(v21, v13);
	if (sign1(dot(cross(v21, nor), p1)) + sign1(dot(cross(v32, nor), p2)) + sign1(dot(cross(v13, nor), p3)) < 2.0f) {
		vec3 e1 = v21 * clamp01(dot(v21, p1) / length2(v21)) - p1;
		vec3 e2 = v32 * clamp01(dot(v32, p2) / length2(v32)) - p2;
		vec3 e3 = v13 * clamp01(dot(v13, p3) / length2(v13)) - p3;
		return std::min(std::min(length2(e1), length2(e2)), length2(e3));
	}
	return dot(nor, p1) * dot(nor, p1) / length2(nor);
}
static vec3 closest_on_segment(vec3 a, vec3 b, vec3 c) {
	float t = dot(c - a, b - a) / dot(b - a, b - a);
	t = std::max(std::min(t, 1.0f), 0.0f);
	return a + t * (b - a);
}
static vec3 tri_closest_point(const Triangle& t, vec3 point) {
	vec3 n = tri_normal(t);
	point = point - dot(n, point - t.a) * n;
	vec3 la = t.a - point, lb = t.b - point, lc = t.c - point;
	vec3 u = cross(lb, lc), v = cross(lc, la), w = cross(la, lb);
	if (dot(u, v) >= 0.0f && dot(u, w) >= 0.0f) return point;
	vec3 c1 = closest_on_segment(t.a, t.b, point), c2 = closest_on_segment(t.b, t.c, point), c3 = closest_on_segment(t.c, t.a, point);
	float m1 = length2(point - c1), m2 = length2(point - c2), m3 = length2(point - c3);
	float mn = std::min(std::min(m1, m2), m3);
	if (mn == m1) return c1;
	if (mn == m2) return c2;
	return c3;
}
void orc_triangle_distance_sq(const float tri[9], const float p[3], float* out) {
	Triangle t = {{tri[0], tri[1], tri[2]}, {tri[3], tri[4], tri[5]}, {tri[6], tri[7], tri[8]}};
	*out = tri_distance_sq(t, {p[0], p[1], p[2]});
}
void orc_triangle_closest_point(const float tri[9], const float p[3], float out[3]) {
	Triangle t = {{tri[0], tri[1], tri[2]}, {tri[3], tri[4], tri[5]}, {tri[6], tri[7], tri[8]}};
	vec3 c = tri_closest_point(t, {p[0], p[1], p[2]});
	out[0] = c.x; out[1] = c.y; out[2] = c.z;
}
// pcg32 in its default state (default_rng_t rng; pcg32.h PCG32_DEFAULT_STATE / PCG32_DEFAULT_STREAM), skip-ahead, two floats
static void default_rng_offset(uint64_t delta, float out[2]) {
	uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
	uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
	while (delta > 0) {
		if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
		cur_plus = (cur_mult + 1) * cur_plus;
		cur_mult *= cur_mult;
		delta /= 2;
	}
	state = acc_mult * state + acc_plus;
	for (int k = 0; k < 2; ++k) {
		uint64_t old = state;
		state = old * 0x5851f42d4c957f2dULL + inc;
		uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
		uint32_t rot = (uint32_t)(old >> 59u);
		uint32_t u = (((xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31))) >> 9) | 0x3f800000u;
		float f; std::memcpy(&f, &u, 4);
		out[k] = f - 1.0f;
	}
}
static float fractf1(float x) { return x - std::floor(x); }
// random_val.cuh:45-54,84-101 with N_DIRS = 32 (epsilon 1.33)
static vec3 fibonacci_dir32(uint32_t i, float ox, float oy) {
	const float epsilon = 1.33f, GOLDEN_RATIO = 1.6180339887498948482045868343656f;
	float px = fractf1((i + epsilon) / (32 - 1 + 2 * epsilon) + ox), py = fractf1(i / GOLDEN_RATIO + oy);
	float cos_theta = -2.0f * px + 1.0f, phi = 2.0f * 3.14159265358979323846f * (py - 0.5f);
	float sin_theta = std::sqrt(std::fmax(1.0f - cos_theta * cos_theta, 0.0f));
	return {sin_theta * std::cos(phi), sin_theta * std::sin(phi), cos_theta};
}
void orc_fibonacci_dir32(uint32_t i, const float offset[2], float out[3]) { vec3 d = fibonacci_dir32(i, offset[0], offset[1]); out[0] = d.x; out[1] = d.y; out[2] = d.z; }
void orc_default_rng_offset(uint64_t advance, float out[2]) { default_rng_offset(advance, out); }
// TriangleBvh::unsigned_distance / signed_distance_watertight (triangle_bvh.cu:285-297) by brute force over every triangle of
// the scene: the minimum does not depend on the BVH, so this is the witness for the product's pruned walk.
void orc_scene_distance(const orc_scene* s, uint32_t n, const float* pos, int32_t mode, float max_distance, int32_t n_threads, float* out, int32_t* tri_out) {
	int nt = n_threads > 1 ? n_threads : 1;
	std::atomic<uint32_t> next{0};
	auto work = [&] {
		for (uint32_t b; (b = next.fetch_add(64)) < n;) {
			for (uint32_t i = b; i < std::min(n, b + 64); ++i) {
				vec3 p = {pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]};
				float best = max_distance * max_distance;
				int bm = -1, bt = -1;
				for (size_t m = 0; m < s->meshes.size(); ++m)
					for (size_t k = 0; k < s->meshes[m].tris.size(); ++k) {
						float d2 = tri_distance_sq(s->meshes[m].tris[k], p);
						if (d2 < best || (d2 == best && bm < 0)) { best = d2; bm = (int)m; bt = (int)k; }
					}
				if (bm < 0) { out[i] = max_distance; if (tri_out) tri_out[i] = -1; continue; }
				float d = std::sqrt(best);
				if (tri_out) tri_out[i] = bt;
				if (mode == 1) {
					vec3 cp = tri_closest_point(s->meshes[bm].tris[bt], p);
					vec3 ns = {0.0f, 0.0f, 0.0f}; float nw = 0.0f;
					for (const Triangle& t : s->meshes[bm].tris) if (tri_distance_sq(t, cp) < 1e-6f) { ns = ns + tri_normal(t); nw += 1.0f; }
					vec3 avg = ns / nw;
					d = std::copysign(d, dot(avg, p - cp));
				} else if (mode == 2) { // signed_distance_raystab (:299-317) with the kernel's rng (:573-574): any escaping stab ray -> outside
					float off[2];
					default_rng_offset((uint64_t)i * 2, off);
					bool inside = true;
					for (uint32_t k = 0; k < 32 && inside; ++k) {
						int mi, ti; float tt; bool touched;
						scene_nearest(*s, false, p, fibonacci_dir32(k, off[0], off[1]), false, mi, ti, tt, touched);
						if (ti < 0) inside = false;
					}
					if (inside) d = -d;
				}
				out[i] = d;
			}
		}
	};
	std::vector<std::thread> th;
	for (int k = 1; k < nt; ++k) th.emplace_back(work);
	work();
	for (auto& t : th) t.join();
}

void orc_scene_trace(const orc_scene* s, const orc_render_params* p, uint32_t n, const float* o, const float* d,
	int32_t* mesh_idx, int32_t* tri_idx, float* t, int32_t brute) {
	int nt = p->n_threads > 1 ? p->n_threads : 1;
	std::atomic<uint32_t> next{0};
	auto work = [&] {
		for (uint32_t b; (b = next.fetch_add(1024)) < n;) {
			for (uint32_t i = b; i < std::min(n, b + 1024); ++i) {
				int mi, ti; float tt; bool touched;
				scene_nearest(*s, p->compat_nearest_aabb_only != 0, {o[3 * i], o[3 * i + 1], o[3 * i + 2]}, {d[3 * i], d[3 * i + 1], d[3 * i + 2]}, brute != 0, mi, ti, tt, touched);
				mesh_idx[i] = mi; tri_idx[i] = ti; t[i] = tt;
			}
		}
	};
	std::vector<std::thread> th;
	for (int i = 1; i < nt; ++i) th.emplace_back(work);
	work();
	for (auto& x : th) x.join();
}
void orc_render_mesh(const orc_scene* s, const orc_render_params* p, float* frame, float* depth) { render_mesh(*s, *p, frame, depth); }

uint64_t orc_render_frame(const orc_scene* scene, const orc_render_params* p, const orc_net_config* c, const uint16_t* params,
	const uint8_t* bitfield, float* frame, float* depth, uint32_t* n_steps_out) {
	size_t n = (size_t)p->width * p->height;
	std::memset(frame, 0, n * 4 * sizeof(float)); // render_buffer.cu:603 clear
	std::memset(depth, 0, n * sizeof(float));
	bool has_mesh = scene && !scene->meshes.empty();
	Net* net = (params && p->envmap && p->env_samples > 0) ? net_new(*c, params) : nullptr;
	if (has_mesh) render_mesh(*scene, *p, frame, depth, net, bitfield);
	if (net) net_delete(net);
	uint64_t total = 0;
	if (params) total = orc_render_nerf(p, c, params, bitfield, frame, depth, n_steps_out, has_mesh ? 1 : 0);
	return total;
}

} // extern "C"
