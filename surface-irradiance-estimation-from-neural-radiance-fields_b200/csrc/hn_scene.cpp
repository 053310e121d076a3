// Scene side of the boundary: data/geometry/*.json scene files, ascii .obj / binary .stl meshes, load_mesh's
// normalisation (src/testbed_geometry_training.cu:2786-2866), and the B200 BVH: 4-wide, 128-byte nodes with padded child
// boxes in SoA form, triangles re-ordered into leaf order carrying their original index. The split rule is the
// reference's (median on the axis of maximum centroid variance, two rounds -> 4 children, src/triangle_bvh.cu:425-508)
// made deterministic with an index tie-break; the tree itself is ours, parity is on (mesh, original triangle, t).
#include "hn_context.h"
#include "hn_json.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <functional>
#include <numeric>
#include <sstream>

namespace hn {

static constexpr uint32_t LEAF_SIZE = 8; // build(tris, 8), src/testbed_geometry_training.cu:2844

static std::string lower_ext(const std::string& path) {
	size_t p = path.find_last_of('.');
	std::string e = p == std::string::npos ? "" : path.substr(p + 1);
	std::transform(e.begin(), e.end(), e.begin(), [](unsigned char c) { return (char)std::tolower(c); });
	return e;
}

// ascii .obj: v / f records, polygons fan-triangulated, negative (relative) indices accepted
static std::vector<float> load_obj_triangles(const std::string& path) {
	std::ifstream f(path);
	if (!f) throw std::runtime_error("cannot open mesh '" + path + "'");
	std::vector<float> verts, out;
	std::string line;
	while (std::getline(f, line)) {
		if (line.size() < 2) continue;
		if (line[0] == 'v' && (line[1] == ' ' || line[1] == '\t')) {
			std::istringstream ss(line.substr(2));
			float x, y, z;
			if (ss >> x >> y >> z) { verts.push_back(x); verts.push_back(y); verts.push_back(z); }
		} else if (line[0] == 'f' && (line[1] == ' ' || line[1] == '\t')) {
			std::istringstream ss(line.substr(2));
			std::string tok;
			std::vector<long> idx;
			while (ss >> tok) {
				long v = std::strtol(tok.c_str(), nullptr, 10); // "v", "v/vt", "v//vn", "v/vt/vn"
				long nv = (long)(verts.size() / 3);
				if (v < 0) v = nv + v; else v = v - 1;
				if (v < 0 || v >= nv) throw std::runtime_error("obj face index out of range in '" + path + "'");
				idx.push_back(v);
			}
			for (size_t k = 2; k < idx.size(); ++k) {
				const long tri[3] = {idx[0], idx[k - 1], idx[k]};
				for (long t : tri) { out.push_back(verts[3 * t]); out.push_back(verts[3 * t + 1]); out.push_back(verts[3 * t + 2]); }
			}
		}
	}
	return out;
}

static std::vector<float> load_stl_triangles(const std::string& path) {
	std::ifstream f(path, std::ios::binary);
	if (!f) throw std::runtime_error("cannot open mesh '" + path + "'");
	char header[80];
	uint32_t n = 0;
	f.read(header, 80);
	f.read((char*)&n, 4);
	std::vector<float> out((size_t)n * 9);
	for (uint32_t i = 0; i < n; ++i) {
		float rec[12];
		uint16_t attr;
		f.read((char*)rec, 48);
		f.read((char*)&attr, 2);
		if (!f) throw std::runtime_error("truncated stl '" + path + "'");
		std::memcpy(&out[(size_t)i * 9], rec + 3, 36);
	}
	return out;
}

void add_mesh(hn_context* ctx, const float* tris, size_t n, const float center[3], bool normalize) {
	if (n == 0) throw std::runtime_error("mesh has no triangles");
	HostMesh m;
	m.tris.resize(n);
	std::memcpy(m.tris.data(), tris, n * sizeof(Tri));
	std::memcpy(m.center, center, 12);
	if (normalize) {
		// load_mesh :2805-2823: box of all vertices, inflated by 0.5% of its diagonal, scaled so the longest side is 1,
		// centred on 0.5, then offset by `center`. Same float operation order as the reference.
		float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
		float* v = &m.tris[0].a[0];
		for (size_t i = 0; i < n * 3; ++i) for (int d = 0; d < 3; ++d) { lo[d] = std::min(lo[d], v[3 * i + d]); hi[d] = std::max(hi[d], v[3 * i + d]); }
		float diag[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
		float len = 0.0f;
		for (int d = 0; d < 3; ++d) len += diag[d] * diag[d];
		const float infl = std::sqrt(len) * 0.005f;
		for (int d = 0; d < 3; ++d) { lo[d] -= infl; hi[d] += infl; diag[d] = hi[d] - lo[d]; }
		m.mesh_scale = std::max(diag[0], std::max(diag[1], diag[2]));
		for (size_t i = 0; i < n * 3; ++i) for (int d = 0; d < 3; ++d) {
			float x = (v[3 * i + d] - lo[d] - 0.5f * diag[d]) / m.mesh_scale + 0.5f;
			v[3 * i + d] = x + center[d];
		}
	}
	ctx->meshes.push_back(std::move(m));
	ctx->scene_built = false;
}

void load_mesh_file(hn_context* ctx, const std::string& path, const float center[3]) {
	std::string ext = lower_ext(path);
	std::vector<float> t;
	if (ext == "obj") t = load_obj_triangles(path);
	else if (ext == "stl") t = load_stl_triangles(path);
	else throw std::runtime_error("mesh data path must be a mesh in ascii .obj or binary .stl format.");
	add_mesh(ctx, t.data(), t.size() / 9, center, true);
}

// ---------------------------------------------------------------- BVH4 build
struct Builder {
	HostMesh& m;
	std::vector<std::array<float, 3>> centroid;
	explicit Builder(HostMesh& mesh) : m(mesh) {
		centroid.resize(m.tris.size());
		for (size_t i = 0; i < m.tris.size(); ++i) for (int d = 0; d < 3; ++d) centroid[i][d] = (m.tris[i].a[d] + m.tris[i].b[d] + m.tris[i].c[d]) / 3.0f;
	}
	void bounds(uint32_t first, uint32_t count, float lo[3], float hi[3]) const {
		for (int d = 0; d < 3; ++d) { lo[d] = INFINITY; hi[d] = -INFINITY; }
		for (uint32_t k = first; k < first + count; ++k) {
			const Tri& t = m.tris[m.order[k]];
			for (int d = 0; d < 3; ++d) {
				lo[d] = std::min(lo[d], std::min(t.a[d], std::min(t.b[d], t.c[d])));
				hi[d] = std::max(hi[d], std::max(t.a[d], std::max(t.b[d], t.c[d])));
			}
		}
	}
	// Binned surface-area heuristic (16 bins per axis over the centroid bounds): the split plane that minimises
	// area(left) * n(left) + area(right) * n(right) over the three axes. Rays that graze a finely tessellated surface visit far fewer
	// nodes than with median splits, and those rays are the critical path of the mesh pass once a frame is sharded over several GPUs.
	// Returns 0 when no plane separates the centroids (the caller falls back to the median split). The hit a traversal reports does not
	// depend on the tree (nearest t, ties by original triangle index), so parity with the brute-force oracle is untouched.
	bool use_sah = true;
	std::vector<std::array<float, 6>> tri_box;
	void ensure_boxes() {
		if (!tri_box.empty()) return;
		tri_box.resize(m.tris.size());
		for (size_t i = 0; i < m.tris.size(); ++i) {
			const Tri& t = m.tris[i];
			for (int d = 0; d < 3; ++d) {
				tri_box[i][d] = std::min(t.a[d], std::min(t.b[d], t.c[d]));
				tri_box[i][3 + d] = std::max(t.a[d], std::max(t.b[d], t.c[d]));
			}
		}
	}
	static float half_area(const float lo[3], const float hi[3]) {
		const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
		return dx * dy + dy * dz + dz * dx;
	}
	uint32_t split_sah(uint32_t first, uint32_t count) {
		constexpr int NB = 16;
		ensure_boxes();
		float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
		for (uint32_t k = first; k < first + count; ++k) for (int d = 0; d < 3; ++d) { const float c = centroid[m.order[k]][d]; clo[d] = std::min(clo[d], c); chi[d] = std::max(chi[d], c); }
		float best_cost = INFINITY; int best_axis = -1, best_plane = 0;
		for (int axis = 0; axis < 3; ++axis) {
			const float ext = chi[axis] - clo[axis];
			if (!(ext > 0.0f)) continue;
			const float scale = (float)NB / ext;
			uint32_t cnt[NB] = {};
			float blo[NB][3], bhi[NB][3];
			for (int b = 0; b < NB; ++b) for (int d = 0; d < 3; ++d) { blo[b][d] = INFINITY; bhi[b][d] = -INFINITY; }
			for (uint32_t k = first; k < first + count; ++k) {
				const uint32_t t = m.order[k];
				const int b = std::min(NB - 1, std::max(0, (int)((centroid[t][axis] - clo[axis]) * scale)));
				++cnt[b];
				for (int d = 0; d < 3; ++d) { blo[b][d] = std::min(blo[b][d], tri_box[t][d]); bhi[b][d] = std::max(bhi[b][d], tri_box[t][3 + d]); }
			}
			// sweep: areas and counts to the left of plane p (bins 0..p-1) and to the right (bins p..NB-1)
			float la[NB], ra[NB]; uint32_t ln[NB], rn[NB];
			float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
			uint32_t n = 0;
			for (int p = 1; p < NB; ++p) {
				for (int d = 0; d < 3; ++d) { lo[d] = std::min(lo[d], blo[p - 1][d]); hi[d] = std::max(hi[d], bhi[p - 1][d]); }
				n += cnt[p - 1]; ln[p] = n; la[p] = n ? half_area(lo, hi) : 0.0f;
			}
			for (int d = 0; d < 3; ++d) { lo[d] = INFINITY; hi[d] = -INFINITY; }
			n = 0;
			for (int p = NB - 1; p >= 1; --p) {
				for (int d = 0; d < 3; ++d) { lo[d] = std::min(lo[d], blo[p][d]); hi[d] = std::max(hi[d], bhi[p][d]); }
				n += cnt[p]; rn[p] = n; ra[p] = n ? half_area(lo, hi) : 0.0f;
			}
			for (int p = 1; p < NB; ++p) {
				if (!ln[p] || !rn[p]) continue;
				const float cost = la[p] * (float)ln[p] + ra[p] * (float)rn[p];
				if (cost < best_cost) { best_cost = cost; best_axis = axis; best_plane = p; }
			}
		}
		if (best_axis < 0) return 0;
		const float scale = 16.0f / (chi[best_axis] - clo[best_axis]);
		auto b = m.order.begin() + first;
		auto mid = std::stable_partition(b, b + count, [&](uint32_t t) {
			return std::min(NB - 1, std::max(0, (int)((centroid[t][best_axis] - clo[best_axis]) * scale))) < best_plane;
		});
		const uint32_t left = (uint32_t)(mid - b);
		return (left == 0 || left == count) ? 0u : left;
	}
	// median split on the axis of maximum centroid variance
	uint32_t split(uint32_t first, uint32_t count) {
		if (use_sah && count > 4) { const uint32_t l = split_sah(first, count); if (l) return l; }
		double mean[3] = {0, 0, 0}, var[3] = {0, 0, 0};
		for (uint32_t k = first; k < first + count; ++k) for (int d = 0; d < 3; ++d) mean[d] += centroid[m.order[k]][d];
		for (int d = 0; d < 3; ++d) mean[d] /= count;
		for (uint32_t k = first; k < first + count; ++k) for (int d = 0; d < 3; ++d) { double x = centroid[m.order[k]][d] - mean[d]; var[d] += x * x; }
		int axis = (var[0] >= var[1] && var[0] >= var[2]) ? 0 : (var[1] >= var[2] ? 1 : 2);
		uint32_t mid = count / 2;
		auto b = m.order.begin() + first;
		std::nth_element(b, b + mid, b + count, [&](uint32_t x, uint32_t y) {
			float cx = centroid[x][axis], cy = centroid[y][axis];
			return cx < cy || (cx == cy && x < y);
		});
		return mid;
	}
	void build_node(uint32_t node_idx, uint32_t first, uint32_t count) {
		struct Range { uint32_t first, count; };
		std::vector<Range> r = {{first, count}};
		for (int round = 0; round < 2; ++round) {
			std::vector<Range> next;
			for (const Range& x : r) {
				if (x.count >= 2 && (x.count > LEAF_SIZE || round == 0 || r.size() == 1)) {
					uint32_t mid = split(x.first, x.count);
					next.push_back({x.first, mid});
					next.push_back({x.first + mid, x.count - mid});
				} else {
					next.push_back(x);
				}
			}
			r.swap(next);
		}
		Node4 node;
		std::memset(&node, 0, sizeof(node));
		std::vector<std::pair<uint32_t, Range>> todo;
		for (size_t i = 0; i < 4; ++i) {
			if (i >= r.size()) {
				node.lo_x[i] = node.lo_y[i] = node.lo_z[i] = INFINITY;
				node.hi_x[i] = node.hi_y[i] = node.hi_z[i] = -INFINITY;
				node.child[i] = 0; node.count[i] = 0;
				continue;
			}
			float lo[3], hi[3];
			bounds(r[i].first, r[i].count, lo, hi);
			// conservative padding: the slab test must never reject a ray the triangle test accepts
			float mag = 0.f;
			for (int d = 0; d < 3; ++d) mag = std::max(mag, std::max(std::fabs(lo[d]), std::fabs(hi[d])));
			const float pad = 1e-5f * mag + 1e-5f;
			node.lo_x[i] = lo[0] - pad; node.lo_y[i] = lo[1] - pad; node.lo_z[i] = lo[2] - pad;
			node.hi_x[i] = hi[0] + pad; node.hi_y[i] = hi[1] + pad; node.hi_z[i] = hi[2] + pad;
			if (r[i].count <= LEAF_SIZE) {
				node.child[i] = -1 - (int32_t)((r[i].first << 4) | r[i].count);
				node.count[i] = (int32_t)r[i].count;
			} else {
				uint32_t child_idx = (uint32_t)m.nodes.size();
				m.nodes.emplace_back();
				node.child[i] = (int32_t)child_idx;
				node.count[i] = -1;
				todo.push_back({child_idx, r[i]});
			}
		}
		m.nodes[node_idx] = node;
		for (auto& t : todo) build_node(t.first, t.second.first, t.second.count);
	}
};

static void build_mesh_bvh(HostMesh& m, bool sah) {
	if (m.tris.size() >= (1u << 27)) throw std::runtime_error("mesh too large for the leaf encoding");
	m.order.resize(m.tris.size());
	std::iota(m.order.begin(), m.order.end(), 0u);
	m.nodes.clear();
	m.nodes.reserve(m.tris.size() / 2 + 4);
	m.nodes.emplace_back();
	Builder b(m);
	b.use_sah = sah;
	b.build_node(0, 0, (uint32_t)m.tris.size());
	// exact mesh box (geometry_bvh.cu:18-35)
	float lo[3], hi[3];
	b.bounds(0, (uint32_t)m.tris.size(), lo, hi);
	for (int d = 0; d < 3; ++d) { m.bb[d] = lo[d]; m.bb[3 + d] = hi[d]; }
}

void build_scene(hn_context* ctx) {
	if (ctx->meshes.empty()) { ctx->scene_built = true; return; }
	size_t n_nodes = 0, n_tris = 0;
	for (HostMesh& m : ctx->meshes) {
		if (m.nodes.empty()) build_mesh_bvh(m, ctx->opt.bvh_sah != 0);
		n_nodes += m.nodes.size();
		n_tris += m.tris.size();
	}
	std::vector<Node4> nodes;
	nodes.reserve(n_nodes);
	std::vector<float4> tris(n_tris * 3);
	std::vector<MeshDev> md(ctx->meshes.size());
	size_t tri_base = 0;
	float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
	for (size_t mi = 0; mi < ctx->meshes.size(); ++mi) {
		HostMesh& m = ctx->meshes[mi];
		const uint32_t node_base = (uint32_t)nodes.size();
		for (Node4 n : m.nodes) {
			for (int i = 0; i < 4; ++i) if (n.count[i] == -1) n.child[i] += (int32_t)node_base; // inner children -> global node index
			nodes.push_back(n);
		}
		for (size_t k = 0; k < m.order.size(); ++k) {
			const Tri& t = m.tris[m.order[k]];
			float w;
			int32_t orig = (int32_t)m.order[k];
			std::memcpy(&w, &orig, 4);
			tris[(tri_base + k) * 3 + 0] = make_float4(t.a[0], t.a[1], t.a[2], w);
			tris[(tri_base + k) * 3 + 1] = make_float4(t.b[0], t.b[1], t.b[2], 0.f);
			tris[(tri_base + k) * 3 + 2] = make_float4(t.c[0], t.c[1], t.c[2], 0.f);
		}
		md[mi].node_root = node_base;
		md[mi].tri_first = (uint32_t)tri_base;
		std::memcpy(md[mi].bb, m.bb, sizeof(m.bb));
		for (int d = 0; d < 3; ++d) { lo[d] = std::min(lo[d], m.bb[d]); hi[d] = std::max(hi[d], m.bb[3 + d]); }
		tri_base += m.tris.size();
	}
	for (int d = 0; d < 3; ++d) { ctx->scene_bb[d] = lo[d]; ctx->scene_bb[3 + d] = hi[d]; }
	ctx->d_nodes.ensure(nodes.size());
	ctx->d_tris.ensure(tris.size());
	ctx->d_meshes.ensure(md.size());
	HN_CUDA(cudaMemcpy(ctx->d_nodes.p, nodes.data(), nodes.size() * sizeof(Node4), cudaMemcpyHostToDevice));
	HN_CUDA(cudaMemcpy(ctx->d_tris.p, tris.data(), tris.size() * sizeof(float4), cudaMemcpyHostToDevice));
	HN_CUDA(cudaMemcpy(ctx->d_meshes.p, md.data(), md.size() * sizeof(MeshDev), cudaMemcpyHostToDevice));
	ctx->d_stack_overflow.ensure(1);
	HN_CUDA(cudaMemset(ctx->d_stack_overflow.p, 0, sizeof(uint32_t)));
	// load_scene :3183-3189: m_aabb = m_render_aabb = root box inflated by 4.0
	if (!ctx->opt.mesh_aabb_user) {
		for (int d = 0; d < 3; ++d) { ctx->opt.mesh_aabb[d] = lo[d] - 4.0f; ctx->opt.mesh_aabb[3 + d] = hi[d] + 4.0f; }
	}
	ctx->scene_built = true;
}

// load_scene (src/testbed_geometry_training.cu:3101-3210)
void load_scene_file(hn_context* ctx, const std::string& path) {
	Json j = JsonParser(read_text_file(path)).parse();
	if (!j.is_object() || !j.contains("geometry")) throw std::runtime_error("Geometry file must contain an array of geometry metadata.");
	const Json& geoms = j.at("geometry");
	std::string dir = ".";
	size_t p = path.find_last_of('/');
	if (p != std::string::npos) dir = path.substr(0, p);
	ctx->meshes.clear();
	for (const Json& g : geoms.arr) {
		std::string type = g.at("type").str;
		std::string mpath = g.at("path").str;
		const Json& c = g.at("center");
		if (c.arr.size() < 3) throw std::runtime_error("geometry 'center' needs 3 numbers");
		float center[3] = {(float)c.arr[0].num, (float)c.arr[1].num, (float)c.arr[2].num};
		// the shipped scene files carry the author's absolute paths; fall back to the file name next to the scene / in objs/
		auto exists = [](const std::string& f) { std::ifstream t(f); return (bool)t; };
		if (!exists(mpath)) {
			size_t q = mpath.find_last_of('/');
			std::string base = q == std::string::npos ? mpath : mpath.substr(q + 1);
			if (exists(dir + "/" + base)) mpath = dir + "/" + base;
			else if (exists(dir + "/objs/" + base)) mpath = dir + "/objs/" + base;
		}
		if (type == "Mesh") {
			load_mesh_file(ctx, mpath, center);
		} else if (type == "Nerf") {
			// load_snapshot(path) (src/testbed_geometry_training.cu:3172): the hybrid scene pins the NeRF to the unit cube
			// (src/testbed.cu:5303-5307). A missing file is tolerated: the NeRF may also enter through hn_set_params.
			if (exists(mpath)) {
				load_snapshot_file(ctx, mpath);
				for (int d = 0; d < 3; ++d) { ctx->opt.train_aabb[d] = 0.0f; ctx->opt.train_aabb[3 + d] = 1.0f; }
			}
		} else {
			throw std::runtime_error("Geometry type must be either 'Mesh' or 'Nerf'.");
		}
	}
	build_scene(ctx);
	// load_scene :3183-3189 runs AFTER the entries are loaded: m_render_aabb = m_aabb = mesh root box inflated by 4, identity
	// crop rotation - so in a Geometry-mode scene the NeRF rays enter at THAT box (not at the snapshot's render box) and the
	// DDA walks from there; the NeRF itself stays pinned to the unit cube (train_aabb above).
	if (!ctx->meshes.empty()) {
		std::memcpy(ctx->opt.render_aabb, ctx->opt.mesh_aabb, sizeof(ctx->opt.render_aabb));
		const float ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
		std::memcpy(ctx->opt.render_aabb_to_local, ident, sizeof(ident));
	}
}

} // namespace hn
