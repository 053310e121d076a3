// NeRF dataset ingestion: transforms.json + PNG images -> the in-memory training set of hn_train.cu.
//   load_nerf                      src/nerf_loader.cu:273-770 (frames, intrinsics, scale / offset / aabb_scale)
//   read_focal_length              src/nerf_loader.cu:243-271 (x_fov | fl_x | camera_angle_x, same for y)
//   NerfDataset::nerf_matrix_to_ngp include/neural-graphics-primitives/nerf_loader.h:101-120
//   read_rgba (Byte images)        include/neural-graphics-primitives/common_device.cuh:797-816 (sRGB bytes, straight alpha)
// The reference decodes images with stb_image; here a PNG reader (8-bit grey / grey+alpha / RGB / RGBA, non-interlaced) sits on
// the DEFLATE decoder the snapshot reader already has, and hn_jpeg.cpp reads baseline / sequential JPEG. One resolution for the whole set; focal length, principal point and lens
// (read_lens, :175-241: OpenCV / fisheye coefficients, f-theta polynomial, latlong / equirectangular) per frame over the file-level
// values, as the reference layers them (:678-696). Depth / ray files, rolling-shutter frame pairs, EXR and JPEG are not on this path.
#include "hn_context.h"
#include "hn_json.h"
#include <cmath>
#include <cstring>
#include <fstream>

namespace hn {

namespace {

uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

struct Image8 { int w = 0, h = 0; std::vector<uint8_t> rgba; };

// PNG (ISO/IEC 15948): signature, IHDR, concatenated IDAT = one zlib stream, per-row filter byte (None/Sub/Up/Average/Paeth)
Image8 decode_png(const std::string& path) {
	const std::string f = read_text_file(path);
	static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
	if (f.size() < 8 + 25 || std::memcmp(f.data(), sig, 8) != 0) throw std::runtime_error("'" + path + "' is not a PNG file");
	const uint8_t* p = (const uint8_t*)f.data();
	size_t pos = 8;
	Image8 img;
	int bit_depth = 0, color_type = 0, interlace = 0;
	std::string idat;
	while (pos + 12 <= f.size()) {
		const uint32_t len = be32(p + pos);
		const std::string type(f.data() + pos + 4, 4);
		if (pos + 12 + len > f.size()) throw std::runtime_error("'" + path + "': truncated PNG chunk");
		const uint8_t* d = p + pos + 8;
		if (type == "IHDR") {
			if (len < 13) throw std::runtime_error("'" + path + "': bad IHDR");
			img.w = (int)be32(d); img.h = (int)be32(d + 4);
			bit_depth = d[8]; color_type = d[9]; interlace = d[12];
		} else if (type == "IDAT") {
			idat.append((const char*)d, len);
		} else if (type == "IEND") {
			break;
		}
		pos += 12 + len;
	}
	if (img.w <= 0 || img.h <= 0) throw std::runtime_error("'" + path + "': missing IHDR");
	if (bit_depth != 8 || interlace != 0 || !(color_type == 0 || color_type == 2 || color_type == 4 || color_type == 6))
		throw std::runtime_error("'" + path + "': only 8-bit non-interlaced grey / RGB (+ alpha) PNGs are supported");
	const int ch = color_type == 0 ? 1 : color_type == 4 ? 2 : color_type == 2 ? 3 : 4;
	const std::string raw = zlib_inflate(idat);
	const size_t stride = (size_t)img.w * ch;
	if (raw.size() < (stride + 1) * (size_t)img.h) throw std::runtime_error("'" + path + "': PNG pixel data too short");
	std::vector<uint8_t> cur(stride), prev(stride, 0);
	img.rgba.resize((size_t)img.w * img.h * 4);
	for (int y = 0; y < img.h; ++y) {
		const uint8_t* row = (const uint8_t*)raw.data() + (stride + 1) * (size_t)y;
		const int filter = row[0];
		for (size_t i = 0; i < stride; ++i) {
			const int a = i >= (size_t)ch ? cur[i - ch] : 0, b = prev[i], c = i >= (size_t)ch ? prev[i - ch] : 0;
			int pred = 0;
			switch (filter) {
				case 0: pred = 0; break;
				case 1: pred = a; break;
				case 2: pred = b; break;
				case 3: pred = (a + b) / 2; break;
				case 4: { const int pp = a + b - c, pa = std::abs(pp - a), pb = std::abs(pp - b), pc = std::abs(pp - c); pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
				default: throw std::runtime_error("'" + path + "': bad PNG filter type");
			}
			cur[i] = (uint8_t)(row[1 + i] + pred);
		}
		uint8_t* out = img.rgba.data() + (size_t)y * img.w * 4;
		for (int x = 0; x < img.w; ++x) {
			const uint8_t* s = cur.data() + (size_t)x * ch;
			if (ch <= 2) { out[4 * x] = out[4 * x + 1] = out[4 * x + 2] = s[0]; out[4 * x + 3] = ch == 2 ? s[1] : 255; }
			else { out[4 * x] = s[0]; out[4 * x + 1] = s[1]; out[4 * x + 2] = s[2]; out[4 * x + 3] = ch == 4 ? s[3] : 255; }
		}
		prev.swap(cur);
	}
	return img;
}

} // namespace
// 8-bit image file -> RGBA8 (stbi_load(path, ..., 4) of the reference's loader): PNG or JPEG, told apart by their signatures
void decode_image_file(const std::string& path, int& w, int& h, std::vector<uint8_t>& rgba) {
	const std::string f = read_text_file(path);
	if (f.size() >= 2 && (uint8_t)f[0] == 0xFF && (uint8_t)f[1] == 0xD8) { decode_jpeg(path, f, w, h, rgba); return; }
	Image8 img = decode_png(path);
	w = img.w; h = img.h; rgba.swap(img.rgba);
}
namespace {

float srgb_to_linear_host(float s) { return s <= 0.04045f ? s / 12.92f : std::pow((s + 0.055f) / 1.055f, 2.4f); }

std::string dirname_of(const std::string& path) {
	const size_t p = path.find_last_of('/');
	return p == std::string::npos ? std::string(".") : path.substr(0, p);
}
bool has_extension(const std::string& s) {
	const size_t slash = s.find_last_of('/'), dot = s.find_last_of('.');
	return dot != std::string::npos && (slash == std::string::npos || dot > slash);
}
std::string lower_ext(const std::string& s) {
	const size_t dot = s.find_last_of('.');
	std::string e = dot == std::string::npos ? std::string() : s.substr(dot + 1);
	for (auto& c : e) c = (char)std::tolower((unsigned char)c);
	return e;
}
// fov_to_focal_length (common_device.cuh): degrees -> pixels
float fov_to_focal(int res, float deg) { return 0.5f * (float)res / std::tan(0.5f * deg * 3.14159265358979323846f / 180.f); }
float focal_from(const Json& j, int res, const char* axis) {
	const std::string a(axis);
	if (j.contains(a + "_fov")) return fov_to_focal(res, (float)j.at(a + "_fov").num);
	if (j.contains("fl_" + a)) return (float)j.at("fl_" + a).num;
	if (j.contains("camera_angle_" + a)) return fov_to_focal(res, (float)j.at("camera_angle_" + a).num * 180.f / 3.14159265358979323846f);
	return 0.f;
}

// read_focal_length (src/nerf_loader.cu:243-271): x_fov | fl_x | camera_angle_x (and y); one axis fills both
bool read_focal(const Json& j, int w, int h, float focal[2]) {
	const float fx = focal_from(j, w, "x"), fy = focal_from(j, h, "y");
	if (fx != 0.f) { focal[0] = focal[1] = fx; if (fy != 0.f) focal[1] = fy; return true; }
	if (fy != 0.f) { focal[0] = focal[1] = fy; return true; }
	return false;
}
// read_lens (src/nerf_loader.cu:175-241): OpenCV k1 k2 p1 p2 (k3 k4 for is_fisheye), principal point cx / w, cy / h, f-theta polynomial,
// latlong / equirectangular flags. A json object without lens keys leaves what the outer level set.
void read_lens(const Json& j, hn::ImageMeta& m) {
	int mode = 0;
	const Json* fe = j.find("is_fisheye");
	const int opencv_mode = (fe && fe->type == Json::Bool && fe->b) ? 4 : 1;
	auto opencv = [&](const char* name, int idx) {
		if (const Json* v = j.find(name)) { m.lens_params[idx] = (float)v->num; if (m.lens_params[idx] != 0.f) mode = opencv_mode; }
	};
	opencv("k1", 0); opencv("k2", 1); opencv("k3", 2); opencv("k4", 3);
	opencv("p1", 2); opencv("p2", 3);
	if (j.contains("cx")) { if (!j.contains("w")) throw std::runtime_error("transforms.json: cx needs w beside it"); m.principal[0] = (float)j.at("cx").num / (float)j.at("w").num; }
	if (j.contains("cy")) { if (!j.contains("h")) throw std::runtime_error("transforms.json: cy needs h beside it"); m.principal[1] = (float)j.at("cy").num / (float)j.at("h").num; }
	if (j.contains("ftheta_p0")) {
		for (int k = 0; k < 5; ++k) m.lens_params[k] = (float)j.at(std::string("ftheta_p") + char('0' + k)).num;
		m.lens_params[5] = (float)j.at("w").num; m.lens_params[6] = (float)j.at("h").num;
		mode = 2;
	}
	if (j.contains("latlong")) mode = 3;
	if (j.contains("equirectangular")) mode = 5;
	if (mode != 0) m.lens_mode = mode; // an outer distortion mode is not overridden by nothing
}

} // namespace

void load_training_data_file(hn_context* ctx, const std::string& path_in) {
	std::string path = path_in;
	{ // a directory stands for its transforms.json (Testbed::load_training_data walks the folder)
		std::ifstream probe(path + "/transforms.json");
		if (probe.good()) path += "/transforms.json";
	}
	const Json root = JsonParser(read_text_file(path)).parse();
	if (!root.is_object() || !root.contains("frames") || !root.at("frames").is_array() || root.at("frames").arr.empty())
		throw std::runtime_error("No training images were found for NeRF training!"); // src/nerf_loader.cu:409-411
	const std::string base = dirname_of(path);
	const auto& frames = root.at("frames").arr;
	float scale = (float)root.value("scale", 0.33), offset[3] = {0.5f, 0.5f, 0.5f}; // NERF_SCALE, nerf_loader.cu:402-403,470-502
	if (const Json* o = root.find("offset")) {
		if (o->is_array() && o->arr.size() == 3) for (int k = 0; k < 3; ++k) offset[k] = (float)o->arr[k].num;
		else if (o->type == Json::Number) offset[0] = offset[1] = offset[2] = (float)o->num;
	}
	hn::ImageMeta global_meta = {};
	global_meta.principal[0] = global_meta.principal[1] = 0.5f;
	read_lens(root, global_meta);
	std::vector<hn::ImageMeta> metas(frames.size(), global_meta);
	std::vector<float> pixels, cams;
	int w = 0, h = 0;
	for (size_t i = 0; i < frames.size(); ++i) {
		const Json& fr = frames[i];
		if (!fr.contains("file_path") || !fr.contains("transform_matrix")) throw std::runtime_error("transforms.json: frame without file_path / transform_matrix");
		std::string fp = fr.at("file_path").str;
		if (!has_extension(fp)) fp += ".png"; // nerf_loader.cu:560-563
		const std::string ext = lower_ext(fp);
		if (ext != "png" && ext != "jpg" && ext != "jpeg") throw std::runtime_error("image '" + fp + "': only PNG and JPEG images are supported on this path (no EXR / binary float images)");
		Image8 img;
		decode_image_file(fp[0] == '/' ? fp : base + "/" + fp, img.w, img.h, img.rgba);
		// per-frame intrinsics and lens over the file-level ones (nerf_loader.cu:678-696)
		bool got_fl = read_focal(root, img.w, img.h, metas[i].focal);
		got_fl |= read_focal(fr, img.w, img.h, metas[i].focal);
		if (!got_fl) throw std::runtime_error("Couldn't read fov."); // nerf_loader.cu:680-682
		read_lens(fr, metas[i]);
		if (i == 0) { w = img.w; h = img.h; pixels.resize(frames.size() * (size_t)w * h * 4); cams.resize(frames.size() * 12); }
		else if (img.w != w || img.h != h) throw std::runtime_error("training images of different resolutions are not supported on this path");
		float* dst = pixels.data() + i * (size_t)w * h * 4;
		for (size_t k = 0; k < (size_t)w * h; ++k) { // read_rgba Byte: sRGB -> linear, alpha kept straight here (hn_train multiplies)
			const uint8_t* s = img.rgba.data() + 4 * k;
			dst[4 * k] = srgb_to_linear_host((float)s[0] * (1.0f / 255.0f)); dst[4 * k + 1] = srgb_to_linear_host((float)s[1] * (1.0f / 255.0f));
			dst[4 * k + 2] = srgb_to_linear_host((float)s[2] * (1.0f / 255.0f)); dst[4 * k + 3] = (float)s[3] * (1.0f / 255.0f);
		}
		// transform_matrix rows -> mat4x3 columns (nerf_loader.cu:700-706), then nerf_matrix_to_ngp
		const Json& m = fr.at("transform_matrix");
		if (!m.is_array() || m.arr.size() < 3) throw std::runtime_error("transforms.json: transform_matrix must have at least 3 rows");
		float col[4][3];
		for (int r = 0; r < 3; ++r) {
			if (!m.arr[r].is_array() || m.arr[r].arr.size() < 4) throw std::runtime_error("transforms.json: transform_matrix rows need 4 entries");
			for (int c = 0; c < 4; ++c) col[c][r] = (float)m.arr[r].arr[c].num;
		}
		for (int r = 0; r < 3; ++r) { col[1][r] *= -1.f; col[2][r] *= -1.f; col[3][r] = col[3][r] * scale + offset[r]; }
		float* cam = cams.data() + 12 * i;
		for (int c = 0; c < 4; ++c) { cam[3 * c + 0] = col[c][1]; cam[3 * c + 1] = col[c][2]; cam[3 * c + 2] = col[c][0]; } // cycle axes xyz <- yzx
	}
	const float focal[2] = {metas[0].focal[0], metas[0].focal[1]}, principal[2] = {metas[0].principal[0], metas[0].principal[1]};
	// aabb_scale drives the boxes, cascades and the cone angle exactly as hn_set_aabb_scale does (load_nerf_post, testbed_nerf.cu:2718-2745)
	const int aabb_scale = (int)root.value("aabb_scale", 1.0);
	if (aabb_scale < 1 || aabb_scale > 128 || (aabb_scale & (aabb_scale - 1))) throw std::runtime_error("transforms.json: aabb_scale must be a power of two in [1, 128]");
	if (aabb_scale != ctx->aabb_scale && ctx->has_net) {
		ctx->aabb_scale = aabb_scale;
		reapply_network_config(ctx);
		for (int d = 0; d < 3; ++d) {
			ctx->opt.train_aabb[d] = ctx->opt.render_aabb[d] = 0.5f - 0.5f * aabb_scale;
			ctx->opt.train_aabb[3 + d] = ctx->opt.render_aabb[3 + d] = 0.5f + 0.5f * aabb_scale;
		}
	}
	train_set_dataset(ctx, pixels.data(), (uint32_t)frames.size(), w, h, cams.data(), focal, principal);
	{
		const size_t n = frames.size();
		std::vector<float> f(2 * n), pp(2 * n), lp(7 * n);
		std::vector<int32_t> lm(n);
		for (size_t i = 0; i < n; ++i) {
			f[2 * i] = metas[i].focal[0]; f[2 * i + 1] = metas[i].focal[1]; pp[2 * i] = metas[i].principal[0]; pp[2 * i + 1] = metas[i].principal[1];
			lm[i] = metas[i].lens_mode;
			for (int k = 0; k < 7; ++k) lp[7 * i + k] = metas[i].lens_params[k];
		}
		train_set_image_metadata(ctx, (uint32_t)n, f.data(), pp.data(), lm.data(), lp.data());
	}
	ctx->train.srgb_space = 1; // byte images live in sRGB: the target is blended there (compute_loss_kernel_train_nerf :1031-1037)
}

} // namespace hn
