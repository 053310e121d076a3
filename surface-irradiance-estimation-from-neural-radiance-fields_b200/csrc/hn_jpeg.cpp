// JPEG reader of the dataset loader: the reference decodes training images with stb_image
// (src/nerf_loader.cu:575-590, stbi_load); real transforms.json sets are mostly .jpg. Baseline and extended sequential DCT JPEG
// (ITU-T T.81: SOF0 / SOF1, 8-bit, Huffman), grey or YCbCr (JFIF) or RGB (Adobe transform 0), any sampling factors up to 4, restart
// intervals. Chroma of h2v1 / h2v2 images is upsampled with the triangle filter libjpeg calls "fancy upsampling" (stb_image's
// resample_row_h_2 / hv_2 do the same), other ratios by replication. Progressive (SOF2), arithmetic-coded, lossless and 12-bit
// files raise an error. Not a copy of stb_image: decoded pixels agree with libjpeg-turbo (PIL) to +-3 of 255 (different IDCT
// arithmetic), which is what tests/test_images.py asserts.
#include "hn_context.h"
#include <cmath>
#include <cstring>

namespace hn {

namespace {

struct Huff {
	// canonical Huffman code (T.81 Annex C / F.2.2.3): maxcode / valptr / mincode per length, values in code order
	int mincode[17], maxcode[18], valptr[17];
	uint8_t vals[256];
	bool set = false;
	void build(const uint8_t counts[16], const uint8_t* symbols, int n) {
		std::memcpy(vals, symbols, (size_t)n);
		int code = 0, k = 0;
		for (int len = 1; len <= 16; ++len) {
			valptr[len] = k;
			mincode[len] = code;
			code += counts[len - 1];
			k += counts[len - 1];
			maxcode[len] = counts[len - 1] ? code - 1 : -1;
			code <<= 1;
		}
		maxcode[17] = 0x7fffffff;
		set = true;
	}
};

struct Component { int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0, pred = 0; int bw = 0, bh = 0; std::vector<uint8_t> plane; int pw = 0, ph = 0; };

struct BitReader {
	const uint8_t* p; const uint8_t* end;
	uint32_t acc = 0; int nbits = 0; bool hit_marker = false;
	BitReader(const uint8_t* b, const uint8_t* e) : p(b), end(e) {}
	void fill() {
		while (nbits <= 24) {
			int byte = 0;
			if (!hit_marker && p < end) {
				byte = *p;
				if (byte == 0xFF) {
					const int next = p + 1 < end ? p[1] : 0xD9;
					if (next == 0x00) p += 2;                 // stuffed zero
					else { hit_marker = true; byte = 0; }      // a marker: feed zeros until the caller consumes it
				} else ++p;
			}
			acc |= (uint32_t)byte << (24 - nbits);
			nbits += 8;
		}
	}
	int bit() { if (nbits < 1) fill(); const int b = (int)(acc >> 31); acc <<= 1; --nbits; return b; }
	int bits(int n) { if (!n) return 0; if (nbits < n) fill(); const int v = (int)(acc >> (32 - n)); acc <<= n; nbits -= n; return v; }
	void reset() { acc = 0; nbits = 0; hit_marker = false; }
};

int decode_symbol(BitReader& br, const Huff& h) {
	int code = 0;
	for (int len = 1; len <= 16; ++len) {
		code = (code << 1) | br.bit();
		if (h.maxcode[len] >= 0 && code <= h.maxcode[len] && code >= h.mincode[len]) return h.vals[h.valptr[len] + code - h.mincode[len]];
	}
	throw std::runtime_error("JPEG: bad Huffman code");
}
int extend(int v, int n) { return n && v < (1 << (n - 1)) ? v - (1 << n) + 1 : v; } // T.81 F.2.2.1

const uint8_t ZIGZAG[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
	35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// separable float inverse DCT (T.81 A.3.3), level shift + clamp
void idct_block(const float in[64], uint8_t* out, int stride) {
	static float C[8][8];
	static bool init = false;
	if (!init) {
		for (int x = 0; x < 8; ++x) for (int u = 0; u < 8; ++u) C[x][u] = (u == 0 ? std::sqrt(0.125f) : 0.5f) * std::cos((2 * x + 1) * u * 3.14159265358979323846f / 16.0f);
		init = true;
	}
	float tmp[64];
	for (int y = 0; y < 8; ++y) for (int x = 0; x < 8; ++x) { float s = 0; for (int u = 0; u < 8; ++u) s += C[x][u] * in[y * 8 + u]; tmp[y * 8 + x] = s; }
	for (int x = 0; x < 8; ++x) for (int y = 0; y < 8; ++y) {
		float s = 0;
		for (int v = 0; v < 8; ++v) s += C[y][v] * tmp[v * 8 + x];
		const int q = (int)std::lrintf(s + 128.0f);
		out[y * stride + x] = (uint8_t)std::min(255, std::max(0, q));
	}
}

} // namespace

void decode_jpeg(const std::string& path, const std::string& f, int& w_out, int& h_out, std::vector<uint8_t>& rgba) {
	const uint8_t* p = (const uint8_t*)f.data();
	const size_t n = f.size();
	auto bad = [&](const std::string& why) { return std::runtime_error("'" + path + "': " + why); };
	if (n < 4 || p[0] != 0xFF || p[1] != 0xD8) throw bad("not a JPEG file");
	uint16_t qt[4][64] = {};
	Huff hdc[4], hac[4];
	std::vector<Component> comp;
	int width = 0, height = 0, restart_interval = 0, adobe_transform = -1;
	bool have_sof = false, decoded = false;
	size_t pos = 2;
	while (pos + 4 <= n && !decoded) {
		if (p[pos] != 0xFF) { ++pos; continue; }
		const int marker = p[pos + 1];
		if (marker == 0xFF) { ++pos; continue; }
		if (marker == 0xD8 || (marker >= 0xD0 && marker <= 0xD7) || marker == 0x01) { pos += 2; continue; }
		if (marker == 0xD9) break;
		const size_t len = ((size_t)p[pos + 2] << 8) | p[pos + 3];
		if (len < 2 || pos + 2 + len > n) throw bad("truncated JPEG segment");
		const uint8_t* d = p + pos + 4;
		const size_t dl = len - 2;
		if (marker == 0xDB) { // DQT
			size_t i = 0;
			while (i < dl) {
				const int pq = d[i] >> 4, tq = d[i] & 15; ++i;
				if (tq > 3 || i + (pq ? 128 : 64) > dl) throw bad("bad quantisation table");
				for (int k = 0; k < 64; ++k) { qt[tq][ZIGZAG[k]] = pq ? (uint16_t)((d[i] << 8) | d[i + 1]) : d[i]; i += pq ? 2 : 1; }
			}
		} else if (marker == 0xC4) { // DHT
			size_t i = 0;
			while (i + 17 <= dl) {
				const int tc = d[i] >> 4, th = d[i] & 15; ++i;
				int total = 0;
				for (int k = 0; k < 16; ++k) total += d[i + k];
				if (th > 3 || tc > 1 || total > 256 || i + 16 + total > dl) throw bad("bad Huffman table");
				(tc ? hac[th] : hdc[th]).build(d + i, d + i + 16, total);
				i += 16 + total;
			}
		} else if (marker == 0xC0 || marker == 0xC1) { // SOF0 / SOF1
			if (dl < 6 || d[0] != 8) throw bad("only 8-bit JPEGs are supported");
			height = (d[1] << 8) | d[2]; width = (d[3] << 8) | d[4];
			const int nc = d[5];
			if (width <= 0 || height <= 0 || !(nc == 1 || nc == 3) || dl < (size_t)(6 + 3 * nc)) throw bad("unsupported JPEG frame (grey and 3-component images only)");
			comp.resize(nc);
			for (int c = 0; c < nc; ++c) {
				comp[c].id = d[6 + 3 * c]; comp[c].h = d[7 + 3 * c] >> 4; comp[c].v = d[7 + 3 * c] & 15; comp[c].tq = d[8 + 3 * c];
				if (comp[c].h < 1 || comp[c].h > 4 || comp[c].v < 1 || comp[c].v > 4 || comp[c].tq > 3) throw bad("bad JPEG sampling factors");
			}
			have_sof = true;
		} else if (marker == 0xC2) throw bad("progressive JPEGs are not supported (re-save as baseline)");
		else if (marker == 0xC3 || (marker >= 0xC5 && marker <= 0xCF && marker != 0xC8 && marker != 0xCC)) throw bad("lossless / hierarchical / arithmetic-coded JPEGs are not supported");
		else if (marker == 0xDD) { if (dl >= 2) restart_interval = (d[0] << 8) | d[1]; }
		else if (marker == 0xEE) { if (dl >= 12 && std::memcmp(d, "Adobe", 5) == 0) adobe_transform = d[11]; }
		else if (marker == 0xDA) { // SOS + entropy-coded data
			if (!have_sof) throw bad("scan before frame header");
			const int ns = d[0];
			if (ns != (int)comp.size() || dl < (size_t)(1 + 2 * ns + 3)) throw bad("non-interleaved multi-scan JPEGs are not supported");
			for (int k = 0; k < ns; ++k) {
				const int cid = d[1 + 2 * k];
				bool found = false;
				for (auto& c : comp) if (c.id == cid) { c.td = d[2 + 2 * k] >> 4; c.ta = d[2 + 2 * k] & 15; found = true; }
				if (!found) throw bad("scan names an unknown component");
			}
			int hmax = 1, vmax = 1;
			for (auto& c : comp) { hmax = std::max(hmax, c.h); vmax = std::max(vmax, c.v); }
			if (comp.size() == 1) { comp[0].h = comp[0].v = 1; hmax = vmax = 1; } // a single-component scan is never interleaved (A.2.2)
			const int mcu_w = 8 * hmax, mcu_h = 8 * vmax, mcus_x = (width + mcu_w - 1) / mcu_w, mcus_y = (height + mcu_h - 1) / mcu_h;
			for (auto& c : comp) {
				if (!hdc[c.td].set || !hac[c.ta].set) throw bad("scan uses a Huffman table that was not defined");
				c.pw = mcus_x * c.h * 8; c.ph = mcus_y * c.v * 8; c.plane.assign((size_t)c.pw * c.ph, 0); c.pred = 0;
			}
			BitReader br(p + pos + 2 + len, p + n);
			int until_restart = restart_interval;
			float block[64];
			for (int my = 0; my < mcus_y; ++my) for (int mx = 0; mx < mcus_x; ++mx) {
				if (restart_interval && until_restart == 0) { // RSTn: byte-align, skip the marker, reset the predictors
					br.reset();
					while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) ++br.p;
					if (br.p + 1 < br.end) br.p += 2;
					for (auto& c : comp) c.pred = 0;
					until_restart = restart_interval;
				}
				for (auto& c : comp) for (int by = 0; by < c.v; ++by) for (int bx = 0; bx < c.h; ++bx) {
					std::memset(block, 0, sizeof(block));
					const int t = decode_symbol(br, hdc[c.td]);
					if (t > 11) throw bad("bad DC category");
					c.pred += extend(br.bits(t), t);
					block[0] = (float)(c.pred * (int)qt[c.tq][0]);
					for (int k = 1; k < 64;) {
						const int rs = decode_symbol(br, hac[c.ta]), r = rs >> 4, s = rs & 15;
						if (s == 0) { if (r == 15) { k += 16; continue; } break; } // ZRL / EOB
						k += r;
						if (k > 63) throw bad("AC coefficient index out of range");
						const int zz = ZIGZAG[k];
						block[zz] = (float)(extend(br.bits(s), s) * (int)qt[c.tq][zz]);
						++k;
					}
					idct_block(block, c.plane.data() + (size_t)((my * c.v + by) * 8) * c.pw + (size_t)(mx * c.h + bx) * 8, c.pw);
				}
				if (restart_interval) --until_restart;
			}
			// ---- planes -> RGBA
			rgba.assign((size_t)width * height * 4, 255);
			std::vector<std::vector<uint8_t>> full(comp.size());
			for (size_t ci = 0; ci < comp.size(); ++ci) {
				Component& c = comp[ci];
				const int fx = hmax / c.h, fy = vmax / c.v;
				if (hmax % c.h || vmax % c.v) throw bad("fractional JPEG sampling ratios are not supported");
				std::vector<uint8_t>& o = full[ci];
				o.resize((size_t)width * height);
				// the component's own extent (T.81 A.1.1): ceil(X * Hi / Hmax) x ceil(Y * Vi / Vmax)
				const int cw = (width * c.h + hmax - 1) / hmax, chh = (height * c.v + vmax - 1) / vmax;
				auto at = [&](int x, int y) { x = std::min(std::max(x, 0), cw - 1); y = std::min(std::max(y, 0), chh - 1); return (int)c.plane[(size_t)y * c.pw + x]; };
				for (int y = 0; y < height; ++y) for (int x = 0; x < width; ++x) {
					int v;
					if (fx == 1 && fy == 1) v = at(x, y);
					else if (fx == 2 && fy <= 2) { // triangle filter: 3/4 nearer sample + 1/4 farther one, per axis
						const int sx = x >> 1, nx = (x & 1) ? sx + 1 : sx - 1;
						if (fy == 1) v = (3 * at(sx, y) + at(nx, y) + 2) >> 2;
						else {
							const int sy = y >> 1, ny = (y & 1) ? sy + 1 : sy - 1;
							const int near_row = 3 * at(sx, sy) + at(nx, sy), far_row = 3 * at(sx, ny) + at(nx, ny);
							v = (3 * near_row + far_row + 8) >> 4;
						}
					} else v = at(x / fx, y / fy);
					o[(size_t)y * width + x] = (uint8_t)v;
				}
			}
			for (size_t i = 0; i < (size_t)width * height; ++i) {
				uint8_t* o = rgba.data() + 4 * i;
				if (comp.size() == 1) { o[0] = o[1] = o[2] = full[0][i]; continue; }
				if (adobe_transform == 0) { o[0] = full[0][i]; o[1] = full[1][i]; o[2] = full[2][i]; continue; } // stored as RGB
				const float Y = (float)full[0][i], cb = (float)full[1][i] - 128.0f, cr = (float)full[2][i] - 128.0f; // JFIF YCbCr (BT.601 full range)
				const float r = Y + 1.402f * cr, g = Y - 0.344136f * cb - 0.714136f * cr, b = Y + 1.772f * cb;
				o[0] = (uint8_t)std::min(255, std::max(0, (int)std::lrintf(r)));
				o[1] = (uint8_t)std::min(255, std::max(0, (int)std::lrintf(g)));
				o[2] = (uint8_t)std::min(255, std::max(0, (int)std::lrintf(b)));
			}
			decoded = true;
		}
		pos += 2 + len;
	}
	if (!decoded) throw bad("no image data in JPEG file");
	w_out = width; h_out = height;
}

} // namespace hn
