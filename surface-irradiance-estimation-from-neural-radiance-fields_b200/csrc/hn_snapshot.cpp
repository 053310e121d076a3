// Snapshots: the reference's `.ingp` / `.msgpack` files (Testbed::save_snapshot / load_snapshot, src/testbed.cu:5219-5488):
// the network config JSON plus a "snapshot" object, serialised as MessagePack (nlohmann::json::to_msgpack), and for `.ingp`
// wrapped in a gzip member by zstr (read: gzip, zlib or plain, auto-detected like zstr::istream). This is how a trained NeRF enters a hybrid scene (load_scene, type "Nerf",
// src/testbed_geometry_training.cu:3172). Self-contained: a DEFLATE decoder (RFC 1950/1951), a MessagePack reader/writer.
#include "hn_context.h"
#include "hn_json.h"
#include <cstring>
#include <cmath>
#include <fstream>

namespace hn {

// ---------------------------------------------------------------- zlib (RFC 1950) / DEFLATE (RFC 1951)
namespace {

struct BitReader {
	const uint8_t* p; size_t n, pos = 0; uint32_t buf = 0; int cnt = 0;
	BitReader(const uint8_t* d, size_t len) : p(d), n(len) {}
	uint32_t bits(int need) {
		while (cnt < need) {
			if (pos >= n) throw std::runtime_error("snapshot: truncated zlib stream");
			buf |= (uint32_t)p[pos++] << cnt; cnt += 8;
		}
		uint32_t v = need ? (buf & ((1u << need) - 1u)) : 0u;
		buf >>= need; cnt -= need;
		return v;
	}
	void align() { buf = 0; cnt = 0; }
};

struct Huffman {
	uint16_t count[16]; uint16_t symbol[288];
	void build(const uint8_t* lengths, int n) {
		std::memset(count, 0, sizeof(count));
		for (int i = 0; i < n; ++i) count[lengths[i]]++;
		count[0] = 0;
		uint16_t offs[16]; offs[1] = 0;
		for (int l = 1; l < 15; ++l) offs[l + 1] = offs[l] + count[l];
		for (int i = 0; i < n; ++i) if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
	}
	int decode(BitReader& br) const { // canonical code, one bit at a time
		int code = 0, first = 0, index = 0;
		for (int len = 1; len <= 15; ++len) {
			code |= (int)br.bits(1);
			int c = count[len];
			if (code - c < first) return symbol[index + (code - first)];
			index += c; first += c; first <<= 1; code <<= 1;
		}
		throw std::runtime_error("snapshot: bad Huffman code");
	}
};

const uint16_t LEN_BASE[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint16_t LEN_EXTRA[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t DIST_BASE[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint16_t DIST_EXTRA[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

void inflate_codes(BitReader& br, const Huffman& lit, const Huffman& dist, std::string& out) {
	for (;;) {
		int sym = lit.decode(br);
		if (sym < 256) { out.push_back((char)sym); continue; }
		if (sym == 256) return;
		sym -= 257;
		if (sym >= 29) throw std::runtime_error("snapshot: bad length symbol");
		size_t len = LEN_BASE[sym] + br.bits(LEN_EXTRA[sym]);
		int ds = dist.decode(br);
		if (ds >= 30) throw std::runtime_error("snapshot: bad distance symbol");
		size_t d = DIST_BASE[ds] + br.bits(DIST_EXTRA[ds]);
		if (d > out.size()) throw std::runtime_error("snapshot: distance too far back");
		size_t from = out.size() - d;
		for (size_t i = 0; i < len; ++i) out.push_back(out[from + i]);
	}
}

uint32_t adler32(const std::string& s) {
	uint32_t a = 1, b = 0;
	size_t i = 0;
	while (i < s.size()) {
		size_t n = std::min<size_t>(5552, s.size() - i);
		for (size_t k = 0; k < n; ++k) { a += (uint8_t)s[i + k]; b += a; }
		a %= 65521u; b %= 65521u; i += n;
	}
	return (b << 16) | a;
}

} // namespace

// raw DEFLATE body (RFC 1951) starting at br; leaves br byte-aligned just behind the last block
static void deflate_body(BitReader& br, std::string& out) {
	bool last;
	do {
		last = br.bits(1) != 0;
		uint32_t type = br.bits(2);
		if (type == 0) {
			br.align();
			if (br.pos + 4 > br.n) throw std::runtime_error("snapshot: truncated stored block");
			uint32_t len = br.p[br.pos] | (br.p[br.pos + 1] << 8), nlen = br.p[br.pos + 2] | (br.p[br.pos + 3] << 8);
			br.pos += 4;
			if ((len ^ 0xffffu) != nlen || br.pos + len > br.n) throw std::runtime_error("snapshot: bad stored block");
			out.append((const char*)br.p + br.pos, len);
			br.pos += len;
		} else if (type == 1) {
			uint8_t l[288];
			for (int i = 0; i < 144; ++i) l[i] = 8;
			for (int i = 144; i < 256; ++i) l[i] = 9;
			for (int i = 256; i < 280; ++i) l[i] = 7;
			for (int i = 280; i < 288; ++i) l[i] = 8;
			Huffman lit, dist;
			lit.build(l, 288);
			uint8_t dl[30];
			for (int i = 0; i < 30; ++i) dl[i] = 5;
			dist.build(dl, 30);
			inflate_codes(br, lit, dist, out);
		} else if (type == 2) {
			int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
			static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
			uint8_t cl[19] = {0};
			for (int i = 0; i < ncode; ++i) cl[order[i]] = (uint8_t)br.bits(3);
			Huffman clh;
			clh.build(cl, 19);
			uint8_t lengths[320] = {0};
			int idx = 0;
			while (idx < nlen + ndist) {
				int sym = clh.decode(br);
				if (sym < 16) { lengths[idx++] = (uint8_t)sym; continue; }
				int rep; uint8_t val = 0;
				if (sym == 16) { if (!idx) throw std::runtime_error("snapshot: bad code lengths"); val = lengths[idx - 1]; rep = 3 + (int)br.bits(2); }
				else if (sym == 17) rep = 3 + (int)br.bits(3);
				else rep = 11 + (int)br.bits(7);
				if (idx + rep > nlen + ndist) throw std::runtime_error("snapshot: bad code lengths");
				while (rep--) lengths[idx++] = val;
			}
			Huffman lit, dist;
			lit.build(lengths, nlen);
			dist.build(lengths + nlen, ndist);
			inflate_codes(br, lit, dist, out);
		} else {
			throw std::runtime_error("snapshot: bad DEFLATE block type");
		}
	} while (!last);
	// bytes already pulled into the bit buffer but not consumed belong to the trailer
	br.pos -= (size_t)(br.cnt / 8);
	br.align();
}

static bool is_zlib_header(const uint8_t* d, size_t n) { return n >= 2 && (d[0] & 0x0f) == 8 && ((d[0] << 8) | d[1]) % 31 == 0 && !(d[1] & 0x20); }

std::string zlib_inflate(const std::string& in) {
	if (in.size() < 6) throw std::runtime_error("snapshot: zlib stream too short");
	const uint8_t* d = (const uint8_t*)in.data();
	if (!is_zlib_header(d, in.size())) throw std::runtime_error("snapshot: not a zlib stream");
	BitReader br(d + 2, in.size() - 2);
	std::string out;
	out.reserve(in.size() * 2);
	deflate_body(br, out);
	if (br.pos + 4 <= br.n) {
		uint32_t want = ((uint32_t)br.p[br.pos] << 24) | ((uint32_t)br.p[br.pos + 1] << 16) | ((uint32_t)br.p[br.pos + 2] << 8) | br.p[br.pos + 3];
		if (want != adler32(out)) throw std::runtime_error("snapshot: zlib checksum mismatch");
	}
	return out;
}

static uint32_t crc32_of(const std::string& s) {
	static uint32_t table[256];
	static bool init = false;
	if (!init) {
		for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i; for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; }
		init = true;
	}
	uint32_t c = 0xFFFFFFFFu;
	for (unsigned char ch : s) c = table[(c ^ ch) & 0xFFu] ^ (c >> 8);
	return c ^ 0xFFFFFFFFu;
}

// gzip member (RFC 1952): what the reference WRITES - Testbed::save_snapshot streams through zstr::ostream, whose
// deflateInit2(..., 15 + 16, ...) selects the gzip container even at Z_NO_COMPRESSION (dependencies/zstr/src/zstr.hpp:107).
std::string gzip_inflate(const std::string& in) {
	const uint8_t* d = (const uint8_t*)in.data();
	const size_t n = in.size();
	if (n < 18 || d[0] != 0x1f || d[1] != 0x8b) throw std::runtime_error("snapshot: not a gzip stream");
	if (d[2] != 8) throw std::runtime_error("snapshot: gzip compression method is not DEFLATE");
	const uint8_t flg = d[3];
	size_t pos = 10;
	auto need = [&](size_t k) { if (pos + k > n) throw std::runtime_error("snapshot: truncated gzip header"); };
	if (flg & 4) { need(2); size_t xlen = d[pos] | (d[pos + 1] << 8); pos += 2; need(xlen); pos += xlen; } // FEXTRA
	if (flg & 8) { while (true) { need(1); if (!d[pos++]) break; } }                                     // FNAME
	if (flg & 16) { while (true) { need(1); if (!d[pos++]) break; } }                                    // FCOMMENT
	if (flg & 2) { need(2); pos += 2; }                                                                  // FHCRC
	BitReader br(d + pos, n - pos);
	std::string out;
	out.reserve(in.size() * 2);
	deflate_body(br, out);
	if (br.pos + 8 > br.n) throw std::runtime_error("snapshot: truncated gzip trailer");
	const uint8_t* t = br.p + br.pos;
	const uint32_t crc = t[0] | (t[1] << 8) | (t[2] << 16) | ((uint32_t)t[3] << 24), isize = t[4] | (t[5] << 8) | (t[6] << 16) | ((uint32_t)t[7] << 24);
	if (crc != crc32_of(out)) throw std::runtime_error("snapshot: gzip CRC32 mismatch");
	if (isize != (uint32_t)out.size()) throw std::runtime_error("snapshot: gzip length mismatch");
	return out;
}

// zstr::istream auto-detection (zstr.hpp:240-262): gzip magic 1f 8b, a valid zlib header, else the bytes pass through as they are
std::string snapshot_unwrap(const std::string& in) {
	const uint8_t* d = (const uint8_t*)in.data();
	if (in.size() >= 2 && d[0] == 0x1f && d[1] == 0x8b) return gzip_inflate(in);
	if (is_zlib_header(d, in.size())) return zlib_inflate(in);
	return in;
}

// gzip member of stored (uncompressed) blocks: the container the reference writes, trivially produced
std::string gzip_store(const std::string& in) {
	std::string out;
	const unsigned char hdr[10] = {0x1f, 0x8b, 8, 0, 0, 0, 0, 0, 0, 3}; // no flags, no mtime, OS = unix
	out.append((const char*)hdr, 10);
	size_t i = 0;
	do {
		size_t k = std::min<size_t>(65535, in.size() - i);
		bool last = i + k >= in.size();
		out.push_back(last ? 1 : 0);
		out.push_back((char)(k & 0xff)); out.push_back((char)(k >> 8));
		out.push_back((char)(~k & 0xff)); out.push_back((char)((~k >> 8) & 0xff));
		out.append(in, i, k);
		i += k;
	} while (i < in.size());
	const uint32_t c = crc32_of(in), sz = (uint32_t)in.size();
	for (int sft = 0; sft < 32; sft += 8) out.push_back((char)((c >> sft) & 0xff));
	for (int sft = 0; sft < 32; sft += 8) out.push_back((char)((sz >> sft) & 0xff));
	return out;
}

// zlib stream of stored (uncompressed) blocks: valid input for zstr / zlib, trivially produced
std::string zlib_store(const std::string& in) {
	std::string out;
	out.push_back((char)0x78); out.push_back((char)0x01);
	size_t i = 0;
	do {
		size_t n = std::min<size_t>(65535, in.size() - i);
		bool last = i + n >= in.size();
		out.push_back(last ? 1 : 0);
		out.push_back((char)(n & 0xff)); out.push_back((char)(n >> 8));
		out.push_back((char)(~n & 0xff)); out.push_back((char)((~n >> 8) & 0xff));
		out.append(in, i, n);
		i += n;
	} while (i < in.size());
	uint32_t a = adler32(in);
	for (int s = 24; s >= 0; s -= 8) out.push_back((char)((a >> s) & 0xff));
	return out;
}

// ---------------------------------------------------------------- MessagePack <-> Json
namespace {

struct MsgReader {
	const uint8_t* p; size_t n, pos = 0;
	uint8_t u8() { if (pos >= n) throw std::runtime_error("snapshot: truncated msgpack"); return p[pos++]; }
	uint64_t be(int bytes) { uint64_t v = 0; for (int i = 0; i < bytes; ++i) v = (v << 8) | u8(); return v; }
	std::string raw(size_t len) { if (pos + len > n) throw std::runtime_error("snapshot: truncated msgpack"); std::string r((const char*)p + pos, len); pos += len; return r; }
	Json num(double v) { Json j; j.type = Json::Number; j.num = v; return j; }
	Json str(size_t len) { Json j; j.type = Json::String; j.str = raw(len); return j; }
	Json bin(size_t len) { Json j; j.type = Json::Binary; j.str = raw(len); return j; }
	Json arr(size_t len) { Json j; j.type = Json::Array; j.arr.reserve(len); for (size_t i = 0; i < len; ++i) j.arr.push_back(value()); return j; }
	Json map(size_t len) {
		Json j; j.type = Json::Object;
		for (size_t i = 0; i < len; ++i) { Json k = value(); if (k.type != Json::String) throw std::runtime_error("snapshot: non-string map key"); j.obj.emplace_back(k.str, value()); }
		return j;
	}
	Json value() {
		const uint8_t t = u8();
		if (t <= 0x7f) return num(t);
		if (t <= 0x8f) return map(t & 0x0f);
		if (t <= 0x9f) return arr(t & 0x0f);
		if (t <= 0xbf) return str(t & 0x1f);
		if (t >= 0xe0) return num((int8_t)t);
		switch (t) {
			case 0xc0: return Json();
			case 0xc2: { Json j; j.type = Json::Bool; j.b = false; return j; }
			case 0xc3: { Json j; j.type = Json::Bool; j.b = true; return j; }
			case 0xc4: return bin(be(1));
			case 0xc5: return bin(be(2));
			case 0xc6: return bin(be(4));
			case 0xc7: { size_t l = be(1); u8(); return bin(l); }
			case 0xc8: { size_t l = be(2); u8(); return bin(l); }
			case 0xc9: { size_t l = be(4); u8(); return bin(l); }
			case 0xca: { uint32_t v = (uint32_t)be(4); float f; std::memcpy(&f, &v, 4); return num(f); }
			case 0xcb: { uint64_t v = be(8); double d; std::memcpy(&d, &v, 8); return num(d); }
			case 0xcc: return num((double)be(1));
			case 0xcd: return num((double)be(2));
			case 0xce: return num((double)be(4));
			case 0xcf: return num((double)be(8));
			case 0xd0: return num((int8_t)be(1));
			case 0xd1: return num((int16_t)be(2));
			case 0xd2: return num((int32_t)be(4));
			case 0xd3: return num((double)(int64_t)be(8));
			case 0xd4: u8(); return bin(1);
			case 0xd5: u8(); return bin(2);
			case 0xd6: u8(); return bin(4);
			case 0xd7: u8(); return bin(8);
			case 0xd8: u8(); return bin(16);
			case 0xd9: return str(be(1));
			case 0xda: return str(be(2));
			case 0xdb: return str(be(4));
			case 0xdc: return arr(be(2));
			case 0xdd: return arr(be(4));
			case 0xde: return map(be(2));
			case 0xdf: return map(be(4));
		}
		throw std::runtime_error("snapshot: unknown msgpack type");
	}
};

void put_be(std::string& o, uint64_t v, int bytes) { for (int i = bytes - 1; i >= 0; --i) o.push_back((char)((v >> (8 * i)) & 0xff)); }
void msg_write(std::string& o, const Json& j) {
	switch (j.type) {
		case Json::Null: o.push_back((char)0xc0); break;
		case Json::Bool: o.push_back((char)(j.b ? 0xc3 : 0xc2)); break;
		case Json::Number: {
			const double v = j.num;
			if (v == std::floor(v) && std::fabs(v) < 9e15) { // integers in their smallest encoding, as nlohmann does
				if (v >= 0) {
					const uint64_t u = (uint64_t)v;
					if (u <= 0x7f) o.push_back((char)u);
					else if (u <= 0xff) { o.push_back((char)0xcc); put_be(o, u, 1); }
					else if (u <= 0xffff) { o.push_back((char)0xcd); put_be(o, u, 2); }
					else if (u <= 0xffffffffull) { o.push_back((char)0xce); put_be(o, u, 4); }
					else { o.push_back((char)0xcf); put_be(o, u, 8); }
				} else {
					const int64_t i = (int64_t)v;
					if (i >= -32) o.push_back((char)(int8_t)i);
					else if (i >= -128) { o.push_back((char)0xd0); put_be(o, (uint64_t)i, 1); }
					else if (i >= -32768) { o.push_back((char)0xd1); put_be(o, (uint64_t)i, 2); }
					else if (i >= -2147483648ll) { o.push_back((char)0xd2); put_be(o, (uint64_t)i, 4); }
					else { o.push_back((char)0xd3); put_be(o, (uint64_t)i, 8); }
				}
			} else if ((double)(float)v == v) { float f = (float)v; uint32_t u; std::memcpy(&u, &f, 4); o.push_back((char)0xca); put_be(o, u, 4); }
			else { uint64_t u; std::memcpy(&u, &v, 8); o.push_back((char)0xcb); put_be(o, u, 8); }
			break;
		}
		case Json::String:
			if (j.str.size() <= 31) o.push_back((char)(0xa0 | j.str.size()));
			else if (j.str.size() <= 0xff) { o.push_back((char)0xd9); put_be(o, j.str.size(), 1); }
			else if (j.str.size() <= 0xffff) { o.push_back((char)0xda); put_be(o, j.str.size(), 2); }
			else { o.push_back((char)0xdb); put_be(o, j.str.size(), 4); }
			o += j.str;
			break;
		case Json::Binary:
			if (j.str.size() <= 0xff) { o.push_back((char)0xc4); put_be(o, j.str.size(), 1); }
			else if (j.str.size() <= 0xffff) { o.push_back((char)0xc5); put_be(o, j.str.size(), 2); }
			else { o.push_back((char)0xc6); put_be(o, j.str.size(), 4); }
			o += j.str;
			break;
		case Json::Array:
			if (j.arr.size() <= 15) o.push_back((char)(0x90 | j.arr.size()));
			else if (j.arr.size() <= 0xffff) { o.push_back((char)0xdc); put_be(o, j.arr.size(), 2); }
			else { o.push_back((char)0xdd); put_be(o, j.arr.size(), 4); }
			for (auto& e : j.arr) msg_write(o, e);
			break;
		case Json::Object:
			if (j.obj.size() <= 15) o.push_back((char)(0x80 | j.obj.size()));
			else if (j.obj.size() <= 0xffff) { o.push_back((char)0xde); put_be(o, j.obj.size(), 2); }
			else { o.push_back((char)0xdf); put_be(o, j.obj.size(), 4); }
			for (auto& kv : j.obj) { Json k; k.type = Json::String; k.str = kv.first; msg_write(o, k); msg_write(o, kv.second); }
			break;
	}
}

Json jnum(double v) { Json j; j.type = Json::Number; j.num = v; return j; }
Json jstr(const std::string& v) { Json j; j.type = Json::String; j.str = v; return j; }
Json jvec(const float* v, int n) { Json j; j.type = Json::Array; for (int i = 0; i < n; ++i) j.arr.push_back(jnum(v[i])); return j; }
void read_vec(const Json& j, float* out, int n) {
	if (j.type != Json::Array || (int)j.arr.size() < n) throw std::runtime_error("snapshot: vector has the wrong size");
	for (int i = 0; i < n; ++i) out[i] = (float)j.arr[i].num;
}
void read_box(const Json& j, float* out6) { read_vec(j.at("min"), out6, 3); read_vec(j.at("max"), out6 + 3, 3); }
bool ends_with(const std::string& s, const char* suffix) { size_t l = std::strlen(suffix); return s.size() >= l && s.compare(s.size() - l, l, suffix) == 0; }

} // namespace

Json msgpack_parse(const std::string& bytes) { MsgReader r{(const uint8_t*)bytes.data(), bytes.size()}; return r.value(); }
std::string msgpack_write(const Json& j) { std::string o; msg_write(o, j); return o; }

// ---------------------------------------------------------------- load / save
// Testbed::load_snapshot (src/testbed.cu:5285-5463), for what the render path needs: network config, parameters
// (Trainer::deserialize, tcnn trainer.h:290-318: "__half" or "float" blob), the fp16 density grid -> occupancy bitfield
// (update_density_grid_mean_and_bitfield), aabb / render_aabb, camera, sun/up direction, exposure, background colour.
void load_snapshot_file(hn_context* ctx, const std::string& path) {
	std::string bytes = read_text_file(path);
	if (ends_with(path, ".ingp")) bytes = snapshot_unwrap(bytes); // zstr::ifstream (src/testbed.cu:255-262): gzip, zlib or plain
	Json cfg = msgpack_parse(bytes);
	bytes.clear(); bytes.shrink_to_fit();
	if (!cfg.is_object() || !cfg.contains("snapshot")) throw std::runtime_error("File '" + path + "' does not contain a snapshot.");
	const Json& snap = cfg.at("snapshot");
	if (snap.value("version", 0.0) < 1.0) throw std::runtime_error("Snapshot uses an old format and can not be loaded.");
	const std::string mode = snap.value("mode", std::string("nerf"));
	if (mode != "nerf" && mode != "Nerf" && mode != "geometry" && mode != "Geometry") throw std::runtime_error("snapshot mode '" + mode + "' is not on the hybrid NeRF + mesh path");
	if (snap.contains("density_grid_size") && (uint32_t)snap.at("density_grid_size").num != GRIDSIZE) throw std::runtime_error("Incompatible grid size.");
	int aabb_scale = 1;
	if (snap.contains("nerf") && snap.at("nerf").contains("aabb_scale")) aabb_scale = (int)snap.at("nerf").at("aabb_scale").num;
	if (aabb_scale < 1 || aabb_scale > 128 || (aabb_scale & (aabb_scale - 1))) throw std::runtime_error("snapshot: bad aabb_scale");
	ctx->aabb_scale = aabb_scale;
	apply_network_config(ctx, cfg); // reset_network(false)
	Options& o = ctx->opt;
	// load_nerf_post: boxes follow aabb_scale (src/testbed_nerf.cu:2722-2727); a hybrid (Geometry) scene pins the NeRF to the
	// unit cube (src/testbed.cu:5303-5307). Explicit boxes in the snapshot override.
	for (int d = 0; d < 3; ++d) {
		o.train_aabb[d] = o.render_aabb[d] = 0.5f - 0.5f * aabb_scale;
		o.train_aabb[3 + d] = o.render_aabb[3 + d] = 0.5f + 0.5f * aabb_scale;
	}
	if (snap.contains("aabb")) read_box(snap.at("aabb"), o.train_aabb);
	if (snap.contains("render_aabb")) read_box(snap.at("render_aabb"), o.render_aabb);
	if (snap.contains("render_aabb_to_local")) { // mat3 as 3 rows (tcnn vec_json.h:40-48) -> column-major option
		const Json& m = snap.at("render_aabb_to_local");
		if (m.type == Json::Array && m.arr.size() == 3) {
			for (int row = 0; row < 3; ++row) { float r3[3]; read_vec(m.arr[row], r3, 3); for (int col = 0; col < 3; ++col) o.render_aabb_to_local[3 * col + row] = r3[col]; }
		}
	}
	if (snap.contains("sun_dir")) read_vec(snap.at("sun_dir"), o.sun_dir, 3);
	if (snap.contains("up_dir")) read_vec(snap.at("up_dir"), o.up_dir, 3);
	o.exposure = (float)snap.value("exposure", (double)o.exposure);
	if (snap.contains("background_color")) read_vec(snap.at("background_color"), o.background_color, 4);
	if (snap.contains("camera")) {
		const Json& cam = snap.at("camera");
		if (cam.contains("matrix")) { // mat4x3 as 3 rows x 4 columns (tcnn vec_json.h:40-48) -> the ABI's 4 columns x vec3
			const Json& m = cam.at("matrix");
			if (m.type != Json::Array || m.arr.size() != 3) throw std::runtime_error("snapshot: camera matrix must have 3 rows");
			for (int row = 0; row < 3; ++row) {
				float r4[4];
				read_vec(m.arr[row], r4, 4);
				for (int col = 0; col < 4; ++col) ctx->camera[3 * col + row] = r4[col];
			}
		}
		o.fov_axis = (float)cam.value("fov_axis", (double)o.fov_axis);
		if (cam.contains("relative_focal_length")) read_vec(cam.at("relative_focal_length"), o.relative_focal_length, 2);
		if (cam.contains("screen_center")) read_vec(cam.at("screen_center"), o.screen_center, 2);
		o.zoom = (float)cam.value("zoom", (double)o.zoom);
	}
	// parameters
	const std::string ptype = snap.value("params_type", std::string("__half"));
	const Json& pb = snap.at("params_binary");
	if (pb.type != Json::Binary) throw std::runtime_error("snapshot: params_binary is not binary");
	if (ptype == "__half") upload_params(ctx, pb.str.data(), pb.str.size() / 2, 0);
	else if (ptype == "float") upload_params(ctx, pb.str.data(), pb.str.size() / 4, 1);
	else throw std::runtime_error("Trainer: snapshot parameters must be of type float of __half");
	// optimizer state (Trainer::deserialize -> Optimizer::deserialize): innermost Adam moments + step counters
	if (snap.contains("optimizer")) {
		const Json* oj = &snap.at("optimizer");
		while (oj->is_object() && oj->contains("nested")) oj = &oj->at("nested");
		if (oj->is_object() && oj->contains("first_moments_binary") && oj->contains("second_moments_binary")) {
			const Json& a1 = oj->at("first_moments_binary"); const Json& a2 = oj->at("second_moments_binary");
			if (a1.type != Json::Binary || a2.type != Json::Binary || a1.str.size() != a2.str.size()) throw std::runtime_error("snapshot: bad optimizer moments");
			const Json* ps = oj->find("param_steps_binary");
			if (ps && (ps->type != Json::Binary || ps->str.size() != a1.str.size())) ps = nullptr;
			train_import_state(ctx, (uint32_t)oj->value("current_step", snap.value("training_step", 0.0)), (const float*)a1.str.data(), (const float*)a2.str.data(),
				ps ? (const uint32_t*)ps->str.data() : nullptr, a1.str.size() / 4);
		}
	}
	// density grid (fp16, Morton order, 128^3 per cascade)
	if (snap.contains("density_grid_binary")) {
		const Json& gb = snap.at("density_grid_binary");
		if (gb.type != Json::Binary) throw std::runtime_error("snapshot: density_grid_binary is not binary");
		const size_t n = gb.str.size() / 2;
		if (n == (size_t)GRID_CELLS * (size_t)(o.max_cascade + 1)) {
			std::vector<float> grid(n);
			const __half* h = (const __half*)gb.str.data();
			for (size_t i = 0; i < n; ++i) grid[i] = __half2float(h[i]);
			upload_density_grid(ctx, grid.data(), (uint32_t)(o.max_cascade + 1));
		} else if (n != 0) {
			throw std::runtime_error("Incompatible number of grid cascades.");
		}
	}
}

// Testbed::save_snapshot (src/testbed.cu:5219-5283), render-relevant members; `.ingp` = zlib stream (stored blocks).
void save_snapshot_file(hn_context* ctx, const std::string& path) {
	if (!ctx->has_net || !ctx->has_params) throw std::runtime_error("nothing to save: load a network config and parameters first");
	Json cfg = msgpack_parse(ctx->config_blob);
	Json& snap = cfg.set("snapshot");
	snap.type = Json::Object;
	snap.set("version") = jnum(1);
	snap.set("mode") = jstr("nerf");
	snap.set("n_params") = jnum((double)ctx->net.n_params);
	snap.set("params_type") = jstr("__half");
	if (ctx->params_host.size() != ctx->net.n_params) { // stale after training steps: read the fp16 inference copy back
		ctx->params_host.resize(ctx->net.n_params);
		HN_CUDA(cudaMemcpy(ctx->params_host.data(), ctx->params.p, ctx->net.n_params * sizeof(__half), cudaMemcpyDeviceToHost));
	}
	Json pb; pb.type = Json::Binary; pb.str.assign((const char*)ctx->params_host.data(), ctx->params_host.size() * sizeof(__half));
	snap.set("params_binary") = std::move(pb);
	snap.set("density_grid_size") = jnum(GRIDSIZE);
	Json gb; gb.type = Json::Binary;
	{
		if (ctx->density_grid_host.empty() && ctx->density_grid.p && ctx->grid_cascades) { // refreshed on the device by training
			ctx->density_grid_host.resize((size_t)GRID_CELLS * ctx->grid_cascades);
			HN_CUDA(cudaMemcpy(ctx->density_grid_host.data(), ctx->density_grid.p, ctx->density_grid_host.size() * sizeof(float), cudaMemcpyDeviceToHost));
		}
		std::vector<__half> h(ctx->density_grid_host.size());
		for (size_t i = 0; i < h.size(); ++i) h[i] = __float2half_rn(ctx->density_grid_host[i]);
		gb.str.assign((const char*)h.data(), h.size() * sizeof(__half));
	}
	snap.set("density_grid_binary") = std::move(gb);
	const Options& o = ctx->opt;
	Json& nerf = snap.set("nerf");
	nerf.type = Json::Object;
	nerf.set("aabb_scale") = jnum(ctx->aabb_scale);
	Json& rgb = nerf.set("rgb");
	rgb.type = Json::Object;
	rgb.set("rays_per_batch") = jnum(ctx->train.ready ? ctx->train.rays_per_batch : 4096); rgb.set("measured_batch_size") = jnum(ctx->train.n_samples_last);
	rgb.set("measured_batch_size_before_compaction") = jnum(ctx->train.n_samples_last);
	snap.set("training_step") = jnum(ctx->train.ready ? ctx->train.step : 0);
	snap.set("loss") = jnum(ctx->train.loss_last);
	if (ctx->opt.snapshot_optimizer_state && ctx->train.ready) {
		// Trainer::serialize(true): the optimizer stack serialises itself outside-in (Ema > ExponentialDecay > Adam for
		// configs/nerf/base.json; tcnn ema.h:190-195, exponential_decay.h:136-142, adam.h:278-286)
		std::vector<float> m1, m2; std::vector<uint32_t> steps;
		train_export_state(ctx, m1, m2, steps);
		auto bin = [](const void* p, size_t bytes) { Json b; b.type = Json::Binary; b.str.assign((const char*)p, bytes); return b; };
		Json adam; adam.type = Json::Object;
		adam.set("current_step") = jnum(ctx->train.step);
		adam.set("base_learning_rate") = jnum(ctx->train_cfg.lr);
		adam.set("first_moments_binary") = bin(m1.data(), m1.size() * 4);
		adam.set("second_moments_binary") = bin(m2.data(), m2.size() * 4);
		adam.set("param_steps_binary") = bin(steps.data(), steps.size() * 4);
		Json cur = std::move(adam);
		// wrap inside-out following the config's nesting
		std::vector<std::string> chain;
		for (const Json* oj = cfg.find("optimizer"); oj && oj->is_object(); oj = oj->find("nested")) chain.push_back(oj->value("otype", std::string("Adam")));
		for (int i = (int)chain.size() - 2; i >= 0; --i) {
			Json w; w.type = Json::Object;
			std::string ot = chain[i];
			for (auto& ch : ot) ch = (char)std::tolower((unsigned char)ch);
			w.set("nested") = std::move(cur);
			if (ot == "ema") w.set("weights_ema_binary") = bin(ctx->params_host.data(), ctx->params_host.size() * sizeof(__half));
			else if (ot == "exponentialdecay") { w.set("learning_rate") = jnum(ctx->train_cfg.lr); w.set("learning_rate_factor") = jnum(ctx->train.lr_last > 0.f ? ctx->train.lr_last / ctx->train_cfg.lr : 1.0); }
			cur = std::move(w);
		}
		snap.set("optimizer") = std::move(cur);
	}
	auto box = [](const float* b) { Json j; j.type = Json::Object; j.set("min") = jvec(b, 3); j.set("max") = jvec(b + 3, 3); return j; };
	snap.set("aabb") = box(o.train_aabb);
	snap.set("render_aabb") = box(o.render_aabb);
	{
		Json m; m.type = Json::Array;
		for (int row = 0; row < 3; ++row) { float r3[3]; for (int col = 0; col < 3; ++col) r3[col] = o.render_aabb_to_local[3 * col + row]; m.arr.push_back(jvec(r3, 3)); }
		snap.set("render_aabb_to_local") = std::move(m);
	}
	snap.set("up_dir") = jvec(o.up_dir, 3);
	snap.set("sun_dir") = jvec(o.sun_dir, 3);
	snap.set("exposure") = jnum(o.exposure);
	snap.set("background_color") = jvec(o.background_color, 4);
	Json& cam = snap.set("camera");
	cam.type = Json::Object;
	Json m; m.type = Json::Array;
	for (int row = 0; row < 3; ++row) { float r4[4]; for (int col = 0; col < 4; ++col) r4[col] = ctx->camera[3 * col + row]; m.arr.push_back(jvec(r4, 4)); }
	cam.set("matrix") = std::move(m);
	cam.set("fov_axis") = jnum(o.fov_axis);
	cam.set("relative_focal_length") = jvec(o.relative_focal_length, 2);
	cam.set("screen_center") = jvec(o.screen_center, 2);
	cam.set("zoom") = jnum(o.zoom);
	std::string bytes = msgpack_write(cfg);
	if (ends_with(path, ".ingp")) bytes = gzip_store(bytes); // the gzip container zstr::ostream produces
	std::ofstream f(path, std::ios::binary);
	if (!f) throw std::runtime_error("cannot write '" + path + "'");
	f.write(bytes.data(), (std::streamsize)bytes.size());
}

} // namespace hn
