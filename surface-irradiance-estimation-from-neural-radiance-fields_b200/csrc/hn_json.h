// Minimal JSON reader for the reference's on-disk formats: network configs (configs/nerf/*.json with "parent" merge,
// src/testbed.cu:86-97,223-317) and scene files (data/geometry/*.json). Comments are accepted, as the reference parses
// with nlohmann::json::parse(f, nullptr, true, /*ignore_comments=*/true).
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include <cstdlib>
#include <cctype>

namespace hn {

struct Json {
	enum Type { Null, Bool, Number, String, Array, Object, Binary } type = Null; // Binary: msgpack bin (snapshots), bytes in `str`
	bool b = false;
	double num = 0.0;
	std::string str;
	std::vector<Json> arr;
	std::vector<std::pair<std::string, Json>> obj; // insertion order kept

	bool is_object() const { return type == Object; }
	bool is_array() const { return type == Array; }
	bool contains(const std::string& k) const { return find(k) != nullptr; }
	const Json* find(const std::string& k) const {
		if (type != Object) return nullptr;
		for (auto& kv : obj) if (kv.first == k) return &kv.second;
		return nullptr;
	}
	Json* find(const std::string& k) {
		if (type != Object) return nullptr;
		for (auto& kv : obj) if (kv.first == k) return &kv.second;
		return nullptr;
	}
	const Json& at(const std::string& k) const {
		const Json* j = find(k);
		if (!j) throw std::runtime_error("json: missing key '" + k + "'");
		return *j;
	}
	Json& set(const std::string& k) {
		if (type != Object) { type = Object; obj.clear(); }
		if (Json* j = find(k)) return *j;
		obj.emplace_back(k, Json());
		return obj.back().second;
	}
	double value(const std::string& k, double def) const { const Json* j = find(k); return (j && j->type == Number) ? j->num : def; }
	std::string value(const std::string& k, const std::string& def) const { const Json* j = find(k); return (j && j->type == String) ? j->str : def; }
};

class JsonParser {
public:
	explicit JsonParser(const std::string& text) : s(text) {}
	Json parse() {
		Json v = parse_value();
		skip_ws();
		if (p != s.size()) fail("trailing characters");
		return v;
	}
private:
	const std::string& s;
	size_t p = 0;
	[[noreturn]] void fail(const std::string& why) const { throw std::runtime_error("json: " + why + " at offset " + std::to_string(p)); }
	void skip_ws() {
		while (p < s.size()) {
			if (std::isspace((unsigned char)s[p])) { ++p; continue; }
			if (s[p] == '/' && p + 1 < s.size() && s[p + 1] == '/') { while (p < s.size() && s[p] != '\n') ++p; continue; }
			if (s[p] == '/' && p + 1 < s.size() && s[p + 1] == '*') {
				p += 2;
				while (p + 1 < s.size() && !(s[p] == '*' && s[p + 1] == '/')) ++p;
				p += 2;
				continue;
			}
			break;
		}
	}
	Json parse_value() {
		skip_ws();
		if (p >= s.size()) fail("unexpected end");
		char c = s[p];
		if (c == '{') return parse_object();
		if (c == '[') return parse_array();
		if (c == '"') { Json j; j.type = Json::String; j.str = parse_string(); return j; }
		if (s.compare(p, 4, "true") == 0) { p += 4; Json j; j.type = Json::Bool; j.b = true; return j; }
		if (s.compare(p, 5, "false") == 0) { p += 5; Json j; j.type = Json::Bool; j.b = false; return j; }
		if (s.compare(p, 4, "null") == 0) { p += 4; return Json(); }
		return parse_number();
	}
	Json parse_number() {
		const char* begin = s.c_str() + p;
		char* end = nullptr;
		double v = std::strtod(begin, &end);
		if (end == begin) fail("bad value");
		p += (size_t)(end - begin);
		Json j; j.type = Json::Number; j.num = v;
		return j;
	}
	std::string parse_string() {
		++p; // opening quote
		std::string out;
		while (p < s.size() && s[p] != '"') {
			char c = s[p++];
			if (c == '\\') {
				if (p >= s.size()) fail("bad escape");
				char e = s[p++];
				switch (e) {
					case 'n': out += '\n'; break;
					case 't': out += '\t'; break;
					case 'r': out += '\r'; break;
					case 'b': out += '\b'; break;
					case 'f': out += '\f'; break;
					case 'u': {
						if (p + 4 > s.size()) fail("bad \\u escape");
						unsigned cp = (unsigned)std::strtoul(s.substr(p, 4).c_str(), nullptr, 16);
						p += 4;
						if (cp < 0x80) out += (char)cp;
						else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
						else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
						break;
					}
					default: out += e; break; // \" \\ \/
				}
			} else {
				out += c;
			}
		}
		if (p >= s.size()) fail("unterminated string");
		++p; // closing quote
		return out;
	}
	Json parse_array() {
		Json j; j.type = Json::Array;
		++p;
		skip_ws();
		if (p < s.size() && s[p] == ']') { ++p; return j; }
		while (true) {
			j.arr.push_back(parse_value());
			skip_ws();
			if (p >= s.size()) fail("unterminated array");
			if (s[p] == ',') { ++p; continue; }
			if (s[p] == ']') { ++p; break; }
			fail("expected , or ]");
		}
		return j;
	}
	Json parse_object() {
		Json j; j.type = Json::Object;
		++p;
		skip_ws();
		if (p < s.size() && s[p] == '}') { ++p; return j; }
		while (true) {
			skip_ws();
			if (p >= s.size() || s[p] != '"') fail("expected key");
			std::string k = parse_string();
			skip_ws();
			if (p >= s.size() || s[p] != ':') fail("expected :");
			++p;
			j.set(k) = parse_value();
			skip_ws();
			if (p >= s.size()) fail("unterminated object");
			if (s[p] == ',') { ++p; continue; }
			if (s[p] == '}') { ++p; break; }
			fail("expected , or }");
		}
		return j;
	}
};

// nlohmann merge_patch semantics used for "parent" configs (src/testbed.cu:86-97): objects merge recursively,
// null deletes, everything else replaces.
inline void json_merge_patch(Json& target, const Json& patch) {
	if (patch.type != Json::Object) { target = patch; return; }
	if (target.type != Json::Object) { target = Json(); target.type = Json::Object; }
	for (auto& kv : patch.obj) {
		if (kv.second.type == Json::Null) {
			for (size_t i = 0; i < target.obj.size(); ++i) if (target.obj[i].first == kv.first) { target.obj.erase(target.obj.begin() + i); break; }
		} else {
			json_merge_patch(target.set(kv.first), kv.second);
		}
	}
}

} // namespace hn
