// synerfgine_b200 — minimal JSON reader for the virtual-scene description (the reference uses nlohmann::json,
// src/synerfgine/engine.cu:21-78). Supports objects, arrays, strings (with \" \\ \/ \b \f \n \r \t \uXXXX escapes),
// numbers, true/false/null, plus // and /* */ comments and trailing commas, which the shipped scene files use.
#pragma once
#include <string>
#include <vector>
#include <map>
#include <memory>
#include <stdexcept>
#include <cstdlib>
#include <cctype>

namespace snbjson {

struct Value;
typedef std::shared_ptr<Value> Ptr;

struct Value {
	enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
	bool b = false;
	double num = 0.0;
	std::string str;
	std::vector<Ptr> arr;
	std::vector<std::pair<std::string, Ptr>> obj;

	bool has(const std::string& k) const { for (auto& kv : obj) if (kv.first == k) return true; return false; }
	const Value& at(const std::string& k) const {
		for (auto& kv : obj) if (kv.first == k) return *kv.second;
		throw std::runtime_error("missing key \"" + k + "\"");
	}
	const Value& at(size_t i) const {
		if (kind != Array || i >= arr.size()) throw std::runtime_error("array index out of range");
		return *arr[i];
	}
	size_t size() const { return kind == Array ? arr.size() : (kind == Object ? obj.size() : 0); }
	double number() const { if (kind == Number) return num; if (kind == Bool) return b ? 1.0 : 0.0; throw std::runtime_error("expected a number"); }
	float f() const { return (float)number(); }
	int i() const { return (int)number(); }
	bool boolean() const { if (kind == Bool) return b; if (kind == Number) return num != 0.0; throw std::runtime_error("expected a boolean"); }
	const std::string& s() const { if (kind != String) throw std::runtime_error("expected a string"); return str; }
};

class Parser {
public:
	explicit Parser(const std::string& text) : t(text), p(0) {}
	Ptr parse() { Ptr v = value(); ws(); if (p != t.size()) fail("trailing characters"); return v; }
private:
	const std::string& t; size_t p;
	[[noreturn]] void fail(const std::string& m) { throw std::runtime_error("JSON: " + m + " at offset " + std::to_string(p)); }
	void ws() {
		for (;;) {
			while (p < t.size() && std::isspace((unsigned char)t[p])) ++p;
			if (p + 1 < t.size() && t[p] == '/' && t[p + 1] == '/') { while (p < t.size() && t[p] != '\n') ++p; continue; }
			if (p + 1 < t.size() && t[p] == '/' && t[p + 1] == '*') { p += 2; while (p + 1 < t.size() && !(t[p] == '*' && t[p + 1] == '/')) ++p; p += 2; continue; }
			break;
		}
	}
	Ptr value() {
		ws();
		if (p >= t.size()) fail("unexpected end");
		const char c = t[p];
		if (c == '{') return object();
		if (c == '[') return array();
		if (c == '"') { Ptr v = std::make_shared<Value>(); v->kind = Value::String; v->str = string(); return v; }
		if (t.compare(p, 4, "true") == 0) { p += 4; Ptr v = std::make_shared<Value>(); v->kind = Value::Bool; v->b = true; return v; }
		if (t.compare(p, 5, "false") == 0) { p += 5; Ptr v = std::make_shared<Value>(); v->kind = Value::Bool; v->b = false; return v; }
		if (t.compare(p, 4, "null") == 0) { p += 4; return std::make_shared<Value>(); }
		return number();
	}
	Ptr number() {
		const char* s = t.c_str() + p; char* e = nullptr;
		const double d = std::strtod(s, &e);
		if (e == s) fail("bad token");
		p += (size_t)(e - s);
		Ptr v = std::make_shared<Value>(); v->kind = Value::Number; v->num = d; return v;
	}
	std::string string() {
		std::string out; ++p;
		while (p < t.size() && t[p] != '"') {
			char c = t[p++];
			if (c == '\\') {
				if (p >= t.size()) fail("bad escape");
				const char e = t[p++];
				switch (e) {
					case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
					case 'b': out += '\b'; break; case 'f': out += '\f'; break;
					case 'u': { if (p + 4 > t.size()) fail("bad \\u"); unsigned cp = (unsigned)std::strtoul(t.substr(p, 4).c_str(), nullptr, 16); p += 4;
						if (cp < 0x80) out += (char)cp; else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
						else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); } break; }
					default: out += e;
				}
			} else out += c;
		}
		if (p >= t.size()) fail("unterminated string");
		++p;
		return out;
	}
	Ptr array() {
		Ptr v = std::make_shared<Value>(); v->kind = Value::Array; ++p;
		for (;;) {
			ws();
			if (p < t.size() && t[p] == ']') { ++p; return v; }
			v->arr.push_back(value());
			ws();
			if (p < t.size() && t[p] == ',') { ++p; continue; }
			if (p < t.size() && t[p] == ']') { ++p; return v; }
			fail("expected , or ]");
		}
	}
	Ptr object() {
		Ptr v = std::make_shared<Value>(); v->kind = Value::Object; ++p;
		for (;;) {
			ws();
			if (p < t.size() && t[p] == '}') { ++p; return v; }
			if (p >= t.size() || t[p] != '"') fail("expected a key");
			std::string k = string();
			ws();
			if (p >= t.size() || t[p] != ':') fail("expected :");
			++p;
			Ptr val = value();
			bool replaced = false;
			for (auto& kv : v->obj) if (kv.first == k) { kv.second = val; replaced = true; }
			if (!replaced) v->obj.emplace_back(k, val);
			ws();
			if (p < t.size() && t[p] == ',') { ++p; continue; }
			if (p < t.size() && t[p] == '}') { ++p; return v; }
			fail("expected , or }");
		}
	}
};

inline Ptr parse(const std::string& text) { return Parser(text).parse(); }

} // namespace snbjson
