// rtc_json.h — a small JSON reader for the two documents the host side parses: config.json
// (src/main.zig:56-69) and the JSON chunk of a glTF 2.0 asset (what libs/zgltf/src/main.zig parses
// with std.json). Values are kept in a tree of Json nodes; numbers as double.
#pragma once
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace rtc_host {

struct Json {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;     // insertion order kept

    const Json* find(const char* key) const
    {
        if (kind != Object) return nullptr;
        for (const auto& kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool has(const char* key) const { const Json* j = find(key); return j && j->kind != Null; }
    const Json& at(const char* key) const
    {
        const Json* j = find(key);
        if (!j) throw std::runtime_error(std::string("missing JSON key '") + key + "'");
        return *j;
    }
    double number(const char* key, double def) const { const Json* j = find(key); return j && j->kind == Number ? j->num : def; }
    long long integer(const char* key, long long def) const { const Json* j = find(key); return j && j->kind == Number ? (long long)j->num : def; }
    std::string string(const char* key, const std::string& def) const { const Json* j = find(key); return j && j->kind == String ? j->str : def; }
    size_t size() const { return kind == Array ? arr.size() : (kind == Object ? obj.size() : 0); }
};

class JsonParser {
public:
    JsonParser(const char* p, size_t n) : p_(p), end_(p + n) {}
    Json parse()
    {
        Json v = value();
        ws();
        if (p_ != end_) fail("trailing characters");
        return v;
    }

private:
    const char* p_;
    const char* end_;
    [[noreturn]] void fail(const char* what) const { throw std::runtime_error(std::string("JSON: ") + what); }
    void ws() { while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_; }
    bool lit(const char* s)
    {
        const size_t n = strlen(s);
        if ((size_t)(end_ - p_) >= n && memcmp(p_, s, n) == 0) { p_ += n; return true; }
        return false;
    }
    static void utf8(std::string& out, unsigned cp)
    {
        if (cp < 0x80) out += (char)cp;
        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
        else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
        else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    }
    unsigned hex4()
    {
        if (end_ - p_ < 4) fail("bad \\u escape");
        unsigned v = 0;
        for (int i = 0; i < 4; ++i) {
            const char c = *p_++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= (unsigned)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (unsigned)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (unsigned)(c - 'A' + 10);
            else fail("bad \\u escape");
        }
        return v;
    }
    std::string string()
    {
        std::string out;
        ++p_;                                   // opening quote
        while (true) {
            if (p_ >= end_) fail("unterminated string");
            const char c = *p_++;
            if (c == '"') break;
            if (c != '\\') { out += c; continue; }
            if (p_ >= end_) fail("unterminated escape");
            const char e = *p_++;
            switch (e) {
                case '"': out += '"'; break;
                case '\\': out += '\\'; break;
                case '/': out += '/'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'n': out += '\n'; break;
                case 'r': out += '\r'; break;
                case 't': out += '\t'; break;
                case 'u': {
                    unsigned cp = hex4();
                    if (cp >= 0xD800 && cp < 0xDC00 && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
                        p_ += 2;
                        const unsigned lo = hex4();
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    }
                    utf8(out, cp);
                    break;
                }
                default: fail("bad escape");
            }
        }
        return out;
    }
    Json value()
    {
        ws();
        if (p_ >= end_) fail("unexpected end");
        Json v;
        const char c = *p_;
        if (c == '{') {
            v.kind = Json::Object;
            ++p_;
            ws();
            if (p_ < end_ && *p_ == '}') { ++p_; return v; }
            while (true) {
                ws();
                if (p_ >= end_ || *p_ != '"') fail("expected a key");
                std::string k = string();
                ws();
                if (p_ >= end_ || *p_ != ':') fail("expected ':'");
                ++p_;
                v.obj.emplace_back(std::move(k), value());
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == '}') { ++p_; break; }
                fail("expected ',' or '}'");
            }
        } else if (c == '[') {
            v.kind = Json::Array;
            ++p_;
            ws();
            if (p_ < end_ && *p_ == ']') { ++p_; return v; }
            while (true) {
                v.arr.push_back(value());
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == ']') { ++p_; break; }
                fail("expected ',' or ']'");
            }
        } else if (c == '"') {
            v.kind = Json::String;
            v.str = string();
        } else if (lit("true")) { v.kind = Json::Bool; v.b = true; }
        else if (lit("false")) { v.kind = Json::Bool; v.b = false; }
        else if (lit("null")) { v.kind = Json::Null; }
        else {
            const char* s = p_;
            while (p_ < end_ && (strchr("+-.eE", *p_) || (*p_ >= '0' && *p_ <= '9'))) ++p_;
            if (s == p_) fail("unexpected character");
            const std::string tok(s, p_);
            char* e = nullptr;
            v.kind = Json::Number;
            v.num = strtod(tok.c_str(), &e);
            if (!e || *e) fail("bad number");
        }
        return v;
    }
};

}  // namespace rtc_host
