// Minimal JSON DOM for the glTF loader (host, load-time only).  The reference parses with
// rapidjson (hw8/src/sceneio.cpp:351-354); numbers are read as double and narrowed to float
// exactly where the reference calls GetFloat().  Own parser => no build-time dependency.
#pragma once
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace b200rt_host {

class Json {
public:
    enum Kind { Null, Bool, Number, String, Array, Object };
    Kind kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    bool has(const char *key) const { return find(key) != nullptr; }
    const Json *find(const char *key) const {
        if (kind != Object) return nullptr;
        for (const auto &kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    // Missing members resolve to a shared Null value instead of UB (reference: rapidjson assert).
    const Json &operator[](const char *key) const {
        const Json *p = find(key);
        return p ? *p : null_value();
    }
    const Json &operator[](int i) const { return (*this)[(size_t)(i < 0 ? ~(size_t)0 : (size_t)i)]; }
    const Json &operator[](size_t i) const { return (kind == Array && i < arr.size()) ? arr[i] : null_value(); }
    size_t size() const { return kind == Array ? arr.size() : 0; }
    bool is_null() const { return kind == Null; }
    bool is_number() const { return kind == Number; }
    float as_float() const { return (float)num; }
    size_t as_uint() const { return num > 0 ? (size_t)num : 0; }

    static const Json &null_value() {
        static const Json n;
        return n;
    }

    // Returns false on malformed input; `err` receives a byte offset.
    static bool parse(const std::string &text, Json &out, size_t &err) {
        Parser p{text.data(), text.data() + text.size(), text.data()};
        p.ws();
        if (!p.value(out, 0)) { err = (size_t)(p.cur - p.begin); return false; }
        p.ws();
        if (p.cur != p.end) { err = (size_t)(p.cur - p.begin); return false; }
        return true;
    }

private:
    struct Parser {
        const char *cur, *end, *begin;
        void ws() { while (cur < end && (*cur == ' ' || *cur == '\n' || *cur == '\t' || *cur == '\r')) ++cur; }
        bool lit(const char *s) {
            size_t n = strlen(s);
            if ((size_t)(end - cur) < n || memcmp(cur, s, n) != 0) return false;
            cur += n;
            return true;
        }
        static void utf8(std::string &o, unsigned cp) {
            if (cp < 0x80) o += (char)cp;
            else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 0x3F)); }
            else { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
        }
        bool string(std::string &o) {
            if (cur >= end || *cur != '"') return false;
            ++cur;
            while (cur < end && *cur != '"') {
                if (*cur == '\\') {
                    if (++cur >= end) return false;
                    switch (*cur) {
                        case 'n': o += '\n'; break; case 't': o += '\t'; break; case 'r': o += '\r'; break;
                        case 'b': o += '\b'; break; case 'f': o += '\f'; break;
                        case 'u': {
                            if (end - cur < 5) return false;
                            char hex[5] = {cur[1], cur[2], cur[3], cur[4], 0};
                            utf8(o, (unsigned)strtoul(hex, nullptr, 16));
                            cur += 4;
                            break;
                        }
                        default: o += *cur;
                    }
                    ++cur;
                } else o += *cur++;
            }
            if (cur >= end) return false;
            ++cur;
            return true;
        }
        bool value(Json &v, int depth) {
            if (depth > 256 || cur >= end) return false;
            char c = *cur;
            if (c == '{') {
                v.kind = Object;
                ++cur; ws();
                if (cur < end && *cur == '}') { ++cur; return true; }
                while (true) {
                    ws();
                    std::string k;
                    if (!string(k)) return false;
                    ws();
                    if (cur >= end || *cur != ':') return false;
                    ++cur; ws();
                    v.obj.emplace_back(std::move(k), Json());
                    if (!value(v.obj.back().second, depth + 1)) return false;
                    ws();
                    if (cur < end && *cur == ',') { ++cur; continue; }
                    if (cur < end && *cur == '}') { ++cur; return true; }
                    return false;
                }
            }
            if (c == '[') {
                v.kind = Array;
                ++cur; ws();
                if (cur < end && *cur == ']') { ++cur; return true; }
                while (true) {
                    ws();
                    v.arr.emplace_back();
                    if (!value(v.arr.back(), depth + 1)) return false;
                    ws();
                    if (cur < end && *cur == ',') { ++cur; continue; }
                    if (cur < end && *cur == ']') { ++cur; return true; }
                    return false;
                }
            }
            if (c == '"') { v.kind = String; return string(v.str); }
            if (c == 't') { v.kind = Bool; v.b = true; return lit("true"); }
            if (c == 'f') { v.kind = Bool; v.b = false; return lit("false"); }
            if (c == 'n') { v.kind = Null; return lit("null"); }
            char *e = nullptr;
            std::string tmp(cur, (size_t)std::min<ptrdiff_t>(end - cur, 64));
            double d = strtod(tmp.c_str(), &e);
            if (e == tmp.c_str()) return false;
            v.kind = Number;
            v.num = d;
            cur += (e - tmp.c_str());
            return true;
        }
    };
};

}  // namespace b200rt_host
