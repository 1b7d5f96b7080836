// Minimal JSON value + recursive-descent parser for mega_ag.json / task_signature.json.
// The reference links nlohmann/json (lib/nlohmann, used by mega_ag_runners/mega_ag.cpp:132-305); the
// graph files only need objects, arrays, strings, booleans, null and numbers.  Integers are kept
// exact as uint64/int64 (moduli are up to 61 bits), everything else as double.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace lsr {

class Json {
public:
    enum Kind { Null, Bool, UInt, Int, Real, String, Array, Object };
    Kind kind = Null;
    bool b = false;
    uint64_t u = 0;
    int64_t i = 0;
    double d = 0.0;
    std::string s;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;  // insertion order kept (graph files are small per object)

    bool is_null() const { return kind == Null; }
    bool is_object() const { return kind == Object; }
    bool is_array() const { return kind == Array; }
    bool is_number() const { return kind == UInt || kind == Int || kind == Real; }
    bool contains(const std::string& k) const { return find(k) != nullptr; }
    const Json* find(const std::string& k) const {
        if (kind != Object) return nullptr;
        for (auto& kv : obj)
            if (kv.first == k) return &kv.second;
        return nullptr;
    }
    const Json& at(const std::string& k) const {
        const Json* p = find(k);
        if (!p) throw std::runtime_error("json: missing key '" + k + "'");
        return *p;
    }
    const Json& operator[](const std::string& k) const { return at(k); }
    const Json& operator[](size_t idx) const {
        if (kind != Array || idx >= arr.size()) throw std::runtime_error("json: array index out of range");
        return arr[idx];
    }
    size_t size() const { return kind == Array ? arr.size() : kind == Object ? obj.size() : 0; }

    bool as_bool() const {
        if (kind != Bool) throw std::runtime_error("json: not a boolean");
        return b;
    }
    uint64_t as_u64() const {
        if (kind == UInt) return u;
        if (kind == Int && i >= 0) return (uint64_t)i;
        if (kind == Real && d >= 0 && d == (double)(uint64_t)d) return (uint64_t)d;
        throw std::runtime_error("json: not an unsigned integer");
    }
    int64_t as_i64() const {
        if (kind == Int) return i;
        if (kind == UInt && u <= (uint64_t)INT64_MAX) return (int64_t)u;
        if (kind == Real && d == (double)(int64_t)d) return (int64_t)d;
        throw std::runtime_error("json: not an integer");
    }
    int as_int() const { return (int)as_i64(); }
    double as_double() const {
        if (kind == Real) return d;
        if (kind == UInt) return (double)u;
        if (kind == Int) return (double)i;
        throw std::runtime_error("json: not a number");
    }
    const std::string& as_string() const {
        if (kind != String) throw std::runtime_error("json: not a string");
        return s;
    }
    std::vector<uint64_t> as_u64_vector() const {
        if (kind != Array) throw std::runtime_error("json: not an array");
        std::vector<uint64_t> v;
        v.reserve(arr.size());
        for (auto& e : arr) v.push_back(e.as_u64());
        return v;
    }

    static Json parse(const std::string& text) {
        Parser p{text.data(), text.data() + text.size()};
        Json v = p.value();
        p.ws();
        if (p.cur != p.end) throw std::runtime_error("json: trailing characters");
        return v;
    }
    static Json parse_file(const std::string& path) {
        std::ifstream f(path, std::ios::binary);
        if (!f.is_open()) throw std::runtime_error("Cannot open JSON file " + path);
        std::stringstream ss;
        ss << f.rdbuf();
        return parse(ss.str());
    }

private:
    struct Parser {
        const char* cur;
        const char* end;
        void ws() {
            while (cur < end && (*cur == ' ' || *cur == '\n' || *cur == '\t' || *cur == '\r')) cur++;
        }
        [[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string("json: ") + what); }
        Json value() {
            ws();
            if (cur >= end) fail("unexpected end");
            switch (*cur) {
                case '{': return object();
                case '[': return array();
                case '"': {
                    Json v;
                    v.kind = String;
                    v.s = string();
                    return v;
                }
                case 't': return literal("true", Bool, true);
                case 'f': return literal("false", Bool, false);
                case 'n': return literal("null", Null, false);
                default: return number();
            }
        }
        Json literal(const char* word, Kind k, bool bv) {
            size_t n = std::strlen(word);
            if ((size_t)(end - cur) < n || std::strncmp(cur, word, n) != 0) fail("bad literal");
            cur += n;
            Json v;
            v.kind = k;
            v.b = bv;
            return v;
        }
        Json number() {
            const char* st = cur;
            bool neg = false, real = false;
            if (cur < end && *cur == '-') {
                neg = true;
                cur++;
            }
            if (cur >= end || !((*cur >= '0' && *cur <= '9'))) fail("bad number");
            while (cur < end && ((*cur >= '0' && *cur <= '9') || *cur == '.' || *cur == 'e' || *cur == 'E' ||
                                 *cur == '+' || *cur == '-')) {
                if (*cur == '.' || *cur == 'e' || *cur == 'E') real = true;
                cur++;
            }
            std::string tok(st, cur);
            Json v;
            if (real || tok.size() > 20) {
                v.kind = Real;
                v.d = std::strtod(tok.c_str(), nullptr);
            } else if (neg) {
                v.kind = Int;
                v.i = std::strtoll(tok.c_str(), nullptr, 10);
            } else {
                v.kind = UInt;
                v.u = std::strtoull(tok.c_str(), nullptr, 10);
            }
            return v;
        }
        std::string string() {
            cur++;  // opening quote
            std::string out;
            while (cur < end && *cur != '"') {
                char c = *cur++;
                if (c != '\\') {
                    out.push_back(c);
                    continue;
                }
                if (cur >= end) fail("bad escape");
                char e = *cur++;
                switch (e) {
                    case 'n': out.push_back('\n'); break;
                    case 't': out.push_back('\t'); break;
                    case 'r': out.push_back('\r'); break;
                    case 'b': out.push_back('\b'); break;
                    case 'f': out.push_back('\f'); break;
                    case 'u': {
                        if (end - cur < 4) fail("bad \\u escape");
                        unsigned cp = (unsigned)std::strtoul(std::string(cur, cur + 4).c_str(), nullptr, 16);
                        cur += 4;
                        if (cp < 0x80) out.push_back((char)cp);
                        else if (cp < 0x800) {
                            out.push_back((char)(0xC0 | (cp >> 6)));
                            out.push_back((char)(0x80 | (cp & 0x3F)));
                        } else {
                            out.push_back((char)(0xE0 | (cp >> 12)));
                            out.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
                            out.push_back((char)(0x80 | (cp & 0x3F)));
                        }
                        break;
                    }
                    default: out.push_back(e); break;  // \" \\ \/
                }
            }
            if (cur >= end) fail("unterminated string");
            cur++;
            return out;
        }
        Json array() {
            cur++;
            Json v;
            v.kind = Array;
            ws();
            if (cur < end && *cur == ']') {
                cur++;
                return v;
            }
            for (;;) {
                v.arr.push_back(value());
                ws();
                if (cur >= end) fail("unterminated array");
                if (*cur == ',') {
                    cur++;
                    continue;
                }
                if (*cur == ']') {
                    cur++;
                    return v;
                }
                fail("expected , or ]");
            }
        }
        Json object() {
            cur++;
            Json v;
            v.kind = Object;
            ws();
            if (cur < end && *cur == '}') {
                cur++;
                return v;
            }
            for (;;) {
                ws();
                if (cur >= end || *cur != '"') fail("expected object key");
                std::string k = string();
                ws();
                if (cur >= end || *cur != ':') fail("expected :");
                cur++;
                v.obj.emplace_back(std::move(k), value());
                ws();
                if (cur >= end) fail("unterminated object");
                if (*cur == ',') {
                    cur++;
                    continue;
                }
                if (*cur == '}') {
                    cur++;
                    return v;
                }
                fail("expected , or }");
            }
        }
    };
};

}  // namespace lsr
