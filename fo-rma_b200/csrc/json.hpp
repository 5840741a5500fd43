// json.hpp — minimal RFC 8259 reader for the scene files (stands in for serde_json at
// basics/scene_loader.rs:3-7).  DOM of vectors; numbers are parsed with strtod and narrowed to
// f32 by the caller, which is what serde_json does for an `f32` field.
#pragma once
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace forma::json {

struct Value {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Value> arr;
    std::vector<std::pair<std::string, Value>> obj;

    const Value* find(const char* key) const {
        if (type != Object) return nullptr;
        for (const auto& kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
};

struct ParseError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

class Parser {
public:
    Parser(const char* s, size_t n) : p_(s), end_(s + n), begin_(s) {}
    Value parse() {
        Value v = value(0);
        ws();
        if (p_ != end_) fail("trailing characters");
        return v;
    }

private:
    const char *p_, *end_, *begin_;
    [[noreturn]] void fail(const char* what) {
        throw ParseError(std::string(what) + " at byte " + std::to_string(p_ - begin_));
    }
    void ws() {
        while (p_ < end_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) ++p_;
    }
    bool lit(const char* s) {
        size_t n = std::strlen(s);
        if ((size_t)(end_ - p_) >= n && std::memcmp(p_, s, n) == 0) { p_ += n; return true; }
        return false;
    }
    Value value(int depth) {
        if (depth > 64) fail("nesting too deep");
        ws();
        if (p_ >= end_) fail("unexpected end");
        Value v;
        char c = *p_;
        if (c == '{') {
            ++p_;
            v.type = Value::Object;
            ws();
            if (p_ < end_ && *p_ == '}') { ++p_; return v; }
            for (;;) {
                ws();
                if (p_ >= end_ || *p_ != '"') fail("expected object key");
                std::string k = string();
                ws();
                if (p_ >= end_ || *p_ != ':') fail("expected ':'");
                ++p_;
                v.obj.emplace_back(std::move(k), value(depth + 1));
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == '}') { ++p_; break; }
                fail("expected ',' or '}'");
            }
        } else if (c == '[') {
            ++p_;
            v.type = Value::Array;
            ws();
            if (p_ < end_ && *p_ == ']') { ++p_; return v; }
            for (;;) {
                v.arr.push_back(value(depth + 1));
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == ']') { ++p_; break; }
                fail("expected ',' or ']'");
            }
        } else if (c == '"') {
            v.type = Value::String;
            v.str = string();
        } else if (lit("true")) {
            v.type = Value::Bool; v.b = true;
        } else if (lit("false")) {
            v.type = Value::Bool; v.b = false;
        } else if (lit("null")) {
            v.type = Value::Null;
        } else if (c == '-' || (c >= '0' && c <= '9')) {
            v.type = Value::Number;
            // copy the token so strtod never runs past a non-terminated buffer
            const char* q = p_;
            while (q < end_ && (std::strchr("+-.eE", *q) || (*q >= '0' && *q <= '9'))) ++q;
            char buf[64];
            size_t n = (size_t)(q - p_);
            if (n == 0 || n >= sizeof buf) fail("bad number");
            std::memcpy(buf, p_, n);
            buf[n] = 0;
            char* e = nullptr;
            v.num = std::strtod(buf, &e);
            if (e != buf + n) fail("bad number");
            p_ = q;
        } else {
            fail("unexpected character");
        }
        return v;
    }
    std::string string() {
        ++p_;  // opening quote
        std::string s;
        while (p_ < end_ && *p_ != '"') {
            char c = *p_++;
            if (c == '\\') {
                if (p_ >= end_) fail("bad escape");
                char e = *p_++;
                switch (e) {
                    case '"': s += '"'; break;
                    case '\\': s += '\\'; break;
                    case '/': s += '/'; break;
                    case 'b': s += '\b'; break;
                    case 'f': s += '\f'; break;
                    case 'n': s += '\n'; break;
                    case 'r': s += '\r'; break;
                    case 't': s += '\t'; break;
                    case 'u': {
                        if (end_ - p_ < 4) fail("bad \\u escape");
                        unsigned cp = 0;
                        for (int i = 0; i < 4; ++i) {
                            char h = *p_++;
                            cp <<= 4;
                            if (h >= '0' && h <= '9') cp |= (unsigned)(h - '0');
                            else if (h >= 'a' && h <= 'f') cp |= (unsigned)(h - 'a' + 10);
                            else if (h >= 'A' && h <= 'F') cp |= (unsigned)(h - 'A' + 10);
                            else fail("bad \\u escape");
                        }
                        if (cp < 0x80) s += (char)cp;
                        else if (cp < 0x800) { s += (char)(0xC0 | (cp >> 6)); s += (char)(0x80 | (cp & 0x3F)); }
                        else { s += (char)(0xE0 | (cp >> 12)); s += (char)(0x80 | ((cp >> 6) & 0x3F)); s += (char)(0x80 | (cp & 0x3F)); }
                        break;
                    }
                    default: fail("bad escape");
                }
            } else {
                s += c;
            }
        }
        if (p_ >= end_) fail("unterminated string");
        ++p_;
        return s;
    }
};

inline Value parse(const char* s, size_t n) { return Parser(s, n).parse(); }

}  // namespace forma::json
