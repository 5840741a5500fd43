// json.hpp — minimal RFC 8259 reader for the scene files (stands in for serde_json at
// basics/scene_loader.rs:3-7).  DOM of vectors; numbers are parsed with std::from_chars (correctly rounded, like
// strtod) and narrowed to f32 by the caller, which is what serde_json does for an `f32` field.
// A Value is 24 bytes: scalars carry no string or vector, strings / arrays / objects keep theirs in a heap box - the
// 100 000-object scene of config D is 4 M values and 37 MB of text, and its load time is what a caller of the JSON
// API pays before the first frame (first version: every number constructed, moved and destroyed a string and two
// vectors, was copied for strtod, and the whole document became a DOM whose destruction alone took 0.4 s: 0.95 s;
// now, with the objects array streamed to the loader (parse_streaming), 0.25 s on the same machine).
#pragma once
#include <charconv>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace forma::json {

struct Value {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false;
    double num = 0.0;
    struct Box {
        std::string str;
        std::vector<Value> arr;
        std::vector<std::pair<std::string, Value>> obj;
    };
    std::unique_ptr<Box> box;  // String / Array / Object only

    Box& make() { if (!box) box.reset(new Box); return *box; }
    const std::string& str() const { static const std::string e; return box ? box->str : e; }
    const std::vector<Value>& arr() const { static const std::vector<Value> e; return box ? box->arr : e; }
    const std::vector<std::pair<std::string, Value>>& obj() const {
        static const std::vector<std::pair<std::string, Value>> e;
        return box ? box->obj : e;
    }
    const Value* find(const char* key) const {
        if (type != Object) return nullptr;
        for (const auto& kv : obj())
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
    // Streaming of ONE top-level array: the elements of root[key] (when it is an array) are handed to `on_item` one at
    // a time and not kept - root[key] comes back as an empty Array.  The scene loader streams "objects": a DOM of
    // config D's 100 000 objects is 1.4 M heap boxes whose destruction alone took 0.4 s.
    template <class F>
    Value parse_streaming(const char* key, F&& on_item) {
        stream_key_ = key;
        on_item_ = [&](Value&& v) { on_item(std::move(v)); };
        return parse();
    }
    Value parse() {
        Value v = value(0);
        ws();
        if (p_ != end_) fail("trailing characters");
        return v;
    }

private:
    const char *p_, *end_, *begin_;
    const char* stream_key_ = nullptr;
    std::function<void(Value&&)> on_item_;
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
            auto& members = v.make().obj;
            members.reserve(8);
            for (;;) {
                ws();
                if (p_ >= end_ || *p_ != '"') fail("expected object key");
                std::string k = string();
                ws();
                if (p_ >= end_ || *p_ != ':') fail("expected ':'");
                ++p_;
                if (depth == 0 && stream_key_ && k == stream_key_) {
                    ws();
                    if (p_ < end_ && *p_ == '[') {  // hand the elements over one by one; keep an empty Array
                        ++p_;
                        Value a;
                        a.type = Value::Array;
                        ws();
                        if (p_ < end_ && *p_ == ']') ++p_;
                        else
                            for (;;) {
                                on_item_(value(depth + 2));
                                ws();
                                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                                if (p_ < end_ && *p_ == ']') { ++p_; break; }
                                fail("expected ',' or ']'");
                            }
                        members.emplace_back(std::move(k), std::move(a));
                        ws();
                        if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                        if (p_ < end_ && *p_ == '}') { ++p_; break; }
                        fail("expected ',' or '}'");
                    }
                }
                members.emplace_back(std::move(k), value(depth + 1));
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
            auto& items = v.make().arr;
            for (;;) {
                items.push_back(value(depth + 1));
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == ']') { ++p_; break; }
                fail("expected ',' or ']'");
            }
        } else if (c == '"') {
            v.type = Value::String;
            v.make().str = string();
        } else if (lit("true")) {
            v.type = Value::Bool; v.b = true;
        } else if (lit("false")) {
            v.type = Value::Bool; v.b = false;
        } else if (lit("null")) {
            v.type = Value::Null;
        } else if (c == '-' || (c >= '0' && c <= '9')) {
            v.type = Value::Number;
            // the token, then std::from_chars on exactly those bytes (it never reads past `q`)
            const char* q = p_;
            while (q < end_ && ((*q >= '0' && *q <= '9') || *q == '.' || *q == '-' || *q == '+' || *q == 'e' || *q == 'E')) ++q;
            if (q == p_ || q - p_ >= 64) fail("bad number");
            const auto r = std::from_chars(p_, q, v.num);
            if (r.ec == std::errc::result_out_of_range) {  // strtod's answer for these: +-inf / 0 (serde_json rejects them; none occur)
                char buf[64];
                std::memcpy(buf, p_, (size_t)(q - p_));
                buf[q - p_] = 0;
                v.num = std::strtod(buf, nullptr);
            } else if (r.ec != std::errc() || r.ptr != q) {
                fail("bad number");
            }
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
