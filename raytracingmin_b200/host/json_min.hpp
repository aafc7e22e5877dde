// json_min.hpp — a small self-contained JSON reader/writer for the host front end.
// The reference parses its scene files with the vendored nlohmann/json
// (src/SettingData.cpp:6-12,129-186); the schema is tiny (objects, arrays, numbers,
// strings), so the front end carries its own parser instead of a third-party tree.
// Numbers are kept as double (nlohmann converts to double/float/int on access the
// same way: value -> static_cast).  Object keys are kept sorted, as nlohmann's
// default std::map does, so "items()" order matches the reference.
#pragma once
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace jsonmin {

struct Value;
using Array = std::vector<Value>;
using Object = std::map<std::string, Value>;

struct Value {
    enum Kind { Null, Bool, Number, String, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0.0;
    bool is_integer = false;  // written without fraction/exponent
    std::string str;
    std::shared_ptr<Array> arr;
    std::shared_ptr<Object> obj;

    Value() = default;
    Value(double v) : kind(Number), num(v) {}
    Value(int v) : kind(Number), num(v), is_integer(true) {}
    Value(const std::string& s) : kind(String), str(s) {}
    static Value array() { Value v; v.kind = Arr; v.arr = std::make_shared<Array>(); return v; }
    static Value object() { Value v; v.kind = Obj; v.obj = std::make_shared<Object>(); return v; }

    bool is_object() const { return kind == Obj; }
    bool is_array() const { return kind == Arr; }
    bool is_number() const { return kind == Number; }
    bool empty() const { return (kind == Obj && obj->empty()) || (kind == Arr && arr->empty()) || kind == Null; }
    bool contains(const std::string& k) const { return kind == Obj && obj->count(k) != 0; }
    const Value& at(const std::string& k) const {
        if (kind != Obj) throw std::runtime_error("json: not an object while looking up '" + k + "'");
        auto it = obj->find(k);
        if (it == obj->end()) throw std::runtime_error("json: missing key '" + k + "'");
        return it->second;
    }
    const Value& at(size_t i) const {
        if (kind != Arr || i >= arr->size()) throw std::runtime_error("json: array index out of range");
        return (*arr)[i];
    }
    size_t size() const { return kind == Arr ? arr->size() : kind == Obj ? obj->size() : 0; }
    double number() const {
        if (kind != Number) throw std::runtime_error("json: number expected");
        return num;
    }
    Value& operator[](const std::string& k) {
        if (kind != Obj) { kind = Obj; obj = std::make_shared<Object>(); }
        return (*obj)[k];
    }
    void push_back(const Value& v) {
        if (kind != Arr) { kind = Arr; arr = std::make_shared<Array>(); }
        arr->push_back(v);
    }
};

class Parser {
public:
    explicit Parser(const std::string& text) : s_(text) {}
    Value parse() {
        Value v = value();
        ws();
        if (i_ != s_.size()) fail("trailing characters");
        return v;
    }

private:
    const std::string& s_;
    size_t i_ = 0;

    [[noreturn]] void fail(const char* what) const {
        throw std::runtime_error(std::string("json parse error at byte ") + std::to_string(i_) + ": " + what);
    }
    void ws() {
        while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\t' || s_[i_] == '\n' || s_[i_] == '\r')) ++i_;
    }
    bool eat(char c) {
        ws();
        if (i_ < s_.size() && s_[i_] == c) { ++i_; return true; }
        return false;
    }
    Value value() {
        ws();
        if (i_ >= s_.size()) fail("unexpected end");
        const char c = s_[i_];
        if (c == '{') return object();
        if (c == '[') return array();
        if (c == '"') return Value(string());
        if (c == 't' || c == 'f' || c == 'n') return literal();
        return number();
    }
    Value literal() {
        Value v;
        if (s_.compare(i_, 4, "true") == 0) { v.kind = Value::Bool; v.b = true; i_ += 4; }
        else if (s_.compare(i_, 5, "false") == 0) { v.kind = Value::Bool; v.b = false; i_ += 5; }
        else if (s_.compare(i_, 4, "null") == 0) { i_ += 4; }
        else fail("bad literal");
        return v;
    }
    Value number() {
        const size_t b = i_;
        bool integer = true;
        if (i_ < s_.size() && (s_[i_] == '-' || s_[i_] == '+')) ++i_;
        while (i_ < s_.size() && (std::isdigit((unsigned char)s_[i_]) || s_[i_] == '.' || s_[i_] == 'e' || s_[i_] == 'E' ||
                                  s_[i_] == '-' || s_[i_] == '+')) {
            if (s_[i_] == '.' || s_[i_] == 'e' || s_[i_] == 'E') integer = false;
            ++i_;
        }
        if (i_ == b) fail("number expected");
        char* end = nullptr;
        const std::string tok = s_.substr(b, i_ - b);
        const double v = std::strtod(tok.c_str(), &end);
        if (end == tok.c_str() || *end != '\0') fail("malformed number");
        Value out(v);
        out.is_integer = integer;
        return out;
    }
    std::string string() {
        if (s_[i_] != '"') fail("string expected");
        ++i_;
        std::string out;
        while (i_ < s_.size() && s_[i_] != '"') {
            char c = s_[i_++];
            if (c == '\\') {
                if (i_ >= s_.size()) fail("bad escape");
                const char e = s_[i_++];
                switch (e) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {
                        if (i_ + 4 > s_.size()) fail("bad \\u escape");
                        const unsigned cp = (unsigned)std::strtoul(s_.substr(i_, 4).c_str(), nullptr, 16);
                        i_ += 4;
                        if (cp < 0x80) out += (char)cp;
                        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                        else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                        break;
                    }
                    default: out += e;  // \" \\ \/
                }
            } else {
                out += c;
            }
        }
        if (i_ >= s_.size()) fail("unterminated string");
        ++i_;
        return out;
    }
    Value array() {
        ++i_;
        Value v = Value::array();
        if (eat(']')) return v;
        for (;;) {
            v.arr->push_back(value());
            if (eat(',')) continue;
            if (eat(']')) return v;
            fail("',' or ']' expected");
        }
    }
    Value object() {
        ++i_;
        Value v = Value::object();
        if (eat('}')) return v;
        for (;;) {
            ws();
            if (i_ >= s_.size() || s_[i_] != '"') fail("key expected");
            const std::string k = string();
            if (!eat(':')) fail("':' expected");
            (*v.obj)[k] = value();
            if (eat(',')) continue;
            if (eat('}')) return v;
            fail("',' or '}' expected");
        }
    }
};

inline Value parse(const std::string& text) { return Parser(text).parse(); }

inline void dump(const Value& v, std::string& out) {
    char buf[64];
    switch (v.kind) {
        case Value::Null: out += "null"; break;
        case Value::Bool: out += v.b ? "true" : "false"; break;
        case Value::Number:
            if (v.is_integer && std::fabs(v.num) < 9e15) std::snprintf(buf, sizeof buf, "%lld", (long long)v.num);
            else {
                std::snprintf(buf, sizeof buf, "%.17g", v.num);
                // shortest representation that round-trips
                for (int prec = 1; prec < 17; ++prec) {
                    char t[64];
                    std::snprintf(t, sizeof t, "%.*g", prec, v.num);
                    if (std::strtod(t, nullptr) == v.num) { std::snprintf(buf, sizeof buf, "%s", t); break; }
                }
                if (!std::strpbrk(buf, ".eEni")) std::snprintf(buf + std::strlen(buf), sizeof buf - std::strlen(buf), ".0");
            }
            out += buf;
            break;
        case Value::String:
            out += '"';
            for (char c : v.str) {
                if (c == '"' || c == '\\') { out += '\\'; out += c; }
                else if (c == '\n') out += "\\n";
                else if (c == '\t') out += "\\t";
                else out += c;
            }
            out += '"';
            break;
        case Value::Arr: {
            out += '[';
            bool first = true;
            for (const auto& e : *v.arr) { if (!first) out += ','; first = false; dump(e, out); }
            out += ']';
            break;
        }
        case Value::Obj: {
            out += '{';
            bool first = true;
            for (const auto& kv : *v.obj) {
                if (!first) out += ',';
                first = false;
                dump(Value(kv.first), out);
                out += ':';
                dump(kv.second, out);
            }
            out += '}';
            break;
        }
    }
}

inline std::string dump(const Value& v) { std::string s; dump(v, s); return s; }

}  // namespace jsonmin
