// Minimal JSON reader/writer for the OptiMo scene format (the reference parses it with json11,
// external/json11; this is an independent small recursive-descent parser, enough for
// {min_frame, max_frame, frames_per_second, joints{...}, ik_handlers[...]}, OptiMo/core_io.cpp:216-250).
#ifndef OPTIMO_B200_JSON_HPP
#define OPTIMO_B200_JSON_HPP

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace optimo {

class Json {
public:
    enum Type { Null, Bool, Number, String, Array, Object };
    Json() : type_(Null) {}
    static Json number(double v) { Json j; j.type_ = Number; j.num_ = v; return j; }
    static Json string(const std::string& s) { Json j; j.type_ = String; j.str_ = s; return j; }
    static Json array() { Json j; j.type_ = Array; return j; }
    static Json object() { Json j; j.type_ = Object; return j; }

    Type type() const { return type_; }
    bool is_null() const { return type_ == Null; }
    double number_value() const { return type_ == Number ? num_ : 0.0; }
    int int_value() const { return static_cast<int>(number_value()); }
    const std::string& string_value() const { return str_; }
    const std::vector<Json>& array_items() const { return arr_; }
    std::vector<Json>& array_items() { return arr_; }
    const Json& operator[](const std::string& key) const
    {
        static const Json null_json;
        const auto it = obj_.find(key);
        return it == obj_.end() ? null_json : it->second;
    }
    const Json& operator[](size_t i) const
    {
        static const Json null_json;
        return i < arr_.size() ? arr_[i] : null_json;
    }
    void set(const std::string& key, const Json& v) { obj_[key] = v; }
    void push_back(const Json& v) { arr_.push_back(v); }

    /// Parses `text`; throws std::runtime_error with a position on malformed input.
    static Json parse(const std::string& text)
    {
        size_t i = 0;
        Json j = parse_value(text, i, 0);
        skip_ws(text, i);
        if (i != text.size()) fail("trailing characters", i);
        return j;
    }

    /// Numbers are written with 17 significant digits so that doubles round-trip (as json11 does).
    std::string dump() const
    {
        std::string out;
        dump_to(out);
        return out;
    }

private:
    Type type_;
    double num_ = 0.0;
    bool bool_ = false;
    std::string str_;
    std::vector<Json> arr_;
    std::map<std::string, Json> obj_;

    [[noreturn]] static void fail(const char* what, size_t pos)
    {
        throw std::runtime_error(std::string("JSON parse error: ") + what + " at offset " + std::to_string(pos));
    }
    static void skip_ws(const std::string& s, size_t& i)
    {
        while (i < s.size() && (s[i] == ' ' || s[i] == '\n' || s[i] == '\r' || s[i] == '\t')) ++i;
    }
    static Json parse_value(const std::string& s, size_t& i, int depth)
    {
        if (depth > 512) fail("nesting too deep", i);
        skip_ws(s, i);
        if (i >= s.size()) fail("unexpected end", i);
        const char c = s[i];
        if (c == '{') {
            Json j = object();
            ++i; skip_ws(s, i);
            if (i < s.size() && s[i] == '}') { ++i; return j; }
            for (;;) {
                skip_ws(s, i);
                if (i >= s.size() || s[i] != '"') fail("expected string key", i);
                const std::string key = parse_string(s, i);
                skip_ws(s, i);
                if (i >= s.size() || s[i] != ':') fail("expected ':'", i);
                ++i;
                j.obj_[key] = parse_value(s, i, depth + 1);
                skip_ws(s, i);
                if (i < s.size() && s[i] == ',') { ++i; continue; }
                if (i < s.size() && s[i] == '}') { ++i; return j; }
                fail("expected ',' or '}'", i);
            }
        }
        if (c == '[') {
            Json j = array();
            ++i; skip_ws(s, i);
            if (i < s.size() && s[i] == ']') { ++i; return j; }
            for (;;) {
                j.arr_.push_back(parse_value(s, i, depth + 1));
                skip_ws(s, i);
                if (i < s.size() && s[i] == ',') { ++i; continue; }
                if (i < s.size() && s[i] == ']') { ++i; return j; }
                fail("expected ',' or ']'", i);
            }
        }
        if (c == '"') return string(parse_string(s, i));
        if (s.compare(i, 4, "true") == 0) { i += 4; Json j; j.type_ = Bool; j.bool_ = true; return j; }
        if (s.compare(i, 5, "false") == 0) { i += 5; Json j; j.type_ = Bool; j.bool_ = false; return j; }
        if (s.compare(i, 4, "null") == 0) { i += 4; return Json(); }
        // number
        const char* begin = s.c_str() + i;
        char* end = nullptr;
        const double v = std::strtod(begin, &end);
        if (end == begin) fail("unexpected character", i);
        i += static_cast<size_t>(end - begin);
        return number(v);
    }
    static std::string parse_string(const std::string& s, size_t& i)
    {
        std::string out;
        ++i;  // opening quote
        while (i < s.size() && s[i] != '"') {
            if (s[i] == '\\') {
                ++i;
                if (i >= s.size()) fail("bad escape", i);
                switch (s[i]) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {
                        if (i + 4 >= s.size()) fail("bad \\u escape", i);
                        const unsigned cp = static_cast<unsigned>(std::strtoul(s.substr(i + 1, 4).c_str(), nullptr, 16));
                        if (cp < 0x80) out += static_cast<char>(cp);
                        else if (cp < 0x800) { out += static_cast<char>(0xC0 | (cp >> 6)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
                        else { out += static_cast<char>(0xE0 | (cp >> 12)); out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
                        i += 4;
                        break;
                    }
                    default: out += s[i];
                }
                ++i;
            } else {
                out += s[i++];
            }
        }
        if (i >= s.size()) fail("unterminated string", i);
        ++i;  // closing quote
        return out;
    }
    void dump_to(std::string& out) const
    {
        switch (type_) {
            case Null: out += "null"; break;
            case Bool: out += bool_ ? "true" : "false"; break;
            case Number: {
                if (std::isfinite(num_)) {
                    char buf[40];
                    if (num_ == std::floor(num_) && std::fabs(num_) < 1e15) std::snprintf(buf, sizeof buf, "%.0f", num_);
                    else std::snprintf(buf, sizeof buf, "%.17g", num_);
                    out += buf;
                } else out += "null";
                break;
            }
            case String: {
                out += '"';
                for (char ch : str_) {
                    if (ch == '"' || ch == '\\') { out += '\\'; out += ch; }
                    else if (ch == '\n') out += "\\n";
                    else out += ch;
                }
                out += '"';
                break;
            }
            case Array: {
                out += '[';
                for (size_t k = 0; k < arr_.size(); ++k) { if (k) out += ", "; arr_[k].dump_to(out); }
                out += ']';
                break;
            }
            case Object: {
                out += '{';
                bool first = true;
                for (const auto& kv : obj_) {
                    if (!first) out += ", ";
                    first = false;
                    out += '"'; out += kv.first; out += "\": ";
                    kv.second.dump_to(out);
                }
                out += '}';
                break;
            }
        }
    }
};

}  // namespace optimo
#endif
