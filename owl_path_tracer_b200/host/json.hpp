// json.hpp — small JSON reader for assets/settings.json and assets/<scene>.json.
// The reference uses nlohmann/json (path_tracer/src/utils/parser.cpp); only what those
// two file formats need is implemented: objects, arrays, strings, numbers, booleans,
// null.  Access errors throw std::runtime_error, as nlohmann's .get<T>() does on a
// missing key or a type mismatch.
#pragma once
#include <cstdlib>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace host {

class Json {
public:
    enum Kind { Null, Bool, Number, String, Array, Object };
    Kind kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    bool is_array() const { return kind == Array; }
    bool is_object() const { return kind == Object; }
    bool contains(const std::string& k) const {
        for (auto& kv : obj)
            if (kv.first == k) return true;
        return false;
    }
    const Json& operator[](const std::string& k) const {
        if (kind != Object) throw std::runtime_error("json: not an object while reading key '" + k + "'");
        for (auto& kv : obj)
            if (kv.first == k) return kv.second;
        throw std::runtime_error("json: missing key '" + k + "'");
    }
    const Json& operator[](size_t i) const {
        if (kind != Array || i >= arr.size()) throw std::runtime_error("json: bad array access");
        return arr[i];
    }
    size_t size() const { return kind == Array ? arr.size() : obj.size(); }
    float as_float() const {
        if (kind != Number) throw std::runtime_error("json: number expected");
        return (float)num;
    }
    int as_int() const {
        if (kind != Number) throw std::runtime_error("json: number expected");
        return (int)num;
    }
    bool as_bool() const {
        if (kind != Bool) throw std::runtime_error("json: boolean expected");
        return b;
    }
    const std::string& as_string() const {
        if (kind != String) throw std::runtime_error("json: string expected");
        return str;
    }

    static Json parse(const std::string& text) {
        size_t p = 0;
        Json j = value(text, p);
        ws(text, p);
        if (p != text.size()) throw std::runtime_error("json: trailing characters");
        return j;
    }
    static Json load(const std::string& path) {
        std::ifstream f(path);
        if (!f) throw std::runtime_error("json: cannot open " + path);
        std::stringstream ss;
        ss << f.rdbuf();
        return parse(ss.str());
    }

private:
    static void ws(const std::string& s, size_t& p) {
        while (p < s.size()) {
            unsigned char c = (unsigned char)s[p];
            if (c == ' ' || c == '\t' || c == '\n' || c == '\r')
                ++p;
            else if (c == 0xEF && p + 2 < s.size() && (unsigned char)s[p + 1] == 0xBB && (unsigned char)s[p + 2] == 0xBF)
                p += 3;  // UTF-8 BOM
            else
                break;
        }
    }
    static std::string string_lit(const std::string& s, size_t& p) {
        std::string out;
        ++p;  // opening quote
        while (p < s.size() && s[p] != '"') {
            char c = s[p++];
            if (c == '\\' && p < s.size()) {
                char e = s[p++];
                switch (e) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {
                        unsigned cp = (unsigned)strtoul(s.substr(p, 4).c_str(), nullptr, 16);
                        p += 4;
                        if (cp < 0x80)
                            out += (char)cp;
                        else if (cp < 0x800) {
                            out += (char)(0xC0 | (cp >> 6));
                            out += (char)(0x80 | (cp & 0x3F));
                        } else {
                            out += (char)(0xE0 | (cp >> 12));
                            out += (char)(0x80 | ((cp >> 6) & 0x3F));
                            out += (char)(0x80 | (cp & 0x3F));
                        }
                        break;
                    }
                    default: out += e;
                }
            } else
                out += c;
        }
        if (p >= s.size()) throw std::runtime_error("json: unterminated string");
        ++p;
        return out;
    }
    static Json value(const std::string& s, size_t& p) {
        ws(s, p);
        if (p >= s.size()) throw std::runtime_error("json: unexpected end");
        Json j;
        char c = s[p];
        if (c == '{') {
            j.kind = Object;
            ++p;
            ws(s, p);
            if (p < s.size() && s[p] == '}') {
                ++p;
                return j;
            }
            for (;;) {
                ws(s, p);
                if (p >= s.size() || s[p] != '"') throw std::runtime_error("json: key expected");
                std::string k = string_lit(s, p);
                ws(s, p);
                if (p >= s.size() || s[p] != ':') throw std::runtime_error("json: ':' expected");
                ++p;
                Json v = value(s, p);
                j.obj.emplace_back(std::move(k), std::move(v));
                ws(s, p);
                if (p < s.size() && s[p] == ',') {
                    ++p;
                    continue;
                }
                if (p < s.size() && s[p] == '}') {
                    ++p;
                    break;
                }
                throw std::runtime_error("json: ',' or '}' expected");
            }
        } else if (c == '[') {
            j.kind = Array;
            ++p;
            ws(s, p);
            if (p < s.size() && s[p] == ']') {
                ++p;
                return j;
            }
            for (;;) {
                j.arr.push_back(value(s, p));
                ws(s, p);
                if (p < s.size() && s[p] == ',') {
                    ++p;
                    continue;
                }
                if (p < s.size() && s[p] == ']') {
                    ++p;
                    break;
                }
                throw std::runtime_error("json: ',' or ']' expected");
            }
        } else if (c == '"') {
            j.kind = String;
            j.str = string_lit(s, p);
        } else if (s.compare(p, 4, "true") == 0) {
            j.kind = Bool;
            j.b = true;
            p += 4;
        } else if (s.compare(p, 5, "false") == 0) {
            j.kind = Bool;
            j.b = false;
            p += 5;
        } else if (s.compare(p, 4, "null") == 0) {
            p += 4;
        } else {
            const char* st = s.c_str() + p;
            char* en = nullptr;
            j.num = strtod(st, &en);
            if (en == st) throw std::runtime_error("json: unexpected character");
            j.kind = Number;
            p += (size_t)(en - st);
        }
        return j;
    }
};

}  // namespace host
