// JsonMin.h — a small JSON reader for the planner's config file (config.json is JSON syntax that the
// reference loads through yaml-cpp, src/ConfigParserYAML.cpp:10-19).  Objects keep file order, which
// the reference relies on when it enumerates a component's boxes (:56-68).
#pragma once

#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace epp {

class Json {
public:
    enum Kind { Null, Bool, Number, String, Array, Object };
    Kind kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    static Json parse(const std::string &text)
    {
        std::size_t i = 0;
        Json v = value(text, i);
        ws(text, i);
        if (i != text.size()) fail("trailing characters", i);
        return v;
    }

    bool has(const std::string &key) const
    {
        for (const auto &kv : obj)
            if (kv.first == key) return true;
        return false;
    }
    // yaml-cpp's operator[] on a missing key yields an invalid node whose .as<T>() throws; same here.
    const Json &operator[](const std::string &key) const
    {
        for (const auto &kv : obj)
            if (kv.first == key) return kv.second;
        throw std::runtime_error("config: missing key \"" + key + "\"");
    }
    const Json &operator[](std::size_t i) const
    {
        if (kind != Array || i >= arr.size()) throw std::runtime_error("config: bad sequence index");
        return arr[i];
    }
    double asDouble() const
    {
        if (kind != Number) throw std::runtime_error("config: bad conversion to double");
        return num;
    }
    int asInt() const { return static_cast<int>(asDouble()); }
    bool asBool() const
    {
        if (kind != Bool) throw std::runtime_error("config: bad conversion to bool");
        return b;
    }
    const std::string &asString() const
    {
        if (kind != String) throw std::runtime_error("config: bad conversion to string");
        return str;
    }

private:
    [[noreturn]] static void fail(const char *what, std::size_t at)
    {
        throw std::runtime_error(std::string("config: JSON parse error: ") + what + " at offset " + std::to_string(at));
    }
    static void ws(const std::string &t, std::size_t &i)
    {
        while (i < t.size() && (t[i] == ' ' || t[i] == '\n' || t[i] == '\t' || t[i] == '\r')) ++i;
    }
    static std::string string_(const std::string &t, std::size_t &i)
    {
        std::string out;
        ++i; // opening quote
        while (i < t.size() && t[i] != '"') {
            if (t[i] == '\\') {
                if (++i >= t.size()) fail("bad escape", i);
                switch (t[i]) {
                case 'n': out += '\n'; break;
                case 't': out += '\t'; break;
                case 'r': out += '\r'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'u': // config keys are ASCII; keep the escape verbatim
                    out += "\\u";
                    break;
                default: out += t[i];
                }
                ++i;
            } else {
                out += t[i++];
            }
        }
        if (i >= t.size()) fail("unterminated string", i);
        ++i;
        return out;
    }
    static Json value(const std::string &t, std::size_t &i)
    {
        ws(t, i);
        if (i >= t.size()) fail("unexpected end", i);
        Json v;
        const char c = t[i];
        if (c == '{') {
            v.kind = Object;
            ++i;
            ws(t, i);
            if (i < t.size() && t[i] == '}') {
                ++i;
                return v;
            }
            for (;;) {
                ws(t, i);
                if (i >= t.size() || t[i] != '"') fail("expected key", i);
                std::string key = string_(t, i);
                ws(t, i);
                if (i >= t.size() || t[i] != ':') fail("expected ':'", i);
                ++i;
                v.obj.emplace_back(std::move(key), value(t, i));
                ws(t, i);
                if (i < t.size() && t[i] == ',') {
                    ++i;
                    continue;
                }
                if (i < t.size() && t[i] == '}') {
                    ++i;
                    return v;
                }
                fail("expected ',' or '}'", i);
            }
        }
        if (c == '[') {
            v.kind = Array;
            ++i;
            ws(t, i);
            if (i < t.size() && t[i] == ']') {
                ++i;
                return v;
            }
            for (;;) {
                v.arr.push_back(value(t, i));
                ws(t, i);
                if (i < t.size() && t[i] == ',') {
                    ++i;
                    continue;
                }
                if (i < t.size() && t[i] == ']') {
                    ++i;
                    return v;
                }
                fail("expected ',' or ']'", i);
            }
        }
        if (c == '"') {
            v.kind = String;
            v.str = string_(t, i);
            return v;
        }
        if (t.compare(i, 4, "true") == 0) {
            v.kind = Bool;
            v.b = true;
            i += 4;
            return v;
        }
        if (t.compare(i, 5, "false") == 0) {
            v.kind = Bool;
            v.b = false;
            i += 5;
            return v;
        }
        if (t.compare(i, 4, "null") == 0) {
            i += 4;
            return v;
        }
        char *end = nullptr;
        v.num = std::strtod(t.c_str() + i, &end); // correctly rounded, like yaml-cpp's stream extraction
        if (end == t.c_str() + i) fail("unexpected token", i);
        v.kind = Number;
        i = static_cast<std::size_t>(end - t.c_str());
        return v;
    }
};

} // namespace epp
