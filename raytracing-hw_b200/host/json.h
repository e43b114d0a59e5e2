// Minimal JSON reader for the glTF front end (numbers via strtod, exactly like the json.hpp that tinygltf
// v2.8.10 uses when TINYGLTF_USE_RAPIDJSON is not a compile definition -- SURVEY.md section 2 row 14).
#pragma once
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace rtb {

struct Json {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    bool is_null() const { return type == Null; }
    bool is_object() const { return type == Object; }
    bool is_array() const { return type == Array; }
    bool is_number() const { return type == Number; }

    const Json *find(const std::string &key) const {
        if (type != Object) return nullptr;
        for (auto &kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool has(const std::string &key) const { return find(key) != nullptr; }
    const Json &at(const std::string &key) const {
        const Json *j = find(key);
        if (!j) throw std::runtime_error("[gltf] missing key \"" + key + "\"");
        return *j;
    }
    const Json &at(size_t i) const {
        if (type != Array || i >= arr.size()) throw std::runtime_error("[gltf] array index out of range");
        return arr[i];
    }
    size_t size() const { return type == Array ? arr.size() : (type == Object ? obj.size() : 0); }
    double number_or(const std::string &key, double dflt) const {
        const Json *j = find(key);
        return (j && j->type == Number) ? j->num : dflt;
    }
    int int_or(const std::string &key, int dflt) const {
        const Json *j = find(key);
        return (j && j->type == Number) ? (int) j->num : dflt;
    }
    std::string string_or(const std::string &key, const std::string &dflt) const {
        const Json *j = find(key);
        return (j && j->type == String) ? j->str : dflt;
    }
    std::vector<double> numbers(const std::string &key) const {
        std::vector<double> out;
        const Json *j = find(key);
        if (j && j->type == Array)
            for (auto &e : j->arr) out.push_back(e.num);
        return out;
    }
};

class JsonParser {
public:
    explicit JsonParser(const std::string &text) : s(text), p(0) {}
    Json parse() {
        Json j = value();
        ws();
        if (p != s.size()) fail("trailing characters");
        return j;
    }

private:
    const std::string &s;
    size_t p;

    [[noreturn]] void fail(const char *what) { throw std::runtime_error(std::string("[json] ") + what + " at byte " + std::to_string(p)); }
    void ws() {
        while (p < s.size() && (s[p] == ' ' || s[p] == '\t' || s[p] == '\n' || s[p] == '\r')) ++p;
    }
    Json value() {
        ws();
        if (p >= s.size()) fail("unexpected end");
        char c = s[p];
        if (c == '{') return object();
        if (c == '[') return array();
        if (c == '"') {
            Json j;
            j.type = Json::String;
            j.str = string();
            return j;
        }
        if (c == 't' || c == 'f') {
            Json j;
            j.type = Json::Bool;
            if (s.compare(p, 4, "true") == 0) { j.b = true; p += 4; }
            else if (s.compare(p, 5, "false") == 0) { j.b = false; p += 5; }
            else fail("bad literal");
            return j;
        }
        if (c == 'n') {
            if (s.compare(p, 4, "null") != 0) fail("bad literal");
            p += 4;
            return Json();
        }
        return number();
    }
    Json number() {
        const char *start = s.c_str() + p;
        char *end = nullptr;
        double v = std::strtod(start, &end);
        if (end == start) fail("bad number");
        p += (size_t) (end - start);
        Json j;
        j.type = Json::Number;
        j.num = v;
        return j;
    }
    std::string string() {
        std::string out;
        ++p;  // opening quote
        while (true) {
            if (p >= s.size()) fail("unterminated string");
            char c = s[p++];
            if (c == '"') break;
            if (c == '\\') {
                if (p >= s.size()) fail("bad escape");
                char e = s[p++];
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
                        if (p + 4 > s.size()) fail("bad \\u escape");
                        unsigned cp = (unsigned) std::strtoul(s.substr(p, 4).c_str(), nullptr, 16);
                        p += 4;
                        if (cp < 0x80) out += (char) cp;
                        else if (cp < 0x800) { out += (char) (0xC0 | (cp >> 6)); out += (char) (0x80 | (cp & 0x3F)); }
                        else { out += (char) (0xE0 | (cp >> 12)); out += (char) (0x80 | ((cp >> 6) & 0x3F)); out += (char) (0x80 | (cp & 0x3F)); }
                        break;
                    }
                    default: fail("bad escape");
                }
            } else {
                out += c;
            }
        }
        return out;
    }
    Json array() {
        Json j;
        j.type = Json::Array;
        ++p;
        ws();
        if (p < s.size() && s[p] == ']') { ++p; return j; }
        while (true) {
            j.arr.push_back(value());
            ws();
            if (p >= s.size()) fail("unterminated array");
            if (s[p] == ',') { ++p; continue; }
            if (s[p] == ']') { ++p; break; }
            fail("expected , or ]");
        }
        return j;
    }
    Json object() {
        Json j;
        j.type = Json::Object;
        ++p;
        ws();
        if (p < s.size() && s[p] == '}') { ++p; return j; }
        while (true) {
            ws();
            if (p >= s.size() || s[p] != '"') fail("expected string key");
            std::string k = string();
            ws();
            if (p >= s.size() || s[p] != ':') fail("expected :");
            ++p;
            j.obj.emplace_back(std::move(k), value());
            ws();
            if (p >= s.size()) fail("unterminated object");
            if (s[p] == ',') { ++p; continue; }
            if (s[p] == '}') { ++p; break; }
            fail("expected , or }");
        }
        return j;
    }
};

}  // namespace rtb
