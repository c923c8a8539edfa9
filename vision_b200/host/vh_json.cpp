#include "vh_json.h"
#include <cstdlib>
#include <fstream>
#include <sstream>

namespace vh {

std::string Json::strip_comments(const std::string &s) {
    std::string out;
    out.reserve(s.size());
    bool in_str = false;
    for (size_t i = 0; i < s.size(); ++i) {
        char c = s[i];
        if (in_str) {
            out.push_back(c);
            if (c == '\\' && i + 1 < s.size()) out.push_back(s[++i]);
            else if (c == '"') in_str = false;
        } else if (c == '"') {
            in_str = true; out.push_back(c);
        } else if (c == '/' && i + 1 < s.size() && s[i + 1] == '/') {
            while (i < s.size() && s[i] != '\n') ++i;
            out.push_back('\n');
        } else if (c == '/' && i + 1 < s.size() && s[i + 1] == '*') {
            size_t e = s.find("*/", i + 2);
            i = (e == std::string::npos) ? s.size() : e + 1;
        } else out.push_back(c);
    }
    return out;
}

namespace {
struct Parser {
    const std::string &s; size_t i = 0;
    explicit Parser(const std::string &t) : s(t) {}
    [[noreturn]] void fail(const char *m) { throw std::runtime_error(std::string("json: ") + m + " at offset " + std::to_string(i)); }
    void ws() { while (i < s.size() && (s[i] == ' ' || s[i] == '\n' || s[i] == '\t' || s[i] == '\r')) ++i; }
    Json value() {
        ws();
        if (i >= s.size()) fail("unexpected end");
        char c = s[i];
        if (c == '{') return object();
        if (c == '[') return array();
        if (c == '"') { Json j; j.kind = Json::String; j.str = string(); return j; }
        if (s.compare(i, 4, "true") == 0) { i += 4; return Json::boolean(true); }
        if (s.compare(i, 5, "false") == 0) { i += 5; return Json::boolean(false); }
        if (s.compare(i, 4, "null") == 0) { i += 4; return Json(); }
        char *end = nullptr;
        double v = std::strtod(s.c_str() + i, &end);
        if (end == s.c_str() + i) fail("bad value");
        i = end - s.c_str();
        return Json::number(v);
    }
    std::string string() {
        std::string o; ++i;
        while (i < s.size() && s[i] != '"') {
            if (s[i] == '\\' && i + 1 < s.size()) {
                char e = s[++i];
                switch (e) {
                    case 'n': o.push_back('\n'); break; case 't': o.push_back('\t'); break;
                    case 'r': o.push_back('\r'); break; case 'b': o.push_back('\b'); break; case 'f': o.push_back('\f'); break;
                    case 'u': { o.push_back('?'); i += 4; break; }
                    default: o.push_back(e);
                }
                ++i;
            } else o.push_back(s[i++]);
        }
        if (i >= s.size()) fail("unterminated string");
        ++i;
        return o;
    }
    Json array() {
        Json j; j.kind = Json::Array; ++i; ws();
        if (i < s.size() && s[i] == ']') { ++i; return j; }
        for (;;) {
            j.arr.push_back(value()); ws();
            if (i < s.size() && s[i] == ',') { ++i; ws(); if (i < s.size() && s[i] == ']') { ++i; return j; } continue; }
            if (i < s.size() && s[i] == ']') { ++i; return j; }
            fail("expected , or ]");
        }
    }
    Json object() {
        Json j; j.kind = Json::Object; ++i; ws();
        if (i < s.size() && s[i] == '}') { ++i; return j; }
        for (;;) {
            ws();
            if (i >= s.size() || s[i] != '"') fail("expected key");
            std::string k = string(); ws();
            if (i >= s.size() || s[i] != ':') fail("expected :");
            ++i;
            j.obj.emplace_back(k, value()); ws();
            if (i < s.size() && s[i] == ',') { ++i; ws(); if (i < s.size() && s[i] == '}') { ++i; return j; } continue; }
            if (i < s.size() && s[i] == '}') { ++i; return j; }
            fail("expected , or }");
        }
    }
};
}  // namespace

Json Json::parse(const std::string &text) {
    Parser p(text);
    Json j = p.value();
    p.ws();
    return j;
}

Json Json::parse_file(const std::string &path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::stringstream ss; ss << f.rdbuf();
    return parse(strip_comments(ss.str()));
}

}  // namespace vh
