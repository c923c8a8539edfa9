// Minimal JSON (with C/C++ comments) for the host layer. Stands in for nlohmann::json as used by the
// reference's scene importer: comment stripping mirrors reference src/base/import/json_util.h:12-110.
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace vh {

class Json {
public:
    enum Kind { Null, Bool, Number, String, Array, Object };
    Kind kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;   // insertion order kept (scene order matters)

    bool is_null() const { return kind == Null; }
    bool is_object() const { return kind == Object; }
    bool is_array() const { return kind == Array; }
    bool is_number() const { return kind == Number; }
    bool is_string() const { return kind == String; }
    bool contains(const std::string &k) const { return find(k) != nullptr; }
    const Json *find(const std::string &k) const {
        if (kind != Object) return nullptr;
        for (auto &kv : obj) if (kv.first == k) return &kv.second;
        return nullptr;
    }
    const Json &at(const std::string &k) const {
        static const Json null_json;
        const Json *j = find(k);
        return j ? *j : null_json;
    }
    void set(const std::string &k, Json v) {
        kind = Object;
        for (auto &kv : obj) if (kv.first == k) { kv.second = std::move(v); return; }
        obj.emplace_back(k, std::move(v));
    }
    static Json number(double v) { Json j; j.kind = Number; j.num = v; return j; }
    static Json string(const std::string &s) { Json j; j.kind = String; j.str = s; return j; }
    static Json boolean(bool v) { Json j; j.kind = Bool; j.b = v; return j; }

    static std::string strip_comments(const std::string &s);
    static Json parse(const std::string &text);          // throws std::runtime_error
    static Json parse_file(const std::string &path);     // strips comments first
};

}  // namespace vh
