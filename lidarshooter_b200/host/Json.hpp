// Json.hpp -- minimal JSON value + parser for the reference's config files.
// The reference parses with jsoncpp, comments allowed (SRC/LidarDevice.cpp:485-492); its shipped configs contain
// `//` comments (config/hesai-pandar-XT-32-lidar_0000.json:78,92) and string values with "//" inside
// ("http://sensoraddr", :4), so comments are skipped by the tokenizer, never by text substitution.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace lidarshooter {
namespace json {

class Value {
public:
    enum Type { Null, Bool, Number, String, Array, Object };
    Value() = default;
    Type type() const { return _type; }
    bool isNull() const { return _type == Null; }
    bool isObject() const { return _type == Object; }
    bool isArray() const { return _type == Array; }
    bool isMember(const std::string &key) const { return _type == Object && _object.count(key) != 0; }
    // object access; a missing key yields a Null value (jsoncpp semantics)
    const Value &operator[](const std::string &key) const;
    const Value &operator[](size_t index) const;
    size_t size() const { return _type == Array ? _array.size() : (_type == Object ? _object.size() : 0); }
    const std::vector<Value> &items() const { return _array; }
    // typed getters with defaults (jsoncpp's get(key, default).asX())
    double asDouble(double dflt = 0.0) const { return _type == Number ? _number : (_type == Bool ? (_bool ? 1.0 : 0.0) : dflt); }
    float asFloat(float dflt = 0.0f) const { return _type == Number ? static_cast<float>(_number) : dflt; }
    int asInt(int dflt = 0) const { return _type == Number ? static_cast<int>(_number) : dflt; }
    unsigned asUInt(unsigned dflt = 0) const { return _type == Number ? static_cast<unsigned>(_number) : dflt; }
    bool asBool(bool dflt = false) const { return _type == Bool ? _bool : (_type == Number ? _number != 0.0 : dflt); }
    std::string asString(const std::string &dflt = "") const { return _type == String ? _string : dflt; }

private:
    friend class Parser;
    Type _type = Null;
    bool _bool = false;
    double _number = 0.0;
    std::string _string;
    std::vector<Value> _array;
    std::map<std::string, Value> _object;
};

// Parses `text`; returns false and fills `error` (with line number) on malformed input.
bool parse(const std::string &text, Value &out, std::string &error);
bool parseFile(const std::string &path, Value &out, std::string &error);

}  // namespace json
}  // namespace lidarshooter
