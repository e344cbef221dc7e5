#include "Json.hpp"

#include <cctype>
#include <cstdlib>
#include <fstream>
#include <sstream>

namespace lidarshooter {
namespace json {

static const Value kNull;

const Value &Value::operator[](const std::string &key) const
{
    if (_type != Object) return kNull;
    auto it = _object.find(key);
    return it == _object.end() ? kNull : it->second;
}

const Value &Value::operator[](size_t index) const
{
    if (_type != Array || index >= _array.size()) return kNull;
    return _array[index];
}

class Parser {
public:
    Parser(const std::string &text) : s(text) {}
    bool run(Value &out, std::string &err)
    {
        try {
            skip();
            out = value();
            skip();
            if (i != s.size()) fail("trailing characters");
            return true;
        } catch (const std::string &e) {
            err = e;
            return false;
        }
    }

private:
    const std::string &s;
    size_t i = 0;
    int line = 1;

    [[noreturn]] void fail(const std::string &what) { throw "json: " + what + " at line " + std::to_string(line); }

    void skip()
    {
        while (i < s.size()) {
            const char c = s[i];
            if (c == '\n') { ++line; ++i; }
            else if (std::isspace(static_cast<unsigned char>(c))) ++i;
            else if (c == '/' && i + 1 < s.size() && s[i + 1] == '/') { while (i < s.size() && s[i] != '\n') ++i; }
            else if (c == '/' && i + 1 < s.size() && s[i + 1] == '*') {
                i += 2;
                while (i + 1 < s.size() && !(s[i] == '*' && s[i + 1] == '/')) { if (s[i] == '\n') ++line; ++i; }
                if (i + 1 >= s.size()) fail("unterminated comment");
                i += 2;
            } else break;
        }
    }

    Value value()
    {
        skip();
        if (i >= s.size()) fail("unexpected end of input");
        const char c = s[i];
        if (c == '{') return object();
        if (c == '[') return array();
        if (c == '"') { Value v; v._type = Value::String; v._string = string(); return v; }
        if (c == 't' || c == 'f' || c == 'n') return literal();
        return number();
    }

    Value object()
    {
        Value v; v._type = Value::Object;
        ++i; skip();
        if (i < s.size() && s[i] == '}') { ++i; return v; }
        while (true) {
            skip();
            if (i >= s.size() || s[i] != '"') fail("expected a member name");
            std::string key = string();
            skip();
            if (i >= s.size() || s[i] != ':') fail("expected ':'");
            ++i;
            v._object[key] = value();
            skip();
            if (i < s.size() && s[i] == ',') { ++i; continue; }
            if (i < s.size() && s[i] == '}') { ++i; return v; }
            fail("expected ',' or '}'");
        }
    }

    Value array()
    {
        Value v; v._type = Value::Array;
        ++i; skip();
        if (i < s.size() && s[i] == ']') { ++i; return v; }
        while (true) {
            v._array.push_back(value());
            skip();
            if (i < s.size() && s[i] == ',') { ++i; continue; }
            if (i < s.size() && s[i] == ']') { ++i; return v; }
            fail("expected ',' or ']'");
        }
    }

    std::string string()
    {
        std::string out;
        ++i;
        while (i < s.size() && s[i] != '"') {
            char c = s[i++];
            if (c == '\\') {
                if (i >= s.size()) fail("bad escape");
                const char e = s[i++];
                switch (e) {
                case 'n': out += '\n'; break;
                case 't': out += '\t'; break;
                case 'r': out += '\r'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'u': {
                    if (i + 4 > s.size()) fail("bad \\u escape");
                    const unsigned cp = static_cast<unsigned>(std::strtoul(s.substr(i, 4).c_str(), nullptr, 16));
                    i += 4;
                    if (cp < 0x80) out += static_cast<char>(cp);
                    else if (cp < 0x800) { out += static_cast<char>(0xC0 | (cp >> 6)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
                    else { out += static_cast<char>(0xE0 | (cp >> 12)); out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
                    break;
                }
                default: out += e;
                }
            } else {
                if (c == '\n') ++line;
                out += c;
            }
        }
        if (i >= s.size()) fail("unterminated string");
        ++i;
        return out;
    }

    Value literal()
    {
        Value v;
        if (s.compare(i, 4, "true") == 0) { v._type = Value::Bool; v._bool = true; i += 4; }
        else if (s.compare(i, 5, "false") == 0) { v._type = Value::Bool; v._bool = false; i += 5; }
        else if (s.compare(i, 4, "null") == 0) { i += 4; }
        else fail("bad literal");
        return v;
    }

    Value number()
    {
        const char *b = s.c_str() + i;
        char *e = nullptr;
        const double d = std::strtod(b, &e);
        if (e == b) fail("bad number");
        i += static_cast<size_t>(e - b);
        Value v; v._type = Value::Number; v._number = d;
        return v;
    }
};

bool parse(const std::string &text, Value &out, std::string &error) { return Parser(text).run(out, error); }

bool parseFile(const std::string &path, Value &out, std::string &error)
{
    std::ifstream f(path);
    if (!f) { error = "cannot open " + path; return false; }
    std::stringstream ss;
    ss << f.rdbuf();
    return parse(ss.str(), out, error);
}

}  // namespace json
}  // namespace lidarshooter
