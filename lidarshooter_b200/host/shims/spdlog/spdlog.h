// shim: the spdlog calls the tracers make (SRC/ITracer.cpp:5-14: get / stdout_color_mt; logger->info/debug/warn/error).
// The shim logger keeps its lines so that tests can read what the adapter logged.
#pragma once
#include <cstdio>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <utility>
#include <vector>
namespace spdlog {
namespace level {
enum level_enum { trace = 0, debug = 1, info = 2, warn = 3, err = 4, critical = 5, off = 6 };
}
class logger {
public:
    explicit logger(std::string name) : _name(std::move(name)) {}
    const std::string &name() const { return _name; }
    void set_level(level::level_enum l) { _level = l; }
    void log(level::level_enum lvl, const std::string &msg)
    {
        std::lock_guard<std::mutex> lk(_m);
        _lines.emplace_back((int)lvl, msg);
        if (lvl >= _level && _echo) std::fprintf(stderr, "[%s] [%d] %s\n", _name.c_str(), (int)lvl, msg.c_str());
    }
    void trace(const std::string &m) { log(level::trace, m); }
    void debug(const std::string &m) { log(level::debug, m); }
    void info(const std::string &m) { log(level::info, m); }
    void warn(const std::string &m) { log(level::warn, m); }
    void error(const std::string &m) { log(level::err, m); }
    // shim-only: what was logged (tests)
    std::vector<std::pair<int, std::string>> lines() { std::lock_guard<std::mutex> lk(_m); return _lines; }
    void set_echo(bool on) { _echo = on; }
private:
    std::string _name;
    level::level_enum _level = level::warn;
    bool _echo = true;
    std::mutex _m;
    std::vector<std::pair<int, std::string>> _lines;
};
inline std::map<std::string, std::shared_ptr<logger>> &registry_() { static std::map<std::string, std::shared_ptr<logger>> r; return r; }
inline std::shared_ptr<logger> get(const std::string &name)
{
    auto it = registry_().find(name);
    return it == registry_().end() ? nullptr : it->second;
}
inline std::shared_ptr<logger> stdout_color_mt(const std::string &name)
{
    auto p = std::make_shared<logger>(name);
    registry_()[name] = p;
    return p;
}
}  // namespace spdlog
