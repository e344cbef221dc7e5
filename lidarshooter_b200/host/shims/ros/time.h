// shim: ros::Time (whole seconds + nanoseconds)
#pragma once
#include <chrono>
#include <cstdint>
namespace ros {
class Time {
public:
    std::uint32_t sec = 0, nsec = 0;
    Time() = default;
    Time(std::uint32_t s, std::uint32_t ns) : sec(s), nsec(ns) {}
    explicit Time(double t)
    {
        const double whole = t < 0.0 ? 0.0 : (double)(std::uint64_t)t;
        double frac = (t < 0.0 ? 0.0 : t - whole) * 1e9 + 0.5;
        sec = (std::uint32_t)whole; nsec = (std::uint32_t)frac;
        if (nsec >= 1000000000u) { nsec -= 1000000000u; ++sec; }
    }
    double toSec() const { return (double)sec + 1e-9 * (double)nsec; }
    static Time now()
    {
        const long long ns = std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::system_clock::now().time_since_epoch()).count();
        return Time((std::uint32_t)(ns / 1000000000LL), (std::uint32_t)(ns % 1000000000LL));
    }
    static void init() {}
};
}  // namespace ros
