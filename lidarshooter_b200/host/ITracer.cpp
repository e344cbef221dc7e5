// ITracer.cpp -- stand-in for the reference's SRC/ITracer.cpp:3-26 (constructor: logger lookup, trace cloud ownership).
// Not part of the drop-in: a lidarshooter build links its own ITracer.cpp.
#include "ITracer.hpp"

namespace lidarshooter {

ITracer::ITracer(LidarDevice::Ptr _sensorConfig, sensor_msgs::PointCloud2::Ptr _traceStorage, std::shared_ptr<spdlog::logger> __logger)
    : _config(_sensorConfig), _geometryCount(0)
{
    // SRC/ITracer.cpp:5-14: take the given logger, else the application's, else create it
    if (__logger == nullptr) {
        _logger = spdlog::get(LIDARSHOOTER_APPLICATION_NAME);
        if (_logger == nullptr) _logger = spdlog::stdout_color_mt(LIDARSHOOTER_APPLICATION_NAME);
    } else {
        _logger = __logger;
    }
    // SRC/ITracer.cpp:16-25: own the cloud when none is given; zero width/height
    _traceCloud = _traceStorage ? _traceStorage : sensor_msgs::PointCloud2::Ptr(new sensor_msgs::PointCloud2());
    _traceCloud->width = 0;
    _traceCloud->height = 0;
}

}  // namespace lidarshooter
