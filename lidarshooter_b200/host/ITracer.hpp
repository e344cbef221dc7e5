// ITracer.hpp -- the tracer interface of the reference, re-declared 1:1 (SRC/ITracer.hpp:29-152; ctor SRC/ITracer.cpp:
// 3-26) so that B200Tracer.{hpp,cpp} -- the file a maintainer drops next to EmbreeTracer.hpp -- compiles here exactly as
// it does against the reference's own ITracer.hpp: same signatures, same third-party type names (Eigen, pcl,
// sensor_msgs, spdlog, RTCGeometryType), same protected `_logger`.  MeshProjector::traceAffineMesh's call sequence
// (updateGeometry* -> commitScene -> traceScene, SRC/MeshProjector.cpp:446-471) works unchanged against it.
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>

#include <Eigen/Dense>
#include <Eigen/Geometry>
#include <embree3/rtcore.h>
#include <pcl/PolygonMesh.h>
#include <sensor_msgs/PointCloud2.h>
#include <spdlog/spdlog.h>
#include <spdlog/sinks/stdout_color_sinks.h>

#include "LidarDevice.hpp"

#ifndef LIDARSHOOTER_APPLICATION_NAME
#define LIDARSHOOTER_APPLICATION_NAME "LiDARShooter"     // SRC/LidarShooter.hpp.in (cmake-configured in the reference)
#endif

namespace lidarshooter {

// error conventions of the reference (SRC/Exceptions.hpp:62-176)
class TraceException : public std::runtime_error {
public:
    TraceException(const std::string &where, const std::string &what, long code)
        : std::runtime_error(where + ": " + what), _code(code) {}
    long getErrorCode() const { return _code; }
private:
    long _code;
};
class BadGeometryException : public std::runtime_error {
public:
    BadGeometryException(const std::string &where, const std::string &what, long code, long expected, long got)
        : std::runtime_error(where + ": " + what + " (expected " + std::to_string(expected) + ", got " + std::to_string(got) + ")"), _code(code) {}
    long getErrorCode() const { return _code; }
private:
    long _code;
};
class MeshNotSetException : public std::runtime_error {
public:
    MeshNotSetException(const std::string &where, const std::string &what, long code)
        : std::runtime_error(where + ": " + what), _code(code) {}
    long getErrorCode() const { return _code; }
private:
    long _code;
};

class ITracer {
public:
    using Ptr = std::shared_ptr<ITracer>;
    using ConstPtr = std::shared_ptr<ITracer const>;

    ITracer(LidarDevice::Ptr _sensorConfig, sensor_msgs::PointCloud2::Ptr _traceStorage = nullptr, std::shared_ptr<spdlog::logger> _logger = nullptr);
    virtual ITracer::Ptr getPtr() = 0;
    virtual ~ITracer() = default;

    virtual int addGeometry(const std::string &_meshName, enum RTCGeometryType _geometryType, int _numVertices, int _numElements) = 0;
    virtual int removeGeometry(const std::string &_meshName) = 0;
    virtual int updateGeometry(const std::string &_meshName, Eigen::Affine3f _transform, pcl::PolygonMesh::Ptr &_mesh) = 0;
    virtual int updateGeometry(const std::string &_meshName, Eigen::Vector3f _translation, Eigen::Vector3f _rotation, pcl::PolygonMesh::Ptr &_mesh) = 0;
    virtual int commitScene() = 0;
    virtual int traceScene(std::uint32_t _frameIndex) = 0;
    virtual long getGeometryCount() const { return _geometryCount; }

    sensor_msgs::PointCloud2::Ptr getTraceCloud() { return _traceCloud; }
    void setTraceCloud(sensor_msgs::PointCloud2::Ptr _traceStorage) { _traceCloud = _traceStorage; }
    LidarDevice::Ptr getSensorConfig() { return _config; }
    void setSensorConfig(LidarDevice::Ptr __config) { _config = __config; }

protected:
    void setGeometryCount(long _count) { _geometryCount = _count; }
    std::shared_ptr<spdlog::logger> _logger;       // SRC/ITracer.hpp:139

private:
    LidarDevice::Ptr _config;
    long _geometryCount;
    sensor_msgs::PointCloud2::Ptr _traceCloud;
};

}  // namespace lidarshooter
