// ITracer.hpp -- the tracer interface of the reference, re-declared 1:1 (SRC/ITracer.hpp:29-152) over this
// repository's stand-in types.  B200Tracer implements it; MeshProjector::traceAffineMesh's call sequence
// (updateGeometry* -> commitScene -> traceScene, SRC/MeshProjector.cpp:446-471) works unchanged against it.
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>

#include "LidarDevice.hpp"
#include "Types.hpp"

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

    ITracer(LidarDevice::Ptr _sensorConfig, PointCloud2::Ptr _traceStorage = nullptr);
    virtual ITracer::Ptr getPtr() = 0;
    virtual ~ITracer() = default;

    virtual int addGeometry(const std::string &_meshName, enum RTCGeometryType _geometryType, int _numVertices, int _numElements) = 0;
    virtual int removeGeometry(const std::string &_meshName) = 0;
    virtual int updateGeometry(const std::string &_meshName, Affine3f _transform, PolygonMesh::Ptr &_mesh) = 0;
    virtual int updateGeometry(const std::string &_meshName, Vector3f _translation, Vector3f _rotation, PolygonMesh::Ptr &_mesh) = 0;
    virtual int commitScene() = 0;
    virtual int traceScene(std::uint32_t _frameIndex) = 0;
    virtual long getGeometryCount() const { return _geometryCount; }

    PointCloud2::Ptr getTraceCloud() { return _traceCloud; }
    void setTraceCloud(PointCloud2::Ptr _traceStorage) { _traceCloud = _traceStorage; }
    LidarDevice::Ptr getSensorConfig() { return _config; }
    void setSensorConfig(LidarDevice::Ptr __config) { _config = __config; }

protected:
    void setGeometryCount(long _count) { _geometryCount = _count; }

private:
    LidarDevice::Ptr _config;
    long _geometryCount;
    PointCloud2::Ptr _traceCloud;
};

}  // namespace lidarshooter
