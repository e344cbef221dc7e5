#include "B200Tracer.hpp"

#include <cstring>

#include "lsb200.h"

namespace lidarshooter {

B200Tracer::Ptr B200Tracer::create(LidarDevice::Ptr _sensorConfig, sensor_msgs::PointCloud2::Ptr _traceStorage,
                                   std::shared_ptr<spdlog::logger> _logger, int _cudaDevice)
{
    return B200Tracer::Ptr(new B200Tracer(_sensorConfig, _traceStorage, _logger, _cudaDevice));
}

// the library's log sink -> the spdlog logger every reference tracer logs to (levels are spdlog's)
void B200Tracer::logThunk(int _level, const char *_message, void *_self)
{
    B200Tracer *self = static_cast<B200Tracer *>(_self);
    if (self->_logger) self->_logger->log(static_cast<spdlog::level::level_enum>(_level), std::string(_message));
}

B200Tracer::B200Tracer(LidarDevice::Ptr _sensorConfig, sensor_msgs::PointCloud2::Ptr _traceStorage,
                       std::shared_ptr<spdlog::logger> __logger, int _cudaDevice)
    : ITracer(_sensorConfig, _traceStorage, __logger)
{
    const int rc = lsb200_create(_cudaDevice, &_ctx);
    if (rc != LSB200_OK || !_ctx) {
        _logger->error("B200Tracer: lsb200_create failed: no usable CUDA device (there is no CPU path)");
        throw TraceException(__FILE__, "lsb200_create failed: no usable CUDA device (B200Tracer has no CPU path)", rc);
    }
    lsb200_set_log_callback(_ctx, &B200Tracer::logThunk, this);
    LidarDevice::Ptr dev = getSensorConfig();
    if (dev) {
        // the four LidarDevice getters of INTEGRATION.md: channel table, azimuth range, quaternion, translation
        float q[4], t[3];
        dev->quaternion(q);
        dev->translation(t);
        const auto &vert = dev->verticalAngles();
        const int src = lsb200_set_sensor(_ctx, 0, vert.data(), static_cast<int>(vert.size()), dev->azimuthBegin(), dev->azimuthEnd(),
                                          static_cast<int>(dev->getScanRayCount()), q, t);
        if (src != LSB200_OK) {
            const std::string msg = lsb200_last_error(_ctx);
            lsb200_destroy(_ctx); _ctx = nullptr;
            throw TraceException(__FILE__, "lsb200_set_sensor failed: " + msg, src);
        }
        _logger->info("B200Tracer: " + dev->getSensorUid() + " on CUDA device " + std::to_string(_cudaDevice) + ", " +
                      std::to_string(dev->getTotalRays()) + " rays per scan");
    }
}

int B200Tracer::setSensorLimits(bool _on)
{
    float lo = -180.0f, hi = 180.0f;
    auto dev = getSensorConfig();
    if (_on) dev->horizontalAngleRange(lo, hi);
    return lsb200_set_sensor_limits(_ctx, 0, _on ? dev->detectionRadius() : 0.0f, lo, hi);
}

int B200Tracer::setIntensity(int _model, bool _filter)
{
    int rc = lsb200_set_option(_ctx, LSB200_OPT_INTENSITY_MODEL, _model);
    if (rc != LSB200_OK) return rc;
    auto dev = getSensorConfig();
    return lsb200_set_sensor_intensity_filter(_ctx, 0, (_filter && dev->useIntensityFiltering()) ? dev->minimumIntensity() : 0.0f);
}

int B200Tracer::setWatertight(bool _on) { return lsb200_set_option(_ctx, LSB200_OPT_WATERTIGHT, _on ? 1 : 0); }

B200Tracer::~B200Tracer()
{
    if (_ctx) lsb200_destroy(_ctx);
}

std::string B200Tracer::lastError() const { return lsb200_last_error(_ctx); }

int B200Tracer::getGeometryId(const std::string &_meshName) const { return lsb200_get_geometry_id(_ctx, _meshName.c_str()); }

RTCGeometryType B200Tracer::getGeometryType(const std::string &_meshName)
{
    auto it = _geoms.find(_meshName);
    if (it == _geoms.end()) throw TraceException(__FILE__, "Geometry key does not exist in geometry types map", 8);
    return it->second.type;
}

long B200Tracer::getVertexCount(const std::string &_meshName)
{
    auto it = _geoms.find(_meshName);
    if (it == _geoms.end()) throw TraceException(__FILE__, "Geometry key does not exist in vertex count map", 1);
    return it->second.nVerts;
}

long B200Tracer::getElementCount(const std::string &_meshName)
{
    auto it = _geoms.find(_meshName);
    if (it == _geoms.end()) throw TraceException(__FILE__, "Geometry key does not exist in element count map", 4);
    return it->second.nElems;
}

int B200Tracer::addGeometry(const std::string &_meshName, enum RTCGeometryType _geometryType, int _numVertices, int _numElements)
{
    // SRC/EmbreeTracer.cpp:115-218: TRIANGLE and QUAD are supported, everything else "returns false"
    if (_geometryType != RTC_GEOMETRY_TYPE_TRIANGLE && _geometryType != RTC_GEOMETRY_TYPE_QUAD) return 0;
    const int id = lsb200_add_geometry(_ctx, _meshName.c_str(), static_cast<int>(_geometryType), _numVertices, _numElements);
    if (id < 0) return id;
    _geoms[_meshName] = Geom{_geometryType, _numVertices, _numElements, false};
    setGeometryCount(getGeometryCount() + 1);
    return id;
}

int B200Tracer::removeGeometry(const std::string &_meshName)
{
    const int id = lsb200_remove_geometry(_ctx, _meshName.c_str());
    if (id < 0) return -1;                                     // SRC/EmbreeTracer.cpp:224-225
    _geoms.erase(_meshName);
    setGeometryCount(getGeometryCount() - 1);
    return id;
}

Eigen::Affine3f B200Tracer::transformFromComponents(Eigen::Vector3f _translation, Eigen::Vector3f _rotation)
{
    const float t[3] = {_translation.x(), _translation.y(), _translation.z()}, r[3] = {_rotation.x(), _rotation.y(), _rotation.z()};
    float rows[12];
    lsb200_affine_from_components(t, r, rows);
    Eigen::Affine3f A = Eigen::Affine3f::Identity();
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 4; ++j) A(i, j) = rows[4 * i + j];
    return A;
}

int B200Tracer::update(const std::string &_meshName, const float A[12], pcl::PolygonMesh::Ptr &_mesh)
{
    if (!_mesh) throw MeshNotSetException(__FILE__, "No mesh data to transform", 1);       // SRC/MeshTransformer.cpp:144-151
    auto it = _geoms.find(_meshName);
    if (it == _geoms.end()) throw TraceException(__FILE__, "Geometry key does not exist in geometry types map", 8);
    Geom &g = it->second;
    const int arity = g.type == RTC_GEOMETRY_TYPE_QUAD ? 4 : 3;
    const size_t meshVerts = _mesh->cloud.point_step ? _mesh->cloud.data.size() / _mesh->cloud.point_step : 0;
    if (static_cast<long>(meshVerts) < g.nVerts || static_cast<long>(_mesh->polygons.size()) < g.nElems)
        return -1;
    const uint32_t *elems = nullptr;
    if (!g.topologySent) {
        // index copy (SRC/MeshTransformer.cpp:486-552), once: the reference repeats it every frame (its TODO at :153)
        _elemScratch.resize(static_cast<size_t>(g.nElems) * arity);
        for (long e = 0; e < g.nElems; ++e) {
            const auto &poly = _mesh->polygons[e].vertices;
            if (static_cast<int>(poly.size()) != arity)
                throw BadGeometryException(__FILE__, "Wrong number of vertices in an element", 1, arity, static_cast<long>(poly.size()));
            for (int k = 0; k < arity; ++k)      // pcl::index_t is signed in PCL 1.12: converted, not memcpy'd
                _elemScratch[static_cast<size_t>(e) * arity + k] = static_cast<uint32_t>(poly[k]);
        }
        elems = _elemScratch.data();
    }
    const int rc = lsb200_update_geometry(_ctx, _meshName.c_str(), _mesh->cloud.data.data(), static_cast<int>(_mesh->cloud.point_step), elems, A);
    if (rc == LSB200_OK) g.topologySent = true;
    return rc;
}

int B200Tracer::updateGeometry(const std::string &_meshName, Eigen::Affine3f _transform, pcl::PolygonMesh::Ptr &_mesh)
{
    float rows[12];                                  // row-major 3x4 [L | t]
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 4; ++j) rows[4 * i + j] = _transform(i, j);
    return update(_meshName, rows, _mesh);
}

int B200Tracer::updateGeometry(const std::string &_meshName, Eigen::Vector3f _translation, Eigen::Vector3f _rotation, pcl::PolygonMesh::Ptr &_mesh)
{
    const float t[3] = {_translation.x(), _translation.y(), _translation.z()}, r[3] = {_rotation.x(), _rotation.y(), _rotation.z()};
    float rows[12];
    lsb200_affine_from_components(t, r, rows);       // T * Rz * Ry * Rx (SRC/MeshTransformer.cpp:467-477)
    return update(_meshName, rows, _mesh);
}

int B200Tracer::commitScene()
{
    return lsb200_commit_scene(_ctx, LSB200_COMMIT_AUTO);
}

int B200Tracer::traceScene(std::uint32_t _frameIndex)
{
    // SRC/EmbreeTracer.cpp:297-367: header, clear, trace, width = number of hits.  The cloud is rewritten in place.
    sensor_msgs::PointCloud2::Ptr cloud = getTraceCloud();
    getSensorConfig()->initMessage(cloud, static_cast<int>(_frameIndex));
    const size_t capacity = getSensorConfig()->getTotalRays();
    cloud->data.resize(capacity * 32);
    uint32_t n = 0;
    uint32_t flags = 0;
    if (_ordered) flags |= LSB200_TRACE_ORDERED;
    if (!_ringChannel) flags |= LSB200_TRACE_RING_ZERO;          // default: the Embree path's ring = 0 (SRC/EmbreeTracer.cpp:333-334)
    if (_worldPoints) flags |= LSB200_TRACE_WORLD_POINTS;
    const int rc = lsb200_trace_scene(_ctx, 0, _frameIndex, cloud->data.data(), capacity, &n, flags);
    cloud->data.resize(static_cast<size_t>(n) * 32);
    cloud->width = n;
    return rc;
}

int B200Tracer::traceRaw(std::vector<float> &_t, std::vector<int32_t> &_tri)
{
    const size_t R = getSensorConfig()->getTotalRays();
    _t.resize(R);
    _tri.resize(R);
    return lsb200_trace_raw(_ctx, 0, _t.data(), _tri.data());
}

}  // namespace lidarshooter
