// B200Tracer.hpp -- the drop-in tracer: `class B200Tracer : public ITracer`, created exactly like the
// reference's back-ends (EmbreeTracer::create / OptixTracer::create, SRC/EmbreeTracer.hpp:51, SRC/OptixTracer.hpp:91)
// and selected by the one line at SRC/mainwindow.cpp:258.  A thin marshaller: every virtual forwards to the C ABI
// of include/lsb200.h; all arithmetic happens in the CUDA library.  No CPU fallback: construction throws when the
// CUDA context cannot be created.
#pragma once
#include <map>
#include <memory>
#include <string>

#include "ITracer.hpp"

struct lsb200_ctx;

namespace lidarshooter {

class B200Tracer : public ITracer, public std::enable_shared_from_this<B200Tracer> {
public:
    using Ptr = std::shared_ptr<B200Tracer>;
    using ConstPtr = std::shared_ptr<B200Tracer const>;

    static B200Tracer::Ptr create(LidarDevice::Ptr _sensorConfig, PointCloud2::Ptr _traceStorage = nullptr, int _cudaDevice = 0);
    ITracer::Ptr getPtr() override { return shared_from_this(); }
    ~B200Tracer() override;

    int getGeometryId(const std::string &_meshName) const;                 // -1 if unknown (SRC/EmbreeTracer.cpp:80-89)
    RTCGeometryType getGeometryType(const std::string &_meshName);         // throws TraceException code 8
    long getVertexCount(const std::string &_meshName);                     // throws TraceException code 1
    long getElementCount(const std::string &_meshName);                    // throws TraceException code 4

    int addGeometry(const std::string &_meshName, enum RTCGeometryType _geometryType, int _numVertices, int _numElements) override;
    int removeGeometry(const std::string &_meshName) override;
    int updateGeometry(const std::string &_meshName, Affine3f _transform, PolygonMesh::Ptr &_mesh) override;
    int updateGeometry(const std::string &_meshName, Vector3f _translation, Vector3f _rotation, PolygonMesh::Ptr &_mesh) override;
    int commitScene() override;
    int traceScene(std::uint32_t _frameIndex) override;

    // extensions that do not exist in the reference
    void setOrderedOutput(bool _on) { _ordered = _on; }       // points in ray-id order (deterministic)
    void setRingZeroCompat(bool _on) { _ringZero = _on; }     // ring = 0 like the Embree path (SRC/EmbreeTracer.cpp:333-334)
    void setWorldFramePoints(bool _on) { _worldPoints = _on; } // emit R p + T (CloudTransformer, SRC/CloudTransformer.cpp:202-265)
    // apply the device config's detection_radius / horizontal_angle_range (lsb200_set_sensor_limits); off = reference behaviour
    int setSensorLimits(bool _on);
    int traceRaw(std::vector<float> &_t, std::vector<int32_t> &_tri);   // per-ray results (parity tests)
    lsb200_ctx *context() { return _ctx; }
    std::string lastError() const;

    static Affine3f transformFromComponents(Vector3f _translation, Vector3f _rotation);   // SRC/MeshTransformer.cpp:467-477

private:
    B200Tracer(LidarDevice::Ptr _sensorConfig, PointCloud2::Ptr _traceStorage, int _cudaDevice);
    struct Geom { RTCGeometryType type; long nVerts, nElems; bool topologySent; };
    int update(const std::string &_meshName, const float A[12], PolygonMesh::Ptr &_mesh);

    lsb200_ctx *_ctx = nullptr;
    std::map<std::string, Geom> _geoms;
    bool _ordered = false, _ringZero = false, _worldPoints = false;
    std::vector<uint32_t> _elemScratch;
};

}  // namespace lidarshooter
