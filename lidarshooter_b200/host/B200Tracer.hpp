// B200Tracer.hpp -- THE drop-in tracer: add this file and B200Tracer.cpp next to EmbreeTracer.hpp in
// ros_ws/src/lidarshooter/src/ and change the one line at SRC/mainwindow.cpp:258 to
//
//     ITracer::Ptr tracer = B200Tracer::create(device, nullptr, loggerTop);
//
// `class B200Tracer : public ITracer`, created exactly like the reference's back-ends (EmbreeTracer::create /
// OptixTracer::create, SRC/EmbreeTracer.hpp:51, SRC/OptixTracer.hpp:91).  A thin marshaller: every virtual forwards to
// the C ABI of include/lsb200.h (liblsb200.so); all arithmetic happens in the CUDA library.  No CPU fallback:
// construction throws when the CUDA context cannot be created.
//
// This same source is what this repository builds and tests (liblsb200_host.so, tests/test_gpu_host.py): here against
// host/shims/ and host/ITracer.hpp + host/LidarDevice.hpp (1:1 re-declarations of the reference's headers), in a
// lidarshooter tree against the real Eigen / PCL / ROS / spdlog / Embree headers and the reference's own ITracer.hpp.
// One requirement on the reference: LidarDevice must expose the numbers it parsed -- see "LidarDevice getters" in
// INTEGRATION.md (four const accessors; the reference keeps them private, SRC/LidarDevice.hpp:302-440).
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "ITracer.hpp"

struct lsb200_ctx;

namespace lidarshooter {

class B200Tracer : public ITracer, public std::enable_shared_from_this<B200Tracer> {
public:
    using Ptr = std::shared_ptr<B200Tracer>;
    using ConstPtr = std::shared_ptr<B200Tracer const>;

    // same parameter list as EmbreeTracer::create (SRC/EmbreeTracer.hpp:51) + an optional CUDA device ordinal
    static B200Tracer::Ptr create(LidarDevice::Ptr _sensorConfig, sensor_msgs::PointCloud2::Ptr _traceStorage = nullptr,
                                  std::shared_ptr<spdlog::logger> _logger = nullptr, int _cudaDevice = 0);
    ITracer::Ptr getPtr() override { return shared_from_this(); }
    ~B200Tracer() override;

    int getGeometryId(const std::string &_meshName) const;                 // -1 if unknown (SRC/EmbreeTracer.cpp:80-89)
    RTCGeometryType getGeometryType(const std::string &_meshName);         // throws TraceException code 8
    long getVertexCount(const std::string &_meshName);                     // throws TraceException code 1
    long getElementCount(const std::string &_meshName);                    // throws TraceException code 4

    int addGeometry(const std::string &_meshName, enum RTCGeometryType _geometryType, int _numVertices, int _numElements) override;
    int removeGeometry(const std::string &_meshName) override;
    int updateGeometry(const std::string &_meshName, Eigen::Affine3f _transform, pcl::PolygonMesh::Ptr &_mesh) override;
    int updateGeometry(const std::string &_meshName, Eigen::Vector3f _translation, Eigen::Vector3f _rotation, pcl::PolygonMesh::Ptr &_mesh) override;
    int commitScene() override;
    int traceScene(std::uint32_t _frameIndex) override;

    // extensions that do not exist in the reference (all off by default: the default cloud bytes are the Embree path's)
    void setOrderedOutput(bool _on) { _ordered = _on; }       // points in ray-id order (deterministic)
    // ring field: the Embree path writes 0 for every point (SRC/EmbreeTracer.cpp:333-334) -- the default here; the OptiX
    // path writes the channel index (SRC/LidarDeviceKernels.cu:51) -- setRingChannel(true)
    void setRingChannel(bool _on) { _ringChannel = _on; }
    void setWorldFramePoints(bool _on) { _worldPoints = _on; } // emit R p + T (CloudTransformer, SRC/CloudTransformer.cpp:202-265)
    // apply the device config's detection_radius / horizontal_angle_range (lsb200_set_sensor_limits); off = reference behaviour
    int setSensorLimits(bool _on);
    // apply the device config's use_intensity_filtering / minimum_intensity with intensity model `_model` (0 = constant 64.0)
    int setIntensity(int _model, bool _filter);
    int setWatertight(bool _on);                               // Woop-Benthin-Wald 2013 intersector instead of Embree-ordered MT
    int traceRaw(std::vector<float> &_t, std::vector<int32_t> &_tri);   // per-ray results (parity tests)
    lsb200_ctx *context() { return _ctx; }
    std::string lastError() const;

    static Eigen::Affine3f transformFromComponents(Eigen::Vector3f _translation, Eigen::Vector3f _rotation);   // SRC/MeshTransformer.cpp:467-477

private:
    B200Tracer(LidarDevice::Ptr _sensorConfig, sensor_msgs::PointCloud2::Ptr _traceStorage, std::shared_ptr<spdlog::logger> _logger, int _cudaDevice);
    static void logThunk(int _level, const char *_message, void *_self);
    struct Geom { RTCGeometryType type; long nVerts, nElems; bool topologySent; };
    int update(const std::string &_meshName, const float A[12], pcl::PolygonMesh::Ptr &_mesh);

    lsb200_ctx *_ctx = nullptr;
    std::map<std::string, Geom> _geoms;
    bool _ordered = false, _ringChannel = false, _worldPoints = false;
    std::vector<uint32_t> _elemScratch;
};

}  // namespace lidarshooter
