// LidarDevice.hpp -- host-side mirror of the reference's LidarDevice as far as the trace path needs it
// (SRC/LidarDevice.hpp, SRC/LidarDevice.cpp): the device JSON schema, the sensor transform, the ray generator
// conventions and the PointCloud2 header.  Same method names and return conventions; no Embree / ROS / Poco.
// Out of scope: HTTP fetch of the transform from the SENSR API (cpp:635-718) -- a config
// that names neither an inline sensorConfig nor an existing sensorConfigFile raises ConfigurationException.
#pragma once
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "Json.hpp"
#include "Types.hpp"

namespace lidarshooter {

class ConfigurationException : public std::runtime_error {
public:
    ConfigurationException(const std::string &file, const std::string &what, int code)
        : std::runtime_error(file + ": " + what), _code(code) {}
    int code() const { return _code; }
private:
    int _code;
};

// One ray as the reference's generator emits it (RTCRayHit fields that the path uses).
struct RayDir {
    float org[3] = {0.f, 0.f, 0.f};
    float dir[3] = {0.f, 0.f, 0.f};
    float tnear = 0.f;
    float tfar = 0.f;
};

class LidarDevice : public std::enable_shared_from_this<LidarDevice> {
public:
    using Ptr = std::shared_ptr<LidarDevice>;
    static Ptr create(const std::string &_config);                                  // SRC/LidarDevice.cpp:26-32
    static Ptr create(const std::string &_config, const std::string &_sensorUid);   // SRC/LidarDevice.cpp:34-40
    // synthetic device without a file (benchmarks): elevations in degrees above horizontal
    static Ptr create(const std::vector<float> &_vertical, float _begin, float _end, unsigned _count,
                      const float _qwxyz[4], const float _txyz[3], const std::string &_sensorUid = "synthetic");
    Ptr getPtr() { return shared_from_this(); }
    float detectionRadius() const { return _device.detectionRadius; }
    bool useIntensityFiltering() const { return _device.useIntensityFiltering; }      // config/hesai-pandar-XT-32-lidar_0000.json:19
    float minimumIntensity() const { return _device.minimumIntensity; }               // config/hesai-pandar-XT-32-lidar_0000.json:20
    void horizontalAngleRange(float &_lo, float &_hi) const { _lo = _device.horizontalAngleRange[0]; _hi = _device.horizontalAngleRange[1]; }

    void initialize(const std::string &_config, const std::string &_sensorUid = "");
    void initMessage(PointCloud2::Ptr _msg, int _frameIndex);                       // SRC/LidarDevice.cpp:94-115

    // stateful generator, horizontal index fastest; returns 1 when the cursor wrapped (SRC/LidarDevice.cpp:160-192, 824-845)
    int nextRay(RayDir &_ray, int *_valid = nullptr);
    int allRays(std::vector<RayDir> &_rays);           // ray id = channel * H + azimuth (the nextRay order)
    void reset();
    void getCurrentIndex(int *_verticalIndex, int *_horizontalIndex);

    void originToSensor(Vector3f &_sensor) const;          // SRC/LidarDevice.cpp:383-391
    void originToSensorInverse(Vector3f &_sensor) const;   // SRC/LidarDevice.cpp:393-401

    unsigned int getTotalRays() const { return _channels.count; }
    unsigned int getTotalChannels() const { return static_cast<unsigned>(_channels.vertical.size()); }
    unsigned int getScanRayCount() const { return _channels.horizontal.count; }
    const std::string &getSensorUid() const { return _device.sensorUid; }
    const std::string &getOutputFolder() const { return _outputFolder; }

    // what the C ABI's lsb200_set_sensor takes.  verticalAngles / azimuthBegin / azimuthEnd / quaternion / translation are
    // THE accessors a lidarshooter tree has to add to its LidarDevice for the drop-in ("LidarDevice getters")
    const std::vector<float> &verticalAngles() const { return _channels.vertical; }
    float azimuthBegin() const { return _channels.horizontal.begin; }
    float azimuthEnd() const { return _channels.horizontal.end; }
    float azimuthStep() const { return _channels.horizontal.step; }
    void quaternion(float q[4]) const;     // (w, x, y, z), not normalised
    void translation(float t[3]) const;    // (base_to_origin.tx, base_to_origin.ty, sensor_to_base.tz)
    const float *rotation() const { return _device.R; }
    const float *rotationInverse() const { return _device.Rinv; }
    const std::vector<PointField> &fields() const { return _message.fields; }
    int pointStep() const { return _message.pointStep; }

private:
    LidarDevice() = default;
    int loadConfiguration(const std::string &_config, const std::string &_sensorUid);
    int loadTransformationFromFile(const std::string &_transformFile);
    int loadTransformationFromJson(const json::Value &_transformJson);
    int advanceRayIndex();
    void updateRotation();

    struct {
        std::string sensorUid, sensorApiUrl;
        unsigned sensorApiPort = 9080;
        float tx = 0.f, ty = 0.f, tz = 0.f, qw = 1.f, qx = 0.f, qy = 0.f, qz = 0.f;
        float R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, Rinv[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        // carried by the reference's config, never applied by it (config/hesai-pandar-XT-32-lidar_0000.json:14-18)
        float detectionRadius = 0.f;                 // 0 = unlimited
        float horizontalAngleRange[2] = {-180.f, 180.f};
        bool useIntensityFiltering = false;
        float minimumIntensity = 0.f;
    } _device;
    struct {
        std::string frameId = "PandarXT-32";
        int pointStep = 32;
        bool isBigendian = false, isDense = true;
        std::vector<PointField> fields;
    } _message;
    struct {
        std::vector<float> vertical;
        struct { float begin = 0.f, end = 360.f, step = 0.f; unsigned count = 128; } horizontal;
        unsigned count = 0;
    } _channels;
    std::string _outputFolder = ".";
    unsigned _verticalIndex = 0, _horizontalIndex = 0;
    std::mutex _rayMutex;
};

// Small fp32 helpers in Eigen's evaluation order (shared with B200Tracer / CloudTransformer)
void quaternionToRotation(const float q_wxyz[4], float R[9]);   // Eigen::Quaternionf::toRotationMatrix, un-normalised
void inverse3x3(const float M[9], float Minv[9]);               // Eigen::Matrix3f::inverse (cofactors)

}  // namespace lidarshooter
