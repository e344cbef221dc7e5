#include "LidarDevice.hpp"

#include <chrono>
#include <cmath>
#include <cstring>
#include <fstream>
#include <limits>

namespace lidarshooter {

void quaternionToRotation(const float q[4], float R[9])
{
    const float w = q[0], x = q[1], y = q[2], z = q[3];
    const float tx = 2.0f * x, ty = 2.0f * y, tz = 2.0f * z;
    const float twx = tx * w, twy = ty * w, twz = tz * w;
    const float txx = tx * x, txy = ty * x, txz = tz * x;
    const float tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R[0] = 1.0f - (tyy + tzz); R[1] = txy - twz;          R[2] = txz + twy;
    R[3] = txy + twz;          R[4] = 1.0f - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy;          R[7] = tyz + twx;          R[8] = 1.0f - (txx + tyy);
}

static inline float cofactor(const float *M, int i, int j)
{
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return M[3 * i1 + j1] * M[3 * i2 + j2] - M[3 * i1 + j2] * M[3 * i2 + j1];
}

void inverse3x3(const float M[9], float Mi[9])
{
    const float c0 = cofactor(M, 0, 0), c1 = cofactor(M, 1, 0), c2 = cofactor(M, 2, 0);
    const float p0 = c0 * M[0], p1 = c1 * M[3], p2 = c2 * M[6];
    const float s12 = p1 + p2;
    const float det = p0 + s12;
    const float invdet = 1.0f / det;
    Mi[0] = c0 * invdet; Mi[1] = c1 * invdet; Mi[2] = c2 * invdet;
    for (int j = 1; j < 3; ++j)
        for (int i = 0; i < 3; ++i) Mi[3 * j + i] = cofactor(M, i, j) * invdet;
}

LidarDevice::Ptr LidarDevice::create(const std::string &_config)
{
    Ptr p(new LidarDevice());
    p->initialize(_config);
    return p;
}

LidarDevice::Ptr LidarDevice::create(const std::string &_config, const std::string &_sensorUid)
{
    Ptr p(new LidarDevice());
    p->initialize(_config, _sensorUid);
    return p;
}

LidarDevice::Ptr LidarDevice::create(const std::vector<float> &_vertical, float _begin, float _end, unsigned _count,
                                     const float q[4], const float t[3], const std::string &_sensorUid)
{
    Ptr p(new LidarDevice());
    p->_device.sensorUid = _sensorUid;
    p->_device.qw = q[0]; p->_device.qx = q[1]; p->_device.qy = q[2]; p->_device.qz = q[3];
    p->_device.tx = t[0]; p->_device.ty = t[1]; p->_device.tz = t[2];
    p->updateRotation();
    p->_channels.vertical = _vertical;
    p->_channels.horizontal.begin = _begin;
    p->_channels.horizontal.end = _end;
    p->_channels.horizontal.count = _count;
    p->_channels.horizontal.step = (_end - _begin) / (_count - 1);
    p->_channels.count = static_cast<unsigned>(_vertical.size()) * _count;
    const char *names[5] = {"x", "y", "z", "intensity", "ring"};
    const uint32_t offs[5] = {0, 4, 8, 16, 20};
    for (int k = 0; k < 5; ++k) {
        PointField f; f.name = names[k]; f.offset = offs[k]; f.datatype = k < 4 ? PointField::FLOAT32 : PointField::UINT16; f.count = 1;
        p->_message.fields.push_back(f);
    }
    p->reset();
    return p;
}

void LidarDevice::initialize(const std::string &_config, const std::string &_sensorUid)
{
    loadConfiguration(_config, _sensorUid);
    reset();
}

void LidarDevice::updateRotation()
{
    const float q[4] = {_device.qw, _device.qx, _device.qy, _device.qz};
    quaternionToRotation(q, _device.R);
    inverse3x3(_device.R, _device.Rinv);
}

int LidarDevice::loadConfiguration(const std::string &_config, const std::string &_sensorUid)
{
    json::Value jsonData;
    std::string errs;
    if (!json::parseFile(_config, jsonData, errs)) throw ConfigurationException(_config, errs, 2);

    // Device (SRC/LidarDevice.cpp:495-554): inline sensorConfig > sensorConfigFile > SENSR API (out of scope)
    if (jsonData.isMember("device")) {
        const json::Value &dev = jsonData["device"];
        if (!_sensorUid.empty()) _device.sensorUid = _sensorUid;
        else if (dev.isMember("sensorUid")) _device.sensorUid = dev["sensorUid"].asString("");
        _device.sensorApiUrl = dev["sensorApiUrl"].asString("localhost");
        _device.sensorApiPort = dev["sensorApiPort"].asUInt(9080);
        if (dev.isMember("sensorConfig")) {
            loadTransformationFromJson(dev["sensorConfig"]);
        } else if (dev.isMember("sensorConfigFile")) {
            const std::string path = dev["sensorConfigFile"].asString("transform-device.json");
            if (!std::ifstream(path).good()) throw ConfigurationException(path, "File not found", 1);
            loadTransformationFromFile(path);
        } else {
            throw ConfigurationException(_config, "no sensorConfig / sensorConfigFile; the SENSR HTTP fetch is out of scope", 3);
        }
    }

    // Message (SRC/LidarDevice.cpp:559-584)
    if (jsonData.isMember("message")) {
        const json::Value &msg = jsonData["message"];
        _message.frameId = msg["frameId"].asString("PandarXT-32");
        _message.pointStep = msg["pointStep"].asInt(32);
        _message.isBigendian = msg["isBigendian"].asBool(false);
        _message.isDense = msg["isDense"].asBool(true);
        _message.fields.clear();
        for (const json::Value &field : msg["pointFields"].items()) {
            PointField pf;
            pf.name = field["name"].asString();
            pf.offset = static_cast<uint32_t>(field["offset"].asInt());
            pf.datatype = static_cast<uint8_t>(field["datatype"].asInt());
            pf.count = static_cast<uint32_t>(field["count"].asInt());
            _message.fields.push_back(pf);
        }
    }

    // Channels (SRC/LidarDevice.cpp:586-620)
    if (jsonData.isMember("channels")) {
        const json::Value &ch = jsonData["channels"];
        _channels.vertical.clear();
        for (const json::Value &angle : ch["vertical"].items()) _channels.vertical.push_back(angle.asFloat());
        if (ch.isMember("horizontal")) {
            const json::Value &hz = ch["horizontal"];
            if (hz.isMember("range")) {
                _channels.horizontal.begin = hz["range"]["begin"].asFloat(0.0f);
                _channels.horizontal.end = hz["range"]["end"].asFloat(360.0f);
            }
            _channels.horizontal.count = hz["count"].asUInt(128);
            // float / unsigned -> float division (SRC/LidarDevice.cpp:611)
            _channels.horizontal.step = (_channels.horizontal.end - _channels.horizontal.begin) / (_channels.horizontal.count - 1);
        }
        _channels.count = static_cast<unsigned>(_channels.vertical.size()) * _channels.horizontal.count;
    }
    if (jsonData.isMember("outputFolder")) _outputFolder = jsonData["outputFolder"].asString(".");
    return 0;
}

int LidarDevice::loadTransformationFromFile(const std::string &_transformFile)
{
    json::Value v;
    std::string errs;
    if (!json::parseFile(_transformFile, v, errs)) throw ConfigurationException(_transformFile, errs, 2);
    return loadTransformationFromJson(v);
}

int LidarDevice::loadTransformationFromJson(const json::Value &t)
{
    // SRC/LidarDevice.cpp:758-822
    if (t.isMember("uid") && _device.sensorUid.empty()) _device.sensorUid = t["uid"].asString("");
    if (t.isMember("base_to_origin")) {
        _device.tx = t["base_to_origin"]["tx"].asFloat(0.0f);
        _device.ty = t["base_to_origin"]["ty"].asFloat(0.0f);
    }
    if (t.isMember("sensor_to_base")) {
        const json::Value &s = t["sensor_to_base"];
        _device.qw = s["qw"].asFloat(0.0f); _device.qx = s["qx"].asFloat(0.0f);
        _device.qy = s["qy"].asFloat(0.0f); _device.qz = s["qz"].asFloat(0.0f);
        _device.tz = s["tz"].asFloat(0.0f);
        updateRotation();
    }
    if (t.isMember("detection_radius")) _device.detectionRadius = t["detection_radius"].asFloat(0.0f);
    if (t.isMember("use_intensity_filtering")) _device.useIntensityFiltering = t["use_intensity_filtering"].asBool(false);
    if (t.isMember("minimum_intensity")) _device.minimumIntensity = t["minimum_intensity"].asFloat(0.0f);
    if (t.isMember("horizontal_angle_range") && t["horizontal_angle_range"].size() == 2) {
        _device.horizontalAngleRange[0] = t["horizontal_angle_range"][size_t(0)].asFloat(-180.0f);
        _device.horizontalAngleRange[1] = t["horizontal_angle_range"][size_t(1)].asFloat(180.0f);
    }
    return 0;
}

void LidarDevice::initMessage(PointCloud2::Ptr _msg, int _frameIndex)
{
    _msg->fields.clear();
    for (const auto &field : _message.fields) _msg->fields.push_back(field);
    _msg->header.frame_id = _message.frameId;
    _msg->header.stamp = ros::Time::now();            // SRC/LidarDevice.cpp:103
    _msg->header.seq = static_cast<uint32_t>(_frameIndex);
    _msg->height = 1;
    _msg->width = 0;
    _msg->point_step = static_cast<uint32_t>(_message.pointStep);
    _msg->row_step = 0 * _message.pointStep;       // the reference leaves it 0 (SRC/LidarDevice.cpp:112)
    _msg->is_bigendian = _message.isBigendian;
    _msg->is_dense = _message.isDense;
}

static inline void rayDirection(float preChi, float prePhi, float dir[3])
{
    // double expression narrowed to float BEFORE the float sin/cos (SRC/LidarDevice.cpp:172-179)
    const float theta = (90.0 - preChi) * M_PI / 180.0;
    const float phi = prePhi * M_PI / 180.0;
    dir[0] = std::sin(theta) * std::cos(phi);
    dir[1] = std::sin(theta) * std::sin(phi);
    dir[2] = std::cos(theta);
}

int LidarDevice::nextRay(RayDir &_ray, int *_valid)
{
    std::lock_guard<std::mutex> rayLock(_rayMutex);
    const float preChi = _channels.vertical[_verticalIndex];
    const float prePhi = _channels.horizontal.begin + _channels.horizontal.step * static_cast<float>(_horizontalIndex);
    rayDirection(preChi, prePhi, _ray.dir);
    _ray.org[0] = _ray.org[1] = _ray.org[2] = 0.f;
    _ray.tnear = 0.f;
    _ray.tfar = std::numeric_limits<float>::infinity();
    if (_valid) *_valid = -1;
    return advanceRayIndex();
}

int LidarDevice::allRays(std::vector<RayDir> &_rays)
{
    _rays.resize(getTotalRays());
    size_t k = 0;
    for (float chi : _channels.vertical)
        for (unsigned h = 0; h < _channels.horizontal.count; ++h, ++k) {
            const float prePhi = _channels.horizontal.begin + _channels.horizontal.step * static_cast<float>(h);
            rayDirection(chi, prePhi, _rays[k].dir);
            _rays[k].org[0] = _rays[k].org[1] = _rays[k].org[2] = 0.f;
            _rays[k].tnear = 0.f;
            _rays[k].tfar = std::numeric_limits<float>::infinity();
        }
    return 0;
}

void LidarDevice::reset() { _verticalIndex = 0; _horizontalIndex = 0; }

void LidarDevice::getCurrentIndex(int *v, int *h) { *v = static_cast<int>(_verticalIndex); *h = static_cast<int>(_horizontalIndex); }

int LidarDevice::advanceRayIndex()
{
    if (_horizontalIndex == _channels.horizontal.count - 1) {
        _horizontalIndex = 0;
        if (_verticalIndex == _channels.vertical.size() - 1) { _verticalIndex = 0; return 1; }
        ++_verticalIndex;
    } else {
        ++_horizontalIndex;
    }
    return 0;
}

static inline void mulMat3(const float *M, const float p[3], float o[3])
{
    for (int r = 0; r < 3; ++r) {
        const float p0 = M[3 * r] * p[0], p1 = M[3 * r + 1] * p[1], p2 = M[3 * r + 2] * p[2];
        const float s12 = p1 + p2;
        o[r] = p0 + s12;
    }
}

void LidarDevice::originToSensor(Vector3f &s) const
{
    const float d[3] = {s[0] - _device.tx, s[1] - _device.ty, s[2] - _device.tz};
    float o[3];
    mulMat3(_device.Rinv, d, o);
    s = Vector3f(o[0], o[1], o[2]);
}

void LidarDevice::originToSensorInverse(Vector3f &s) const
{
    float o[3];
    mulMat3(_device.R, s.data(), o);
    s = Vector3f(o[0] + _device.tx, o[1] + _device.ty, o[2] + _device.tz);
}

void LidarDevice::quaternion(float q[4]) const { q[0] = _device.qw; q[1] = _device.qx; q[2] = _device.qy; q[3] = _device.qz; }
void LidarDevice::translation(float t[3]) const { t[0] = _device.tx; t[1] = _device.ty; t[2] = _device.tz; }

}  // namespace lidarshooter
