// CloudConverter.hpp -- decoding of the traced PointCloud2 payload: the 32-byte record written by the compaction
// kernel / the reference's XYZIRBytes (SRC/XYZIRBytes.cpp:24-40, SRC/XYZIRPoint.hpp) and the PointCloud2 -> xyz array
// conversion the viewer side does (SRC/CloudConverter.hpp:30-90).  Sensor -> world transform of a cloud
// (CloudTransformer::applyInverseTransform with an identity mesh transform, SRC/CloudTransformer.cpp:202-265) is
// done ON THE DEVICE by the compaction kernel (B200Tracer::setWorldFramePoints); applyInverseTransform() below is
// the host restatement used to check it.
#pragma once
#include <cstring>
#include <vector>

#include "LidarDevice.hpp"
#include "Types.hpp"

namespace lidarshooter {

#pragma pack(push, 1)
struct XYZIRPoint {          // 32 bytes
    float xPos, yPos, zPos;
    uint8_t pad0[4];
    float intensity;
    int32_t ring;
    uint8_t pad1[8];
};
#pragma pack(pop)
static_assert(sizeof(XYZIRPoint) == 32, "point record must be 32 bytes");

struct PointXYZ { float x, y, z; };

class CloudConverter {
public:
    explicit CloudConverter(PointCloud2::Ptr __cloud) : _cloud(__cloud) {}
    // PointCloud2 -> xyz (offsets taken from the fields, like SRC/CloudConverter.hpp:41-70)
    void to(std::vector<PointXYZ> &_output) const
    {
        uint32_t ox = 0, oy = 4, oz = 8;
        for (const auto &f : _cloud->fields) { if (f.name == "x") ox = f.offset; if (f.name == "y") oy = f.offset; if (f.name == "z") oz = f.offset; }
        const size_t n = static_cast<size_t>(_cloud->width) * (_cloud->height ? _cloud->height : 1);
        _output.resize(n);
        for (size_t i = 0; i < n; ++i) {
            const uint8_t *p = _cloud->data.data() + i * _cloud->point_step;
            std::memcpy(&_output[i].x, p + ox, 4); std::memcpy(&_output[i].y, p + oy, 4); std::memcpy(&_output[i].z, p + oz, 4);
        }
    }
private:
    PointCloud2::Ptr _cloud;
};

// sensor frame -> world frame in place: p <- R p + T (LidarDevice::originToSensorInverse per point)
inline void applyInverseTransform(PointCloud2 &cloud, const LidarDevice &device)
{
    const size_t n = cloud.point_step ? cloud.data.size() / cloud.point_step : 0;
    for (size_t i = 0; i < n; ++i) {
        float *p = reinterpret_cast<float *>(cloud.data.data() + i * cloud.point_step);
        Vector3f v(p[0], p[1], p[2]);
        device.originToSensorInverse(v);
        p[0] = v[0]; p[1] = v[1]; p[2] = v[2];
    }
}

}  // namespace lidarshooter
