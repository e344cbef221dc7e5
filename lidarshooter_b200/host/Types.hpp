// Types.hpp -- the few data types the trace path exchanges with its caller, shaped like the ones the reference
// uses so that B200Tracer's signatures read like EmbreeTracer's (SRC/ITracer.hpp:50-94):
//   Eigen::Vector3f / Eigen::Affine3f  -> lidarshooter::Vector3f / Affine3f   (Eigen is not in this image)
//   pcl::PolygonMesh                   -> lidarshooter::PolygonMesh           (cloud blob + polygons)
//   sensor_msgs::PointCloud2           -> lidarshooter::PointCloud2
//   RTCGeometryType                    -> lidarshooter::RTCGeometryType       (same enumerator values as Embree 3)
// INTEGRATION.md shows the adapter compiled against the real headers; these stand-ins have the same field names.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace lidarshooter {

enum RTCGeometryType { RTC_GEOMETRY_TYPE_TRIANGLE = 0, RTC_GEOMETRY_TYPE_QUAD = 1, RTC_GEOMETRY_TYPE_GRID = 2 };

struct Vector3f {
    float v[3] = {0.f, 0.f, 0.f};
    Vector3f() = default;
    Vector3f(float x, float y, float z) : v{x, y, z} {}
    float x() const { return v[0]; }
    float y() const { return v[1]; }
    float z() const { return v[2]; }
    float &operator[](int i) { return v[i]; }
    float operator[](int i) const { return v[i]; }
    static Vector3f Zero() { return Vector3f(); }
};

// Row-major 3x4 [L | t], the layout lsb200_update_geometry takes.
struct Affine3f {
    float m[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    static Affine3f Identity() { return Affine3f(); }
    const float *data() const { return m; }
};

struct PointField {
    std::string name;
    uint32_t offset = 0;
    uint8_t datatype = 0;
    uint32_t count = 0;
    enum { INT8 = 1, UINT8 = 2, INT16 = 3, UINT16 = 4, INT32 = 5, UINT32 = 6, FLOAT32 = 7, FLOAT64 = 8 };
};

// ros::Time: whole seconds + nanoseconds (a double cannot hold a present-day stamp to the nanosecond)
struct Time {
    uint32_t sec = 0, nsec = 0;
    Time() = default;
    Time(uint32_t s, uint32_t ns) : sec(s), nsec(ns) {}
    explicit Time(double t)
    {
        const double whole = t < 0.0 ? 0.0 : (double)(uint64_t)t;
        double frac = (t < 0.0 ? 0.0 : t - whole) * 1e9 + 0.5;
        sec = (uint32_t)whole; nsec = (uint32_t)frac;
        if (nsec >= 1000000000u) { nsec -= 1000000000u; ++sec; }
    }
    double toSec() const { return (double)sec + 1e-9 * (double)nsec; }
};

struct Header {
    uint32_t seq = 0;
    Time stamp;                // the reference uses ros::Time::now()
    std::string frame_id;
};

struct PointCloud2 {
    using Ptr = std::shared_ptr<PointCloud2>;
    Header header;
    uint32_t height = 0, width = 0;
    std::vector<PointField> fields;
    bool is_bigendian = false;
    uint32_t point_step = 0, row_step = 0;
    std::vector<uint8_t> data;
    bool is_dense = false;
};

// pcl::PCLPointCloud2 + pcl::Vertices, as far as MeshTransformer reads them (SRC/MeshTransformer.cpp:173-194,
// 486-552): xyz float32 at offset 0 of every point_step-byte record; polygons[k].vertices = 3 (or 4) indices.
struct Vertices {
    std::vector<uint32_t> vertices;
};
struct PolygonMesh {
    using Ptr = std::shared_ptr<PolygonMesh>;
    struct Cloud {
        uint32_t width = 0, height = 1, point_step = 16;
        std::vector<uint8_t> data;
    } cloud;
    std::vector<Vertices> polygons;
    size_t vertexCount() const { return cloud.point_step ? cloud.data.size() / cloud.point_step : 0; }
};

}  // namespace lidarshooter
