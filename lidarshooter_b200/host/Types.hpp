// Types.hpp -- the third-party types the trace path exchanges with its caller, under their REAL names:
//   Eigen::Vector3f / Eigen::Affine3f, pcl::PolygonMesh, sensor_msgs::PointCloud2, RTCGeometryType, spdlog::logger.
// In the reference's build these come from Eigen 3, PCL 1.12, ROS 1, Embree 3.13.4 and spdlog; none of them is
// installed in this image, so the Makefile puts host/shims/ on the include path (minimal stand-ins with the same
// names and members, host/shims/README.md).  `make SHIMS=` compiles the same sources against the real libraries.
// The aliases below keep this repository's own stand-ins for reference classes (LidarDevice, AffineMesh, ...) short;
// the drop-in adapter itself (B200Tracer.hpp / .cpp, ITracer.hpp) spells the real names out.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include <Eigen/Dense>
#include <Eigen/Geometry>
#include <embree3/rtcore.h>
#include <pcl/PolygonMesh.h>
#include <ros/time.h>
#include <sensor_msgs/PointCloud2.h>
#include <spdlog/spdlog.h>

namespace lidarshooter {

using Vector3f = Eigen::Vector3f;
using Affine3f = Eigen::Affine3f;
using PointCloud2 = sensor_msgs::PointCloud2;
using PointField = sensor_msgs::PointField;
using PolygonMesh = pcl::PolygonMesh;
using Time = ros::Time;

// number of vertex records in a pcl::PCLPointCloud2 blob (xyz float32 at offset 0 of every point_step-byte record,
// SRC/MeshTransformer.cpp:173-194)
inline size_t vertexCount(const pcl::PolygonMesh &m) { return m.cloud.point_step ? m.cloud.data.size() / m.cloud.point_step : 0; }

// Eigen::Affine3f <-> the row-major 3x4 [L | t] that lsb200_update_geometry takes
inline void affineToRows(const Eigen::Affine3f &A, float rows[12])
{
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) rows[4 * r + c] = A(r, c);
}
inline Eigen::Affine3f affineFromRows(const float rows[12])
{
    Eigen::Affine3f A = Eigen::Affine3f::Identity();
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) A(r, c) = rows[4 * r + c];
    return A;
}

}  // namespace lidarshooter
