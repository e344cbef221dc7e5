// CloudSerializer.hpp -- ROS-1 wire format of sensor_msgs/PointCloud2 without ROS.
// The reference publishes the traced cloud through roscpp (SRC/MeshProjector.cpp:380-390); a consumer that is not a
// ROS node (a file, a socket, the SENSR ingest) needs the same bytes.  Layout (little-endian, as roscpp's
// ros::serialization writes it):
//   Header   : uint32 seq | uint32 stamp.sec | uint32 stamp.nsec | uint32 len + frame_id bytes
//   uint32 height | uint32 width
//   fields   : uint32 n | n x { uint32 len + name | uint32 offset | uint8 datatype | uint32 count }
//   uint8 is_bigendian | uint32 point_step | uint32 row_step
//   data     : uint32 len + bytes
//   uint8 is_dense
// The header fields are the ones LidarDevice::initMessage sets (SRC/LidarDevice.cpp:94-115).  The reference leaves
// row_step = 0 (SRC/LidarDevice.cpp:112); fixRowStep writes width * point_step instead, which is what consumers
// that validate the message expect.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "Types.hpp"

namespace lidarshooter {

// Appends the serialised message to `out`; returns the number of bytes appended.
size_t serializePointCloud2(const PointCloud2 &cloud, std::vector<uint8_t> &out, bool fixRowStep = false);
// Exact size serializePointCloud2 will append.
size_t serializedLength(const PointCloud2 &cloud);
// Parses one message; returns the bytes consumed, 0 when the buffer is truncated or malformed.
size_t deserializePointCloud2(const uint8_t *buf, size_t len, PointCloud2 &cloud);
// Appends "uint32 length + message" to a file (a plain length-prefixed stream of clouds); false on I/O failure.
bool appendCloudToFile(const std::string &path, const PointCloud2 &cloud, bool fixRowStep = false);

}  // namespace lidarshooter
