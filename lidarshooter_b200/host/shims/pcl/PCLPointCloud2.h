// shim: pcl::PCLPointCloud2 as far as MeshTransformer reads it (SRC/MeshTransformer.cpp:173-194)
#pragma once
#include <cstdint>
#include <string>
#include <vector>
namespace pcl {
struct PCLPointField {
    std::string name;
    std::uint32_t offset = 0;
    std::uint8_t datatype = 0;
    std::uint32_t count = 0;
};
struct PCLPointCloud2 {
    std::uint32_t height = 1, width = 0;
    std::vector<PCLPointField> fields;
    std::uint8_t is_bigendian = 0;
    std::uint32_t point_step = 16, row_step = 0;
    std::vector<std::uint8_t> data;
    std::uint8_t is_dense = 0;
};
}  // namespace pcl
