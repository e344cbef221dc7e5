// shim: pcl::Vertices (PCL 1.12: `Indices vertices`, index_t = std::int32_t unless PCL_INDEX_* says otherwise)
#pragma once
#include <cstdint>
#include <vector>
namespace pcl {
using index_t = std::int32_t;
using Indices = std::vector<index_t>;
struct Vertices {
    Indices vertices;
};
}  // namespace pcl
