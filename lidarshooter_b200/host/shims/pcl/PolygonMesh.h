// shim: pcl::PolygonMesh
#pragma once
#include <memory>
#include <vector>
#include "PCLPointCloud2.h"
#include "Vertices.h"
namespace pcl {
struct PolygonMesh {
    using Ptr = std::shared_ptr<PolygonMesh>;
    using ConstPtr = std::shared_ptr<const PolygonMesh>;
    PCLPointCloud2 cloud;
    std::vector<Vertices> polygons;
};
}  // namespace pcl
