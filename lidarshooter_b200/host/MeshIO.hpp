// MeshIO.hpp -- mesh file readers for the formats the reference ships (mesh/*.stl binary STL, mesh/*.obj).
// The reference loads STL through pcl::io::loadPolygonFileSTL (SRC/mainwindow.cpp:146; SRC/test/EmbreeTracer_test.cpp:40),
// i.e. vtkSTLReader with point merging: exact-duplicate vertices are welded in first-occurrence order, which is what
// yields 98 vertices / 162 triangles for ground.stl (SRC/test/EmbreeTracer_test.cpp:89-90).
#pragma once
#include <string>

#include "Types.hpp"

namespace lidarshooter {
namespace io {

// Returns the number of triangles read, < 0 on error.  point_step of the produced cloud is 16 (x y z pad), like
// a pcl::PointXYZ blob.
int loadPolygonFileSTL(const std::string &_path, PolygonMesh &_mesh);
int loadPolygonFileOBJ(const std::string &_path, PolygonMesh &_mesh);
int loadPolygonFile(const std::string &_path, PolygonMesh &_mesh);   // by extension

// Builds a mesh from packed arrays (n_verts x 3 floats, n_elems x verts_per_elem indices)
void meshFromArrays(const float *_xyz, size_t _nVerts, const uint32_t *_elems, size_t _nElems, int _vertsPerElem,
                    PolygonMesh &_mesh);

}  // namespace io
}  // namespace lidarshooter
