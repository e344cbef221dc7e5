#include "MeshIO.hpp"

#include <array>
#include <cstring>
#include <fstream>
#include <sstream>
#include <unordered_map>
#include <vector>

namespace lidarshooter {
namespace io {

namespace {

struct KeyHash {
    size_t operator()(const std::array<uint32_t, 3> &k) const
    {
        uint64_t h = 1469598103934665603ull;
        for (uint32_t w : k) { h ^= w; h *= 1099511628211ull; }
        return static_cast<size_t>(h);
    }
};

// Welds exact duplicates (bit pattern) in first-occurrence order.
class Welder {
public:
    uint32_t add(const float p[3])
    {
        std::array<uint32_t, 3> k;
        std::memcpy(k.data(), p, 12);
        for (auto &w : k) if (w == 0x80000000u) w = 0;      // -0.0 == +0.0
        auto it = _index.find(k);
        if (it != _index.end()) return it->second;
        const uint32_t id = static_cast<uint32_t>(_xyz.size() / 3);
        _index.emplace(k, id);
        _xyz.insert(_xyz.end(), p, p + 3);
        return id;
    }
    const std::vector<float> &xyz() const { return _xyz; }
private:
    std::unordered_map<std::array<uint32_t, 3>, uint32_t, KeyHash> _index;
    std::vector<float> _xyz;
};

void fillCloud(const std::vector<float> &xyz, PolygonMesh &mesh)
{
    const size_t n = xyz.size() / 3;
    mesh.cloud.point_step = 16;
    mesh.cloud.width = static_cast<uint32_t>(n);
    mesh.cloud.height = 1;
    mesh.cloud.data.assign(n * 16, 0);
    for (size_t i = 0; i < n; ++i) {
        std::memcpy(mesh.cloud.data.data() + 16 * i, &xyz[3 * i], 12);
        const float one = 1.0f;
        std::memcpy(mesh.cloud.data.data() + 16 * i + 12, &one, 4);     // pcl::PointXYZ padding holds 1.0
    }
}

int loadAsciiSTL(const std::string &text, PolygonMesh &mesh)
{
    std::istringstream in(text);
    std::string tok;
    Welder w;
    std::vector<uint32_t> idx;
    while (in >> tok) {
        if (tok == "vertex") {
            float p[3];
            if (!(in >> p[0] >> p[1] >> p[2])) return -2;
            idx.push_back(w.add(p));
        }
    }
    if (idx.size() % 3) return -2;
    fillCloud(w.xyz(), mesh);
    mesh.polygons.resize(idx.size() / 3);
    for (size_t t = 0; t < idx.size() / 3; ++t) mesh.polygons[t].vertices = {(pcl::index_t)idx[3 * t], (pcl::index_t)idx[3 * t + 1], (pcl::index_t)idx[3 * t + 2]};
    return static_cast<int>(idx.size() / 3);
}

}  // namespace

int loadPolygonFileSTL(const std::string &_path, PolygonMesh &_mesh)
{
    std::ifstream f(_path, std::ios::binary);
    if (!f) return -1;
    std::string raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    _mesh = PolygonMesh();
    if (raw.size() < 84) return raw.compare(0, 5, "solid") == 0 ? loadAsciiSTL(raw, _mesh) : -2;
    uint32_t n = 0;
    std::memcpy(&n, raw.data() + 80, 4);
    if (raw.size() != 84 + static_cast<size_t>(n) * 50) {
        // not a well-formed binary STL: try the ASCII dialect
        return raw.compare(0, 5, "solid") == 0 ? loadAsciiSTL(raw, _mesh) : -2;
    }
    Welder w;
    _mesh.polygons.resize(n);
    for (uint32_t t = 0; t < n; ++t) {
        const char *rec = raw.data() + 84 + static_cast<size_t>(t) * 50 + 12;     // skip the facet normal
        _mesh.polygons[t].vertices.resize(3);
        for (int k = 0; k < 3; ++k) {
            float p[3];
            std::memcpy(p, rec + 12 * k, 12);
            _mesh.polygons[t].vertices[k] = w.add(p);
        }
    }
    fillCloud(w.xyz(), _mesh);
    return static_cast<int>(n);
}

int loadPolygonFileOBJ(const std::string &_path, PolygonMesh &_mesh)
{
    std::ifstream f(_path);
    if (!f) return -1;
    _mesh = PolygonMesh();
    std::vector<float> xyz;
    std::string line;
    while (std::getline(f, line)) {
        if (line.size() < 2) continue;
        if (line[0] == 'v' && (line[1] == ' ' || line[1] == '\t')) {
            std::istringstream in(line.substr(2));
            float p[3];
            if (!(in >> p[0] >> p[1] >> p[2])) return -2;
            xyz.insert(xyz.end(), p, p + 3);
        } else if (line[0] == 'f' && (line[1] == ' ' || line[1] == '\t')) {
            // "f a//na b//nb c//nc" (Meshlab), "f a/ta/na ...", "f a b c"; polygons are fanned into triangles
            std::istringstream in(line.substr(2));
            std::string tok;
            std::vector<uint32_t> ids;
            while (in >> tok) {
                const long v = std::strtol(tok.c_str(), nullptr, 10);
                if (v == 0) return -2;
                const long nv = static_cast<long>(xyz.size() / 3);
                ids.push_back(static_cast<uint32_t>(v > 0 ? v - 1 : nv + v));
            }
            for (size_t k = 2; k < ids.size(); ++k) {
                pcl::Vertices poly;
                poly.vertices = {(pcl::index_t)ids[0], (pcl::index_t)ids[k - 1], (pcl::index_t)ids[k]};
                _mesh.polygons.push_back(poly);
            }
        }
    }
    fillCloud(xyz, _mesh);
    return static_cast<int>(_mesh.polygons.size());
}

int loadPolygonFile(const std::string &_path, PolygonMesh &_mesh)
{
    const size_t dot = _path.rfind('.');
    std::string ext = dot == std::string::npos ? "" : _path.substr(dot + 1);
    for (auto &c : ext) c = static_cast<char>(std::tolower(static_cast<unsigned char>(c)));
    if (ext == "stl") return loadPolygonFileSTL(_path, _mesh);
    if (ext == "obj") return loadPolygonFileOBJ(_path, _mesh);
    return -3;
}

void meshFromArrays(const float *_xyz, size_t _nVerts, const uint32_t *_elems, size_t _nElems, int _vertsPerElem,
                    PolygonMesh &_mesh)
{
    _mesh = PolygonMesh();
    std::vector<float> xyz(_xyz, _xyz + 3 * _nVerts);
    fillCloud(xyz, _mesh);
    _mesh.polygons.resize(_nElems);
    for (size_t e = 0; e < _nElems; ++e)
        _mesh.polygons[e].vertices.assign(_elems + e * _vertsPerElem, _elems + (e + 1) * _vertsPerElem);
}

}  // namespace io
}  // namespace lidarshooter
