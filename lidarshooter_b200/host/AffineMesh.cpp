#include "AffineMesh.hpp"

#include <cmath>

#include "Json.hpp"
#include "lsb200.h"

namespace lidarshooter {

Vector3f AffineMesh::transformToGlobal(Vector3f d) const
{
    // Translation(0) * Rz * Ry * Rx, the same composition (and code) as MeshTransformer::transformFromComponents
    const float zero[3] = {0.f, 0.f, 0.f};
    float A[12];
    lsb200_affine_from_components(zero, _angular.data(), A);
    Vector3f o;
    for (int r = 0; r < 3; ++r) {
        const float p0 = A[4 * r] * d[0], p1 = A[4 * r + 1] * d[1], p2 = A[4 * r + 2] * d[2];
        const float s12 = p1 + p2;
        o[r] = A[4 * r + 3] + (p0 + s12);
    }
    return o;
}

void AffineMesh::applyTwist(const Vector3f &l, const Vector3f &a)
{
    const Vector3f g = transformToGlobal(l);
    for (int k = 0; k < 3; ++k) { _linear[k] += g[k]; _angular[k] += a[k]; }
}

int Trajectory::load(const std::string &_path)
{
    json::Value v;
    std::string err;
    if (!json::parseFile(_path, v, err)) return -1;
    if (!v.isMember("trajectory") || !v["trajectory"].isArray()) return -2;
    _segments.clear();
    for (const json::Value &s : v["trajectory"].items()) {
        Segment seg;
        seg.dt = s["dt"].asFloat(0.f);
        for (int k = 0; k < 3; ++k) { seg.linear[k] = s["linear"][k].asFloat(0.f); seg.angular[k] = s["angular"][k].asFloat(0.f); }
        _segments.push_back(seg);
    }
    return static_cast<int>(_segments.size());
}

size_t Trajectory::tickCount(float period) const
{
    size_t n = 0;
    for (const auto &s : _segments) n += static_cast<size_t>(std::lround(s.dt / period));
    return n;
}

void Trajectory::applyTick(AffineMesh &_mesh, size_t k, float period) const
{
    size_t first = 0;
    for (const auto &s : _segments) {
        const size_t n = static_cast<size_t>(std::lround(s.dt / period));
        if (k < first + n) {
            _mesh.applyTwist(Vector3f(s.linear[0] * period, s.linear[1] * period, s.linear[2] * period),
                             Vector3f(s.angular[0] * period, s.angular[1] * period, s.angular[2] * period));
            return;
        }
        first += n;
    }
}

}  // namespace lidarshooter
