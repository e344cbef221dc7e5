// AffineMesh.hpp -- a mesh plus its accumulated linear / angular displacement, and the headless trajectory player.
// Mirrors the state and the integration rule of the reference's AffineMesh (SRC/AffineMesh.hpp:68-85,
// SRC/AffineMesh.cpp:108-128, 275-284) without ROS: joystickCallback(Twist) becomes applyTwist(linear, angular).
// Trajectory reads config/trajectory.json ({"trajectory": [{dt, linear[3], angular[3]}]}) -- the reference ships the
// file but has no consumer for it; the semantics here (one Twist of linear*period / angular*period per trace tick,
// period 0.1 s = MeshProjector's tracePeriod, SRC/MeshProjector.hpp:70) are defined in SURVEY.md 8(d) C2.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "Types.hpp"

namespace lidarshooter {

class AffineMesh {
public:
    using Ptr = std::shared_ptr<AffineMesh>;
    static Ptr create(const std::string &__name, PolygonMesh::Ptr __mesh = nullptr) { return Ptr(new AffineMesh(__name, __mesh)); }

    const std::string &getName() const { return _name; }
    PolygonMesh::Ptr &getMesh() { return _mesh; }
    void setMesh(PolygonMesh::Ptr __mesh) { _mesh = __mesh; }
    Vector3f getLinearDisplacementConst() const { return _linear; }
    Vector3f getAngularDisplacementConst() const { return _angular; }
    void setLinearDisplacement(const Vector3f &l) { _linear = l; }
    void setAngularDisplacement(const Vector3f &a) { _angular = a; }
    void resetLinearDisplacement() { _linear = Vector3f::Zero(); }
    void resetAngularDisplacement() { _angular = Vector3f::Zero(); }

    // Rz(ang.z) Ry(ang.y) Rx(ang.x) * displacement (SRC/AffineMesh.cpp:275-284)
    Vector3f transformToGlobal(Vector3f _displacement) const;
    // linear += transformToGlobal(linear twist); angular += angular twist (SRC/AffineMesh.cpp:108-128)
    void applyTwist(const Vector3f &_linear, const Vector3f &_angular);

private:
    AffineMesh(const std::string &__name, PolygonMesh::Ptr __mesh) : _name(__name), _mesh(__mesh) {}
    std::string _name;
    PolygonMesh::Ptr _mesh;
    Vector3f _linear, _angular;
};

class Trajectory {
public:
    struct Segment { float dt; Vector3f linear, angular; };
    // returns the number of segments, < 0 on error
    int load(const std::string &_path);
    void add(float dt, Vector3f linear, Vector3f angular) { _segments.push_back(Segment{dt, linear, angular}); }
    const std::vector<Segment> &segments() const { return _segments; }
    // number of ticks of `period` seconds over all segments
    size_t tickCount(float period = 0.1f) const;
    // applies tick k (0-based) to the mesh: the Twist of the segment that contains it, scaled by the period
    void applyTick(AffineMesh &_mesh, size_t k, float period = 0.1f) const;

private:
    std::vector<Segment> _segments;
};

}  // namespace lidarshooter
