// capi_host.cpp -- C flattening of the C++ host layer (include/lsb200_host.h) for the ctypes-driven tests.
#include "lsb200_host.h"

#include <cstdio>
#include <cstring>
#include <exception>

#include "AffineMesh.hpp"
#include "B200Tracer.hpp"
#include "CloudConverter.hpp"
#include "CloudSerializer.hpp"
#include "LidarDevice.hpp"
#include "MeshIO.hpp"

using namespace lidarshooter;

struct lsbh_device { LidarDevice::Ptr p; };
struct lsbh_mesh { PolygonMesh::Ptr p; };
struct lsbh_tracer { B200Tracer::Ptr p; std::shared_ptr<spdlog::logger> log; };

static void set_err(char *err, size_t n, const char *msg) { if (err && n) std::snprintf(err, n, "%s", msg); }

template <class F> static int guarded(F f)
{
    try { return f(); }
    catch (const TraceException &e) { return -100 - (int)e.getErrorCode(); }
    catch (const BadGeometryException &e) { return -200 - (int)e.getErrorCode(); }
    catch (const MeshNotSetException &e) { return -300 - (int)e.getErrorCode(); }
    catch (...) { return -999; }
}
extern "C" {

lsbh_device *lsbh_device_create(const char *path, const char *uid, char *err, size_t n)
{
    try {
        auto *d = new lsbh_device();
        d->p = (uid && *uid) ? LidarDevice::create(path, uid) : LidarDevice::create(path);
        return d;
    } catch (const std::exception &e) { set_err(err, n, e.what()); return nullptr; }
}
lsbh_device *lsbh_device_create_synthetic(const float *vertical, int nv, float b, float e, unsigned count, const float q[4], const float t[3])
{
    auto *d = new lsbh_device();
    d->p = LidarDevice::create(std::vector<float>(vertical, vertical + nv), b, e, count, q, t);
    return d;
}
void lsbh_device_destroy(lsbh_device *d) { delete d; }
const char *lsbh_device_sensor_uid(const lsbh_device *d) { return d->p->getSensorUid().c_str(); }
unsigned lsbh_device_total_rays(const lsbh_device *d) { return d->p->getTotalRays(); }
unsigned lsbh_device_total_channels(const lsbh_device *d) { return d->p->getTotalChannels(); }
unsigned lsbh_device_scan_ray_count(const lsbh_device *d) { return d->p->getScanRayCount(); }
int lsbh_device_transform(const lsbh_device *d, float q[4], float t[3], float R[9], float Rinv[9])
{
    d->p->quaternion(q); d->p->translation(t);
    std::memcpy(R, d->p->rotation(), 36); std::memcpy(Rinv, d->p->rotationInverse(), 36);
    return 0;
}
int lsbh_device_channels(const lsbh_device *d, float *vertical, int cap, float *b, float *e, float *step)
{
    const auto &v = d->p->verticalAngles();
    for (int i = 0; i < cap && i < (int)v.size(); ++i) vertical[i] = v[i];
    *b = d->p->azimuthBegin(); *e = d->p->azimuthEnd(); *step = d->p->azimuthStep();
    return (int)v.size();
}
int lsbh_device_next_ray(lsbh_device *d, float dir[3], int *vi, int *hi)
{
    RayDir r; int valid = 0;
    const int state = d->p->nextRay(r, &valid);
    std::memcpy(dir, r.dir, 12);
    d->p->getCurrentIndex(vi, hi);
    return state;
}
void lsbh_device_reset(lsbh_device *d) { d->p->reset(); }
int lsbh_device_all_rays(lsbh_device *d, float *dirs, size_t cap)
{
    std::vector<RayDir> rays;
    d->p->allRays(rays);
    if (cap < rays.size()) return -1;
    for (size_t i = 0; i < rays.size(); ++i) std::memcpy(dirs + 3 * i, rays[i].dir, 12);
    return (int)rays.size();
}
void lsbh_device_origin_to_sensor(const lsbh_device *d, float p[3], int inverse)
{
    Vector3f v(p[0], p[1], p[2]);
    if (inverse) d->p->originToSensorInverse(v); else d->p->originToSensor(v);
    p[0] = v[0]; p[1] = v[1]; p[2] = v[2];
}
int lsbh_device_init_message(lsbh_device *d, int frame, lsb200_cloud_header *out, char *frame_id, size_t fl, char *fields, size_t fsl)
{
    PointCloud2::Ptr c(new PointCloud2());
    c->width = 77;      // must be reset to 0 by initMessage
    d->p->initMessage(c, frame);
    out->seq = c->header.seq; out->height = c->height; out->width = c->width; out->point_step = c->point_step;
    out->row_step = c->row_step; out->is_bigendian = c->is_bigendian; out->is_dense = c->is_dense;
    out->n_fields = (uint8_t)c->fields.size(); out->reserved = 0;
    if (frame_id && fl) std::snprintf(frame_id, fl, "%s", c->header.frame_id.c_str());
    if (fields && fsl) {
        std::string s;
        for (auto &f : c->fields) s += f.name + ":" + std::to_string(f.offset) + ":" + std::to_string((int)f.datatype) + ":" + std::to_string(f.count) + ";";
        std::snprintf(fields, fsl, "%s", s.c_str());
    }
    return 0;
}

lsbh_mesh *lsbh_mesh_load(const char *path)
{
    auto *m = new lsbh_mesh();
    m->p = PolygonMesh::Ptr(new PolygonMesh());
    if (io::loadPolygonFile(path, *m->p) < 0) { delete m; return nullptr; }
    return m;
}
lsbh_mesh *lsbh_mesh_from_arrays(const float *xyz, size_t nv, const uint32_t *elems, size_t ne, int vpe)
{
    auto *m = new lsbh_mesh();
    m->p = PolygonMesh::Ptr(new PolygonMesh());
    io::meshFromArrays(xyz, nv, elems, ne, vpe, *m->p);
    return m;
}
void lsbh_mesh_destroy(lsbh_mesh *m) { delete m; }
int lsbh_mesh_counts(const lsbh_mesh *m, int *nv, int *np, int *step)
{
    *nv = (int)vertexCount(*m->p); *np = (int)m->p->polygons.size(); *step = (int)m->p->cloud.point_step;
    return 0;
}
int lsbh_mesh_copy(const lsbh_mesh *m, float *xyz, uint32_t *elems, int vpe)
{
    const size_t nv = vertexCount(*m->p);
    for (size_t i = 0; i < nv; ++i) std::memcpy(xyz + 3 * i, m->p->cloud.data.data() + i * m->p->cloud.point_step, 12);
    for (size_t e = 0; e < m->p->polygons.size(); ++e) {
        if ((int)m->p->polygons[e].vertices.size() != vpe) return -1;
        for (int k = 0; k < vpe; ++k) elems[e * vpe + k] = static_cast<uint32_t>(m->p->polygons[e].vertices[k]);
    }
    return 0;
}

lsbh_tracer *lsbh_tracer_create(lsbh_device *d, int dev, char *err, size_t n)
{
    try {
        auto *t = new lsbh_tracer();
        t->log = std::make_shared<spdlog::logger>("lsbh_tracer");
        t->log->set_echo(false);
        t->p = B200Tracer::create(d->p, nullptr, t->log, dev);
        return t;
    } catch (const std::exception &e) { set_err(err, n, e.what()); return nullptr; }
}
void lsbh_tracer_destroy(lsbh_tracer *t) { delete t; }
int lsbh_tracer_add_geometry(lsbh_tracer *t, const char *name, int type, int nv, int ne)
{
    return t->p->addGeometry(name, static_cast<RTCGeometryType>(type), nv, ne);
}
int lsbh_tracer_remove_geometry(lsbh_tracer *t, const char *name) { return t->p->removeGeometry(name); }
long lsbh_tracer_geometry_count(const lsbh_tracer *t) { return t->p->getGeometryCount(); }
int lsbh_tracer_geometry_id(const lsbh_tracer *t, const char *name) { return t->p->getGeometryId(name); }

int lsbh_tracer_update_geometry(lsbh_tracer *t, const char *name, lsbh_mesh *mesh, const float A[12])
{
    return guarded([&] {
        const Affine3f a = affineFromRows(A);
        PolygonMesh::Ptr m = mesh ? mesh->p : nullptr;
        return t->p->updateGeometry(name, a, m);
    });
}
int lsbh_tracer_update_geometry_euler(lsbh_tracer *t, const char *name, lsbh_mesh *mesh, const float tr[3], const float rot[3])
{
    return guarded([&] {
        PolygonMesh::Ptr m = mesh ? mesh->p : nullptr;
        return t->p->updateGeometry(name, Vector3f(tr[0], tr[1], tr[2]), Vector3f(rot[0], rot[1], rot[2]), m);
    });
}
int lsbh_tracer_commit_scene(lsbh_tracer *t) { return t->p->commitScene(); }
int lsbh_tracer_trace_scene(lsbh_tracer *t, uint32_t frame) { return t->p->traceScene(frame); }
int lsbh_tracer_set_flags(lsbh_tracer *t, int ordered, int ring_zero, int world)
{
    t->p->setOrderedOutput(ordered != 0); t->p->setRingChannel(ring_zero == 0); t->p->setWorldFramePoints(world != 0);
    return 0;
}
int lsbh_tracer_cloud(lsbh_tracer *t, lsb200_cloud_header *h, const uint8_t **data, size_t *n)
{
    PointCloud2::Ptr c = t->p->getTraceCloud();
    h->seq = c->header.seq; h->height = c->height; h->width = c->width; h->point_step = c->point_step; h->row_step = c->row_step;
    h->is_bigendian = c->is_bigendian; h->is_dense = c->is_dense; h->n_fields = (uint8_t)c->fields.size(); h->reserved = 0;
    *data = c->data.data(); *n = c->data.size();
    return 0;
}
int lsbh_tracer_set_sensor_limits(lsbh_tracer *t, int on) { return t->p->setSensorLimits(on != 0); }
int lsbh_tracer_set_intensity(lsbh_tracer *t, int model, int filter) { return t->p->setIntensity(model, filter != 0); }
int lsbh_tracer_set_watertight(lsbh_tracer *t, int on) { return t->p->setWatertight(on != 0); }
// what the tracer logged through the std::shared_ptr<spdlog::logger> it was created with (shim logger keeps its lines)
int lsbh_tracer_log_count(lsbh_tracer *t) { return (int)t->log->lines().size(); }
int lsbh_tracer_log_line(lsbh_tracer *t, int i, int *level, char *buf, size_t n)
{
    const auto lines = t->log->lines();
    if (i < 0 || (size_t)i >= lines.size()) return -1;
    *level = lines[i].first;
    std::snprintf(buf, n, "%s", lines[i].second.c_str());
    return 0;
}
int lsbh_device_limits(const lsbh_device *d, float *radius, float *az_lo, float *az_hi)
{
    *radius = d->p->detectionRadius();
    d->p->horizontalAngleRange(*az_lo, *az_hi);
    return 0;
}
long lsbh_tracer_serialize_cloud(lsbh_tracer *t, double stamp, int fix_row_step, uint8_t *out, size_t capacity)
{
    PointCloud2::Ptr c = t->p->getTraceCloud();
    if (stamp >= 0.0) c->header.stamp = Time(stamp);
    std::vector<uint8_t> msg;
    const size_t n = serializePointCloud2(*c, msg, fix_row_step != 0);
    if (n != serializedLength(*c)) return -2;
    if (out && n <= capacity) std::memcpy(out, msg.data(), n);
    return (long)n;
}
long lsbh_cloud_reserialize(const uint8_t *msg, size_t len, uint8_t *out, size_t capacity, lsb200_cloud_header *hdr, double *stamp)
{
    PointCloud2 c;
    const size_t used = deserializePointCloud2(msg, len, c);
    if (used == 0) return -1;
    if (hdr) {
        hdr->seq = c.header.seq; hdr->height = c.height; hdr->width = c.width; hdr->point_step = c.point_step; hdr->row_step = c.row_step;
        hdr->is_bigendian = c.is_bigendian; hdr->is_dense = c.is_dense; hdr->n_fields = (uint8_t)c.fields.size(); hdr->reserved = 0;
    }
    if (stamp) *stamp = c.header.stamp.toSec();
    std::vector<uint8_t> again;
    const size_t n = serializePointCloud2(c, again, false);
    if (out && n <= capacity) std::memcpy(out, again.data(), n);
    return (long)n;
}
int lsbh_tracer_append_cloud_to_file(lsbh_tracer *t, const char *path, int fix_row_step)
{
    return appendCloudToFile(path, *t->p->getTraceCloud(), fix_row_step != 0) ? 0 : -1;
}
int lsbh_tracer_trace_raw(lsbh_tracer *t, float *t_out, int32_t *tri_out)
{
    std::vector<float> tt; std::vector<int32_t> tri;
    const int rc = t->p->traceRaw(tt, tri);
    std::memcpy(t_out, tt.data(), tt.size() * 4); std::memcpy(tri_out, tri.data(), tri.size() * 4);
    return rc;
}

int lsbh_trajectory_poses(const char *path, float period, float *poses6, int cap)
{
    Trajectory tr;
    if (tr.load(path) < 0) return -1;
    AffineMesh::Ptr am = AffineMesh::create("mesh");
    const size_t n = tr.tickCount(period);
    for (size_t k = 0; k < n; ++k) {
        tr.applyTick(*am, k, period);
        if ((int)k < cap) {
            const Vector3f l = am->getLinearDisplacementConst(), a = am->getAngularDisplacementConst();
            for (int j = 0; j < 3; ++j) { poses6[6 * k + j] = l[j]; poses6[6 * k + 3 + j] = a[j]; }
        }
    }
    return (int)n;
}

}  // extern "C"
