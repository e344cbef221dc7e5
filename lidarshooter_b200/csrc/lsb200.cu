// lsb200.cu -- context, memory management and the C ABI (include/lsb200.h) of liblsb200.so.
// Host side is C++17; every entry point is extern "C", takes plain pointers and sizes, catches everything.
#include "../../include/lsb200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include <sched.h>
#include <sys/syscall.h>
#include <unistd.h>

#include "common.cuh"
#include "radix_sort.cuh"
#include "build.cuh"
#include "trace.cuh"
#include <nvtx3/nvToolsExt.h>

using namespace lsb;

namespace {

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    // grows geometrically; contents are NOT preserved unless keep == true
    cudaError_t reserve(size_t bytes, cudaStream_t s, bool keep = false, bool zero_new = false)
    {
        if (bytes <= cap) return cudaSuccess;
        size_t ncap = std::max(bytes, cap + cap / 2);
        ncap = (ncap + 255) & ~size_t(255);
        void *np = nullptr;
        cudaError_t e = cudaMalloc(&np, ncap);
        if (e != cudaSuccess) return e;
        if (zero_new) { e = cudaMemsetAsync(np, 0, ncap, s); if (e != cudaSuccess) return e; }
        if (keep && p && cap) { e = cudaMemcpyAsync(np, p, cap, cudaMemcpyDeviceToDevice, s); if (e != cudaSuccess) return e; }
        if (p) { cudaStreamSynchronize(s); cudaFree(p); }
        p = np; cap = ncap;
        return cudaSuccess;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct Geometry {
    std::string name;
    int id = -1, type = 0, n_verts = 0, n_elems = 0, verts_per_elem = 3, stride = 12;
    DevBuf verts, elems;
    bool has_verts = false, has_elems = false;
    float A[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    int n_tris() const { return verts_per_elem == 4 ? 2 * n_elems : n_elems; }
};

struct Sensor {
    bool set = false;
    int V = 0, H = 0;
    float R[9], Rinv[9], T[3], q[4];
    DevBuf tables;                     // sin_t[V] cos_t[V] sin_p[H] cos_p[H]
    DevBuf az_mask;                    // [H] bytes, only while an azimuth window is set
    bool has_mask = false;
    float max_range = INFINITY;
    std::vector<float> az_deg;         // azimuth of every step, as the ray generator computes it
    DevBuf t, tri, points, blocks;     // per-ray results, compacted points, ordered-compaction scratch
    DevBuf points_b, t_b;              // second output slot: lsb200_run_frames overlaps its D2H with the next frame
    std::vector<float> elev, h_tab;    // the set_sensor arguments (lsb200_clone_scene) and the host copy of the tables
    float az_begin = 0.f, az_end = 0.f, lim_range = 0.f, lim_lo = -180.f, lim_hi = 180.f;
    float min_intensity = 0.f;         // lsb200_set_sensor_intensity_filter (0 = off)
    DevBuf inten;                      // per-ray intensity (only with an intensity model / filter)
    DevBuf t2, tri2;                   // second return per ray (LSB200_OPT_RETURNS = 2)
    int n_rays() const { return V * H; }
};

}  // namespace

struct lsb200_ctx {
    int device = 0;
    cudaStream_t stream = nullptr, own_stream = nullptr;
    int frame_mode = LSB200_FRAME_SENSOR;
    int opt_traversal = 6, opt_leaf = 2, opt_top = 1024, opt_timing = 0, opt_ctas_per_sm = 0, opt_stats = 0, opt_morton_uniform = 0, opt_morton_layout = 0, opt_morton_aspect = 60, opt_chain = 1, opt_tile_w_log2 = 3, opt_small_build = 1, opt_fit_rmq = 0, opt_fused_tree = 1, opt_lane_refill = 0, opt_tile_entries = 10, opt_tile_visits = 0, opt_tile_min_units = 4096, opt_super_tiles = 0, opt_watertight = 0, opt_intensity = 0, opt_returns = 1, opt_small_trace_units = 512;
    int sm_count = kSmCount;
    std::map<int, Geometry> geoms;           // ascending geometry id == concatenation order
    std::map<std::string, int> by_name;
    bool topology_dirty = true, committed = false, top_valid = false;
    int built_leaf = 0, built_top = 0, built_watertight = 0, built_fused_tree = 0;
    std::vector<Sensor> sensors;
    bool sensors_dirty = true;
    // scene arrays
    DevBuf d_geoms, d_xverts, d_bounds, d_keys, d_vals, d_keys_tmp, d_vals_tmp, d_sort_scratch, d_tris, d_leaf_ids,
        d_nodes, d_node_parent, d_leaf_parent, d_rendezvous, d_tree_count, d_flags, d_misc /* n_top, work[2] */, d_sensors, d_sensors_b, d_counts, d_sensor_xf, d_counters, d_unit_oct, d_unit_list, d_oct_hist, d_chains, d_chain_counts, d_plan, d_tile_lists, d_tile_counts, d_sup_lists, d_sup_counts, d_tile_iv, d_sup_iv, d_unit_iv;
    std::vector<GeomDesc> h_geoms;
    std::vector<SensorDev> h_sensors, h_sensors_b;
    cudaStream_t copy_stream = nullptr, upload_stream = nullptr;
    cudaEvent_t ev_traced[2] = {}, ev_copied[2] = {};
    cudaEvent_t ev_xform = nullptr, ev_uploaded = nullptr;   // lsb200_run_frames: vertex upload of frame f+1 overlaps frame f after its transform
    // octant queues of the traversal (variant 0): valid for (units_first, units_count, units_epoch, units_world)
    int64_t sensors_epoch = 0, units_epoch = -1;
    int units_first = -1, units_count = -1, units_world = -1;
    int64_t units_build_id = 0;            // bumped whenever the queues are rebuilt: their order is not reproducible (atomics),
                                           // so everything indexed by queue slot (tile intervals, tile lists) dies with them
    int64_t iv_units_id = -1, tile_units_id = -1;
    int64_t commit_epoch = 0, chain_commit_epoch = -1, chain_sensors_epoch = -1;
    // tile entry lists: valid for one BVH, one set of sensor poses, one set of octant queues and one list size
    int64_t tile_commit_epoch = -1, tile_sensors_epoch = -1;
    int tile_first = -1, tile_count = -1, tile_cap = -1, tile_visits = -1;
    // direction intervals of the tiles / supertiles: valid for one set of sensor tables + poses and one traced range
    int64_t iv_sensors_epoch = -1;
    int iv_first = -1, iv_count = -1, iv_world = -1, iv_super = -1;
    float h_xf[12];
    uint32_t *h_count = nullptr;                                 // pinned
    int n_verts = 0, n_tris = 0;
    lsb200_stats stats{};
    cudaEvent_t ev[12] = {};
    // per-device kernel state (function attributes / occupancy are per device, never per process)
    bool trace_attr_set[32] = {};
    int trace_occ[32] = {};
    size_t trace_occ_smem[32] = {};
    lsb200_log_fn log_fn = nullptr;
    void *log_user = nullptr;
    std::string err;
    std::mutex mtx;
};

namespace {

// process-wide default sink (contexts inherit it at creation; also used when there is no context yet)
lsb200_log_fn g_log_fn = nullptr;
void *g_log_user = nullptr;
std::mutex g_log_mtx;

void emit_log(lsb200_ctx *c, int level, const char *msg)
{
    lsb200_log_fn fn = c ? c->log_fn : nullptr;
    void *user = c ? c->log_user : nullptr;
    if (!fn) { std::lock_guard<std::mutex> lk(g_log_mtx); fn = g_log_fn; user = g_log_user; }
    if (fn) fn(level, msg, user);
}

void logf(lsb200_ctx *c, int level, const char *fmt, ...)
{
    if (!(c && c->log_fn) && !g_log_fn) return;          // nobody listens: do not format
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    emit_log(c, level, buf);
}

// every error path of the ABI goes through here: last_error text + one LSB200_LOG_ERROR line to the log sink
int fail(lsb200_ctx *c, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    if (c) c->err = buf;
    emit_log(c, LSB200_LOG_ERROR, buf);
    return code;
}

#define LSB_CU(c, call)                                                                                  \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return fail((c), LSB200_ERR_CUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
    } while (0)

// ---- small host linear algebra in the reference's evaluation order (oracle/lsb_oracle.c mirrors it) -------
inline float dot3e(float a0, float b0, float a1, float b1, float a2, float b2)
{
    const float p0 = a0 * b0, p1 = a1 * b1, p2 = a2 * b2;
    const float s12 = p1 + p2;
    return p0 + s12;
}
void mat3_mul(const float *A, const float *B, float *C)
{
    float T[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            T[3 * i + j] = dot3e(A[3 * i], B[j], A[3 * i + 1], B[3 + j], A[3 * i + 2], B[6 + j]);
    std::memcpy(C, T, sizeof(T));
}
// Eigen::Quaternionf::toRotationMatrix(), no normalisation (SRC/LidarDevice.cpp:808-812)
void quat_to_rotation(const float q[4], float R[9])
{
    const float w = q[0], x = q[1], y = q[2], z = q[3];
    const float tx = 2.0f * x, ty = 2.0f * y, tz = 2.0f * z;
    const float twx = tx * w, twy = ty * w, twz = tz * w;
    const float txx = tx * x, txy = ty * x, txz = tz * x;
    const float tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R[0] = 1.0f - (tyy + tzz); R[1] = txy - twz;          R[2] = txz + twy;
    R[3] = txy + twz;          R[4] = 1.0f - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy;          R[7] = tyz + twx;          R[8] = 1.0f - (txx + tyy);
}
inline float cof3(const float *M, int i, int j)
{
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return M[3 * i1 + j1] * M[3 * i2 + j2] - M[3 * i1 + j2] * M[3 * i2 + j1];
}
// Eigen::Matrix3f::inverse() (SRC/LidarDevice.cpp:813): cofactors / determinant
void inverse3(const float M[9], float Mi[9])
{
    const float c0 = cof3(M, 0, 0), c1 = cof3(M, 1, 0), c2 = cof3(M, 2, 0);
    const float det = dot3e(c0, M[0], c1, M[3], c2, M[6]);
    const float invdet = 1.0f / det;
    Mi[0] = c0 * invdet; Mi[1] = c1 * invdet; Mi[2] = c2 * invdet;
    for (int j = 1; j < 3; ++j)
        for (int i = 0; i < 3; ++i) Mi[3 * j + i] = cof3(M, i, j) * invdet;
}

Geometry *find_geom(lsb200_ctx *c, const char *name)
{
    if (!name) return nullptr;
    auto it = c->by_name.find(name);
    if (it == c->by_name.end()) return nullptr;
    return &c->geoms[it->second];
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

void count_launch(lsb200_ctx *c, int n = 1) { c->stats.kernel_launches += n; }

SceneDev scene_dev(lsb200_ctx *c)
{
    SceneDev S{};
    S.geoms = c->d_geoms.as<GeomDesc>(); S.n_geoms = (int)c->geoms.size();
    S.n_verts = c->n_verts; S.n_tris = c->n_tris;
    S.xverts = c->d_xverts.as<float4>(); S.bounds = c->d_bounds.as<uint32_t>();
    S.keys = c->d_keys.as<uint32_t>(); S.vals = c->d_vals.as<uint32_t>();
    S.tris = c->d_tris.as<float4>(); S.leaf_ids = c->d_leaf_ids.as<int32_t>();
    S.nodes = c->d_nodes.as<float4>(); S.node_parent = c->d_node_parent.as<int32_t>();
    S.leaf_parent = c->d_leaf_parent.as<int32_t>(); S.flags = c->d_flags.as<uint32_t>();
    S.top_cap = c->built_top; S.leaf_max = c->built_leaf;
    S.morton_uniform = c->opt_morton_uniform;
    S.morton_bxy = c->opt_morton_layout >> 4; S.morton_bz = c->opt_morton_layout & 15;
    S.morton_alpha = 0.01f * (float)c->opt_morton_aspect;
    S.plan = c->d_plan.as<MortonPlan>();
    S.watertight = c->built_watertight;
    return S;
}

int upload_geometry(lsb200_ctx *c, Geometry &g, const void *verts, int stride, const uint32_t *elems, cudaStream_t on = nullptr)
{
    const cudaStream_t ust = on ? on : c->stream;
    if (verts) {
        if (stride < 12 || (stride & 3)) return fail(c, LSB200_ERR_INVALID, "vertex stride %d must be >= 12 and a multiple of 4", stride);
        const size_t bytes = (size_t)g.n_verts * (size_t)stride;
        LSB_CU(c, g.verts.reserve(std::max<size_t>(bytes, 16), ust));
        if (bytes) LSB_CU(c, cudaMemcpyAsync(g.verts.p, verts, bytes, cudaMemcpyHostToDevice, ust));
        g.stride = stride; g.has_verts = true;
    }
    if (elems) {
        const size_t bytes = (size_t)g.n_elems * g.verts_per_elem * sizeof(uint32_t);
        LSB_CU(c, g.elems.reserve(std::max<size_t>(bytes, 16), c->stream));
        if (bytes) LSB_CU(c, cudaMemcpyAsync(g.elems.p, elems, bytes, cudaMemcpyHostToDevice, c->stream));
        g.has_elems = true;
        c->topology_dirty = true;
    }
    return 0;
}

}  // namespace

// =============================================================================================================
//  C ABI
// =============================================================================================================
extern "C" {

const char *lsb200_version(void) { return "lsb200 0.1 (sm_100a)"; }

const char *lsb200_last_error(const lsb200_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int lsb200_create(int cuda_device, lsb200_ctx **out)
{
    if (!out) return LSB200_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0)                            // no CPU fallback, by design
        return fail(nullptr, LSB200_ERR_CUDA, "lsb200_create: no CUDA device (this library has no CPU path)");
    if (cuda_device < 0 || cuda_device >= n) return fail(nullptr, LSB200_ERR_INVALID, "lsb200_create: CUDA device %d of %d", cuda_device, n);
    lsb200_ctx *c = new (std::nothrow) lsb200_ctx();
    if (!c) return LSB200_ERR_NOMEM;
    c->device = cuda_device;
    DeviceGuard dg(cuda_device);
    cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, cuda_device);
    // kernels that need more than 48 KB of dynamic shared memory: the opt-in is a per-DEVICE attribute
    if (rsort::prepare_device() != cudaSuccess ||
#ifdef LSB_LAB
        cudaFuncSetAttribute(k_fit_rmq, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RmqTable)) != cudaSuccess ||
#endif
        cudaFuncSetAttribute(k_build_small, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TreeSmem<kSmallThreads>)) != cudaSuccess) {
        delete c; return LSB200_ERR_CUDA;
    }
    if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return LSB200_ERR_CUDA; }
    c->stream = c->own_stream;
    for (auto &e : c->ev) cudaEventCreate(&e);
    cudaMallocHost((void **)&c->h_count, (16 + 2 * kMaxSensors) * sizeof(uint32_t));
    cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&c->upload_stream, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&c->ev_xform, cudaEventDisableTiming); cudaEventCreateWithFlags(&c->ev_uploaded, cudaEventDisableTiming);
    for (int k = 0; k < 2; ++k) { cudaEventCreateWithFlags(&c->ev_traced[k], cudaEventDisableTiming); cudaEventCreateWithFlags(&c->ev_copied[k], cudaEventDisableTiming); }
    if (c->d_bounds.reserve(256, c->stream) != cudaSuccess || c->d_misc.reserve(256, c->stream, false, true) != cudaSuccess ||
        c->d_plan.reserve(256, c->stream, false, true) != cudaSuccess ||
        c->d_sensor_xf.reserve(256, c->stream) != cudaSuccess) { delete c; return LSB200_ERR_CUDA; }
    c->sensors.resize(1);
    { std::lock_guard<std::mutex> lk(g_log_mtx); c->log_fn = g_log_fn; c->log_user = g_log_user; }
    logf(c, LSB200_LOG_INFO, "lsb200: context created on CUDA device %d (%d SMs)", cuda_device, c->sm_count);
    *out = c;
    return LSB200_OK;
}

int lsb200_destroy(lsb200_ctx *c)
{
    if (!c) return LSB200_ERR_INVALID;
    {
        DeviceGuard dg(c->device);
        cudaStreamSynchronize(c->stream);
        for (auto &kv : c->geoms) { kv.second.verts.release(); kv.second.elems.release(); }
        for (auto &s : c->sensors) { s.tables.release(); s.az_mask.release(); s.t.release(); s.tri.release(); s.points.release(); s.points_b.release(); s.t_b.release(); s.blocks.release(); s.inten.release(); s.t2.release(); s.tri2.release(); }
        DevBuf *bufs[] = {&c->d_geoms, &c->d_xverts, &c->d_bounds, &c->d_keys, &c->d_vals, &c->d_keys_tmp, &c->d_vals_tmp,
                          &c->d_sort_scratch, &c->d_tris, &c->d_leaf_ids, &c->d_nodes, &c->d_node_parent, &c->d_leaf_parent, &c->d_rendezvous, &c->d_tree_count,
                          &c->d_flags, &c->d_misc, &c->d_sensors, &c->d_sensors_b, &c->d_counts, &c->d_sensor_xf, &c->d_counters, &c->d_unit_oct, &c->d_unit_list, &c->d_oct_hist, &c->d_chains, &c->d_chain_counts, &c->d_plan, &c->d_tile_lists, &c->d_tile_counts, &c->d_sup_lists, &c->d_sup_counts, &c->d_tile_iv, &c->d_sup_iv, &c->d_unit_iv};
        for (auto *b : bufs) b->release();
        if (c->h_count) cudaFreeHost(c->h_count);
        for (auto &e : c->ev) if (e) cudaEventDestroy(e);
        for (int k = 0; k < 2; ++k) { if (c->ev_traced[k]) cudaEventDestroy(c->ev_traced[k]); if (c->ev_copied[k]) cudaEventDestroy(c->ev_copied[k]); }
        if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
        if (c->upload_stream) { cudaStreamSynchronize(c->upload_stream); cudaStreamDestroy(c->upload_stream); }
        if (c->ev_xform) cudaEventDestroy(c->ev_xform);
        if (c->ev_uploaded) cudaEventDestroy(c->ev_uploaded);
        if (c->own_stream) cudaStreamDestroy(c->own_stream);
    }
    delete c;
    return LSB200_OK;
}

int lsb200_set_stream(lsb200_ctx *c, void *cuda_stream)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    DeviceGuard dg(c->device);
    cudaStreamSynchronize(c->stream);
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return LSB200_OK;
}

int lsb200_set_frame_mode(lsb200_ctx *c, int mode)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (mode != LSB200_FRAME_SENSOR && mode != LSB200_FRAME_WORLD) return fail(c, LSB200_ERR_INVALID, "bad frame mode %d", mode);
    if (mode != c->frame_mode) { c->frame_mode = mode; c->committed = false; }
    return LSB200_OK;
}

int lsb200_set_option(lsb200_ctx *c, int option, int value)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    switch (option) {
#ifdef LSB_LAB
    case LSB200_OPT_TRAVERSAL: if (value < 0 || value > 7) break; c->opt_traversal = value; return LSB200_OK;
    case LSB200_OPT_TRACE_STATS: c->opt_stats = value ? 1 : 0; return LSB200_OK;
    case LSB200_OPT_FIT_RMQ: c->opt_fit_rmq = value ? 1 : 0; return LSB200_OK;
    case LSB200_OPT_SUPER_TILES: c->opt_super_tiles = value ? 1 : 0; c->tile_commit_epoch = -1; return LSB200_OK;
#else
    // product build: the shipped traversal (6) and brute force (3); the evidence variants, the instrumented launches,
    // the range-query fit and the two-level lists live in liblsb200_lab.so (make lab)
    case LSB200_OPT_TRAVERSAL: if (value != 6 && value != 3) return fail(c, LSB200_ERR_INVALID, "traversal variant %d is built into liblsb200_lab.so only", value); c->opt_traversal = value; return LSB200_OK;
    case LSB200_OPT_TRACE_STATS: case LSB200_OPT_FIT_RMQ: case LSB200_OPT_SUPER_TILES:
        if (value) return fail(c, LSB200_ERR_INVALID, "option %d is built into liblsb200_lab.so only", option);
        return LSB200_OK;
#endif
    case LSB200_OPT_ORIGIN_CHAIN: c->opt_chain = value ? 1 : 0; return LSB200_OK;
    case LSB200_OPT_TILE_ENTRIES: if (value < 0 || value > kMaxTileEntries || value == 1) break; c->opt_tile_entries = value; c->tile_commit_epoch = -1; return LSB200_OK;
    case LSB200_OPT_TILE_MIN_UNITS: if (value < 0) break; c->opt_tile_min_units = value; return LSB200_OK;
    case LSB200_OPT_TILE_VISITS: if (value < 0 || value > 4096) break; c->opt_tile_visits = value; c->tile_commit_epoch = -1; return LSB200_OK;
    case LSB200_OPT_SMALL_BUILD: c->opt_small_build = value ? 1 : 0; c->topology_dirty = true; return LSB200_OK;
    case LSB200_OPT_SMALL_TRACE_UNITS: if (value < 0) break; c->opt_small_trace_units = value; return LSB200_OK;
    case LSB200_OPT_RETURNS: if (value < 1 || value > 2) break; c->opt_returns = value; return LSB200_OK;
    case LSB200_OPT_LANE_REFILL:
        if (value < 0 || value > 15) break;
#if !LSB_REFILL
        if (value) return fail(c, LSB200_ERR_INVALID, "lane refill is built into liblsb200_lab.so only (measured slower, see DESIGN.md)");
#endif
        c->opt_lane_refill = value; return LSB200_OK;
    case LSB200_OPT_FUSED_TREE: c->opt_fused_tree = value ? 1 : 0; c->topology_dirty = true; return LSB200_OK;
    case LSB200_OPT_TILE_WIDTH: {
        int w = 0;
        while (w <= 5 && (1 << w) != value) ++w;
        if (w > 5) break;
        c->opt_tile_w_log2 = w; c->sensors_dirty = true; ++c->sensors_epoch; return LSB200_OK;
    }
    case LSB200_OPT_MORTON_UNIFORM: c->opt_morton_uniform = value ? 1 : 0; c->topology_dirty = true; return LSB200_OK;
    case LSB200_OPT_MORTON_ASPECT: if (value < 0 || value > 200) break; c->opt_morton_aspect = value; c->topology_dirty = true; return LSB200_OK;
    case LSB200_OPT_MORTON_LAYOUT: {
        const int bxy = value >> 4, bz = value & 15;
        if (value != 0 && (value < 0 || bxy < 1 || bxy > 16 || bz > 10 || bz > bxy || 2 * bxy + bz > 32)) break;
        c->opt_morton_layout = value; c->topology_dirty = true; return LSB200_OK;
    }
    case LSB200_OPT_LEAF_SIZE: if (value < 1 || value > kMaxLeaf) break; c->opt_leaf = value; c->topology_dirty = true; return LSB200_OK;
    case LSB200_OPT_TOP_NODES: if (value < 0 || value > 3072) break; c->opt_top = value; c->topology_dirty = true; return LSB200_OK;
    case LSB200_OPT_STAGE_TIMING: c->opt_timing = value ? 1 : 0; return LSB200_OK;
    case LSB200_OPT_WATERTIGHT: c->opt_watertight = value ? 1 : 0; c->committed = false; return LSB200_OK;   // records change layout: re-commit
    case LSB200_OPT_INTENSITY_MODEL: if (value < 0 || value > 1) break; c->opt_intensity = value; return LSB200_OK;
    case LSB200_OPT_CTAS_PER_SM: if (value < 0 || value > 32) break; c->opt_ctas_per_sm = value; return LSB200_OK;
    default: break;
    }
    return fail(c, LSB200_ERR_INVALID, "bad option %d = %d", option, value);
}

int lsb200_set_sensor(lsb200_ctx *c, int idx, const float *elev_deg, int V, float az_begin, float az_end, int H,
                      const float q_wxyz[4], const float t_xyz[3])
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (idx < 0 || idx >= kMaxSensors || !elev_deg || V <= 0 || H <= 0 || !q_wxyz || !t_xyz)
        return fail(c, LSB200_ERR_INVALID, "set_sensor: bad arguments");
    if ((int64_t)V * H > (int64_t)1 << 30) return fail(c, LSB200_ERR_INVALID, "set_sensor: too many rays");
    DeviceGuard dg(c->device);
    if ((size_t)idx >= c->sensors.size()) c->sensors.resize(idx + 1);
    Sensor &s = c->sensors[idx];
    s.V = V; s.H = H;
    std::memcpy(s.q, q_wxyz, 16); std::memcpy(s.T, t_xyz, 12);
    quat_to_rotation(s.q, s.R);
    inverse3(s.R, s.Rinv);
    // tables with the reference's narrowing: theta, phi are double expressions narrowed to float BEFORE the
    // float sin/cos (SRC/LidarDevice.cpp:305-318); step = (end - begin) / (count - 1) (SRC/LidarDevice.cpp:611)
    std::vector<float> tab(2 * (size_t)V + 2 * (size_t)H);
    for (int v = 0; v < V; ++v) {
        const float theta = (float)((90.0 - (double)elev_deg[v]) * M_PI / 180.0);
        tab[v] = sinf(theta); tab[V + v] = cosf(theta);
    }
    const float step = (az_end - az_begin) / (float)((unsigned)H - 1u);
    for (int h = 0; h < H; ++h) {
        const float prePhi = az_begin + step * (float)h;
        const float phi = (float)((double)prePhi * M_PI / 180.0);
        tab[2 * V + h] = sinf(phi); tab[2 * V + H + h] = cosf(phi);
    }
    s.az_deg.resize(H);
    for (int h = 0; h < H; ++h) s.az_deg[h] = az_begin + step * (float)h;
    s.has_mask = false; s.max_range = INFINITY;
    s.elev.assign(elev_deg, elev_deg + V); s.az_begin = az_begin; s.az_end = az_end;
    s.lim_range = 0.f; s.lim_lo = -180.f; s.lim_hi = 180.f;
    s.h_tab = tab; s.min_intensity = 0.f;
    LSB_CU(c, s.tables.reserve(tab.size() * 4, c->stream));
    LSB_CU(c, cudaMemcpyAsync(s.tables.p, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, c->stream));
    LSB_CU(c, cudaStreamSynchronize(c->stream));     // tab is a local
    s.set = true;
    c->sensors_dirty = true;
    ++c->sensors_epoch;
    if (idx == 0 && c->frame_mode == LSB200_FRAME_SENSOR) c->committed = false;   // scene lives in this sensor's frame
    return LSB200_OK;
}

// pose-only update of a sensor that is already set: no table upload, no synchronisation (a vehicle-mounted LiDAR moves
// every frame; its ray tables do not).  Everything derived from the pose -- octant queues, tile intervals, origin chains,
// tile lists -- is rebuilt by the next trace.
int lsb200_set_sensor_pose(lsb200_ctx *c, int idx, const float q_wxyz[4], const float t_xyz[3])
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set || !q_wxyz || !t_xyz)
        return fail(c, LSB200_ERR_INVALID, "set_sensor_pose: sensor %d not set", idx);
    Sensor &s = c->sensors[idx];
    std::memcpy(s.q, q_wxyz, 16); std::memcpy(s.T, t_xyz, 12);
    quat_to_rotation(s.q, s.R);
    inverse3(s.R, s.Rinv);
    c->sensors_dirty = true;
    ++c->sensors_epoch;
    if (idx == 0 && c->frame_mode == LSB200_FRAME_SENSOR) c->committed = false;   // scene lives in this sensor's frame
    return LSB200_OK;
}

int lsb200_set_sensor_limits(lsb200_ctx *c, int idx, float max_range, float az_min_deg, float az_max_deg)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set)
        return fail(c, LSB200_ERR_INVALID, "set_sensor_limits: sensor %d not set", idx);
    if (!(az_min_deg <= az_max_deg)) return fail(c, LSB200_ERR_INVALID, "set_sensor_limits: empty azimuth window");
    DeviceGuard dg(c->device);
    Sensor &s = c->sensors[idx];
    s.lim_range = max_range; s.lim_lo = az_min_deg; s.lim_hi = az_max_deg;
    s.max_range = (max_range > 0.0f) ? max_range : INFINITY;
    s.has_mask = !(az_min_deg <= -180.0f && az_max_deg >= 180.0f);
    if (s.has_mask) {
        std::vector<uint8_t> m((size_t)s.H);
        for (int h = 0; h < s.H; ++h) {
            float w = s.az_deg[h];                       // wrapped into (-180, 180]
            while (w > 180.0f) w -= 360.0f;
            while (w <= -180.0f) w += 360.0f;
            m[h] = (w >= az_min_deg && w <= az_max_deg) ? 1 : 0;
        }
        LSB_CU(c, s.az_mask.reserve(m.size(), c->stream));
        LSB_CU(c, cudaMemcpyAsync(s.az_mask.p, m.data(), m.size(), cudaMemcpyHostToDevice, c->stream));
        LSB_CU(c, cudaStreamSynchronize(c->stream));
    }
    c->sensors_dirty = true;
    c->tile_commit_epoch = -1;          // the tile lists bake the detection radius in
    return LSB200_OK;
}

int lsb200_set_sensor_intensity_filter(lsb200_ctx *c, int idx, float min_intensity)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set)
        return fail(c, LSB200_ERR_INVALID, "set_sensor_intensity_filter: sensor %d not set", idx);
    if (!(min_intensity >= 0.0f)) return fail(c, LSB200_ERR_INVALID, "set_sensor_intensity_filter: negative threshold");
    c->sensors[idx].min_intensity = min_intensity;
    c->sensors_dirty = true;
    return LSB200_OK;
}

int lsb200_get_ray_count(const lsb200_ctx *c, int idx)
{
    if (!c || idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set) return LSB200_ERR_INVALID;
    return c->sensors[idx].n_rays();
}

int lsb200_add_geometry(lsb200_ctx *c, const char *name, int type, int n_vertices, int n_elements)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (!name || n_vertices < 0 || n_elements < 0) return fail(c, LSB200_ERR_INVALID, "add_geometry: bad arguments");
    if (type != LSB200_GEOMETRY_TRIANGLE && type != LSB200_GEOMETRY_QUAD)
        return fail(c, LSB200_ERR_INVALID, "add_geometry: unsupported geometry type %d", type);
    if (c->by_name.count(name)) return fail(c, LSB200_ERR_INVALID, "add_geometry: '%s' already exists", name);
    int id = 0;
    while (c->geoms.count(id)) ++id;           // lowest free id, like rtcAttachGeometry
    Geometry &g = c->geoms[id];
    g.name = name; g.id = id; g.type = type; g.n_verts = n_vertices; g.n_elems = n_elements;
    g.verts_per_elem = (type == LSB200_GEOMETRY_QUAD) ? 4 : 3;
    c->by_name[name] = id;
    c->topology_dirty = true;
    c->committed = false;
    return id;
}

int lsb200_get_geometry_count(const lsb200_ctx *c) { return c ? (int)c->geoms.size() : LSB200_ERR_INVALID; }

int lsb200_get_geometry_id(const lsb200_ctx *c, const char *name)
{
    if (!c || !name) return LSB200_ERR_INVALID;
    auto it = c->by_name.find(name);
    return it == c->by_name.end() ? -1 : it->second;
}

void lsb200_affine_from_components(const float t[3], const float r[3], float A[12])
{
    // Translation(t) * Rz * Ry * Rx  (SRC/MeshTransformer.cpp:467-477; Eigen AngleAxisf about the unit axes)
    const float cx = cosf(r[0]), sx = sinf(r[0]), cy = cosf(r[1]), sy = sinf(r[1]), cz = cosf(r[2]), sz = sinf(r[2]);
    const float Rx[9] = {1, 0, 0, 0, cx, -sx, 0, sx, cx};
    const float Ry[9] = {cy, 0, sy, 0, 1, 0, -sy, 0, cy};
    const float Rz[9] = {cz, -sz, 0, sz, cz, 0, 0, 0, 1};
    float Rzy[9], L[9];
    mat3_mul(Rz, Ry, Rzy);
    mat3_mul(Rzy, Rx, L);
    for (int i = 0; i < 3; ++i) { A[4 * i] = L[3 * i]; A[4 * i + 1] = L[3 * i + 1]; A[4 * i + 2] = L[3 * i + 2]; A[4 * i + 3] = t[i]; }
}

int lsb200_update_geometry(lsb200_ctx *c, const char *name, const void *verts, int stride, const uint32_t *elems,
                           const float A[12])
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    Geometry *g = find_geom(c, name);
    if (!g) return fail(c, LSB200_ERR_INVALID, "update_geometry: unknown geometry '%s'", name ? name : "(null)");
    if (!A) return fail(c, LSB200_ERR_INVALID, "update_geometry: null transform");
    DeviceGuard dg(c->device);
    int rc = upload_geometry(c, *g, verts, stride, elems);
    if (rc) return rc;
    std::memcpy(g->A, A, sizeof(g->A));
    return LSB200_OK;
}

int lsb200_update_geometry_euler(lsb200_ctx *c, const char *name, const void *verts, int stride, const uint32_t *elems,
                                 const float t[3], const float r[3])
{
    if (!c || !t || !r) return LSB200_ERR_INVALID;
    float A[12];
    lsb200_affine_from_components(t, r, A);
    return lsb200_update_geometry(c, name, verts, stride, elems, A);
}

int lsb200_update_transform(lsb200_ctx *c, const char *name, const float A[12])
{
    return lsb200_update_geometry(c, name, nullptr, 0, nullptr, A);
}

// ---- commit ----------------------------------------------------------------------------------------------
static int commit_locked(lsb200_ctx *c, int mode)
{
    DeviceGuard dg(c->device);
    int64_t nv = 0, nt = 0;
    for (auto &kv : c->geoms) {
        const Geometry &g = kv.second;
        if (g.n_verts > 0 && g.n_tris() > 0 && !(g.has_verts && g.has_elems))
            return fail(c, LSB200_ERR_STATE, "commit_scene: geometry '%s' was never given vertices/elements", g.name.c_str());
        nv += g.n_verts; nt += g.n_tris();
    }
    if (c->geoms.empty() || nt == 0 || nv == 0) {
        c->committed = false; c->n_tris = 0; c->n_verts = 0;
        return fail(c, LSB200_ERR_EMPTY_SCENE, "commit_scene: empty scene");
    }
    if (nt >= (1 << kLeafShift) || nv >= (int64_t)1 << 31) return fail(c, LSB200_ERR_INVALID, "commit_scene: scene too large");
    if (c->frame_mode == LSB200_FRAME_SENSOR && !c->sensors[0].set)
        return fail(c, LSB200_ERR_STATE, "commit_scene: sensor 0 not set (sensor-frame mode)");
    // a refit needs a hierarchy that was built for exactly this set of geometries
    const bool rebuild = (mode == LSB200_COMMIT_REBUILD) || c->topology_dirty;
    cudaStream_t st = c->stream;
    c->n_verts = (int)nv; c->n_tris = (int)nt;
    if (rebuild) { c->built_leaf = c->opt_leaf; c->built_top = c->opt_top; c->built_fused_tree = (c->opt_fused_tree && !c->opt_fit_rmq && nt > 1) ? 1 : 0; }
    c->built_watertight = c->opt_watertight;        // the leaf records are rewritten by every commit, rebuild or refit

    // geometry descriptor table (small pageable copy: staged by the runtime before the call returns)
    const size_t ng = c->geoms.size();
    c->h_geoms.resize(ng);
    {
        size_t k = 0; int voff = 0, toff = 0;
        for (auto &kv : c->geoms) {
            Geometry &g = kv.second;
            GeomDesc &d = c->h_geoms[k++];
            d.verts = g.verts.as<uint8_t>(); d.elems = g.elems.as<uint32_t>();
            std::memcpy(d.A, g.A, sizeof(d.A));
            d.stride = g.stride; d.verts_per_elem = g.verts_per_elem;
            d.vert_off = voff; d.n_verts = g.n_verts; d.tri_off = toff; d.n_tris = g.n_tris();
            voff += g.n_verts; toff += g.n_tris();
        }
    }
    // small scenes: one fused kernel, descriptors passed as kernel parameters (build.cuh: k_build_small)
    const bool small = c->opt_small_build && nt <= kSmallMaxTris && ng <= (size_t)kSmallMaxGeoms;
    if (!small) {
        LSB_CU(c, c->d_geoms.reserve(ng * sizeof(GeomDesc), st));
        LSB_CU(c, cudaMemcpyAsync(c->d_geoms.p, c->h_geoms.data(), ng * sizeof(GeomDesc), cudaMemcpyHostToDevice, st));
    }

    const size_t T = (size_t)nt;
    LSB_CU(c, c->d_xverts.reserve((size_t)nv * sizeof(float4), st));
    LSB_CU(c, c->d_tris.reserve(T * 48, st));
    LSB_CU(c, c->d_leaf_ids.reserve(T * 4, st));
    if (rebuild) {
        LSB_CU(c, c->d_keys.reserve(T * 4, st)); LSB_CU(c, c->d_vals.reserve(T * 4, st));
        LSB_CU(c, c->d_keys_tmp.reserve(T * 4, st)); LSB_CU(c, c->d_vals_tmp.reserve(T * 4, st));
        if (!small) LSB_CU(c, c->d_sort_scratch.reserve(rsort::scratch_words(nt) * 4, st));
        LSB_CU(c, c->d_nodes.reserve(((size_t)c->built_top + T) * 64, st));
        LSB_CU(c, c->d_node_parent.reserve(T * 4, st)); LSB_CU(c, c->d_leaf_parent.reserve(T * 4, st));
        if (small || c->built_fused_tree) { LSB_CU(c, c->d_rendezvous.reserve(T * 64, st)); LSB_CU(c, c->d_tree_count.reserve(16, st)); }
        LSB_CU(c, c->d_flags.reserve(T * 4, st, false, true));
    }
    SceneDev S = scene_dev(c);
    const bool timing = c->opt_timing != 0;
    int evi = 0;
    // stage boundary: CUDA event (LSB200_OPT_STAGE_TIMING) + NVTX range of the stage that starts here (nullptr ends the last one)
    struct NvtxScope { NvtxScope(const char *n) { nvtxRangePushA(n); } ~NvtxScope() { nvtxRangePop(); } } nvtx_commit("lsb200:commit");
    struct NvtxStage { bool open = false; ~NvtxStage() { if (open) nvtxRangePop(); } } nvtx_stage;      // closed on every return path
    auto mark = [&](const char *next_stage) {
        if (timing) cudaEventRecord(c->ev[evi++], st);
        if (nvtx_stage.open) { nvtxRangePop(); nvtx_stage.open = false; }
        if (next_stage) { nvtxRangePushA(next_stage); nvtx_stage.open = true; }
    };

    if (small) {
        SmallArgs args;
        std::memcpy(args.geoms, c->h_geoms.data(), ng * sizeof(GeomDesc));
        if (c->frame_mode == LSB200_FRAME_SENSOR) { std::memcpy(args.RinvT, c->sensors[0].Rinv, 36); std::memcpy(args.RinvT + 9, c->sensors[0].T, 12); }
        else std::memset(args.RinvT, 0, sizeof(args.RinvT));
        args.keys_tmp = c->d_keys_tmp.as<uint32_t>(); args.vals_tmp = c->d_vals_tmp.as<uint32_t>();
        // origin chains of the sensors folded into the same kernel when there are few of them
        args.n_origins = 0; args.chains = nullptr; args.chain_counts = nullptr;
        const size_t ns = c->sensors.size();
        bool all_set = ns > 0;
        for (const Sensor &sn : c->sensors) all_set = all_set && sn.set;
        if (c->opt_chain && all_set && ns <= 8 && c->opt_traversal != 3) {
            LSB_CU(c, c->d_chains.reserve(ns * 9 * 2 * kMaxChain * sizeof(float4), st));
            LSB_CU(c, c->d_chain_counts.reserve(ns * 9 * 4, st));
            args.n_origins = (int)ns; args.chains = c->d_chains.as<float4>(); args.chain_counts = c->d_chain_counts.as<int>();
            for (size_t i = 0; i < ns; ++i)
                for (int k = 0; k < 3; ++k) args.origins[i][k] = c->frame_mode == LSB200_FRAME_WORLD ? c->sensors[i].T[k] : 0.0f;
        }
        mark("lsb200:build_small");
        // one cluster of kClusterCtas CTAs (__cluster_dims__ on the kernel)
        args.rendezvous = c->d_rendezvous.as<float4>();
        args.work = reinterpret_cast<uint32_t *>(c->d_leaf_parent.as<int32_t>()); args.work_count = c->d_tree_count.as<uint32_t>();
        k_build_small<<<kClusterCtas, kSmallThreads, sizeof(TreeSmem<kSmallThreads>), st>>>(S, args, c->frame_mode == LSB200_FRAME_SENSOR, rebuild ? 1 : 0);
        count_launch(c);
        cudaEventRecord(c->ev_xform, st);
        mark(nullptr);
        c->top_valid = false;
        if (c->built_top > 0 && c->opt_traversal == 4) {
            k_top_level<<<1, 1024, (size_t)c->built_top * sizeof(int), st>>>(S, c->d_misc.as<int>());
            count_launch(c);
            c->top_valid = true;
        }
        LSB_CU(c, cudaGetLastError());
        c->stats.last_commit_was_rebuild = rebuild ? 1 : 0;
        if (timing) {
            LSB_CU(c, cudaStreamSynchronize(st));
            float ms = 0;
            cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]);
            c->stats.ms_transform = c->stats.ms_morton = c->stats.ms_sort = c->stats.ms_hierarchy = c->stats.ms_toplevel = 0;
            c->stats.ms_fit = ms;                 // the fused kernel, reported as one stage
        }
        c->topology_dirty = false;
        c->committed = true;
        ++c->commit_epoch;
        if (args.n_origins > 0) { c->chain_commit_epoch = c->commit_epoch; c->chain_sensors_epoch = c->sensors_epoch; }
        logf(c, LSB200_LOG_DEBUG, "lsb200: commit (%s, fused small build) of %lld triangles / %lld vertices in %zu geometries",
             rebuild ? "rebuild" : "refit", (long long)nt, (long long)nv, ng);
        return LSB200_OK;
    }

    // (1) transform + scene bounds
    if (c->frame_mode == LSB200_FRAME_SENSOR) {
        std::memcpy(c->h_xf, c->sensors[0].Rinv, 36); std::memcpy(c->h_xf + 9, c->sensors[0].T, 12);
        LSB_CU(c, cudaMemcpyAsync(c->d_sensor_xf.p, c->h_xf, 48, cudaMemcpyHostToDevice, st));
    }
    LSB_CU(c, cudaMemsetAsync(S.bounds, 0xFF, 12, st));
    LSB_CU(c, cudaMemsetAsync(S.bounds + 3, 0x00, 12, st));
    mark("lsb200:transform");
    k_transform<<<(unsigned)((nv + kXformBlock - 1) / kXformBlock), 256, 0, st>>>(S, c->frame_mode == LSB200_FRAME_SENSOR, c->d_sensor_xf.as<float>());
    count_launch(c);
    cudaEventRecord(c->ev_xform, st);        // the raw vertex blobs have been read: the next frame's upload may overwrite them
    mark(rebuild ? "lsb200:morton" : "lsb200:tree");
    if (rebuild) {
        // (2) Morton codes  (3) radix sort  (4) Karras hierarchy
        k_morton_plan<<<1, 1, 0, st>>>(S);
        k_morton<<<(unsigned)((nt + kMortonBlock - 1) / kMortonBlock), 256, 0, st>>>(S);
        count_launch(c, 2);
        mark("lsb200:sort");
        count_launch(c, rsort::sort_pairs(S.keys, S.vals, c->d_keys_tmp.as<uint32_t>(), c->d_vals_tmp.as<uint32_t>(), nt,
                                          c->d_sort_scratch.as<uint32_t>(), st));
        mark("lsb200:hierarchy");
        if (!c->built_fused_tree) {
            if (nt > 1) k_hierarchy<<<(unsigned)((nt - 1 + 255) / 256), 256, 0, st>>>(S, S.keys);
            else k_single_leaf_root<<<1, 1, 0, st>>>(S);
            count_launch(c);
        }
        mark("lsb200:tree");
    }
    // (5) leaf records + bottom-up boxes; with the fused tree pass (k_tree, the default) the hierarchy comes out of the same launch
    if (c->built_fused_tree) {
        // phase 2's work list reuses the (otherwise idle) leaf-parent array; its counter is the first word of d_tree_count
        uint32_t *work = reinterpret_cast<uint32_t *>(c->d_leaf_parent.as<int32_t>()), *work_count = c->d_tree_count.as<uint32_t>();
        LSB_CU(c, cudaMemsetAsync(work_count, 0, 4, st));
        k_tree<<<(unsigned)((nt + kTreeBlock - 1) / kTreeBlock), kTreeBlock, 0, st>>>(S, S.keys, S.vals, c->d_rendezvous.as<float4>(), work, work_count);
        // one thread per parked subtree (a block parks ~2 log2(kTreeBlock) of them at most in practice; grid-stride beyond that)
        k_tree_top<<<(unsigned)std::max<int64_t>(kSmCount, ((nt + kTreeBlock - 1) / kTreeBlock) * 16 / 128), 128, 0, st>>>(S, c->d_rendezvous.as<float4>(), work, work_count);
        count_launch(c);
    } else
#ifdef LSB_LAB
    if (c->opt_fit_rmq) {
        k_fit_rmq<<<(unsigned)((nt + kFitBlock - 1) / kFitBlock), kFitBlock, sizeof(RmqTable), st>>>(S, S.vals);
    } else
#endif
    {
        k_fit<<<(unsigned)((nt + kFitBlock - 1) / kFitBlock), kFitBlock, 0, st>>>(S, S.vals);
    }
    count_launch(c);
    mark("lsb200:toplevel");
    // (6) top-level block for the TMA-staged traversal
    c->top_valid = false;
    if (c->built_top > 0 && c->opt_traversal == 4) {       // only the TMA-staged variant reads the top block
        k_top_level<<<1, 1024, (size_t)c->built_top * sizeof(int), st>>>(S, c->d_misc.as<int>());
        count_launch(c);
        c->top_valid = true;
    }
    mark(nullptr);
    LSB_CU(c, cudaGetLastError());
    c->stats.last_commit_was_rebuild = rebuild ? 1 : 0;
    if (timing) {
        LSB_CU(c, cudaStreamSynchronize(st));
        float ms = 0; int k = 0;
        cudaEventElapsedTime(&ms, c->ev[k], c->ev[k + 1]); c->stats.ms_transform = ms; ++k;
        if (rebuild) {
            cudaEventElapsedTime(&ms, c->ev[k], c->ev[k + 1]); c->stats.ms_morton = ms; ++k;
            cudaEventElapsedTime(&ms, c->ev[k], c->ev[k + 1]); c->stats.ms_sort = ms; ++k;
            cudaEventElapsedTime(&ms, c->ev[k], c->ev[k + 1]); c->stats.ms_hierarchy = ms; ++k;
        } else { c->stats.ms_morton = c->stats.ms_sort = c->stats.ms_hierarchy = 0; }
        cudaEventElapsedTime(&ms, c->ev[k], c->ev[k + 1]); c->stats.ms_fit = ms; ++k;
        cudaEventElapsedTime(&ms, c->ev[k], c->ev[k + 1]); c->stats.ms_toplevel = ms; ++k;
    }
    c->topology_dirty = false;
    c->committed = true;
    ++c->commit_epoch;
    logf(c, LSB200_LOG_DEBUG, "lsb200: commit (%s) of %lld triangles / %lld vertices in %zu geometries", rebuild ? "rebuild" : "refit",
         (long long)nt, (long long)nv, ng);
    return LSB200_OK;
}

int lsb200_commit_scene(lsb200_ctx *c, int mode)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (mode < 0 || mode > 2) return fail(c, LSB200_ERR_INVALID, "commit_scene: bad mode %d", mode);
    return commit_locked(c, mode);
}

int lsb200_remove_geometry(lsb200_ctx *c, const char *name)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    Geometry *g = find_geom(c, name);
    if (!g) return fail(c, -1, "remove_geometry: unknown geometry '%s'", name ? name : "(null)");
    DeviceGuard dg(c->device);
    const int id = g->id;
    cudaStreamSynchronize(c->stream);
    g->verts.release(); g->elems.release();
    c->by_name.erase(g->name);
    c->geoms.erase(id);
    c->topology_dirty = true;
    c->committed = false;
    // the reference re-commits inside removeGeometry (SRC/EmbreeTracer.cpp:252); an empty scene stays uncommitted
    if (!c->geoms.empty()) {
        bool ready = true;
        for (auto &kv : c->geoms) ready = ready && kv.second.has_verts && kv.second.has_elems;
        if (ready && (c->frame_mode == LSB200_FRAME_WORLD || c->sensors[0].set)) commit_locked(c, LSB200_COMMIT_REBUILD);
    } else { c->n_tris = 0; c->n_verts = 0; }
    return id;
}

}  // extern "C"

// ---- trace -------------------------------------------------------------------------------------------------
static size_t record_bytes(uint32_t flags)
{
    return (flags & LSB200_TRACE_RANGE_IMAGE) ? 4 : ((flags & LSB200_TRACE_XYZI16) ? 16 : 32);
}

// (re)allocates the per-sensor result buffers of [first, first+count) and refreshes the device sensor tables.
// slot 0 = the context's ordinary output buffers; slot 1 = the second set used by lsb200_run_frames.
static int prepare_sensors(lsb200_ctx *c, int first, int count, int slot = 0)
{
    cudaStream_t st = c->stream;
    const size_t ns = c->sensors.size();
    LSB_CU(c, c->d_counts.reserve((size_t)2 * kMaxSensors * 4, st, true, true));
    bool dirty = c->sensors_dirty || c->h_sensors.size() != ns;
    for (int i = first; i < first + count; ++i) {
        Sensor &s = c->sensors[i];
        if (!s.set) return fail(c, LSB200_ERR_STATE, "trace: sensor %d not set", i);
        const size_t R = (size_t)s.n_rays();
        const size_t nblk = (R + kCompactRays - 1) / kCompactRays;
        if (s.t.cap < R * 4 || s.tri.cap < R * 4 || s.points.cap < R * 32 || s.blocks.cap < nblk * 4) dirty = true;
        if (slot == 1 && (s.points_b.cap < R * 32 || s.t_b.cap < R * 4)) dirty = true;
        LSB_CU(c, s.t.reserve(R * 4, st)); LSB_CU(c, s.tri.reserve(R * 4, st));
        LSB_CU(c, s.points.reserve(R * 32, st)); LSB_CU(c, s.blocks.reserve(std::max<size_t>(nblk, 4) * 4, st));
        if (slot == 1) { LSB_CU(c, s.points_b.reserve(R * 32, st)); LSB_CU(c, s.t_b.reserve(R * 4, st)); }
        if (c->opt_returns == 2) { LSB_CU(c, s.t2.reserve(R * 4, st)); LSB_CU(c, s.tri2.reserve(R * 4, st)); }
        if (c->opt_intensity != 0 || s.min_intensity > 0.0f) {
            if (s.inten.cap < R * 4) dirty = true;
            LSB_CU(c, s.inten.reserve(R * 4, st));
        }
    }
    if (!dirty) return 0;
    c->h_sensors.assign(ns, SensorDev{});
    int unit = 0, sup = 0;
    const int tw = 1 << c->opt_tile_w_log2, th = 32 >> c->opt_tile_w_log2;
    for (size_t i = 0; i < ns; ++i) {
        Sensor &s = c->sensors[i];
        SensorDev &d = c->h_sensors[i];
        d.unit_base = unit; d.sup_base = sup; d.sup_w = 1;
        if (!s.set) continue;
        const float *tab = s.tables.as<float>();
        d.sin_t = tab; d.cos_t = tab + s.V; d.sin_p = tab + 2 * s.V; d.cos_p = tab + 2 * s.V + s.H;
        std::memcpy(d.R, s.R, 36); std::memcpy(d.Rinv, s.Rinv, 36); std::memcpy(d.T, s.T, 12);
        d.V = s.V; d.H = s.H; d.n_rays = s.n_rays();
        d.max_range = s.max_range; d.az_mask = s.has_mask ? s.az_mask.as<uint8_t>() : nullptr;
        d.intensity = s.inten.as<float>(); d.min_intensity = s.min_intensity;
        d.t2_out = s.t2.as<float>(); d.tri2_out = s.tri2.as<int32_t>();
        d.t_out = s.t.as<float>(); d.tri_out = s.tri.as<int32_t>(); d.points = s.points.as<uint8_t>();
        d.count = c->d_counts.as<uint32_t>() + i; d.block_counts = s.blocks.as<uint32_t>();
        d.tiles_h = (s.H + tw - 1) / tw;
        const int tile_rows = (s.V + th - 1) / th;
        unit += tile_rows * d.tiles_h;
        d.sup_base = sup; d.sup_w = (d.tiles_h + kSuperTiles - 1) / kSuperTiles;
        sup += d.sup_w * ((tile_rows + kSuperTiles - 1) / kSuperTiles);
    }
    c->h_sensors_b = c->h_sensors;
    for (size_t i = 0; i < ns; ++i) {
        c->h_sensors_b[i].points = c->sensors[i].points_b.as<uint8_t>();      // null until slot 1 is first used
        if (c->sensors[i].t_b.p) c->h_sensors_b[i].t_out = c->sensors[i].t_b.as<float>();
        c->h_sensors_b[i].count = c->d_counts.as<uint32_t>() + kMaxSensors + i;
    }
    LSB_CU(c, c->d_sensors.reserve(ns * sizeof(SensorDev), st));
    LSB_CU(c, c->d_sensors_b.reserve(ns * sizeof(SensorDev), st));
    LSB_CU(c, cudaMemcpyAsync(c->d_sensors.p, c->h_sensors.data(), ns * sizeof(SensorDev), cudaMemcpyHostToDevice, st));
    LSB_CU(c, cudaMemcpyAsync(c->d_sensors_b.p, c->h_sensors_b.data(), ns * sizeof(SensorDev), cudaMemcpyHostToDevice, st));
    c->sensors_dirty = false;
    return 0;
}

template <int MODE, bool STATS, bool WT = false>
static int launch_trace(lsb200_ctx *c, const TraceParams &P, int n_units)
{
    constexpr bool TOP = (MODE == 4);
    const size_t smem = TOP ? (size_t)P.top_cap * 64 : (MODE == 6 ? (size_t)LSB_SMEM_STACK * kTraceThreads * sizeof(int2) : 0);
    // function attributes and occupancy are per DEVICE, so they are cached per context (one slot per instantiation)
    constexpr int kSlot = MODE * 2 + (STATS ? 1 : 0) + (WT ? 16 : 0);
    if ((TOP || (MODE == 6 && LSB_SMEM_STACK)) && !c->trace_attr_set[kSlot]) {
        LSB_CU(c, cudaFuncSetAttribute(k_trace<MODE, STATS, WT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        c->trace_attr_set[kSlot] = true;
    }
    int grid;
    if (MODE == 2) {
        grid = (n_units + kTraceThreads / 32 - 1) / (kTraceThreads / 32);
    } else {
        int per_sm = c->opt_ctas_per_sm;
        int &occ = c->trace_occ[kSlot]; size_t &occ_smem = c->trace_occ_smem[kSlot];   // one query per instantiation and smem size
        if (per_sm <= 0) {
            if (occ <= 0 || occ_smem != smem) {
                LSB_CU(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_trace<MODE, STATS, WT>, kTraceThreads, smem));
                occ_smem = smem;
            }
            per_sm = occ;
        }
        if (per_sm <= 0) return fail(c, LSB200_ERR_CUDA, "trace kernel does not fit on an SM (smem %zu)", smem);
        grid = c->sm_count * per_sm;
        grid = std::min(grid, (n_units + kTraceThreads / 32 - 1) / (kTraceThreads / 32));
        grid = std::max(grid, 1);
    }
    k_trace<MODE, STATS, WT><<<grid, kTraceThreads, smem, c->stream>>>(P);
    c->stats.trace_grid = grid; c->stats.trace_block = kTraceThreads;
    count_launch(c);
    return 0;
}

template <bool STATS>
static int launch_trace_mode(lsb200_ctx *c, const TraceParams &P, int n_units, int mode)
{
    if (c->built_watertight) {             // watertight records: the shipped traversal and brute force only
        if (mode == 3) return launch_trace<3, false, true>(c, P, n_units);
        if (mode == 6) return launch_trace<6, false, true>(c, P, n_units);
        return fail(c, LSB200_ERR_STATE, "traversal variant %d has no watertight build (use 6 or 3)", mode);
    }
    switch (mode) {
#ifdef LSB_LAB
    case 0: return launch_trace<0, STATS>(c, P, n_units);
    case 1: return launch_trace<1, STATS>(c, P, n_units);
    case 2: return launch_trace<2, STATS>(c, P, n_units);
    case 4: return launch_trace<4, STATS>(c, P, n_units);
    case 5: return launch_trace<5, STATS>(c, P, n_units);
    case 7: return launch_trace<7, STATS>(c, P, n_units);
#endif
    case 3: return launch_trace<3, STATS>(c, P, n_units);
    default: return launch_trace<6, STATS>(c, P, n_units);
    }
}

static int trace_locked(lsb200_ctx *c, int first, int count, uint32_t flags, int slot = 0)
{
    if (first < 0 || count <= 0 || (size_t)(first + count) > c->sensors.size())
        return fail(c, LSB200_ERR_INVALID, "trace: bad sensor range [%d, %d)", first, first + count);
    if (c->frame_mode == LSB200_FRAME_SENSOR && (first != 0 || count != 1))
        return fail(c, LSB200_ERR_INVALID, "trace: sensor-frame contexts hold exactly one sensor (index 0)");
    DeviceGuard dg(c->device);
    cudaStream_t st = c->stream;
    struct NvtxScope { NvtxScope(const char *n) { nvtxRangePushA(n); } ~NvtxScope() { nvtxRangePop(); } } nvtx_trace("lsb200:trace");
    if (flags & LSB200_TRACE_RANGE_IMAGE) flags |= LSB200_TRACE_NO_POINTS;     // the per-ray t array IS the output
    if (int rc = prepare_sensors(c, first, count, slot)) return rc;
    uint32_t *d_cnt = c->d_counts.as<uint32_t>() + (slot ? kMaxSensors : 0);
    const SensorDev *d_sens = slot ? c->d_sensors_b.as<SensorDev>() : c->d_sensors.as<SensorDev>();
    if (c->geoms.empty() || c->n_tris == 0) {
        // empty scene: an initialised, empty cloud and -1 (SRC/OptixTracer.cpp:279-288)
        LSB_CU(c, cudaMemsetAsync(d_cnt + first, 0, (size_t)count * 4, st));
        for (int i = first; i < first + count; ++i) {
            Sensor &s = c->sensors[i];
            LSB_CU(c, cudaMemsetAsync(s.tri.p, 0xFF, (size_t)s.n_rays() * 4, st));          // -1
            k_fill_u32<<<(unsigned)((s.n_rays() + 255) / 256), 256, 0, st>>>(s.t.as<uint32_t>(), 0x7f800000u, s.n_rays());   // +inf
            count_launch(c);
        }
        return fail(c, LSB200_ERR_EMPTY_SCENE, "trace_scene: empty scene");
    }
    if (!c->committed) return fail(c, LSB200_ERR_STATE, "trace_scene: scene not committed");
    TraceParams P{};
    P.nodes = c->d_nodes.as<float4>(); P.tris = c->d_tris.as<float4>(); P.leaf_ids = c->d_leaf_ids.as<int32_t>();
    P.bounds = c->d_bounds.as<uint32_t>(); P.n_top = c->d_misc.as<int>();
    P.sensors = d_sens; P.n_sensors = (int)c->sensors.size();
    P.top_cap = c->built_top; P.n_tris = c->n_tris;
    P.world_frame = c->frame_mode == LSB200_FRAME_WORLD;
    P.tile_w_log2 = c->opt_tile_w_log2;
    // an intensity model other than the reference's constant, or an intensity filter, is a post-pass over the per-ray
    // results (k_intensity); the points are then emitted by the ordered compaction, never by the traversal's own append
    bool need_intensity = c->opt_intensity != 0;
    for (int i = first; i < first + count; ++i) need_intensity = need_intensity || c->sensors[i].min_intensity > 0.0f;
    const bool want_points = !(flags & LSB200_TRACE_NO_POINTS);
    if (need_intensity) flags |= LSB200_TRACE_NO_POINTS;
    P.flags = flags;
    P.work = c->d_misc.as<uint32_t>() + 4;
    // work units of the batch: [unit_base(first), unit_base(last) + units(last))
    const int unit_lo = c->h_sensors[first].unit_base;
    const SensorDev &last = c->h_sensors[first + count - 1];
    const int unit_hi = last.unit_base + ((last.V + (32 >> c->opt_tile_w_log2) - 1) / (32 >> c->opt_tile_w_log2)) * last.tiles_h;
    P.unit_lo = unit_lo; P.n_units = unit_hi;
    const bool timing = c->opt_timing != 0;
    const bool unordered_points = !(flags & (LSB200_TRACE_ORDERED | LSB200_TRACE_NO_POINTS));
    if (unordered_points) LSB_CU(c, cudaMemsetAsync(d_cnt + first, 0, (size_t)count * 4, st));
    if (timing) cudaEventRecord(c->ev[8], st);
    int mode = c->opt_traversal;
    int rc = 0;
    bool tile_timed = false;
    if (mode == 4 && c->built_top > 0 && !c->top_valid) {  // variant switched on after the commit
        k_top_level<<<1, 1024, (size_t)c->built_top * sizeof(int), st>>>(scene_dev(c), c->d_misc.as<int>());
        count_launch(c);
        c->top_valid = true;
    }
    const int n_units = unit_hi - unit_lo;
    // small launches (an XT-32 scan is 150 units): one thread per ray from the root, no queues / chains / lists (k_trace_small)
    // (not when per-tile entry lists were asked for at this size -- LSB200_OPT_TILE_MIN_UNITS lowered by the caller)
    const bool lists_wanted = c->opt_chain && c->opt_tile_entries >= 2 && n_units >= c->opt_tile_min_units;
    const bool small_trace = mode == 6 && !c->opt_stats && n_units < c->opt_small_trace_units && !lists_wanted;
    if (c->opt_chain && mode != 3) {
        // origin chains: one list per sensor, valid for this BVH and these sensor poses
        const size_t ns = c->sensors.size();
        if (c->chain_commit_epoch != c->commit_epoch || c->chain_sensors_epoch != c->sensors_epoch) {
            LSB_CU(c, c->d_chains.reserve(ns * 9 * 2 * kMaxChain * sizeof(float4), st));
            LSB_CU(c, c->d_chain_counts.reserve(ns * 9 * 4, st));
            // both sensor tables hold the same poses; the chain kernel only reads T
            k_origin_chain<<<(unsigned)((ns * 9 + 63) / 64), 64, 0, st>>>(P.nodes, P.top_cap, c->d_sensors.as<SensorDev>(), (int)ns,
                                                                         P.world_frame, P.bounds, c->d_chains.as<float4>(),
                                                                         c->d_chain_counts.as<int>());
            count_launch(c);
            c->chain_commit_epoch = c->commit_epoch; c->chain_sensors_epoch = c->sensors_epoch;
        }
        P.chains = c->d_chains.as<float4>();
        P.chain_counts = c->d_chain_counts.as<int>();
    }
    if (small_trace) {
        int max_r = 0;
        for (int i = first; i < first + count; ++i) max_r = std::max(max_r, c->sensors[i].n_rays());
        // rays per warp: as few as keeps the launch within ~32 warps per SM (an XT-32 scan: one ray per warp)
        int spread = 0;
        while (spread < 5 && ((int64_t)max_r * count << (spread + 1)) <= (int64_t)c->sm_count * 32 * 32) ++spread;
        dim3 grid((unsigned)((((int64_t)max_r << spread) + 127) / 128), (unsigned)count);
        if (c->built_watertight) k_trace_small<true><<<grid, 128, 0, st>>>(P, first, spread);
        else k_trace_small<false><<<grid, 128, 0, st>>>(P, first, spread);
        count_launch(c);
        c->stats.trace_grid = (int)(grid.x * grid.y); c->stats.trace_block = 128;
    }
    if ((mode == 0 || mode == 6 || mode == 7) && !small_trace) {
        // octant queues: depend on the sensor table only, rebuilt when it (or the traced range) changes
        const int wf = P.world_frame ? 1 : 0;
        if (c->units_epoch != c->sensors_epoch || c->units_first != first || c->units_count != count || c->units_world != wf) {
            LSB_CU(c, c->d_unit_oct.reserve((size_t)n_units, st));
            LSB_CU(c, c->d_unit_list.reserve((size_t)n_units * sizeof(int4), st));
            LSB_CU(c, c->d_oct_hist.reserve(64 * 4, st));        // [0, 9) class sizes, [16, 25) front cursors, [32, 41) back cursors
            LSB_CU(c, cudaMemsetAsync(c->d_oct_hist.p, 0, 64 * 4, st));
            const unsigned blocks = (unsigned)((n_units + 255) / 256);
            // when this launch will use per-tile entry lists, the same pass over the rays also yields the tiles' direction
            // intervals (k_tile_intervals' job): a moving sensor rebuilds queues and intervals every frame
            const bool with_iv = mode >= 6 && c->opt_chain && c->opt_tile_entries >= 2 && n_units >= c->opt_tile_min_units && !c->opt_super_tiles;
            if (with_iv) {
                LSB_CU(c, c->d_unit_iv.reserve((size_t)n_units * 2 * sizeof(float4), st));
                LSB_CU(c, c->d_tile_iv.reserve((size_t)n_units * 2 * sizeof(float4), st));
            }
            k_classify_units<<<blocks, 256, 0, st>>>(d_sens, P.n_sensors, unit_lo, unit_hi, wf, P.tile_w_log2, c->d_unit_oct.as<uint8_t>(),
                                                     c->d_oct_hist.as<uint32_t>(), with_iv ? c->d_unit_iv.as<float4>() : nullptr);
            k_scatter_units<<<blocks, 256, 0, st>>>(d_sens, P.n_sensors, unit_lo, unit_hi, P.tile_w_log2, c->d_unit_oct.as<uint8_t>(),
                                                    c->d_oct_hist.as<uint32_t>(), c->d_oct_hist.as<uint32_t>() + 16,
                                                    c->d_unit_list.as<int4>(), with_iv ? c->d_unit_iv.as<float4>() : nullptr,
                                                    with_iv ? c->d_tile_iv.as<float4>() : nullptr);
            count_launch(c, 2);
            c->units_epoch = c->sensors_epoch; c->units_first = first; c->units_count = count; c->units_world = wf;
            ++c->units_build_id;
            if (with_iv) {      // the intervals of this queue order are in place: k_tile_intervals is not needed
                c->iv_sensors_epoch = c->sensors_epoch; c->iv_first = first; c->iv_count = count; c->iv_world = wf; c->iv_super = 0; c->iv_units_id = c->units_build_id;
            }
        }
        P.unit_list = c->d_unit_list.as<int4>();
        P.oct_hist = c->d_oct_hist.as<uint32_t>();
    }
    if (c->opt_stats) {
        LSB_CU(c, c->d_counters.reserve(kCntCount * sizeof(unsigned long long), st));
        LSB_CU(c, cudaMemsetAsync(c->d_counters.p, 0, kCntCount * sizeof(unsigned long long), st));
        P.counters = c->d_counters.as<unsigned long long>();
    }
    // per-tile entry lists (k_tile_entries): worth their launch from a few thousand 32-ray tiles on
    if (mode >= 6 && c->opt_chain && c->opt_tile_entries >= 2 && n_units >= c->opt_tile_min_units && !small_trace) {
        const int cap = c->opt_tile_entries;
        const int max_visits = c->opt_tile_visits > 0 ? c->opt_tile_visits : (16 * cap) / 5;
        P.tile_cap = cap;
        P.refill = c->opt_lane_refill;
        const bool stale = c->tile_commit_epoch != c->commit_epoch || c->tile_sensors_epoch != c->sensors_epoch || c->tile_first != first ||
                           c->tile_count != count || c->tile_cap != cap || c->tile_visits != max_visits || c->tile_units_id != c->units_build_id;
        if (stale || c->opt_stats) {
            LSB_CU(c, c->d_tile_lists.reserve((size_t)n_units * cap * 2 * sizeof(float4), st));
            LSB_CU(c, c->d_tile_counts.reserve((size_t)n_units, st));
        }
        P.tile_lists = c->d_tile_lists.as<float4>();
        P.tile_counts = c->d_tile_counts.as<uint8_t>();
        tile_timed = false;
        // supertiles of the traced sensors: [sup_base(first), sup_base(last) + supertiles(last))
        const SensorDev &hs_last = c->h_sensors[first + count - 1];
        const int last_rows = (hs_last.V + (32 >> c->opt_tile_w_log2) - 1) / (32 >> c->opt_tile_w_log2);
        const int sup_lo = c->h_sensors[first].sup_base;
        const int sup_hi = hs_last.sup_base + hs_last.sup_w * ((last_rows + kSuperTiles - 1) / kSuperTiles);
        {
            const int wf = P.world_frame ? 1 : 0;
            const bool iv_stale = c->iv_sensors_epoch != c->sensors_epoch || c->iv_first != first || c->iv_count != count || c->iv_world != wf ||
                                  c->iv_super != c->opt_super_tiles || c->iv_units_id != c->units_build_id;
            if (iv_stale) {
                LSB_CU(c, c->d_tile_iv.reserve((size_t)n_units * 2 * sizeof(float4), st));
                if (c->opt_super_tiles) LSB_CU(c, c->d_sup_iv.reserve((size_t)sup_hi * 2 * sizeof(float4), st));
            }
            P.tile_iv = c->d_tile_iv.as<float4>();
            P.sup_iv = c->opt_super_tiles ? c->d_sup_iv.as<float4>() : nullptr;
            if (iv_stale) {
                k_tile_intervals<<<(unsigned)((n_units + 127) / 128), 128, 0, st>>>(P);
                count_launch(c);
#ifdef LSB_LAB
                if (c->opt_super_tiles) {
                    k_super_intervals<<<(unsigned)((sup_hi - sup_lo + 3) / 4), 128, 0, st>>>(P, first, count, sup_lo, sup_hi);
                    count_launch(c);
                }
#endif
                c->iv_sensors_epoch = c->sensors_epoch; c->iv_first = first; c->iv_count = count; c->iv_world = wf; c->iv_super = c->opt_super_tiles; c->iv_units_id = c->units_build_id;
            }
        }
        if (c->opt_super_tiles) {
            if (stale || c->opt_stats) {
                LSB_CU(c, c->d_sup_lists.reserve((size_t)sup_hi * kMaxTileEntries * 2 * sizeof(float4), st));
                LSB_CU(c, c->d_sup_counts.reserve((size_t)sup_hi, st));
            }
            P.sup_lists = c->d_sup_lists.as<float4>();
            P.sup_counts = c->d_sup_counts.as<uint8_t>();
        }
        if (stale || c->opt_stats) {
            const unsigned blocks = (unsigned)((n_units + kTileBlock - 1) / kTileBlock);
            const size_t tile_stk_bytes = LSB_TILE_SMEM_STACK ? (size_t)P.tile_cap * kTileBlock * sizeof(int4) : 0;   // <= 32 KB
            if (timing) { cudaEventRecord(c->ev[11], st); tile_timed = true; }
#ifdef LSB_LAB
            if (c->opt_super_tiles) {
                const unsigned sblocks = (unsigned)((sup_hi - sup_lo + 127) / 128);
                if (c->opt_stats) k_super_entries<true><<<sblocks, 128, 0, st>>>(P, first, count, sup_lo, sup_hi, 8 * kMaxTileEntries);
                else k_super_entries<false><<<sblocks, 128, 0, st>>>(P, first, count, sup_lo, sup_hi, 8 * kMaxTileEntries);
                count_launch(c);
            }
            if (c->opt_stats) k_tile_entries<true><<<blocks, kTileBlock, tile_stk_bytes, st>>>(P, max_visits);
            else
#endif
            k_tile_entries<false><<<blocks, kTileBlock, tile_stk_bytes, st>>>(P, max_visits);
            count_launch(c);
            c->tile_commit_epoch = c->commit_epoch; c->tile_sensors_epoch = c->sensors_epoch; c->tile_first = first; c->tile_count = count;
            c->tile_cap = cap; c->tile_visits = max_visits; c->tile_units_id = c->units_build_id;
            if (timing) cudaEventRecord(c->ev[7], st);
        }
    }
    if (small_trace) rc = 0;
    else
#ifdef LSB_LAB
    if (c->opt_stats) rc = launch_trace_mode<true>(c, P, n_units, mode);
    else
#endif
    rc = launch_trace_mode<false>(c, P, n_units, mode);
    if (rc) return rc;
    if (timing) cudaEventRecord(c->ev[9], st);
    int max_rays = 0;
    for (int i = first; i < first + count; ++i) max_rays = std::max(max_rays, c->sensors[i].n_rays());
    if (c->opt_returns == 2) {           // second return of every ray: the next surface behind its first hit (k_second_return)
        dim3 grid((unsigned)((max_rays + 255) / 256), (unsigned)count);
        if (c->built_watertight) k_second_return<true><<<grid, 256, 0, st>>>(P, first);
        else k_second_return<false><<<grid, 256, 0, st>>>(P, first);
        count_launch(c);
    }
    if (need_intensity) {
        const bool small = c->opt_small_build && c->n_tris <= kSmallMaxTris && c->geoms.size() <= (size_t)kSmallMaxGeoms;
        if (small) {        // the fused small build takes its descriptors as kernel parameters: the post-pass needs them in memory
            LSB_CU(c, c->d_geoms.reserve(c->h_geoms.size() * sizeof(GeomDesc), st));
            LSB_CU(c, cudaMemcpyAsync(c->d_geoms.p, c->h_geoms.data(), c->h_geoms.size() * sizeof(GeomDesc), cudaMemcpyHostToDevice, st));
        }
        dim3 grid((unsigned)((max_rays + 255) / 256), (unsigned)count);
        k_intensity<<<grid, 256, 0, st>>>(scene_dev(c), d_sens, first, P.world_frame, c->opt_intensity);
        count_launch(c);
    }
    if (want_points && ((flags & LSB200_TRACE_ORDERED) || need_intensity)) {
        dim3 grid((unsigned)((max_rays + kCompactRays - 1) / kCompactRays), (unsigned)count);
        const SensorDev *sd = d_sens;
        k_count_hits<<<grid, kCompactThreads, 0, st>>>(sd, first);
        k_scan_counts<<<count, 256, 0, st>>>(sd, first);
        k_scatter_hits<<<grid, kCompactThreads, 0, st>>>(sd, first, (flags & ~(uint32_t)LSB200_TRACE_NO_POINTS) | (need_intensity ? kFlagIntensityBuffer : 0u));
        count_launch(c, 3);
    }
    if (timing) {
        cudaEventRecord(c->ev[10], st);
        LSB_CU(c, cudaStreamSynchronize(st));
        cudaEventElapsedTime(&c->stats.ms_trace, c->ev[8], c->ev[9]);
        cudaEventElapsedTime(&c->stats.ms_compact, c->ev[9], c->ev[10]);
        c->stats.ms_tile_lists = 0.0f;
        if (tile_timed) cudaEventElapsedTime(&c->stats.ms_tile_lists, c->ev[11], c->ev[7]);
    }
    LSB_CU(c, cudaGetLastError());
    return LSB200_OK;
}

static int fetch_locked(lsb200_ctx *c, int idx, void *out, size_t capacity, uint32_t *n_pts, uint32_t flags)
{
    if (idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set) return fail(c, LSB200_ERR_INVALID, "fetch: bad sensor %d", idx);
    if (!c->d_counts.p || !c->sensors[idx].points.p) return fail(c, LSB200_ERR_STATE, "fetch: sensor %d was never traced", idx);
    DeviceGuard dg(c->device);
    cudaStream_t st = c->stream;
    if (flags & LSB200_TRACE_RANGE_IMAGE) {
        // dense range image: one f32 per ray in ray-id order (+inf = no return); the "point count" is the ray count
        const uint32_t R = (uint32_t)c->sensors[idx].n_rays();
        if (n_pts) *n_pts = R;
        if (out && capacity < R) return fail(c, LSB200_ERR_CAPACITY, "fetch: range image of %u rays, capacity %zu", R, capacity);
        if (out) {
            LSB_CU(c, cudaMemcpyAsync(out, c->sensors[idx].t.p, (size_t)R * 4, cudaMemcpyDeviceToHost, st));
            LSB_CU(c, cudaStreamSynchronize(st));
        }
        return LSB200_OK;
    }
    LSB_CU(c, cudaMemcpyAsync(c->h_count, c->d_counts.as<uint32_t>() + idx, 4, cudaMemcpyDeviceToHost, st));
    // small clouds (an XT-32 scan: 4 800 rays): the records are copied speculatively -- as many as could exist -- right behind
    // the count, so the call pays ONE host synchronisation instead of two; records past *n_pts in `out` are then unspecified
    const size_t spec = std::min<size_t>((size_t)c->sensors[idx].n_rays(), capacity);
    if (out && spec && spec * record_bytes(flags) <= (size_t)256 * 1024) {
        LSB_CU(c, cudaMemcpyAsync(out, c->sensors[idx].points.p, spec * record_bytes(flags), cudaMemcpyDeviceToHost, st));
        LSB_CU(c, cudaStreamSynchronize(st));
        const uint32_t n1 = c->h_count[0];
        if (n_pts) *n_pts = n1;
        if (n1 > capacity) return fail(c, LSB200_ERR_CAPACITY, "fetch: %u points, capacity %zu", n1, capacity);
        return LSB200_OK;
    }
    LSB_CU(c, cudaStreamSynchronize(st));
    const uint32_t n = c->h_count[0];
    if (n_pts) *n_pts = n;
    const size_t ncopy = std::min<size_t>(n, capacity);
    if (out && ncopy) {
        LSB_CU(c, cudaMemcpyAsync(out, c->sensors[idx].points.p, ncopy * record_bytes(flags), cudaMemcpyDeviceToHost, st));
        LSB_CU(c, cudaStreamSynchronize(st));
    }
    if (n > capacity && out) return fail(c, LSB200_ERR_CAPACITY, "fetch: %u points, capacity %zu", n, capacity);
    return LSB200_OK;
}

extern "C" {

int lsb200_trace_device(lsb200_ctx *c, int idx, uint32_t flags)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    return trace_locked(c, idx, 1, flags);
}

int lsb200_trace_batch(lsb200_ctx *c, int first, int count, uint32_t flags)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    return trace_locked(c, first, count, flags);
}

int lsb200_trace_scene(lsb200_ctx *c, int idx, uint32_t frame, void *out, size_t capacity, uint32_t *n_pts, uint32_t flags)
{
    (void)frame;     // the frame index only goes into the PointCloud2 header (lsb200_init_message)
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (n_pts) *n_pts = 0;
    flags &= ~(uint32_t)LSB200_TRACE_NO_POINTS;
    int rc = trace_locked(c, idx, 1, flags);
    if (rc) return rc;
    return fetch_locked(c, idx, out, capacity, n_pts, flags);
}

int lsb200_fetch_points(lsb200_ctx *c, int idx, void *out, size_t capacity, uint32_t *n_pts, uint32_t flags)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    return fetch_locked(c, idx, out, capacity, n_pts, flags);
}

int lsb200_fetch_batch(lsb200_ctx *c, int first, int count, void *const *outs, const size_t *caps, uint32_t *n_pts, uint32_t flags)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (first < 0 || count <= 0 || count > 16 * 256 || (size_t)(first + count) > c->sensors.size() || !outs || !caps || !n_pts)
        return fail(c, LSB200_ERR_INVALID, "fetch_batch: bad arguments");
    if (!c->d_counts.p) return fail(c, LSB200_ERR_STATE, "fetch_batch: nothing traced yet");
    DeviceGuard dg(c->device);
    cudaStream_t st = c->stream;
    int rc = LSB200_OK;
    if (flags & LSB200_TRACE_RANGE_IMAGE) {
        for (int i = 0; i < count; ++i) {
            const Sensor &s = c->sensors[first + i];
            if (!s.set || !s.t.p) return fail(c, LSB200_ERR_STATE, "fetch_batch: sensor %d was never traced", first + i);
            n_pts[i] = (uint32_t)s.n_rays();
            if (outs[i] && caps[i] < n_pts[i]) { rc = fail(c, LSB200_ERR_CAPACITY, "fetch_batch: sensor %d: range image of %u rays, capacity %zu", first + i, n_pts[i], caps[i]); continue; }
            if (outs[i]) LSB_CU(c, cudaMemcpyAsync(outs[i], s.t.p, (size_t)n_pts[i] * 4, cudaMemcpyDeviceToHost, st));
        }
        LSB_CU(c, cudaStreamSynchronize(st));
        return rc;
    }
    // one copy for all hit counts, one synchronisation, then all point copies back to back
    LSB_CU(c, cudaMemcpyAsync(n_pts, c->d_counts.as<uint32_t>() + first, (size_t)count * 4, cudaMemcpyDeviceToHost, st));
    LSB_CU(c, cudaStreamSynchronize(st));
    for (int i = 0; i < count; ++i) {
        const Sensor &s = c->sensors[first + i];
        if (!s.set || !s.points.p) return fail(c, LSB200_ERR_STATE, "fetch_batch: sensor %d was never traced", first + i);
        const size_t ncopy = std::min<size_t>(n_pts[i], caps[i]);
        if (n_pts[i] > caps[i]) rc = fail(c, LSB200_ERR_CAPACITY, "fetch_batch: sensor %d has %u points, capacity %zu", first + i, n_pts[i], caps[i]);
        if (outs[i] && ncopy) LSB_CU(c, cudaMemcpyAsync(outs[i], s.points.p, ncopy * record_bytes(flags), cudaMemcpyDeviceToHost, st));
    }
    LSB_CU(c, cudaStreamSynchronize(st));
    return rc;
}

// Pipelined frames: frame f's point clouds travel to the host on the copy stream while frame f+1 is being
// transformed, rebuilt and traced on the compute stream (two output slots, events between the streams).
int lsb200_run_frames(lsb200_ctx *c, const lsb200_frame *frames, int n_frames, int first, int count, int commit_mode,
                      uint32_t flags)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (!frames || n_frames <= 0 || first < 0 || count <= 0 || (size_t)(first + count) > c->sensors.size() || count > kMaxSensors)
        return fail(c, LSB200_ERR_INVALID, "run_frames: bad arguments");
    if (commit_mode < 0 || commit_mode > 2) return fail(c, LSB200_ERR_INVALID, "run_frames: bad commit mode %d", commit_mode);
    for (int f = 0; f < n_frames; ++f)
        if (!frames[f].out_points || !frames[f].capacities || !frames[f].n_pts)
            return fail(c, LSB200_ERR_INVALID, "run_frames: frame %d has null output arrays", f);
    DeviceGuard dg(c->device);
    cudaStream_t st = c->stream, cs = c->copy_stream;
    flags &= ~(uint32_t)LSB200_TRACE_NO_POINTS;
    const size_t rec = record_bytes(flags);
    int result = LSB200_OK;
    bool slot_used[2] = {false, false};
    bool pre_uploaded = false;          // the vertex blobs of the frame about to be committed are already on their way (upload stream)

    auto finish = [&](int f) -> int {             // frame f is traced (or being traced): counts, then the clouds
        const int slot = f & 1;
        LSB_CU(c, cudaEventSynchronize(c->ev_traced[slot]));
        const uint32_t *hc = c->h_count + 16 + (size_t)slot * kMaxSensors;
        for (int i = 0; i < count; ++i) {
            const Sensor &s = c->sensors[first + i];
            const bool range_image = (flags & LSB200_TRACE_RANGE_IMAGE) != 0;
            const uint32_t n = range_image ? (uint32_t)s.n_rays() : hc[first + i];
            frames[f].n_pts[i] = n;
            const size_t cap = frames[f].capacities[i];
            if (n > cap) result = fail(c, LSB200_ERR_CAPACITY, "run_frames: frame %d sensor %d has %u points, capacity %zu", f, first + i, n, cap);
            const size_t ncopy = range_image ? (n > cap ? 0 : n) : std::min<size_t>(n, cap);
            const void *src = range_image ? (slot ? s.t_b.p : s.t.p) : (slot ? s.points_b.p : s.points.p);
            if (frames[f].out_points[i] && ncopy)
                LSB_CU(c, cudaMemcpyAsync(frames[f].out_points[i], src, ncopy * rec, cudaMemcpyDeviceToHost, cs));
        }
        LSB_CU(c, cudaEventRecord(c->ev_copied[slot], cs));
        return 0;
    };
    auto body = [&]() -> int {
        for (int f = 0; f < n_frames; ++f) {
            const int slot = f & 1;
            if (slot_used[slot]) LSB_CU(c, cudaStreamWaitEvent(st, c->ev_copied[slot], 0));   // frame f-2 has left the slot
            size_t k = 0;
            for (auto &kv : c->geoms) {
                Geometry &g = kv.second;
                const void *v = (frames[f].verts && frames[f].verts[k]) ? frames[f].verts[k] : nullptr;
                // frame 0 uploads in order; later frames were uploaded on the upload stream while frame f-1 was traced
                if (v && !pre_uploaded) { if (int rc = upload_geometry(c, g, v, g.stride, nullptr)) return rc; }
                if (frames[f].affines) std::memcpy(g.A, frames[f].affines + 12 * k, sizeof(g.A));
                ++k;
            }
            if (pre_uploaded) LSB_CU(c, cudaStreamWaitEvent(st, c->ev_uploaded, 0));
            pre_uploaded = false;
            if (int rc = commit_locked(c, commit_mode)) return rc;
            if (int rc = trace_locked(c, first, count, flags, slot)) return rc;
            // vertex blobs of frame f+1: host -> device on the upload stream as soon as this frame's transform has read the
            // buffers (ev_xform), i.e. under this frame's sort / tree / trace -- only when no buffer has to grow
            if (f + 1 < n_frames && frames[f + 1].verts) {
                bool fits = true, any = false;
                size_t j = 0;
                for (auto &kv : c->geoms) {
                    Geometry &g = kv.second;
                    if (frames[f + 1].verts[j]) { any = true; fits = fits && g.has_verts && g.verts.cap >= (size_t)g.n_verts * (size_t)g.stride; }
                    ++j;
                }
                if (any && fits) {
                    LSB_CU(c, cudaStreamWaitEvent(c->upload_stream, c->ev_xform, 0));
                    j = 0;
                    for (auto &kv : c->geoms) {
                        Geometry &g = kv.second;
                        if (frames[f + 1].verts[j]) { if (int rc = upload_geometry(c, g, frames[f + 1].verts[j], g.stride, nullptr, c->upload_stream)) return rc; }
                        ++j;
                    }
                    LSB_CU(c, cudaEventRecord(c->ev_uploaded, c->upload_stream));
                    pre_uploaded = true;
                }
            }
            LSB_CU(c, cudaMemcpyAsync(c->h_count + 16 + (size_t)slot * kMaxSensors + first,
                                      c->d_counts.as<uint32_t>() + (size_t)slot * kMaxSensors + first, (size_t)count * 4,
                                      cudaMemcpyDeviceToHost, st));
            LSB_CU(c, cudaEventRecord(c->ev_traced[slot], st));
            slot_used[slot] = true;
            if (f > 0) { if (int rc = finish(f - 1)) return rc; }
        }
        return finish(n_frames - 1);
    };
    const int rc = body();
    cudaStreamSynchronize(cs);
    cudaStreamSynchronize(c->upload_stream);
    cudaStreamSynchronize(st);
    if (rc) return rc;
    return result;
}

int lsb200_trace_raw(lsb200_ctx *c, int idx, float *t_out, int32_t *tri_out)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    int rc = trace_locked(c, idx, 1, LSB200_TRACE_NO_POINTS);
    if (rc && rc != LSB200_ERR_EMPTY_SCENE) return rc;
    DeviceGuard dg(c->device);
    Sensor &s = c->sensors[idx];
    const size_t R = (size_t)s.n_rays();
    if (rc == LSB200_ERR_EMPTY_SCENE) {
        for (size_t i = 0; i < R; ++i) { if (t_out) t_out[i] = INFINITY; if (tri_out) tri_out[i] = -1; }
        return rc;
    }
    if (t_out) LSB_CU(c, cudaMemcpyAsync(t_out, s.t.p, R * 4, cudaMemcpyDeviceToHost, c->stream));
    if (tri_out) LSB_CU(c, cudaMemcpyAsync(tri_out, s.tri.p, R * 4, cudaMemcpyDeviceToHost, c->stream));
    LSB_CU(c, cudaStreamSynchronize(c->stream));
    return LSB200_OK;
}

int lsb200_fetch_second_returns(lsb200_ctx *c, int idx, float *t2_out, int32_t *tri2_out)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set) return fail(c, LSB200_ERR_INVALID, "bad sensor %d", idx);
    if (c->opt_returns != 2) return fail(c, LSB200_ERR_STATE, "fetch_second_returns: LSB200_OPT_RETURNS is not 2");
    Sensor &s = c->sensors[idx];
    if (!s.t2.p || !s.tri2.p) return fail(c, LSB200_ERR_STATE, "fetch_second_returns: sensor %d has not been traced with two returns", idx);
    DeviceGuard dg(c->device);
    const size_t R = (size_t)s.n_rays();
    if (t2_out) LSB_CU(c, cudaMemcpyAsync(t2_out, s.t2.p, R * 4, cudaMemcpyDeviceToHost, c->stream));
    if (tri2_out) LSB_CU(c, cudaMemcpyAsync(tri2_out, s.tri2.p, R * 4, cudaMemcpyDeviceToHost, c->stream));
    LSB_CU(c, cudaStreamSynchronize(c->stream));
    return LSB200_OK;
}

int lsb200_device_results(lsb200_ctx *c, int idx, void **d_points, uint32_t **d_count, float **d_t, int32_t **d_tri)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set) return fail(c, LSB200_ERR_INVALID, "bad sensor %d", idx);
    Sensor &s = c->sensors[idx];
    if (d_points) *d_points = s.points.p;
    if (d_count) *d_count = c->d_counts.p ? c->d_counts.as<uint32_t>() + idx : nullptr;
    if (d_t) *d_t = s.t.as<float>();
    if (d_tri) *d_tri = s.tri.as<int32_t>();
    return LSB200_OK;
}

int lsb200_fetch_intensity(lsb200_ctx *c, int idx, float *out)
{
    if (!c || !out) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set) return fail(c, LSB200_ERR_INVALID, "fetch_intensity: bad sensor %d", idx);
    Sensor &s = c->sensors[idx];
    if (!s.inten.p) return fail(c, LSB200_ERR_STATE, "fetch_intensity: no intensity model / filter was active in the last trace of sensor %d", idx);
    DeviceGuard dg(c->device);
    LSB_CU(c, cudaMemcpyAsync(out, s.inten.p, (size_t)s.n_rays() * 4, cudaMemcpyDeviceToHost, c->stream));
    LSB_CU(c, cudaStreamSynchronize(c->stream));
    return LSB200_OK;
}

int lsb200_synchronize(lsb200_ctx *c)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    DeviceGuard dg(c->device);
    LSB_CU(c, cudaStreamSynchronize(c->stream));
    return LSB200_OK;
}

int lsb200_init_message(const lsb200_ctx *c, uint32_t frame, lsb200_cloud_header *h)
{
    if (!c || !h) return LSB200_ERR_INVALID;
    std::memset(h, 0, sizeof(*h));
    h->seq = frame; h->height = 1; h->width = 0; h->point_step = 32; h->row_step = 0;
    h->is_bigendian = 0; h->is_dense = 1; h->n_fields = 5;
    return LSB200_OK;
}

int lsb200_get_stats(lsb200_ctx *c, lsb200_stats *out)
{
    if (!c || !out) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    DeviceGuard dg(c->device);
    c->stats.n_vertices = c->n_verts; c->stats.n_triangles = c->n_tris;
    c->stats.n_nodes = c->n_tris > 0 ? std::max(c->n_tris - 1, 1) : 0;
    c->stats.n_leaves = c->n_tris;
    c->stats.top_nodes = 0;
    if (c->committed) {
        uint32_t b[6]; int ntop = 0;
        LSB_CU(c, cudaMemcpyAsync(b, c->d_bounds.p, 24, cudaMemcpyDeviceToHost, c->stream));
        LSB_CU(c, cudaMemcpyAsync(&ntop, c->d_misc.p, 4, cudaMemcpyDeviceToHost, c->stream));
        LSB_CU(c, cudaStreamSynchronize(c->stream));
        for (int k = 0; k < 3; ++k) { c->stats.scene_lo[k] = ord2f(b[k]); c->stats.scene_hi[k] = ord2f(b[3 + k]); }
        c->stats.top_nodes = (c->built_top > 0 && c->top_valid) ? ntop : 0;
    }
    *out = c->stats;
    return LSB200_OK;
}

int lsb200_get_trace_counters(lsb200_ctx *c, uint64_t *out, int n)
{
    if (!c || !out || n <= 0) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (!c->d_counters.p) return fail(c, LSB200_ERR_STATE, "trace counters: no LSB200_OPT_TRACE_STATS launch yet");
    DeviceGuard dg(c->device);
    unsigned long long tmp[kCntCount];
    LSB_CU(c, cudaMemcpyAsync(tmp, c->d_counters.p, sizeof(tmp), cudaMemcpyDeviceToHost, c->stream));
    LSB_CU(c, cudaStreamSynchronize(c->stream));
    for (int k = 0; k < n; ++k) out[k] = k < kCntCount ? (uint64_t)tmp[k] : 0;
    return LSB200_OK;
}

int lsb200_reset_counters(lsb200_ctx *c)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    c->stats.kernel_launches = 0;
    return LSB200_OK;
}

int lsb200_debug_copy_bvh(lsb200_ctx *c, float *nodes16, int64_t node_cap, int32_t *leaf_ids, int64_t tri_cap, int32_t *root)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (!c->committed) return fail(c, LSB200_ERR_STATE, "debug_copy_bvh: not committed");
    DeviceGuard dg(c->device);
    const int64_t nn = std::max(c->n_tris - 1, 1);
    if (nodes16) {
        if (node_cap < nn) return fail(c, LSB200_ERR_CAPACITY, "debug_copy_bvh: need %lld nodes", (long long)nn);
        LSB_CU(c, cudaMemcpyAsync(nodes16, c->d_nodes.as<float4>() + 4 * (size_t)c->built_top, (size_t)nn * 64, cudaMemcpyDeviceToHost, c->stream));
    }
    if (leaf_ids) {
        if (tri_cap < c->n_tris) return fail(c, LSB200_ERR_CAPACITY, "debug_copy_bvh: need %d ids", c->n_tris);
        LSB_CU(c, cudaMemcpyAsync(leaf_ids, c->d_leaf_ids.p, (size_t)c->n_tris * 4, cudaMemcpyDeviceToHost, c->stream));
    }
    LSB_CU(c, cudaStreamSynchronize(c->stream));
    if (root) *root = c->built_top;     // unified index of Karras node 0; child refs >= top_cap are (ref - top_cap)
    return LSB200_OK;
}

int lsb200_debug_sort_pairs(lsb200_ctx *c, uint32_t *keys, uint32_t *vals, int64_t n)
{
    if (!c || !keys || !vals || n < 0) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (n == 0) return LSB200_OK;
    DeviceGuard dg(c->device);
    DevBuf k, v, kt, vt, sc;
    cudaStream_t st = c->stream;
    int rc = LSB200_OK;
    auto body = [&]() -> int {
        LSB_CU(c, k.reserve(n * 4, st)); LSB_CU(c, v.reserve(n * 4, st)); LSB_CU(c, kt.reserve(n * 4, st));
        LSB_CU(c, vt.reserve(n * 4, st)); LSB_CU(c, sc.reserve(rsort::scratch_words(n) * 4, st));
        LSB_CU(c, cudaMemcpyAsync(k.p, keys, n * 4, cudaMemcpyHostToDevice, st));
        LSB_CU(c, cudaMemcpyAsync(v.p, vals, n * 4, cudaMemcpyHostToDevice, st));
        count_launch(c, rsort::sort_pairs(k.as<uint32_t>(), v.as<uint32_t>(), kt.as<uint32_t>(), vt.as<uint32_t>(), n, sc.as<uint32_t>(), st));
        LSB_CU(c, cudaMemcpyAsync(keys, k.p, n * 4, cudaMemcpyDeviceToHost, st));
        LSB_CU(c, cudaMemcpyAsync(vals, v.p, n * 4, cudaMemcpyDeviceToHost, st));
        LSB_CU(c, cudaStreamSynchronize(st));
        LSB_CU(c, cudaGetLastError());
        return LSB200_OK;
    };
    rc = body();
    cudaStreamSynchronize(st);
    k.release(); v.release(); kt.release(); vt.release(); sc.release();
    return rc;
}

#ifdef LSB_SORT_PROF
// A/B instrumentation only (tools/sort_phase_prof.py): reads and clears the phase counters of k_onesweep
LSB200_API int lsb200_debug_sort_prof(unsigned long long *out8)
{
    unsigned long long z[8] = {};
    if (cudaMemcpyFromSymbol(out8, rsort::g_sort_prof, sizeof(z)) != cudaSuccess) return LSB200_ERR_CUDA;
    if (cudaMemcpyToSymbol(rsort::g_sort_prof, z, sizeof(z)) != cudaSuccess) return LSB200_ERR_CUDA;
    return LSB200_OK;
}
#endif

int lsb200_debug_copy_vertices(lsb200_ctx *c, float *xyz, int64_t cap)
{
    if (!c || !xyz) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (!c->committed) return fail(c, LSB200_ERR_STATE, "debug_copy_vertices: not committed");
    if (cap < c->n_verts) return fail(c, LSB200_ERR_CAPACITY, "debug_copy_vertices: need %d", c->n_verts);
    DeviceGuard dg(c->device);
    std::vector<float> tmp((size_t)c->n_verts * 4);
    LSB_CU(c, cudaMemcpyAsync(tmp.data(), c->d_xverts.p, tmp.size() * 4, cudaMemcpyDeviceToHost, c->stream));
    LSB_CU(c, cudaStreamSynchronize(c->stream));
    for (int i = 0; i < c->n_verts; ++i) { xyz[3 * i] = tmp[4 * i]; xyz[3 * i + 1] = tmp[4 * i + 1]; xyz[3 * i + 2] = tmp[4 * i + 2]; }
    return LSB200_OK;
}

}  // extern "C"

// =============================================================================================================
//  Round-2 additions to the boundary: log sink, pinned host memory placed next to the GPU, range-image expander,
//  scene cloning and the multi-GPU batch entry
// =============================================================================================================
namespace {

// NUMA node of a CUDA device from sysfs (/sys/bus/pci/devices/<bdf>/numa_node); -1 when unknown
int device_numa_node(int cuda_device)
{
    char bdf[32] = {0};
    if (cudaDeviceGetPCIBusId(bdf, sizeof(bdf), cuda_device) != cudaSuccess) return -1;
    for (char *p = bdf; *p; ++p) if (*p >= 'A' && *p <= 'Z') *p = (char)(*p - 'A' + 'a');
    char path[128];
    snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/numa_node", bdf);
    FILE *f = fopen(path, "r");
    if (!f) return -1;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    return node;
}

// CPUs of a NUMA node ("0-15,32-47") -> cpu_set_t; false when the node is unknown
bool node_cpus(int node, cpu_set_t *set)
{
    char path[128];
    snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", node);
    FILE *f = fopen(path, "r");
    if (!f) return false;
    char buf[4096] = {0};
    const bool ok = fgets(buf, sizeof(buf), f) != nullptr;
    fclose(f);
    if (!ok) return false;
    CPU_ZERO(set);
    int n = 0;
    for (char *p = buf; *p;) {
        char *e;
        long a = strtol(p, &e, 10);
        if (e == p) break;
        long b = a;
        if (*e == '-') { p = e + 1; b = strtol(p, &e, 10); }
        for (long k = a; k <= b && k < CPU_SETSIZE; ++k) { CPU_SET((int)k, set); ++n; }
        p = (*e == ',') ? e + 1 : e;
        if (*e != ',') break;
    }
    return n > 0;
}

}  // namespace

extern "C" {

int lsb200_set_log_callback(lsb200_ctx *c, lsb200_log_fn fn, void *user)
{
    if (!c) {                                   // process-wide default: inherited by contexts created afterwards
        std::lock_guard<std::mutex> lk(g_log_mtx);
        g_log_fn = fn; g_log_user = user;
        return LSB200_OK;
    }
    std::lock_guard<std::mutex> lk(c->mtx);
    c->log_fn = fn; c->log_user = user;
    return LSB200_OK;
}

// Pinned host memory for the output clouds / vertex blobs, allocated while the calling thread is bound to the CPUs
// (and, through the kernel's local-allocation policy, the memory) of the NUMA node the context's GPU hangs off, so
// the D2H copies of several GPUs do not all land on one socket's memory controller.
int lsb200_alloc_host(lsb200_ctx *c, size_t bytes, void **out)
{
    if (!c || !out) return LSB200_ERR_INVALID;
    *out = nullptr;
    std::lock_guard<std::mutex> lk(c->mtx);
    DeviceGuard dg(c->device);
    cpu_set_t old_set, node_set;
    bool bound = false;
    const int node = device_numa_node(c->device);
    if (node >= 0 && sched_getaffinity(0, sizeof(old_set), &old_set) == 0 && node_cpus(node, &node_set)) {
        cpu_set_t both;
        CPU_AND(&both, &old_set, &node_set);     // never leave the cpuset the process was given
        if (CPU_COUNT(&both) > 0 && sched_setaffinity(0, sizeof(both), &both) == 0) bound = true;
    }
    void *p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e == cudaSuccess && bytes) std::memset(p, 0, bytes);       // first touch while bound
    if (bound) sched_setaffinity(0, sizeof(old_set), &old_set);
    if (e != cudaSuccess) return fail(c, LSB200_ERR_NOMEM, "alloc_host: cudaHostAlloc(%zu) -> %s", bytes, cudaGetErrorString(e));
    logf(c, LSB200_LOG_DEBUG, "lsb200: %zu pinned host bytes for device %d (NUMA node %d, %s)", bytes, c->device, node,
         bound ? "thread bound during allocation" : "no binding");
    *out = p;
    return LSB200_OK;
}

int lsb200_free_host(lsb200_ctx *c, void *p)
{
    if (!p) return LSB200_OK;
    if (cudaFreeHost(p) != cudaSuccess) return fail(c, LSB200_ERR_CUDA, "free_host failed");
    return LSB200_OK;
}

int lsb200_device_numa_node(const lsb200_ctx *c) { return c ? device_numa_node(c->device) : LSB200_ERR_INVALID; }

// Range image -> point records on the HOST, for consumers that asked for LSB200_TRACE_RANGE_IMAGE (4 bytes per ray over
// PCIe instead of 16 / 32 per hit) but still want the reference's records: the same arithmetic as the device
// (point = t * d_sensor with d_sensor from the host tables, SRC/EmbreeTracer.cpp:331-333), hits in ray-id order, so the
// bytes equal those of an LSB200_TRACE_ORDERED trace.
int lsb200_expand_range_image(lsb200_ctx *c, int idx, const float *t, void *out_points, size_t capacity, uint32_t *n_pts, uint32_t flags)
{
    if (!c || !t) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (idx < 0 || (size_t)idx >= c->sensors.size() || !c->sensors[idx].set) return fail(c, LSB200_ERR_INVALID, "expand_range_image: bad sensor %d", idx);
    const Sensor &s = c->sensors[idx];
    const int V = s.V, H = s.H;
    const float *st = s.h_tab.data(), *ct = st + V, *sp = st + 2 * V, *cp = st + 2 * V + H;
    const bool xyzi16 = (flags & LSB200_TRACE_XYZI16) != 0, world = (flags & LSB200_TRACE_WORLD_POINTS) != 0;
    const bool ring0 = (flags & LSB200_TRACE_RING_ZERO) != 0;
    uint8_t *out = static_cast<uint8_t *>(out_points);
    size_t n = 0;
    for (int v = 0; v < V; ++v) {
        for (int h = 0; h < H; ++h) {
            const float tt = t[(size_t)v * H + h];
            if (!(tt <= 3.402823466e+38f)) continue;               // +inf / NaN: no return
            if (n < capacity && out) {
                const float dsx = st[v] * cp[h], dsy = st[v] * sp[h], dsz = ct[v];
                float x = tt * dsx, y = tt * dsy, z = tt * dsz;
                if (world) {
                    const float wx = (s.R[0] * x + (s.R[1] * y + s.R[2] * z)) + s.T[0];
                    const float wy = (s.R[3] * x + (s.R[4] * y + s.R[5] * z)) + s.T[1];
                    const float wz = (s.R[6] * x + (s.R[7] * y + s.R[8] * z)) + s.T[2];
                    x = wx; y = wy; z = wz;
                }
                if (xyzi16) {
                    const float rec[4] = {x, y, z, 64.0f};
                    std::memcpy(out + 16 * n, rec, 16);
                } else {
                    float rec[8] = {x, y, z, 0.0f, 64.0f, 0.0f, 0.0f, 0.0f};
                    const int32_t ring = ring0 ? 0 : v;
                    std::memcpy(&rec[5], &ring, 4);
                    std::memcpy(out + 32 * n, rec, 32);
                }
            }
            ++n;
        }
    }
    if (n_pts) *n_pts = (uint32_t)n;
    if (out && n > capacity) return fail(c, LSB200_ERR_CAPACITY, "expand_range_image: %zu points, capacity %zu", n, capacity);
    return LSB200_OK;
}

// A second context on another GPU holding the same scene: options, sensors (+ limits), every geometry's vertex and
// index data (device-to-device copies over NVLink / PCIe peer access) and transforms.  The scene is replicated per
// GPU, as north_star asks; nothing is shared afterwards.
int lsb200_clone_scene(lsb200_ctx *src, int cuda_device, lsb200_ctx **out)
{
    if (!src || !out) return LSB200_ERR_INVALID;
    *out = nullptr;
    lsb200_ctx *dst = nullptr;
    int rc = lsb200_create(cuda_device, &dst);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(src->mtx);
    auto bail = [&](int code) { lsb200_destroy(dst); return code; };
    dst->frame_mode = src->frame_mode;
    dst->opt_traversal = src->opt_traversal; dst->opt_leaf = src->opt_leaf; dst->opt_top = src->opt_top;
    dst->opt_ctas_per_sm = src->opt_ctas_per_sm; dst->opt_morton_uniform = src->opt_morton_uniform;
    dst->opt_morton_layout = src->opt_morton_layout; dst->opt_morton_aspect = src->opt_morton_aspect; dst->opt_chain = src->opt_chain;
    dst->opt_tile_w_log2 = src->opt_tile_w_log2; dst->opt_small_build = src->opt_small_build; dst->opt_tile_entries = src->opt_tile_entries;
    dst->opt_tile_visits = src->opt_tile_visits; dst->opt_tile_min_units = src->opt_tile_min_units;
    dst->opt_watertight = src->opt_watertight; dst->opt_intensity = src->opt_intensity; dst->opt_fused_tree = src->opt_fused_tree; dst->opt_lane_refill = src->opt_lane_refill; dst->opt_returns = src->opt_returns; dst->opt_small_trace_units = src->opt_small_trace_units;
    dst->log_fn = src->log_fn; dst->log_user = src->log_user;
    for (size_t i = 0; i < src->sensors.size(); ++i) {
        const Sensor &s = src->sensors[i];
        if (!s.set) continue;
        rc = lsb200_set_sensor(dst, (int)i, s.elev.data(), s.V, s.az_begin, s.az_end, s.H, s.q, s.T);
        if (rc) return bail(rc);
        if (s.lim_range > 0.0f || s.has_mask) { rc = lsb200_set_sensor_limits(dst, (int)i, s.lim_range, s.lim_lo, s.lim_hi); if (rc) return bail(rc); }
        if (s.min_intensity > 0.0f) lsb200_set_sensor_intensity_filter(dst, (int)i, s.min_intensity);
    }
    {
        DeviceGuard dgs(src->device);
        cudaStreamSynchronize(src->stream);         // pending uploads of the source have landed
    }
    for (auto &kv : src->geoms) {
        const Geometry &g = kv.second;
        // ids are assigned lowest-free-first on both sides and the source's ids may have holes: add under the same id
        Geometry &d = dst->geoms[g.id];
        d.name = g.name; d.id = g.id; d.type = g.type; d.n_verts = g.n_verts; d.n_elems = g.n_elems;
        d.verts_per_elem = g.verts_per_elem; d.stride = g.stride;
        std::memcpy(d.A, g.A, sizeof(d.A));
        dst->by_name[g.name] = g.id;
        DeviceGuard dgd(dst->device);
        if (g.has_verts) {
            const size_t bytes = (size_t)g.n_verts * (size_t)g.stride;
            if (d.verts.reserve(std::max<size_t>(bytes, 16), dst->stream) != cudaSuccess) return bail(LSB200_ERR_NOMEM);
            if (bytes && cudaMemcpyPeerAsync(d.verts.p, dst->device, g.verts.p, src->device, bytes, dst->stream) != cudaSuccess) return bail(LSB200_ERR_CUDA);
            d.has_verts = true;
        }
        if (g.has_elems) {
            const size_t bytes = (size_t)g.n_elems * g.verts_per_elem * sizeof(uint32_t);
            if (d.elems.reserve(std::max<size_t>(bytes, 16), dst->stream) != cudaSuccess) return bail(LSB200_ERR_NOMEM);
            if (bytes && cudaMemcpyPeerAsync(d.elems.p, dst->device, g.elems.p, src->device, bytes, dst->stream) != cudaSuccess) return bail(LSB200_ERR_CUDA);
            d.has_elems = true;
        }
    }
    {
        DeviceGuard dgd(dst->device);
        if (cudaStreamSynchronize(dst->stream) != cudaSuccess) return bail(LSB200_ERR_CUDA);
    }
    dst->topology_dirty = true; dst->committed = false;
    *out = dst;
    return LSB200_OK;
}

// Multi-GPU batch: the C entry for a host that owns the whole box (the reference runs one tracer per LiDAR device in
// one process, SRC/mainwindow.cpp:256-274).  One context per GPU (lsb200_create + scene, or lsb200_clone_scene), one
// host thread per context, each running lsb200_run_frames on its share.  No data-path collective: the (device, frame)
// scans are independent.
//   LSB200_SHARD_FRAMES   context g takes frames g, g + n, ... (each GPU rebuilds only its own frames' BVH: config 5)
//   LSB200_SHARD_SENSORS  every context takes every frame and traces a contiguous share of the sensors (a frame is
//                         finished as early as possible, every GPU pays the commit: config 4's latency view)
int lsb200_run_batch(lsb200_ctx *const *ctxs, int n_ctx, const lsb200_frame *frames, int n_frames, int first, int count,
                     int commit_mode, uint32_t flags, int sharding)
{
    if (!ctxs || n_ctx <= 0 || !frames || n_frames <= 0 || count <= 0) return LSB200_ERR_INVALID;
    for (int g = 0; g < n_ctx; ++g) if (!ctxs[g]) return LSB200_ERR_INVALID;
    if (sharding != LSB200_SHARD_FRAMES && sharding != LSB200_SHARD_SENSORS) return fail(ctxs[0], LSB200_ERR_INVALID, "run_batch: bad sharding %d", sharding);
    std::vector<int> rcs((size_t)n_ctx, LSB200_OK);
    std::vector<std::vector<lsb200_frame>> share((size_t)n_ctx);
    std::vector<int> s_first((size_t)n_ctx, first), s_count((size_t)n_ctx, count);
    for (int g = 0; g < n_ctx; ++g) {
        if (sharding == LSB200_SHARD_FRAMES) {
            for (int f = g; f < n_frames; f += n_ctx) share[g].push_back(frames[f]);
        } else {
            const int lo = (int)((long long)count * g / n_ctx), hi = (int)((long long)count * (g + 1) / n_ctx);
            s_first[g] = first + lo; s_count[g] = hi - lo;
            if (hi > lo)
                for (int f = 0; f < n_frames; ++f) {
                    lsb200_frame fr = frames[f];
                    fr.out_points = frames[f].out_points ? frames[f].out_points + lo : nullptr;
                    fr.capacities = frames[f].capacities ? frames[f].capacities + lo : nullptr;
                    fr.n_pts = frames[f].n_pts ? frames[f].n_pts + lo : nullptr;
                    share[g].push_back(fr);
                }
        }
    }
    std::vector<std::thread> threads;
    for (int g = 0; g < n_ctx; ++g) {
        if (share[g].empty()) continue;
        threads.emplace_back([&, g]() {
            rcs[g] = lsb200_run_frames(ctxs[g], share[g].data(), (int)share[g].size(), s_first[g], s_count[g], commit_mode, flags);
        });
    }
    for (auto &t : threads) t.join();
    for (int g = 0; g < n_ctx; ++g) if (rcs[g] != LSB200_OK) return rcs[g];
    return LSB200_OK;
}

}  // extern "C"

// =============================================================================================================
//  Allocation-free sort (no counterpart in the reference).
//  The radix sort's scratch in a file-scope __device__ array: no cudaMalloc / cudaFree on this path.
// =============================================================================================================
namespace {

constexpr int64_t kStaticSortTiles = ((int64_t)LSB200_SORT_STATIC_MAX_PAIRS + rsort::kTile - 1) / rsort::kTile;
// rsort::scratch_words(LSB200_SORT_STATIC_MAX_PAIRS), as a constant expression
constexpr size_t kStaticSortWords = (size_t)rsort::kPasses * rsort::kRadix + 16 + (size_t)rsort::kPasses * kStaticSortTiles * rsort::kRadix;

// Held by every GPU that loads the library, until the process exits (2.0 MiB).  In HOST code this name is only a
// stand-in for the symbol, never a device address: kernels get the pointer cudaGetSymbolAddress returns.
static __device__ __align__(256) uint32_t g_static_sort_scratch[kStaticSortWords];

constexpr int kMaxStaticDevices = 64;
uint32_t *g_static_sort_ptr[kMaxStaticDevices] = {};   // the array's address on each device (it differs per device)
std::mutex g_static_sort_mtx;

// Set-up, once per device, with that device current (the caller holds a DeviceGuard).
int static_sort_scratch(lsb200_ctx *c, uint32_t **out)
{
    if (c->device < 0 || c->device >= kMaxStaticDevices) return fail(c, LSB200_ERR_INVALID, "sort_pairs_static: device %d out of range", c->device);
    std::lock_guard<std::mutex> lk(g_static_sort_mtx);
    if (!g_static_sort_ptr[c->device]) {
        void *p = nullptr;
        LSB_CU(c, cudaGetSymbolAddress(&p, g_static_sort_scratch));
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255u)) return fail(c, LSB200_ERR_CUDA, "sort_pairs_static: bad symbol address %p", p);
        g_static_sort_ptr[c->device] = static_cast<uint32_t *>(p);
    }
    *out = g_static_sort_ptr[c->device];
    return LSB200_OK;
}

}  // namespace

extern "C" int lsb200_sort_pairs_static(lsb200_ctx *c, uint32_t *d_keys, uint32_t *d_vals, uint32_t *d_keys_tmp, uint32_t *d_vals_tmp, int64_t n)
{
    if (!c) return LSB200_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mtx);
    if (n < 0) return fail(c, LSB200_ERR_INVALID, "sort_pairs_static: n = %lld", (long long)n);
    if (n > (int64_t)LSB200_SORT_STATIC_MAX_PAIRS)
        return fail(c, LSB200_ERR_CAPACITY, "sort_pairs_static: %lld pairs exceed LSB200_SORT_STATIC_MAX_PAIRS = %d (the static scratch array is sized for that many)",
                    (long long)n, (int)LSB200_SORT_STATIC_MAX_PAIRS);
    if (n > 0 && (!d_keys || !d_vals || !d_keys_tmp || !d_vals_tmp)) return fail(c, LSB200_ERR_INVALID, "sort_pairs_static: null device pointer");
    DeviceGuard dg(c->device);
    uint32_t *scratch = nullptr;
    const int rc = static_sort_scratch(c, &scratch);
    if (rc != LSB200_OK || n == 0) return rc;
    // sort_pairs clears the words it uses (histograms, tickets, status) on the stream before its first kernel
    count_launch(c, rsort::sort_pairs(d_keys, d_vals, d_keys_tmp, d_vals_tmp, n, scratch, c->stream));
    LSB_CU(c, cudaGetLastError());
    return LSB200_OK;
}
