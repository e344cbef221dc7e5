/*
 * lsb200.h -- C ABI of the B200-native LiDAR ray-casting backend (liblsb200.so).
 *
 * This is the drop-in boundary for lidarshooter's trace step.  Each entry point replaces one call of the
 * reference's tracer interface (reference tree /root/reference, SRC = ros_ws/src/lidarshooter/src); the C++
 * adapter `lidarshooter::B200Tracer : ITracer` (lidarshooter_b200/host/B200Tracer.hpp) is a thin marshaller
 * over these functions.  Plain C types only, no exceptions cross, every function returns an lsb200_status
 * (>= 0 ok, < 0 error; lsb200_last_error() gives the text).
 *
 * Frames.  LSB200_FRAME_SENSOR reproduces the reference's arithmetic: every vertex is moved into the sensor
 * frame (SRC/MeshTransformer.cpp:173-194, SRC/LidarDevice.cpp:383-391), rays start at 0.  One context == one
 * LiDAR device, like one EmbreeTracer.  LSB200_FRAME_WORLD keeps the scene in the world frame, builds ONE
 * BVH per frame and moves the rays instead (o = T, d = R d_s), so many sensors share a context.
 *
 * Per-ray results: t (f32, +inf on a miss) and the global triangle id (i32, -1 on a miss).  The global
 * triangle id is the position of the triangle in the concatenation of all geometries in ascending
 * geometry-id order.  Closest hit = lexicographic minimum of (t, global triangle id).
 */
#ifndef LSB200_H
#define LSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define LSB200_API
#else
#define LSB200_API __attribute__((visibility("default")))
#endif

typedef struct lsb200_ctx lsb200_ctx;

typedef enum lsb200_status {
    LSB200_OK = 0,
    LSB200_ERR_INVALID = -1,      /* bad argument / unknown geometry (SRC/EmbreeTracer.cpp:224-225 returns -1) */
    LSB200_ERR_EMPTY_SCENE = -1,  /* commit/trace on an empty scene return -1 (SRC/OptixTracer.cpp:263-288) */
    LSB200_ERR_CUDA = -2,
    LSB200_ERR_CAPACITY = -3,     /* caller's output buffer too small; *n_pts still holds the hit count */
    LSB200_ERR_STATE = -4,        /* e.g. trace before commit, sensor not set */
    LSB200_ERR_NOMEM = -5
} lsb200_status;

/* geometry element types; values match Embree's RTCGeometryType used by ITracer::addGeometry
 * (SRC/ITracer.hpp:50; SRC/EmbreeTracer.cpp:115-201): TRIANGLE = 0, QUAD = 1. */
enum { LSB200_GEOMETRY_TRIANGLE = 0, LSB200_GEOMETRY_QUAD = 1 };

enum { LSB200_FRAME_SENSOR = 0, LSB200_FRAME_WORLD = 1 };

/* lsb200_commit_scene modes: AUTO = rebuild when topology (add/remove/index data) changed since the last
 * build, otherwise refit (what OptiX's BUILD/UPDATE switch does, SRC/OptixTracer.cpp:529-535). */
enum { LSB200_COMMIT_AUTO = 0, LSB200_COMMIT_REBUILD = 1, LSB200_COMMIT_REFIT = 2 };

/* lsb200_trace_* flags */
enum {
    LSB200_TRACE_ORDERED = 1u << 0,    /* points emitted in ray-id order (scan); default: warp-aggregated atomics */
    LSB200_TRACE_RING_ZERO = 1u << 1,  /* ring = 0 for every point (Embree-path quirk, SRC/EmbreeTracer.cpp:333-334) */
    LSB200_TRACE_XYZI16 = 1u << 2,     /* 16-byte x,y,z,intensity records instead of the 32-byte PointCloud2 record */
    LSB200_TRACE_NO_POINTS = 1u << 3,  /* per-ray t / triangle only (lsb200_trace_raw uses this) */
    LSB200_TRACE_WORLD_POINTS = 1u << 4, /* emit points in the world frame: R p + T (SRC/LidarDevice.cpp:393-401) */
    LSB200_TRACE_RANGE_IMAGE = 1u << 5  /* output = the dense range image: one f32 range per ray in ray-id order (+inf = no
                                           return), 4 bytes per RAY over PCIe instead of 16 / 32 per hit -- what a physical
                                           LiDAR's packets carry.  *n_pts = ray count.  lsb200_expand_range_image turns it
                                           into the reference's records on the host (same bytes as LSB200_TRACE_ORDERED). */
};

/* traversal kernel variants (lsb200_set_option LSB200_OPT_TRAVERSAL) -- for the ncu before/after evidence */
enum {
    LSB200_OPT_TRAVERSAL = 1,      /* 6 (default) = persistent CTAs, 32-ray units queued by ray octant, sign-specialised
                                      slab test, while-while loop (node visits and leaf tests separated); 0 = the same
                                      with leaf tests inside the node loop; 1 = persistent, generic slab test, units in
                                      table order; 2 = one thread per ray (no persistent fetch); 3 = brute force (all
                                      triangles); 4 = variant 1 + top BVH levels staged into shared memory by TMA bulk
                                      copy; 5 = variant 1, while-while; 7 = variant 6 with node visits and leaf tests in ONE loop
                                      (fewer dependent round trips per warp, more instructions: measured 3 % slower).
                                      Variants 6 and 7 start the rays from per-tile entry lists (LSB200_OPT_TILE_ENTRIES) and
                                      prefetch the next unit's ticket / descriptor / list while the current one is traced */
    LSB200_OPT_LEAF_SIZE = 2,      /* max triangles per leaf, 1..8 (default 2) */
    LSB200_OPT_TOP_NODES = 3,      /* number of BVH nodes staged into shared memory by TMA bulk copy (default 1024) */
    LSB200_OPT_STAGE_TIMING = 4,   /* 1 = record CUDA events around every stage (lsb200_get_stats) */
    LSB200_OPT_CTAS_PER_SM = 5,    /* persistent grid = 148 * this (default 0 = occupancy query) */
    LSB200_OPT_TRACE_STATS = 6,    /* 1 = instrumented traversal: per-launch node / leaf / triangle counters */
    LSB200_OPT_MORTON_UNIFORM = 7, /* 1 = Morton codes on cubic cells (one scale for x, y, z); default 0 = per-axis */
    LSB200_OPT_ORIGIN_CHAIN = 8,   /* 1 (default) = per-sensor origin chain: rays start from the list of subtrees that hang off
                                      the root-to-origin chain instead of from the root; 0 = every ray starts at the root */
    LSB200_OPT_TILE_WIDTH = 9,     /* a 32-ray work unit is a tile of W azimuth steps x 32/W channels of one device's
                                      table; W = 1, 2, 4, 8 (default), 16 or 32 */
    LSB200_OPT_MORTON_LAYOUT = 11, /* Morton bits: (bits of x and of y) << 4 | bits of z, 2 * bxy + bz <= 32, bz <= min(bxy, 10);
                                      the extra x / y bits are interleaved above the 3-way part.  0 = the default */
    LSB200_OPT_MORTON_ASPECT = 12, /* 100 * alpha: the next Morton bit goes to the axis with the largest (cell size)^alpha.
                                      0 = round robin x, y, z; 100 = longest side first (cubic cells); default 60 */
    LSB200_OPT_FIT_RMQ = 13,       /* 1 = block-local node boxes by range queries over the block's leaf boxes (k_fit_rmq; measured
                                      slower: 0.86 vs 0.54 ms at 10 M triangles); 0 (default) = bottom-up climb (k_fit) */
    LSB200_OPT_TILE_ENTRIES = 14,  /* entries of the per-tile entry lists (0 = off, 2..16; default 10): one thread per 32-ray tile walks
                                      the part of the BVH descent that the tile's rays share, once, and lists the subtrees / leaves
                                      the tile's pyramid can reach; the rays start from that list (variant 6, >= 4096 tiles) */
    LSB200_OPT_TILE_VISITS = 15,   /* node visits a tile may spend on its list (0 = default: 3.2 x entries) */
    LSB200_OPT_SUPER_TILES = 17,   /* 1 = two-level lists: a tile's walk starts from the list of its supertile (4 x 4 tiles, one walk
                                      per supertile) instead of its sensor's origin chain: 26 -> 9 node visits per tile walk, but
                                      measured slower overall (1.90 vs 1.74 ms per trace stage); 0 (default) = off */
    LSB200_OPT_TILE_MIN_UNITS = 16, /* tile lists are built for launches of at least this many 32-ray tiles (default 4096: below
                                      that, e.g. an XT-32 with 150 tiles, the extra launch costs more than it saves) */
    LSB200_OPT_SMALL_BUILD = 10,   /* 1 (default) = scenes of <= 16384 triangles in <= 16 geometries are committed by ONE
                                      fused single-CTA kernel (transform, Morton, sort, hierarchy, fit); 0 = always the
                                      multi-kernel pipeline */
    LSB200_OPT_WATERTIGHT = 18,    /* 0 (default) = Embree-ordered Moeller-Trumbore, the reference's intersector (drop-in parity;
                                      SRC/EmbreeTracer.cpp:472-480 -> Embree 3.13.4 triangle_intersector_moeller.h); 1 = watertight
                                      Woop-Benthin-Wald 2013 shear / scale test (edge functions on the sheared vertices, exact
                                      double-precision fallback on a zero edge function): a ray can never slip between two
                                      triangles that share an edge.  Bit-exact against the oracle in the same mode. */
    LSB200_OPT_INTENSITY_MODEL = 19, /* 0 (default) = constant 64.0, what the reference writes (SRC/EmbreeTracer.cpp:333); 1 =
                                      Lambertian cos-incidence: 255 * |Ng . d| / (|Ng| |d|) from the hit triangle's normal
                                      (SURVEY 8f row 4).  Models other than 0 (and an intensity filter) emit the points through the
                                      ordered compaction. */
    LSB200_OPT_LANE_REFILL = 21,   /* ray compaction inside the traversal: bit mask of the queue groups that are drained ray by ray
                                      (a lane whose ray has ended takes the next ray of the queue: ballot + rank, unit lists in a
                                      shared-memory ring) instead of one 32-ray unit per warp at a time.  bit 0 = sign-mixed units,
                                      1 = upward octants, 2 = heavy (grazing) fronts of the downward octants, 3 = their light
                                      remainders.  Built into liblsb200_lab.so only; default 0 (off): measured slower on the 10 M-triangle city, see DESIGN.md / profiles/design_choices_r02.md --
                                      the traversal is bound by instruction issue and dependent round trips per unit, and the refill
                                      loop costs more instructions than the idle lanes it fills.  Same per-ray results either way. */
    LSB200_OPT_RETURNS = 22,       /* returns per ray: 1 (default, what the reference produces) or 2 = dual return -- every trace also
                                      finds, per ray, the NEXT surface behind the first hit: the smallest (t, triangle id) greater
                                      than the first hit's (the any-hit multi-return the reference only plans, SRC/LidarShooter.hpp.in:
                                      87-93).  Read with lsb200_fetch_second_returns.  The cloud itself is unchanged. */
    LSB200_OPT_SMALL_TRACE_UNITS = 23, /* trace launches of fewer 32-ray units than this (default 512 = 16 384 rays; an XT-32 scan has 150)
                                      run one thread per ray from the root (k_trace_small) instead of the persistent queue kernel,
                                      whose fixed costs dominate at that size; 0 = never.  Same per-ray results. */
    LSB200_OPT_FUSED_TREE = 20     /* 1 (default) = hierarchy and node boxes come out of ONE bottom-up pass over the sorted codes
                                      (k_tree: Apetrei 2014 with Karras node numbering, block-local rendezvous in shared memory);
                                      0 = the two-kernel form (k_hierarchy: Karras 2012 top-down searches, then k_fit).  Same node
                                      array bit for bit; kept for the before / after evidence. */
};

/* log levels of lsb200_set_log_callback == spdlog::level (the reference logs through spdlog, SRC/ITracer.cpp:5-14) */
enum { LSB200_LOG_TRACE = 0, LSB200_LOG_DEBUG = 1, LSB200_LOG_INFO = 2, LSB200_LOG_WARN = 3, LSB200_LOG_ERROR = 4 };
enum { LSB200_SHARD_FRAMES = 0, LSB200_SHARD_SENSORS = 1 };

/* ---- life cycle -------------------------------------------------------------------------------------- */
/* replaces EmbreeTracer::create / rtcNewDevice + rtcNewScene (SRC/EmbreeTracer.cpp:20-34, 482-486) */
LSB200_API int lsb200_create(int cuda_device, lsb200_ctx **out);
LSB200_API int lsb200_destroy(lsb200_ctx *ctx);
/* all work of this context is enqueued on `cuda_stream` (a cudaStream_t; NULL = the context's own stream) */
LSB200_API int lsb200_set_stream(lsb200_ctx *ctx, void *cuda_stream);
LSB200_API int lsb200_set_frame_mode(lsb200_ctx *ctx, int frame_mode);
LSB200_API int lsb200_set_option(lsb200_ctx *ctx, int option, int value);
LSB200_API const char *lsb200_last_error(const lsb200_ctx *ctx);
/* Log sink: replaces the std::shared_ptr<spdlog::logger> every reference tracer takes (SRC/EmbreeTracer.hpp:51,
 * SRC/ITracer.cpp:5-14; OptiX's own log callback: SRC/OptixTracer.cpp:432-441).  Every error of the ABI (the text of
 * lsb200_last_error) is reported at LSB200_LOG_ERROR, life-cycle events at INFO, per-commit lines at DEBUG.  ctx == NULL
 * sets the process-wide default that contexts created afterwards inherit (and that lsb200_create's own failures use).
 * The callback may be invoked from any thread that calls into the ABI; fn == NULL switches logging off. */
typedef void (*lsb200_log_fn)(int level, const char *message, void *user);
LSB200_API int lsb200_set_log_callback(lsb200_ctx *ctx, lsb200_log_fn fn, void *user);
LSB200_API const char *lsb200_version(void);

/* ---- sensor: what LidarDevice holds for the path (SRC/LidarDevice.cpp:598-617 channels, 788-816 transform) ---
 * elev_deg[V] degrees above horizontal; azimuth table az[h] = az_begin + step * h with
 * step = (az_end - az_begin) / (H - 1) in fp32; q_wxyz is NOT normalised (SRC/LidarDevice.cpp:808-813);
 * t_xyz = (base_to_origin.tx, base_to_origin.ty, sensor_to_base.tz).  The sin/cos tables are computed on the
 * host with the reference's own narrowing (double -> float before sinf/cosf, SRC/LidarDevice.cpp:312-318).
 * sensor_index 0 is the context's device; indices > 0 are allowed in LSB200_FRAME_WORLD only. */
LSB200_API int lsb200_set_sensor(lsb200_ctx *ctx, int sensor_index, const float *elev_deg, int V, float az_begin,
                                 float az_end, int H, const float q_wxyz[4], const float t_xyz[3]);
/* pose-only update of a sensor that is already set (same q / t conventions): no table upload, no synchronisation -- for
 * sensors that move every frame (the reference re-reads the pose only at start-up, SRC/LidarDevice.cpp:788-816) */
LSB200_API int lsb200_set_sensor_pose(lsb200_ctx *ctx, int sensor_index, const float q_wxyz[4], const float t_xyz[3]);
/* Sensor limits the reference carries in its device config but never applies (config/hesai-pandar-XT-32-lidar_0000.json:
 * 14-18 "detection_radius", "horizontal_angle_range").  Semantics defined here: a ray whose closest
 * hit lies beyond max_range returns nothing (max_range <= 0: unlimited); an azimuth step whose angle, wrapped into
 * (-180, 180], lies outside [az_min_deg, az_max_deg] is not fired.  Cleared by lsb200_set_sensor.  Off by default, so
 * the drop-in behaviour equals the reference's. */
LSB200_API int lsb200_set_sensor_limits(lsb200_ctx *ctx, int sensor_index, float max_range, float az_min_deg, float az_max_deg);
/* use_intensity_filtering / minimum_intensity of the device config (config/hesai-pandar-XT-32-lidar_0000.json:19-20; carried
 * but never applied by the reference): a return whose intensity is below min_intensity is dropped (no point, t = +inf,
 * triangle = -1).  0 = off (default).  With the constant model (64.0) only thresholds above 64 drop anything. */
LSB200_API int lsb200_set_sensor_intensity_filter(lsb200_ctx *ctx, int sensor_index, float min_intensity);
LSB200_API int lsb200_get_ray_count(const lsb200_ctx *ctx, int sensor_index);

/* ---- geometry: ITracer::addGeometry / removeGeometry / updateGeometry (SRC/ITracer.hpp:50-85) ------------- */
/* replaces EmbreeTracer::addGeometry (SRC/EmbreeTracer.cpp:115-218).  Returns the geometry id (>= 0), lowest
 * free id first like rtcAttachGeometry; unsupported geometry_type -> LSB200_ERR_INVALID. */
LSB200_API int lsb200_add_geometry(lsb200_ctx *ctx, const char *name, int geometry_type, int n_vertices,
                                   int n_elements);
/* replaces EmbreeTracer::removeGeometry (SRC/EmbreeTracer.cpp:220-260): returns the removed id, -1 if unknown;
 * the scene is re-committed like the reference does (:252) when any geometry is left. */
LSB200_API int lsb200_remove_geometry(lsb200_ctx *ctx, const char *name);
LSB200_API int lsb200_get_geometry_count(const lsb200_ctx *ctx);
LSB200_API int lsb200_get_geometry_id(const lsb200_ctx *ctx, const char *name);
/* replaces EmbreeTracer::updateGeometry(name, Affine3f, mesh) (SRC/EmbreeTracer.cpp:262-274) and
 * MeshTransformer::transformIntoBuffer (SRC/MeshTransformer.cpp:142-205).  verts: n_vertices records of
 * vert_stride_bytes, xyz float32 at offset 0 (pcl::PCLPointCloud2 blob).  elems: n_elements x 3 (or x 4)
 * uint32; NULL = topology unchanged since the last call (the reference re-copies it every frame,
 * SRC/MeshTransformer.cpp:486-552).  affine3x4 row-major [L | t].
 * Buffer lifetime: pageable host memory is staged by the CUDA runtime before the call returns and may be reused at
 * once.  PINNED host memory (lsb200_alloc_host, cudaHostAlloc) is copied by DMA on the context's stream: it must stay
 * untouched until lsb200_synchronize(ctx) -- or any call that returns results to the host (lsb200_trace_scene,
 * lsb200_fetch_points, lsb200_run_frames) -- has returned. */
LSB200_API int lsb200_update_geometry(lsb200_ctx *ctx, const char *name, const void *verts, int vert_stride_bytes,
                                      const uint32_t *elems, const float affine3x4[12]);
/* replaces updateGeometry(name, translation, eulerXYZ, mesh): A = T * Rz * Ry * Rx (SRC/MeshTransformer.cpp:467-477) */
LSB200_API int lsb200_update_geometry_euler(lsb200_ctx *ctx, const char *name, const void *verts,
                                            int vert_stride_bytes, const uint32_t *elems, const float translation[3],
                                            const float euler_xyz[3]);
/* vertices already resident on the device from an earlier update: only the transform changes (no H2D copy) */
LSB200_API int lsb200_update_transform(lsb200_ctx *ctx, const char *name, const float affine3x4[12]);
LSB200_API void lsb200_affine_from_components(const float translation[3], const float euler_xyz[3], float affine3x4[12]);

/* ---- commit: replaces EmbreeTracer::commitScene / rtcCommitScene (SRC/EmbreeTracer.cpp:290-295) ----------
 * transform kernel -> Morton codes -> radix sort -> Karras hierarchy -> bottom-up AABB fit (rebuild), or
 * transform -> leaf + AABB refit (refit).  Empty scene -> -1. */
LSB200_API int lsb200_commit_scene(lsb200_ctx *ctx, int mode);

/* ---- trace: replaces EmbreeTracer::traceScene (SRC/EmbreeTracer.cpp:297-367) ---------------------------
 * out_points: HOST buffer of capacity_pts records (32 B, or 16 B with LSB200_TRACE_XYZI16); *n_pts = hits.
 * Record (SRC/XYZIRBytes.cpp:24-40): x y z f32 | 4 zero bytes | intensity f32 = 64.0 | ring i32 | 8 zero bytes;
 * point = t * d_sensor.  Empty scene: *n_pts = 0 and -1 (SRC/OptixTracer.cpp:279-288).  Only the first *n_pts records of
 * out_points are meaningful; the rest of the buffer is unspecified after the call (small clouds are copied speculatively). */
LSB200_API int lsb200_trace_scene(lsb200_ctx *ctx, int sensor_index, uint32_t frame_index, void *out_points,
                                  size_t capacity_pts, uint32_t *n_pts, uint32_t flags);
/* same, results stay in device memory (lsb200_device_results) -- no synchronisation, no copies */
LSB200_API int lsb200_trace_device(lsb200_ctx *ctx, int sensor_index, uint32_t flags);
/* one launch for the sensors [first_sensor, first_sensor + n_sensors) (LSB200_FRAME_WORLD): the batch form of
 * lsb200_trace_device, used when many LiDAR devices share one scene */
LSB200_API int lsb200_trace_batch(lsb200_ctx *ctx, int first_sensor, int n_sensors, uint32_t flags);
/* per-ray results for parity tests: t_out[R], tri_out[R] host buffers */
LSB200_API int lsb200_trace_raw(lsb200_ctx *ctx, int sensor_index, float *t_out, int32_t *tri_out);
/* second return of every ray of the last trace of `sensor_index` (LSB200_OPT_RETURNS = 2): t2[R] (+inf = none) and the
 * triangle id tri2[R] (-1 = none); either pointer may be NULL.  Rays without a first hit have no second one. */
LSB200_API int lsb200_fetch_second_returns(lsb200_ctx *ctx, int sensor_index, float *t2_out, int32_t *tri2_out);
/* device pointers of the last trace of `sensor_index`: points, point count (uint32), t, triangle id */
LSB200_API int lsb200_device_results(lsb200_ctx *ctx, int sensor_index, void **d_points, uint32_t **d_count,
                                     float **d_t, int32_t **d_tri);
/* copies the device point buffer of sensor_index to the host (count first, then count records) */
LSB200_API int lsb200_fetch_points(lsb200_ctx *ctx, int sensor_index, void *out_points, size_t capacity_pts,
                                   uint32_t *n_pts, uint32_t flags);
/* batch form: the hit counts of n_sensors sensors in one copy, then every point buffer; outs[i] / caps[i] / n_pts[i]
 * belong to sensor first_sensor + i (host arrays; pinned memory makes the copies asynchronous until the final sync) */
LSB200_API int lsb200_fetch_batch(lsb200_ctx *ctx, int first_sensor, int n_sensors, void *const *outs, const size_t *caps,
                                  uint32_t *n_pts, uint32_t flags);
/* host-side expansion of a range image (LSB200_TRACE_RANGE_IMAGE) of `sensor_index` into point records: flags =
 * XYZI16 / RING_ZERO / WORLD_POINTS as for a trace; hits in ray-id order; *n_pts = hits */
LSB200_API int lsb200_expand_range_image(lsb200_ctx *ctx, int sensor_index, const float *t, void *out_points, size_t capacity_pts,
                                         uint32_t *n_pts, uint32_t flags);
/* per-ray intensity of the last trace (LSB200_OPT_INTENSITY_MODEL != 0): intensity_out[R] host buffer, 0 for a miss */
LSB200_API int lsb200_fetch_intensity(lsb200_ctx *ctx, int sensor_index, float *intensity_out);
LSB200_API int lsb200_synchronize(lsb200_ctx *ctx);
/* Pinned host memory for output clouds / vertex blobs, allocated on the NUMA node the context's GPU is attached to (the
 * calling thread is bound to that node's CPUs for the allocation and the first touch): with several GPUs in one box the
 * D2H streams then spread over the sockets' memory controllers instead of converging on one. */
LSB200_API int lsb200_alloc_host(lsb200_ctx *ctx, size_t bytes, void **out);
LSB200_API int lsb200_free_host(lsb200_ctx *ctx, void *ptr);
LSB200_API int lsb200_device_numa_node(const lsb200_ctx *ctx);
/* ---- pipelined frames: the throughput form of the per-frame loop of MeshProjector::traceMeshWrapper
 * (SRC/MeshProjector.cpp:446-471: updateGeometry for every mesh, commitScene, traceScene) over a batch of
 * trajectory frames.  For each frame: optional vertex re-upload, per-geometry affine, commit, one trace launch
 * for the sensors [first_sensor, first_sensor + n_sensors), then the hit counts and the point clouds go to the
 * caller's HOST buffers on a second stream while the next frame is already being built and traced (two device
 * output slots).  Results are identical to calling the per-frame entry points in sequence.  Pinned host buffers
 * are needed for the overlap (pageable ones still work, serialised by the runtime). */
typedef struct lsb200_frame {
    const float *affines;          /* [n_geometries][12], ascending geometry id; NULL = transforms unchanged */
    const void *const *verts;      /* [n_geometries] host vertex blobs to re-upload (stride as first given); NULL
                                      array or NULL entry = vertices already resident */
    void *const *out_points;       /* [n_sensors] host buffers of capacities[i] records */
    const size_t *capacities;      /* [n_sensors] */
    uint32_t *n_pts;               /* [n_sensors] out: hit counts */
} lsb200_frame;
LSB200_API int lsb200_run_frames(lsb200_ctx *ctx, const lsb200_frame *frames, int n_frames, int first_sensor,
                                 int n_sensors, int commit_mode, uint32_t flags);

/* ---- multi-GPU: one process, one context per GPU.  The reference runs one tracer per LiDAR device in one process
 * (SRC/mainwindow.cpp:256-274) and has no multi-GPU code; this is the entry a C++ host uses to drive a whole box.
 * lsb200_clone_scene: a new context on `cuda_device` holding a replica of src's scene (options, sensors, limits, every
 * geometry's vertices / indices copied device to device, transforms).
 * lsb200_run_batch: lsb200_run_frames on n_ctx contexts at once, one host thread each; LSB200_SHARD_FRAMES deals the
 * frames round-robin (context g: frames g, g + n_ctx, ...; give every frame its affines), LSB200_SHARD_SENSORS gives
 * every context every frame and a contiguous share of the sensors.  No data-path collective. */
LSB200_API int lsb200_clone_scene(lsb200_ctx *src, int cuda_device, lsb200_ctx **out);
LSB200_API int lsb200_run_batch(lsb200_ctx *const *ctxs, int n_ctx, const lsb200_frame *frames, int n_frames, int first_sensor,
                                int n_sensors, int commit_mode, uint32_t flags, int sharding);


/* ---- PointCloud2 header: replaces LidarDevice::initMessage (SRC/LidarDevice.cpp:94-115) ------------------ */
typedef struct lsb200_cloud_header {
    uint32_t seq;          /* = frame index */
    uint32_t height;       /* 1 */
    uint32_t width;        /* number of points */
    uint32_t point_step;   /* 32 */
    uint32_t row_step;     /* 0, as the reference leaves it (SRC/LidarDevice.cpp:112) */
    uint8_t is_bigendian;  /* 0 */
    uint8_t is_dense;      /* 1 */
    uint8_t n_fields;      /* 5: x y z intensity ring */
    uint8_t reserved;
} lsb200_cloud_header;
LSB200_API int lsb200_init_message(const lsb200_ctx *ctx, uint32_t frame_index, lsb200_cloud_header *out);

/* ---- diagnostics ----------------------------------------------------------------------------------------- */
typedef struct lsb200_stats {
    int64_t n_vertices, n_triangles, n_nodes, n_leaves;
    int32_t last_commit_was_rebuild;
    int32_t top_nodes;             /* nodes staged into shared memory */
    int32_t trace_grid, trace_block;
    /* milliseconds of the last commit / trace (CUDA events; valid with LSB200_OPT_STAGE_TIMING = 1) */
    float ms_transform, ms_morton, ms_sort, ms_hierarchy, ms_fit, ms_toplevel, ms_trace, ms_compact;
    /* launches issued since the context was created / since lsb200_reset_counters */
    int64_t kernel_launches;
    float scene_lo[3], scene_hi[3];
    float ms_tile_lists;           /* k_tile_entries alone (its time is part of ms_trace); 0 when the lists were reused / off */
    int32_t reserved0;
} lsb200_stats;
LSB200_API int lsb200_get_stats(lsb200_ctx *ctx, lsb200_stats *out);
LSB200_API int lsb200_reset_counters(lsb200_ctx *ctx);
/* totals of the last LSB200_OPT_TRACE_STATS launch: out[0..n) = rays, nodes visited, leaves visited, triangles tested,
 * 32-ray units with one octant, units with mixed octants, deepest stack, hits, culled pops, origin-chain box tests */
LSB200_API int lsb200_get_trace_counters(lsb200_ctx *ctx, uint64_t *out, int n);
/* copies the BVH to the host for structural tests: nodes (n_nodes x 16 floats), leaf-ordered triangle ids.  Only the
 * records a child reference leads to are defined: the fused tree build does not write the records of nodes whose whole
 * range fits a leaf (their parents reference them as leaves); walk from the root (capi.reachable_nodes in the tests). */
LSB200_API int lsb200_debug_copy_bvh(lsb200_ctx *ctx, float *nodes16, int64_t node_capacity, int32_t *leaf_tri_ids,
                                     int64_t tri_capacity, int32_t *root);
/* device sort self-test entry (parity tests of the radix sort): sorts n (key,value) pairs held in HOST arrays */
LSB200_API int lsb200_debug_sort_pairs(lsb200_ctx *ctx, uint32_t *keys, uint32_t *vals, int64_t n);
/* transformed vertices (n x 3 f32, concatenation order) of the last commit, for bit-exact transform tests */
LSB200_API int lsb200_debug_copy_vertices(lsb200_ctx *ctx, float *xyz, int64_t capacity_vertices);

/* ---- allocation-free sort --------------------------------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same device radix sort
 * as lsb200_debug_sort_pairs (the sort of the BVH build), on DEVICE arrays the caller owns: d_keys / d_vals hold the n
 * pairs on entry and the pairs in ascending key order (stable) on return; d_keys_tmp / d_vals_tmp are the caller's
 * ping-pong buffers of n words each (contents destroyed).  The sort's scratch (digit histograms, tile tickets, look-back
 * status words) lives in a file-scope __device__ array of the library sized for LSB200_SORT_STATIC_MAX_PAIRS pairs
 * (2.0 MiB, held by every GPU that loads the library until the process exits): the call issues no cudaMalloc / cudaFree
 * and no allocating library routine.  n > LSB200_SORT_STATIC_MAX_PAIRS returns LSB200_ERR_CAPACITY and
 * lsb200_last_error names the limit.  The work is enqueued on the context's stream and NOT synchronised
 * (lsb200_synchronize, or order your own stream work behind it).  The first call on a device resolves the array's
 * device address (cudaGetSymbolAddress, once per device; n = 0 does only that) -- make it before a capture begins.
 * Every call on a device shares the one array (it is cleared at the start of each call): calls that overlap on the GPU
 * -- two contexts, or two streams of one device -- race; serialise them. */
#define LSB200_SORT_STATIC_MAX_PAIRS (1 << 22)
LSB200_API int lsb200_sort_pairs_static(lsb200_ctx *ctx, uint32_t *d_keys, uint32_t *d_vals, uint32_t *d_keys_tmp,
                                        uint32_t *d_vals_tmp, int64_t n);

#ifdef __cplusplus
}
#endif
#endif /* LSB200_H */
