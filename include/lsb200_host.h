/*
 * lsb200_host.h -- C entry points of the C++ host layer (liblsb200_host.so): the reference-shaped classes
 * lidarshooter::LidarDevice / B200Tracer / AffineMesh / Trajectory and the mesh readers, flattened to C so that
 * the Python test-suite (ctypes) can drive them.  Each function names the reference interface it stands for
 * (SRC = ros_ws/src/lidarshooter/src of the reference tree).  Handles are opaque; no exceptions cross.
 */
#ifndef LSB200_HOST_H
#define LSB200_HOST_H

#include <stddef.h>
#include <stdint.h>

#include "lsb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lsbh_device lsbh_device;
typedef struct lsbh_mesh lsbh_mesh;
typedef struct lsbh_tracer lsbh_tracer;

/* LidarDevice::create(config[, sensorUid]) (SRC/LidarDevice.cpp:26-40, 482-633); NULL + err text on failure */
LSB200_API lsbh_device *lsbh_device_create(const char *config_path, const char *sensor_uid, char *err, size_t err_len);
LSB200_API lsbh_device *lsbh_device_create_synthetic(const float *vertical, int n_vertical, float begin, float end,
                                                     unsigned count, const float q_wxyz[4], const float t_xyz[3]);
LSB200_API void lsbh_device_destroy(lsbh_device *d);
LSB200_API const char *lsbh_device_sensor_uid(const lsbh_device *d);            /* getSensorUid */
LSB200_API unsigned lsbh_device_total_rays(const lsbh_device *d);               /* getTotalRays */
LSB200_API unsigned lsbh_device_total_channels(const lsbh_device *d);           /* getTotalChannels */
LSB200_API unsigned lsbh_device_scan_ray_count(const lsbh_device *d);           /* getScanRayCount */
LSB200_API int lsbh_device_transform(const lsbh_device *d, float q_wxyz[4], float t_xyz[3], float R[9], float Rinv[9]);
LSB200_API int lsbh_device_channels(const lsbh_device *d, float *vertical, int capacity, float *begin, float *end, float *step);
/* nextRay: writes the direction, returns 1 when the cursor wrapped (SRC/LidarDevice.cpp:160-192, 824-845) */
LSB200_API int lsbh_device_next_ray(lsbh_device *d, float dir[3], int *vertical_index, int *horizontal_index);
LSB200_API void lsbh_device_reset(lsbh_device *d);
LSB200_API int lsbh_device_all_rays(lsbh_device *d, float *dirs_xyz, size_t capacity_rays);
/* originToSensor / originToSensorInverse (SRC/LidarDevice.cpp:383-401) */
LSB200_API void lsbh_device_origin_to_sensor(const lsbh_device *d, float p[3], int inverse);
/* initMessage (SRC/LidarDevice.cpp:94-115): header fields + frame id + field table "name:offset:datatype:count;" */
LSB200_API int lsbh_device_init_message(lsbh_device *d, int frame_index, lsb200_cloud_header *out, char *frame_id,
                                        size_t frame_id_len, char *fields, size_t fields_len);

/* pcl::io::loadPolygonFileSTL and the OBJ twin (SRC/mainwindow.cpp:146; SRC/test/EmbreeTracer_test.cpp:40) */
LSB200_API lsbh_mesh *lsbh_mesh_load(const char *path);
LSB200_API lsbh_mesh *lsbh_mesh_from_arrays(const float *xyz, size_t n_verts, const uint32_t *elems, size_t n_elems, int verts_per_elem);
LSB200_API void lsbh_mesh_destroy(lsbh_mesh *m);
LSB200_API int lsbh_mesh_counts(const lsbh_mesh *m, int *n_vertices, int *n_polygons, int *point_step);
LSB200_API int lsbh_mesh_copy(const lsbh_mesh *m, float *xyz, uint32_t *elems, int verts_per_elem);

/* B200Tracer : ITracer (SRC/ITracer.hpp:50-94) */
LSB200_API lsbh_tracer *lsbh_tracer_create(lsbh_device *d, int cuda_device, char *err, size_t err_len);
LSB200_API void lsbh_tracer_destroy(lsbh_tracer *t);
LSB200_API int lsbh_tracer_add_geometry(lsbh_tracer *t, const char *name, int rtc_geometry_type, int n_vertices, int n_elements);
LSB200_API int lsbh_tracer_remove_geometry(lsbh_tracer *t, const char *name);
LSB200_API long lsbh_tracer_geometry_count(const lsbh_tracer *t);
LSB200_API int lsbh_tracer_geometry_id(const lsbh_tracer *t, const char *name);
/* updateGeometry(name, Affine3f, mesh) / updateGeometry(name, translation, rotation, mesh); -100 - code = exception */
LSB200_API int lsbh_tracer_update_geometry(lsbh_tracer *t, const char *name, lsbh_mesh *mesh, const float affine3x4[12]);
LSB200_API int lsbh_tracer_update_geometry_euler(lsbh_tracer *t, const char *name, lsbh_mesh *mesh, const float translation[3],
                                                 const float rotation[3]);
LSB200_API int lsbh_tracer_commit_scene(lsbh_tracer *t);
LSB200_API int lsbh_tracer_trace_scene(lsbh_tracer *t, uint32_t frame_index);
/* ring_zero != 0 (the tracer's default): ring = 0 like the Embree path (SRC/EmbreeTracer.cpp:333-334); 0: ring = channel */
LSB200_API int lsbh_tracer_set_flags(lsbh_tracer *t, int ordered, int ring_zero, int world_points);
LSB200_API int lsbh_tracer_set_intensity(lsbh_tracer *t, int model, int apply_config_filter);
LSB200_API int lsbh_tracer_set_watertight(lsbh_tracer *t, int on);
/* lines the tracer wrote to the spdlog logger it was created with (SRC/EmbreeTracer.hpp:51: third create() argument) */
LSB200_API int lsbh_tracer_log_count(lsbh_tracer *t);
LSB200_API int lsbh_tracer_log_line(lsbh_tracer *t, int index, int *level, char *buf, size_t capacity);
/* getTraceCloud(): header + payload of the cloud the tracer rewrites in place */
LSB200_API int lsbh_tracer_cloud(lsbh_tracer *t, lsb200_cloud_header *hdr, const uint8_t **data, size_t *n_bytes);
LSB200_API int lsbh_tracer_trace_raw(lsbh_tracer *t, float *t_out, int32_t *tri_out);
/* detection_radius / horizontal_angle_range of the device config (config/hesai-pandar-XT-32-lidar_0000.json:14-18) and
 * the switch that applies them through lsb200_set_sensor_limits (off = the reference's behaviour) */
LSB200_API int lsbh_device_limits(const lsbh_device *d, float *radius, float *az_lo, float *az_hi);
LSB200_API int lsbh_tracer_set_sensor_limits(lsbh_tracer *t, int on);
/* ROS-1 wire bytes of the traced sensor_msgs/PointCloud2 (what SRC/MeshProjector.cpp:380-390 publishes through roscpp),
 * host/CloudSerializer.hpp.  stamp < 0 keeps the cloud's own stamp.  Returns the message length (also when it exceeds
 * `capacity`, in which case nothing is written), < 0 on error. */
LSB200_API long lsbh_tracer_serialize_cloud(lsbh_tracer *t, double stamp, int fix_row_step, uint8_t *out, size_t capacity);
/* parse + serialise again (round-trip check of the reader); fills hdr / stamp from the parsed message */
LSB200_API long lsbh_cloud_reserialize(const uint8_t *msg, size_t len, uint8_t *out, size_t capacity, lsb200_cloud_header *hdr,
                                       double *stamp);
/* appends "uint32 length + message" to a file */
LSB200_API int lsbh_tracer_append_cloud_to_file(lsbh_tracer *t, const char *path, int fix_row_step);

/* AffineMesh + Trajectory (SRC/AffineMesh.cpp:108-128, 275-284; config/trajectory.json): poses[k] = lin xyz, ang xyz
 * after tick k.  Returns the number of ticks (may exceed capacity; only `capacity` are written), < 0 on error. */
LSB200_API int lsbh_trajectory_poses(const char *trajectory_json, float period, float *poses6, int capacity);

#ifdef __cplusplus
}
#endif
#endif
