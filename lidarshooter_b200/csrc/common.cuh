// common.cuh -- shared device/host helpers of liblsb200 (B200 / sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

namespace lsb {

constexpr int kSmCount = 148;           // B200: 2 dies x 74 SMs
constexpr int kMaxLeaf = 8;             // leaf reference encodes count-1 in 3 bits
constexpr int kLeafShift = 28;          // leaf ref: ~(first | (count-1) << 28)
constexpr int kMaxSensors = 4096;

// One triangle in leaf order: 12 floats = 3 x float4 (48 B), the Embree-ordered Moeller-Trumbore operands
//   q0 = (v0.x, v0.y, v0.z, e1.x)   q1 = (e1.y, e1.z, e2.x, e2.y)   q2 = (e2.z, Ng.x, Ng.y, Ng.z)
// with e1 = v0 - v1, e2 = v2 - v0, Ng = cross(e2, e1)  (oracle/lsb_oracle.c: tri_setup, lso_intersect).
//
// One BVH node: 16 floats = 4 x float4 (64 B, Aila-Laine layout), both child boxes and both child references
//   q0 = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)   q1 = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
//   q2 = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)   q3 = (child0, child1, first, last) as int bits
// child >= 0: node index; child < 0: leaf reference ~(first | (count-1) << 28) into the leaf-ordered triangles.

struct GeomDesc {            // one per attached geometry, ascending geometry id
    const uint8_t *verts;    // raw vertex blob as uploaded (xyz f32 at offset 0 of each record)
    const uint32_t *elems;   // 3 (or 4) indices per element
    float A[12];             // row-major 3x4 affine [L | t]
    int stride;              // bytes per vertex record
    int verts_per_elem;      // 3 or 4
    int vert_off, n_verts;   // position in the concatenated vertex array
    int tri_off, n_tris;     // position in the concatenated (global) triangle numbering
};

struct SensorDev {           // one per LiDAR device
    const float *sin_t, *cos_t;   // [V]  per channel
    const float *sin_p, *cos_p;   // [H]  per azimuth step
    float R[9];                   // sensor -> world rotation (row-major)
    float Rinv[9];
    float T[3];
    int V, H, n_rays;
    int unit_base;                // first 32-ray work unit of this sensor in a batch
    int tiles_h;                  // work units per row of tiles: ceil(H / tile width)
    int sup_base, sup_w;          // first supertile (4 x 4 tiles) of this sensor, supertiles per row
    float max_range;              // hits beyond this range are dropped (+inf: none) -- lsb200_set_sensor_limits
    const uint8_t *az_mask;       // [H] 1 = azimuth step fired; nullptr = all
    float *t_out;                 // [n_rays]
    int32_t *tri_out;             // [n_rays]
    uint8_t *points;              // [n_rays] records
    uint32_t *count;              // hit count
    uint32_t *block_counts;       // ordered compaction scratch
    float *intensity;             // [n_rays] per-ray intensity (k_intensity; only with an intensity model / filter)
    float min_intensity;          // returns below it are dropped (0: none) -- lsb200_set_sensor_intensity_filter
    float *t2_out;                // [n_rays] second return (LSB200_OPT_RETURNS = 2): the next surface behind the first hit
    int32_t *tri2_out;            // [n_rays]
};

__host__ __device__ inline uint32_t f2ord(float f)
{
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ inline float ord2f(uint32_t u)
{
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemask_lt()
{
    unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m;
}

// Binary search of the geometry that owns global triangle g / global vertex j (descs sorted by offset).
__device__ __forceinline__ int find_geom_by_tri(const GeomDesc *d, int n, int g)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (d[mid].tri_off <= g) lo = mid; else hi = mid - 1; }
    return lo;
}
__device__ __forceinline__ int find_geom_by_vert(const GeomDesc *d, int n, int j)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (d[mid].vert_off <= j) lo = mid; else hi = mid - 1; }
    return lo;
}

}  // namespace lsb
