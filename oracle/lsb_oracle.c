/*
 * lsb_oracle.c -- CPU ORACLE for the lidarshooter trace path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
 * load this library.  The product (lidarshooter_b200/, liblsb200.so) never links, imports or
 * calls anything in oracle/.
 *
 * What it is: a plain-C restatement of the arithmetic of the reference's Embree CPU path
 * (/root/reference, SRC = ros_ws/src/lidarshooter/src):
 *
 *   lso_quat_to_rotation / lso_inverse3      SRC/LidarDevice.cpp:808-813 (Eigen Quaternionf::toRotationMatrix,
 *                                            Matrix3f::inverse; quaternion is NOT normalised)
 *   lso_affine_from_components               SRC/MeshTransformer.cpp:467-477  (T * Rz * Ry * Rx)
 *   lso_transform_vertices                   SRC/MeshTransformer.cpp:173-194 + SRC/LidarDevice.cpp:383-391
 *   lso_ray_tables / lso_ray_dirs            SRC/LidarDevice.cpp:294-342, 598-617, 824-845 (ray id = channel*H + az)
 *   lso_intersect (Moeller-Trumbore)         Embree v3.13.4 kernels/geometry/triangle_intersector_moeller.h
 *                                            (third-party, NOT in /root/reference; pinned by
 *                                            .devcontainer/Dockerfile:24-27; call site SRC/EmbreeTracer.cpp:472-480)
 *   lso_trace_brute / lso_trace_bvh          SRC/EmbreeTracer.cpp:297-367 (closest hit per ray, tnear 0, tfar inf)
 *   lso_pack_points                          SRC/XYZIRBytes.cpp:24-40, SRC/EmbreeTracer.cpp:331-335
 *
 * PARITY STATUS: the reference's own tests pin only point COUNTS at this boundary (1668 / 1781 / 0,
 * SRC/test/EmbreeTracer_test.cpp:134, SRC/test/OptixTracer_test.cpp:168,307); tests/test_oracle.py checks
 * those.  Per-ray hit mask / triangle id / range are "parity unpinned" against Embree itself: Embree is a
 * closed binary dependency that is absent here, its t = T*rcp(|den|) uses the hardware reciprocal estimate and
 * its dot/cross use FMA on AVX2+ hosts, so no CPU restatement can be bit-identical to it.  Deviations chosen
 * here, all deterministic: non-fused IEEE fp32 everywhere (compile with -ffp-contract=off), t = T/|den| with a
 * true division, closest hit = lexicographic minimum of (t, global triangle id).
 *
 * Everything is IEEE-754 binary32 with round-to-nearest and no contraction, so the CUDA kernels (which
 * use __fmul_rn/__fadd_rn/__fsub_rn/__fdiv_rn in the same order) can match bit for bit.
 */
#define _GNU_SOURCE 1
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#if defined(__GNUC__)
#define LSO_API __attribute__((visibility("default")))
#else
#define LSO_API
#endif

/* ------------------------------------------------------------------------------------------------
 * small fixed-size linear algebra, Eigen 3.4 evaluation order (best effort; Eigen is absent here).
 * Eigen's unrolled 3-term reduction is a0 + (a1 + a2) (redux_novec_unroller halves the range).
 * ---------------------------------------------------------------------------------------------- */
static inline float dot3_eigen(float a0, float b0, float a1, float b1, float a2, float b2)
{
    float p0 = a0 * b0, p1 = a1 * b1, p2 = a2 * b2;
    float s12 = p1 + p2;
    return p0 + s12;
}

/* Row-major 3x3 times vector, Eigen coefficient order. */
static inline void mat3_mul_vec(const float *M, const float *p, float *o)
{
    float x = dot3_eigen(M[0], p[0], M[1], p[1], M[2], p[2]);
    float y = dot3_eigen(M[3], p[0], M[4], p[1], M[5], p[2]);
    float z = dot3_eigen(M[6], p[0], M[7], p[1], M[8], p[2]);
    o[0] = x; o[1] = y; o[2] = z;
}

static void mat3_mul_mat3(const float *A, const float *B, float *C)
{
    float T[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            T[3 * i + j] = dot3_eigen(A[3 * i + 0], B[0 + j], A[3 * i + 1], B[3 + j], A[3 * i + 2], B[6 + j]);
    memcpy(C, T, sizeof(T));
}

/* Eigen::Quaternionf::toRotationMatrix(), q = (w,x,y,z), no normalisation (SRC/LidarDevice.cpp:812). */
LSO_API void lso_quat_to_rotation(const float q[4], float R[9])
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

/* Eigen::Matrix3f::inverse() (SRC/LidarDevice.cpp:813), restated from Eigen 3.4 LU/InverseImpl.h
 * compute_inverse_size3_helper: cofactor(i,j) = m(i1,j1)*m(i2,j2) - m(i1,j2)*m(i2,j1) with i1=(i+1)%3 ...,
 * det = sum_i cofactor(i,0)*m(i,0), inverse(j,i) = cofactor(i,j) / det (as a multiply by 1/det). */
static inline float cof3(const float *M, int i, int j)
{
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return M[3 * i1 + j1] * M[3 * i2 + j2] - M[3 * i1 + j2] * M[3 * i2 + j1];
}
LSO_API void lso_inverse3(const float M[9], float Minv[9])
{
    const float c0 = cof3(M, 0, 0), c1 = cof3(M, 1, 0), c2 = cof3(M, 2, 0);
    const float det = dot3_eigen(c0, M[0], c1, M[3], c2, M[6]);
    const float invdet = 1.0f / det;
    Minv[0] = c0 * invdet; Minv[1] = c1 * invdet; Minv[2] = c2 * invdet;
    for (int j = 1; j < 3; ++j)
        for (int i = 0; i < 3; ++i)
            Minv[3 * j + i] = cof3(M, i, j) * invdet;
}

/* Eigen::AngleAxisf(angle, UnitX/Y/Z).toRotationMatrix() products: A = Translation(t)*Rz*Ry*Rx
 * (SRC/MeshTransformer.cpp:467-477).  A is row-major 3x4: [L | t]. */
LSO_API void lso_affine_from_components(const float t[3], const float r[3], float A[12])
{
    const float cx = cosf(r[0]), sx = sinf(r[0]);
    const float cy = cosf(r[1]), sy = sinf(r[1]);
    const float cz = cosf(r[2]), sz = sinf(r[2]);
    const float Rx[9] = {1, 0, 0, 0, cx, -sx, 0, sx, cx};
    const float Ry[9] = {cy, 0, sy, 0, 1, 0, -sy, 0, cy};
    const float Rz[9] = {cz, -sz, 0, sz, cz, 0, 0, 0, 1};
    float Rzy[9], L[9];
    mat3_mul_mat3(Rz, Ry, Rzy);
    mat3_mul_mat3(Rzy, Rx, L);
    for (int i = 0; i < 3; ++i) {
        A[4 * i + 0] = L[3 * i + 0];
        A[4 * i + 1] = L[3 * i + 1];
        A[4 * i + 2] = L[3 * i + 2];
        A[4 * i + 3] = t[i];
    }
}

/* Per-vertex: p1 = t + L*p  (Eigen Transform*vector: res = translation; res += linear*p), then
 * originToSensor: p2 = Rinv * (p1 - T) (SRC/LidarDevice.cpp:383-391).
 * mode 0 = sensor frame (reference arithmetic), mode 1 = world frame (stop after p1).
 * verts: byte blob, xyz float32 at offset 0 of each `stride`-byte record (pcl::PCLPointCloud2 layout). */
LSO_API void lso_transform_vertices(const uint8_t *verts, int64_t n, int stride, const float A[12],
                                    const float Rinv[9], const float T[3], int mode, float *out_xyz)
{
    const float L[9] = {A[0], A[1], A[2], A[4], A[5], A[6], A[8], A[9], A[10]};
#pragma omp parallel for schedule(static) if (n > 65536)
    for (int64_t j = 0; j < n; ++j) {
        float p[3], q[3];
        memcpy(p, verts + (size_t)j * (size_t)stride, 12);
        mat3_mul_vec(L, p, q);
        q[0] = A[3] + q[0]; q[1] = A[7] + q[1]; q[2] = A[11] + q[2];
        if (mode == 0) {
            float d[3] = {q[0] - T[0], q[1] - T[1], q[2] - T[2]};
            mat3_mul_vec(Rinv, d, q);
        }
        out_xyz[3 * j + 0] = q[0]; out_xyz[3 * j + 1] = q[1]; out_xyz[3 * j + 2] = q[2];
    }
}

/* ------------------------------------------------------------------------------------------------
 * Ray generation (SRC/LidarDevice.cpp:294-342).  theta/phi are computed in double from the float
 * angle and narrowed to float BEFORE the float sin/cos, exactly like the reference.
 * ---------------------------------------------------------------------------------------------- */
LSO_API float lso_azimuth_step(float begin, float end, unsigned count)
{
    /* SRC/LidarDevice.cpp:611: float / float(unsigned) */
    return (end - begin) / (float)(count - 1u);
}

LSO_API void lso_ray_tables(const float *elev_deg, int V, float az_begin, float az_step, int H,
                            float *sin_t, float *cos_t, float *sin_p, float *cos_p)
{
    for (int v = 0; v < V; ++v) {
        float theta = (float)((90.0 - (double)elev_deg[v]) * M_PI / 180.0);
        sin_t[v] = sinf(theta);
        cos_t[v] = cosf(theta);
    }
    for (int h = 0; h < H; ++h) {
        float prePhi = az_begin + az_step * (float)h;
        float phi = (float)((double)prePhi * M_PI / 180.0);
        sin_p[h] = sinf(phi);
        cos_p[h] = cosf(phi);
    }
}

LSO_API void lso_ray_dirs(const float *sin_t, const float *cos_t, int V, const float *sin_p, const float *cos_p,
                          int H, float *dirs)
{
#pragma omp parallel for schedule(static) if ((int64_t)V * H > 65536)
    for (int v = 0; v < V; ++v)
        for (int h = 0; h < H; ++h) {
            float *d = dirs + 3 * ((size_t)v * H + h);
            d[0] = sin_t[v] * cos_p[h];
            d[1] = sin_t[v] * sin_p[h];
            d[2] = cos_t[v];
        }
}

/* World-frame mode: d_w = R * d_s, row dot products evaluated left to right. */
LSO_API void lso_rotate_dirs(const float R[9], const float *dirs_s, int64_t n, float *dirs_w)
{
#pragma omp parallel for schedule(static) if (n > 65536)
    for (int64_t i = 0; i < n; ++i) {
        const float *d = dirs_s + 3 * i;
        float x = (R[0] * d[0] + R[1] * d[1]) + R[2] * d[2];
        float y = (R[3] * d[0] + R[4] * d[1]) + R[5] * d[2];
        float z = (R[6] * d[0] + R[7] * d[1]) + R[8] * d[2];
        dirs_w[3 * i + 0] = x; dirs_w[3 * i + 1] = y; dirs_w[3 * i + 2] = z;
    }
}

/* ------------------------------------------------------------------------------------------------
 * Embree-ordered Moeller-Trumbore (triangle stored as v0, e1 = v0-v1, e2 = v2-v0, Ng = cross(e2,e1)).
 * Returns 1 and *t_out when the ray (org, dir, tnear = 0, tfar = +inf) hits.  No back-face culling.
 * ---------------------------------------------------------------------------------------------- */
typedef struct { float v0[3], e1[3], e2[3]; float v1[3], v2[3]; /* raw vertices: watertight mode */ } lso_tri;

static inline void cross3(const float *a, const float *b, float *o)
{
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
static inline float dot3(const float *a, const float *b)
{
    return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2];
}

static inline void tri_setup(const float *a, const float *b, const float *c, lso_tri *t)
{
    for (int k = 0; k < 3; ++k) {
        t->v0[k] = a[k];
        t->e1[k] = a[k] - b[k];
        t->e2[k] = c[k] - a[k];
        t->v1[k] = b[k];
        t->v2[k] = c[k];
    }
}

static inline int lso_intersect(const lso_tri *tri, const float *org, const float *dir, float *t_out)
{
    float Ng[3], C[3], R[3];
    cross3(tri->e2, tri->e1, Ng);
    C[0] = tri->v0[0] - org[0]; C[1] = tri->v0[1] - org[1]; C[2] = tri->v0[2] - org[2];
    cross3(C, dir, R);
    const float den = dot3(Ng, dir);
    const float absDen = fabsf(den);
    const int neg = signbit(den) != 0;
    float U = dot3(R, tri->e2); if (neg) U = -U;
    float V = dot3(R, tri->e1); if (neg) V = -V;
    if (!(den != 0.0f)) return 0;
    if (!(U >= 0.0f)) return 0;
    if (!(V >= 0.0f)) return 0;
    if (!(U + V <= absDen)) return 0;
    float T = dot3(Ng, C); if (neg) T = -T;
    if (!(absDen * 0.0f < T)) return 0;      /* tnear = 0, strict */
    const float t = T / absDen;
    if (!(t <= FLT_MAX)) return 0;           /* tfar = +inf; rejects inf/NaN */
    *t_out = t;
    return 1;
}

/* ------------------------------------------------------------------------------------------------
 * Optional WATERTIGHT mode (not the reference's intersector; north_star asks for it): Woop, Benthin, Wald 2013,
 * "Watertight Ray/Triangle Intersection", on the triangle's three original vertices.  Ray-space: dominant
 * axis kz, shear (Sx, Sy), scale Sz; edge functions U V W on the sheared vertices, recomputed with exact
 * double products when one of them is 0; no back-face culling; tnear = 0 (strict), tfar = +inf.
 * Same operation order as lidarshooter_b200/csrc/trace.cuh: intersect_tri_wt (bit-exact parity in this mode).
 * lso_set_watertight(1) switches lso_trace_brute / lso_trace_bvh to it.
 * ---------------------------------------------------------------------------------------------- */
static int g_watertight = 0;
LSO_API void lso_set_watertight(int on) { g_watertight = on ? 1 : 0; }
LSO_API int lso_get_watertight(void) { return g_watertight; }

static inline int lso_intersect_wt(const lso_tri *tri, const float *org, const float *dir, float *t_out)
{
    const float ax = fabsf(dir[0]), ay = fabsf(dir[1]), az = fabsf(dir[2]);
    int kz = (ax > ay) ? (ax > az ? 0 : 2) : (ay > az ? 1 : 2);
    int kx = kz == 2 ? 0 : kz + 1, ky = kx == 2 ? 0 : kx + 1;
    const float dkz = dir[kz];
    if (dkz < 0.0f) { const int tmp = kx; kx = ky; ky = tmp; }
    const float Sx = dir[kx] / dkz, Sy = dir[ky] / dkz, Sz = 1.0f / dkz;
    float A[3], B[3], C[3];
    for (int k = 0; k < 3; ++k) { A[k] = tri->v0[k] - org[k]; B[k] = tri->v1[k] - org[k]; C[k] = tri->v2[k] - org[k]; }
    const float Ax = A[kx] - Sx * A[kz], Ay = A[ky] - Sy * A[kz];
    const float Bx = B[kx] - Sx * B[kz], By = B[ky] - Sy * B[kz];
    const float Cx = C[kx] - Sx * C[kz], Cy = C[ky] - Sy * C[kz];
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return 0;
    const float det = (U + V) + W;
    if (!(det != 0.0f)) return 0;
    const float Az = Sz * A[kz], Bz = Sz * B[kz], Cz = Sz * C[kz];
    const float T = (U * Az + V * Bz) + W * Cz;
    const float t = T / det;
    if (!(t > 0.0f)) return 0;
    if (!(t <= FLT_MAX)) return 0;
    *t_out = t;
    return 1;
}

static inline int lso_intersect_any(const lso_tri *tri, const float *org, const float *dir, float *t_out)
{
    return g_watertight ? lso_intersect_wt(tri, org, dir, t_out) : lso_intersect(tri, org, dir, t_out);
}

/* Intensity models of include/lsb200.h (LSB200_OPT_INTENSITY_MODEL; semantics defined there -- the reference writes the
 * constant 64.0, SRC/EmbreeTracer.cpp:333, and never applies its config's minimum_intensity): model 0 = 64.0,
 * model 1 = 255 |Ng . d| / (|Ng| |d|) with Ng = cross(e2, e1) of the hit triangle.  Returns below min_intensity become
 * misses (t = +inf, tri = -1, intensity 0).  Same fp32 operation order as trace.cuh: k_intensity. */
LSO_API void lso_intensity(const float *verts, const uint32_t *tris, const float *dirs, int64_t nR, int model,
                           float min_intensity, float *t, int32_t *tri, float *intensity_out)
{
    for (int64_t r = 0; r < nR; ++r) {
        float I = 0.0f;
        if (tri[r] >= 0) {
            I = 64.0f;
            if (model == 1) {
                lso_tri T;
                const uint32_t *tv = tris + 3 * (size_t)tri[r];
                tri_setup(verts + 3 * (size_t)tv[0], verts + 3 * (size_t)tv[1], verts + 3 * (size_t)tv[2], &T);
                float Ng[3];
                cross3(T.e2, T.e1, Ng);
                const float *d = dirs + 3 * r;
                const float den = dot3(Ng, d), nn = dot3(Ng, Ng), dd = dot3(d, d);
                const float len = sqrtf(nn * dd);
                I = len > 0.0f ? (255.0f * fabsf(den)) / len : 0.0f;
            }
            if (I < min_intensity) { t[r] = INFINITY; tri[r] = -1; I = 0.0f; }
        }
        intensity_out[r] = I;
    }
}

/* Single-triangle entry for unit tests. */
LSO_API int lso_intersect_one(const float a[3], const float b[3], const float c[3], const float org[3],
                              const float dir[3], float *t_out)
{
    lso_tri tri; tri_setup(a, b, c, &tri);
    return lso_intersect_any(&tri, org, dir, t_out);
}

/* Brute force O(R*T) closest hit; misses -> t = +inf, tri = -1.  dirs: nR x 3.  orgs: per-ray origins
 * when org_stride == 3, a single shared origin when org_stride == 0. */
LSO_API void lso_trace_brute(const float *verts, const uint32_t *tris, int64_t nT, const float *orgs, int org_stride,
                             const float *dirs, int64_t nR, float *t_out, int32_t *tri_out)
{
    lso_tri *T = (lso_tri *)malloc(sizeof(lso_tri) * (size_t)(nT > 0 ? nT : 1));
    for (int64_t i = 0; i < nT; ++i)
        tri_setup(verts + 3 * (size_t)tris[3 * i], verts + 3 * (size_t)tris[3 * i + 1],
                  verts + 3 * (size_t)tris[3 * i + 2], &T[i]);
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t r = 0; r < nR; ++r) {
        const float *o = orgs + (size_t)org_stride * r;
        const float *d = dirs + 3 * r;
        float best = INFINITY; int32_t bi = -1;
        for (int64_t i = 0; i < nT; ++i) {
            float t;
            if (lso_intersect_any(&T[i], o, d, &t) && t < best) { best = t; bi = (int32_t)i; }
            /* ascending i with strict '<' == lexicographic min of (t, i) */
        }
        t_out[r] = best; tri_out[r] = bi;
    }
    free(T);
}

/* ------------------------------------------------------------------------------------------------
 * CPU BVH (binned SAH, binary, <= 4 triangles per leaf): the stand-in for rtcCommitScene +
 * rtcIntersect (SRC/EmbreeTracer.cpp:290-295, 472-480) at sizes brute force cannot reach, and the CPU
 * baseline that bench.py times.  Boxes are padded so traversal can never reject a triangle that
 * lso_intersect would accept; results are identical to lso_trace_brute (tests/test_oracle.py).
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    float lo[3]; int32_t left;    /* internal: index of left child (right = left+1); leaf: first triangle slot */
    float hi[3]; int32_t count;   /* 0 = internal, >0 = leaf triangle count */
} lso_node;

typedef struct lso_bvh {
    lso_node *nodes; int64_t n_nodes;
    lso_tri *tris;           /* in leaf order */
    int32_t *tri_ids;        /* leaf order -> global triangle id */
    int64_t n_tris;
    int64_t node_cap;
} lso_bvh;

typedef struct { float lo[3], hi[3], c[3]; } prim_box;

#define LSO_BINS 16
#define LSO_LEAF 4

static inline float fminf3(float a, float b) { return a < b ? a : b; }
static inline float fmaxf3(float a, float b) { return a > b ? a : b; }

static inline float half_area(const float *lo, const float *hi)
{
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

typedef struct {
    lso_bvh *bvh;
    prim_box *pb;
    int32_t *idx;
    int64_t next_node;   /* atomic */
} build_ctx;

static int64_t alloc_nodes(build_ctx *bc, int n)
{
    int64_t r;
#pragma omp atomic capture
    { r = bc->next_node; bc->next_node += n; }
    return r;
}

static void build_rec(build_ctx *bc, int64_t node_i, int64_t first, int64_t count, int depth)
{
    lso_node *node = &bc->bvh->nodes[node_i];
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int64_t i = first; i < first + count; ++i) {
        const prim_box *p = &bc->pb[bc->idx[i]];
        for (int k = 0; k < 3; ++k) {
            lo[k] = fminf3(lo[k], p->lo[k]); hi[k] = fmaxf3(hi[k], p->hi[k]);
            clo[k] = fminf3(clo[k], p->c[k]); chi[k] = fmaxf3(chi[k], p->c[k]);
        }
    }
    for (int k = 0; k < 3; ++k) { node->lo[k] = lo[k]; node->hi[k] = hi[k]; }
    if (count <= LSO_LEAF) { node->left = (int32_t)first; node->count = (int32_t)count; return; }

    /* binned SAH over the widest-centroid axis first, all three axes evaluated */
    int best_axis = -1, best_bin = -1; float best_cost = INFINITY;
    for (int ax = 0; ax < 3; ++ax) {
        float ext = chi[ax] - clo[ax];
        if (!(ext > 0.0f)) continue;
        float scale = (float)LSO_BINS / ext;
        int cnt[LSO_BINS]; float blo[LSO_BINS][3], bhi[LSO_BINS][3];
        for (int b = 0; b < LSO_BINS; ++b) {
            cnt[b] = 0;
            for (int k = 0; k < 3; ++k) { blo[b][k] = INFINITY; bhi[b][k] = -INFINITY; }
        }
        for (int64_t i = first; i < first + count; ++i) {
            const prim_box *p = &bc->pb[bc->idx[i]];
            int b = (int)((p->c[ax] - clo[ax]) * scale);
            if (b >= LSO_BINS) b = LSO_BINS - 1;
            if (b < 0) b = 0;
            cnt[b]++;
            for (int k = 0; k < 3; ++k) { blo[b][k] = fminf3(blo[b][k], p->lo[k]); bhi[b][k] = fmaxf3(bhi[b][k], p->hi[k]); }
        }
        float rarea[LSO_BINS]; int rcnt[LSO_BINS];
        float alo[3] = {INFINITY, INFINITY, INFINITY}, ahi[3] = {-INFINITY, -INFINITY, -INFINITY}; int ac = 0;
        for (int b = LSO_BINS - 1; b > 0; --b) {
            for (int k = 0; k < 3; ++k) { alo[k] = fminf3(alo[k], blo[b][k]); ahi[k] = fmaxf3(ahi[k], bhi[b][k]); }
            ac += cnt[b]; rcnt[b] = ac; rarea[b] = ac ? half_area(alo, ahi) : 0.0f;
        }
        for (int k = 0; k < 3; ++k) { alo[k] = INFINITY; ahi[k] = -INFINITY; }
        ac = 0;
        for (int b = 0; b < LSO_BINS - 1; ++b) {
            for (int k = 0; k < 3; ++k) { alo[k] = fminf3(alo[k], blo[b][k]); ahi[k] = fmaxf3(ahi[k], bhi[b][k]); }
            ac += cnt[b];
            if (ac == 0 || rcnt[b + 1] == 0) continue;
            float cost = half_area(alo, ahi) * (float)ac + rarea[b + 1] * (float)rcnt[b + 1];
            if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = b; }
        }
    }
    int64_t mid;
    if (best_axis < 0) {
        mid = first + count / 2;   /* all centroids coincide: median split in index order */
    } else {
        float ext = chi[best_axis] - clo[best_axis];
        float scale = (float)LSO_BINS / ext;
        int64_t i = first, j = first + count - 1;
        while (i <= j) {
            const prim_box *p = &bc->pb[bc->idx[i]];
            int b = (int)((p->c[best_axis] - clo[best_axis]) * scale);
            if (b >= LSO_BINS) b = LSO_BINS - 1;
            if (b < 0) b = 0;
            if (b <= best_bin) ++i;
            else { int32_t tmp = bc->idx[i]; bc->idx[i] = bc->idx[j]; bc->idx[j] = tmp; --j; }
        }
        mid = i;
        if (mid == first || mid == first + count) mid = first + count / 2;
    }
    int64_t l = alloc_nodes(bc, 2);
    node->left = (int32_t)l; node->count = 0;
    int64_t lc = mid - first, rc = count - lc;
    if (count > 8192 && depth < 24) {
#pragma omp task default(shared) firstprivate(l, first, lc, depth)
        build_rec(bc, l, first, lc, depth + 1);
#pragma omp task default(shared) firstprivate(l, mid, rc, depth)
        build_rec(bc, l + 1, mid, rc, depth + 1);
#pragma omp taskwait
    } else {
        build_rec(bc, l, first, lc, depth + 1);
        build_rec(bc, l + 1, mid, rc, depth + 1);
    }
}

LSO_API lso_bvh *lso_bvh_build(const float *verts, const uint32_t *tris, int64_t nT)
{
    lso_bvh *bvh = (lso_bvh *)calloc(1, sizeof(lso_bvh));
    bvh->n_tris = nT;
    if (nT <= 0) return bvh;
    prim_box *pb = (prim_box *)malloc(sizeof(prim_box) * (size_t)nT);
    int32_t *idx = (int32_t *)malloc(sizeof(int32_t) * (size_t)nT);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nT; ++i) {
        const float *a = verts + 3 * (size_t)tris[3 * i], *b = verts + 3 * (size_t)tris[3 * i + 1],
                    *c = verts + 3 * (size_t)tris[3 * i + 2];
        prim_box *p = &pb[i];
        for (int k = 0; k < 3; ++k) {
            float lo = fminf3(a[k], fminf3(b[k], c[k])), hi = fmaxf3(a[k], fmaxf3(b[k], c[k]));
            /* pad: Moeller-Trumbore in fp32 may accept rays passing a few ulp outside the exact box */
            float m = fmaxf3(fabsf(lo), fabsf(hi));
            float pad = m * 4e-6f + 1e-30f;
            p->lo[k] = lo - pad; p->hi[k] = hi + pad;
            p->c[k] = 0.5f * (lo + hi);
        }
        idx[i] = (int32_t)i;
    }
    bvh->node_cap = 2 * nT + 2;
    bvh->nodes = (lso_node *)malloc(sizeof(lso_node) * (size_t)bvh->node_cap);
    build_ctx bc = {bvh, pb, idx, 1};
#pragma omp parallel
#pragma omp single nowait
    build_rec(&bc, 0, 0, nT, 0);
    bvh->n_nodes = bc.next_node;
    bvh->tris = (lso_tri *)malloc(sizeof(lso_tri) * (size_t)nT);
    bvh->tri_ids = idx;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nT; ++i) {
        int32_t g = idx[i];
        tri_setup(verts + 3 * (size_t)tris[3 * g], verts + 3 * (size_t)tris[3 * g + 1],
                  verts + 3 * (size_t)tris[3 * g + 2], &bvh->tris[i]);
    }
    free(pb);
    return bvh;
}

LSO_API void lso_bvh_free(lso_bvh *bvh)
{
    if (!bvh) return;
    free(bvh->nodes); free(bvh->tris); free(bvh->tri_ids); free(bvh);
}

LSO_API int64_t lso_bvh_node_count(const lso_bvh *bvh) { return bvh ? bvh->n_nodes : 0; }

static inline int slab(const lso_node *n, const float *o, const float *inv, float tbest, float *tnear_out)
{
    float t0 = 0.0f, t1 = tbest;
    for (int k = 0; k < 3; ++k) {
        float a = (n->lo[k] - o[k]) * inv[k];
        float b = (n->hi[k] - o[k]) * inv[k];
        float mn = a < b ? a : b, mx = a < b ? b : a;   /* NaN (0*inf) falls through to 'b'/'a' harmlessly below */
        if (mn != mn) mn = -INFINITY;
        if (mx != mx) mx = INFINITY;
        if (mn > t0) t0 = mn;
        if (mx < t1) t1 = mx;
    }
    *tnear_out = t0;
    return t0 <= t1 * 1.0000005f;
}

static void trace_one(const lso_bvh *bvh, const float *o, const float *d, float *t_out, int32_t *tri_out)
{
    float best = INFINITY; int32_t bi = -1;
    if (bvh->n_tris > 0) {
        float inv[3] = {1.0f / d[0], 1.0f / d[1], 1.0f / d[2]};
        int32_t stack[128]; int sp = 0;
        stack[sp++] = 0;
        while (sp) {
            const lso_node *n = &bvh->nodes[stack[--sp]];
            float tn;
            if (!slab(n, o, inv, best, &tn)) continue;
            for (;;) {
                if (n->count) {
                    for (int32_t i = n->left; i < n->left + n->count; ++i) {
                        float t;
                        if (lso_intersect_any(&bvh->tris[i], o, d, &t)) {
                            int32_t g = bvh->tri_ids[i];
                            if (t < best || (t == best && g < bi)) { best = t; bi = g; }
                        }
                    }
                    break;
                }
                const lso_node *l = &bvh->nodes[n->left], *r = l + 1;
                float tl, tr;
                int hl = slab(l, o, inv, best, &tl), hr = slab(r, o, inv, best, &tr);
                if (hl && hr) {
                    if (tr < tl) { const lso_node *tmp = l; l = r; r = tmp; }
                    if (sp < 127) stack[sp++] = (int32_t)(r - bvh->nodes);
                    n = l;
                } else if (hl) n = l;
                else if (hr) n = r;
                else break;
            }
        }
    }
    *t_out = best; *tri_out = bi;
}

LSO_API void lso_trace_bvh(const lso_bvh *bvh, const float *orgs, int org_stride, const float *dirs, int64_t nR,
                           float *t_out, int32_t *tri_out)
{
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t r = 0; r < nR; ++r)
        trace_one(bvh, orgs + (size_t)org_stride * r, dirs + 3 * r, &t_out[r], &tri_out[r]);
}

/* Second return (include/lsb200.h: LSB200_OPT_RETURNS = 2; the reference only plans it, SRC/LidarShooter.hpp.in:87-93):
 * per ray with a first hit (t1, id1), the lexicographically smallest (t, triangle id) greater than it -- the next
 * surface behind the first one.  Same walk and padding as trace_one; the lower bound applies to triangles only. */
LSO_API void lso_trace_bvh_second(const lso_bvh *bvh, const float *orgs, int org_stride, const float *dirs, int64_t nR,
                                  const float *t1, const int32_t *tri1, float *t2_out, int32_t *tri2_out)
{
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t r = 0; r < nR; ++r) {
        float best = INFINITY; int32_t bi = -1;
        if (tri1[r] >= 0 && bvh->n_tris > 0) {
            const float *o = orgs + (size_t)org_stride * r, *d = dirs + 3 * r;
            const float lo_t = t1[r]; const int32_t lo_i = tri1[r];
            float inv[3] = {1.0f / d[0], 1.0f / d[1], 1.0f / d[2]};
            int32_t stack[128]; int sp = 0;
            stack[sp++] = 0;
            while (sp) {
                const lso_node *n = &bvh->nodes[stack[--sp]];
                float tn;
                if (!slab(n, o, inv, best, &tn)) continue;
                if (n->count) {
                    for (int32_t i = n->left; i < n->left + n->count; ++i) {
                        float t;
                        if (!lso_intersect_any(&bvh->tris[i], o, d, &t)) continue;
                        const int32_t g = bvh->tri_ids[i];
                        if (!(t > lo_t || (t == lo_t && g > lo_i))) continue;
                        if (t < best || (t == best && (bi < 0 || g < bi))) { best = t; bi = g; }
                    }
                } else if (sp < 126) { stack[sp++] = n->left; stack[sp++] = n->left + 1; }
            }
        }
        t2_out[r] = best; tri2_out[r] = bi;
    }
}

/* ------------------------------------------------------------------------------------------------
 * PACKET form of lso_trace_bvh: 8 consecutive rays (AVX2) walk the BVH together with one shared stack, the way
 * rtcIntersect16 walks a ray packet (SRC/EmbreeTracer.cpp:297-367, 472-480: the reference traces 16-ray
 * packets).  It exists to give bench.py a CPU baseline closer to what Embree would do than the one-ray-at-a-time
 * walk above.  Per lane the arithmetic is lso_intersect's, operation for operation (separate mul / add / sub,
 * IEEE division, no FMA: the function is compiled for AVX2 without FMA), a node is entered when ANY lane's slab
 * test passes (a superset of what each ray alone would enter) and closest hit = min (t, triangle id) does not
 * depend on the order: results are bit-identical to lso_trace_bvh (tests/test_oracle.py).
 * Default (Moeller-Trumbore) mode only; returns 0 without touching the outputs when AVX2 is missing or the
 * watertight mode is on -- the caller then uses lso_trace_bvh.
 * ---------------------------------------------------------------------------------------------- */
#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>
#define LSO_AVX2 __attribute__((target("avx2")))

LSO_AVX2 static inline __m256 dot3_v(__m256 a0, __m256 a1, __m256 a2, __m256 b0, __m256 b1, __m256 b2)
{
    return _mm256_add_ps(_mm256_add_ps(_mm256_mul_ps(a0, b0), _mm256_mul_ps(a1, b1)), _mm256_mul_ps(a2, b2));
}

/* slab() for 8 rays: bit i of the result = ray i enters the box; tnear per lane in *tn */
LSO_AVX2 static inline int slab_v(const lso_node *n, const __m256 *o, const __m256 *inv, __m256 tbest, __m256 *tn)
{
    const __m256 ninf = _mm256_set1_ps(-INFINITY), pinf = _mm256_set1_ps(INFINITY);
    __m256 t0 = _mm256_setzero_ps(), t1 = tbest;
    for (int k = 0; k < 3; ++k) {
        const __m256 a = _mm256_mul_ps(_mm256_sub_ps(_mm256_set1_ps(n->lo[k]), o[k]), inv[k]);
        const __m256 b = _mm256_mul_ps(_mm256_sub_ps(_mm256_set1_ps(n->hi[k]), o[k]), inv[k]);
        __m256 mn = _mm256_min_ps(a, b);          /* a < b ? a : b  (b when unordered) */
        __m256 mx = _mm256_max_ps(b, a);          /* a < b ? b : a  (a when unordered) */
        mn = _mm256_blendv_ps(mn, ninf, _mm256_cmp_ps(mn, mn, _CMP_UNORD_Q));
        mx = _mm256_blendv_ps(mx, pinf, _mm256_cmp_ps(mx, mx, _CMP_UNORD_Q));
        t0 = _mm256_blendv_ps(t0, mn, _mm256_cmp_ps(mn, t0, _CMP_GT_OQ));
        t1 = _mm256_blendv_ps(t1, mx, _mm256_cmp_ps(mx, t1, _CMP_LT_OQ));
    }
    *tn = t0;
    return _mm256_movemask_ps(_mm256_cmp_ps(t0, _mm256_mul_ps(t1, _mm256_set1_ps(1.0000005f)), _CMP_LE_OQ));
}

LSO_AVX2 static void trace_packet8(const lso_bvh *bvh, const float *orgs, int org_stride, const float *dirs, int64_t r0, int cnt,
                                   float *t_out, int32_t *tri_out)
{
    float tmp[6][8];
    for (int l = 0; l < 8; ++l) {
        const int64_t r = r0 + (l < cnt ? l : cnt - 1);          /* ragged tail: repeat the last ray */
        const float *o = orgs + (size_t)org_stride * r, *d = dirs + 3 * r;
        for (int k = 0; k < 3; ++k) { tmp[k][l] = o[k]; tmp[3 + k][l] = d[k]; }
    }
    __m256 o[3], d[3], inv[3];
    for (int k = 0; k < 3; ++k) {
        o[k] = _mm256_loadu_ps(tmp[k]); d[k] = _mm256_loadu_ps(tmp[3 + k]);
        inv[k] = _mm256_div_ps(_mm256_set1_ps(1.0f), d[k]);
    }
    __m256 best = _mm256_set1_ps(INFINITY);
    __m256i bi = _mm256_set1_epi32(-1);
    const __m256 sign = _mm256_set1_ps(-0.0f), zero = _mm256_setzero_ps(), fmax = _mm256_set1_ps(FLT_MAX);
    int32_t stack[128]; int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const lso_node *n = &bvh->nodes[stack[--sp]];
        __m256 tn;
        if (!slab_v(n, o, inv, best, &tn)) continue;
        for (;;) {
            if (n->count) {
                for (int32_t i = n->left; i < n->left + n->count; ++i) {
                    const lso_tri *tri = &bvh->tris[i];
                    float Ng[3];
                    cross3(tri->e2, tri->e1, Ng);
                    const __m256 C0 = _mm256_sub_ps(_mm256_set1_ps(tri->v0[0]), o[0]);
                    const __m256 C1 = _mm256_sub_ps(_mm256_set1_ps(tri->v0[1]), o[1]);
                    const __m256 C2 = _mm256_sub_ps(_mm256_set1_ps(tri->v0[2]), o[2]);
                    const __m256 R0 = _mm256_sub_ps(_mm256_mul_ps(C1, d[2]), _mm256_mul_ps(C2, d[1]));
                    const __m256 R1 = _mm256_sub_ps(_mm256_mul_ps(C2, d[0]), _mm256_mul_ps(C0, d[2]));
                    const __m256 R2 = _mm256_sub_ps(_mm256_mul_ps(C0, d[1]), _mm256_mul_ps(C1, d[0]));
                    const __m256 n0 = _mm256_set1_ps(Ng[0]), n1 = _mm256_set1_ps(Ng[1]), n2 = _mm256_set1_ps(Ng[2]);
                    const __m256 den = dot3_v(n0, n1, n2, d[0], d[1], d[2]);
                    const __m256 absDen = _mm256_andnot_ps(sign, den), sgn = _mm256_and_ps(sign, den);
                    const __m256 U = _mm256_xor_ps(dot3_v(R0, R1, R2, _mm256_set1_ps(tri->e2[0]), _mm256_set1_ps(tri->e2[1]), _mm256_set1_ps(tri->e2[2])), sgn);
                    const __m256 V = _mm256_xor_ps(dot3_v(R0, R1, R2, _mm256_set1_ps(tri->e1[0]), _mm256_set1_ps(tri->e1[1]), _mm256_set1_ps(tri->e1[2])), sgn);
                    __m256 ok = _mm256_cmp_ps(den, zero, _CMP_NEQ_UQ);
                    ok = _mm256_and_ps(ok, _mm256_cmp_ps(U, zero, _CMP_GE_OQ));
                    ok = _mm256_and_ps(ok, _mm256_cmp_ps(V, zero, _CMP_GE_OQ));
                    ok = _mm256_and_ps(ok, _mm256_cmp_ps(_mm256_add_ps(U, V), absDen, _CMP_LE_OQ));
                    if (!_mm256_movemask_ps(ok)) continue;
                    const __m256 T = _mm256_xor_ps(dot3_v(n0, n1, n2, C0, C1, C2), sgn);
                    ok = _mm256_and_ps(ok, _mm256_cmp_ps(_mm256_mul_ps(absDen, zero), T, _CMP_LT_OQ));
                    const __m256 t = _mm256_div_ps(T, absDen);
                    ok = _mm256_and_ps(ok, _mm256_cmp_ps(t, fmax, _CMP_LE_OQ));
                    const __m256i g = _mm256_set1_epi32(bvh->tri_ids[i]);
                    const __m256 closer = _mm256_cmp_ps(t, best, _CMP_LT_OQ);
                    const __m256 tie = _mm256_and_ps(_mm256_cmp_ps(t, best, _CMP_EQ_OQ), _mm256_castsi256_ps(_mm256_cmpgt_epi32(bi, g)));
                    const __m256 take = _mm256_and_ps(ok, _mm256_or_ps(closer, tie));
                    best = _mm256_blendv_ps(best, t, take);
                    bi = _mm256_castps_si256(_mm256_blendv_ps(_mm256_castsi256_ps(bi), _mm256_castsi256_ps(g), take));
                }
                break;
            }
            const lso_node *l = &bvh->nodes[n->left], *r = l + 1;
            __m256 tl, tr;
            const int hl = slab_v(l, o, inv, best, &tl), hr = slab_v(r, o, inv, best, &tr);
            if (hl && hr) {
                /* near child = the one most of the lanes that enter both reach first */
                const int both = hl & hr;
                const int r_first = _mm256_movemask_ps(_mm256_cmp_ps(tr, tl, _CMP_LT_OQ)) & both;
                if (2 * __builtin_popcount(r_first) > __builtin_popcount(both)) { const lso_node *x = l; l = r; r = x; }
                if (sp < 127) stack[sp++] = (int32_t)(r - bvh->nodes);
                n = l;
            } else if (hl) n = l;
            else if (hr) n = r;
            else break;
        }
    }
    float tb[8]; int32_t ib[8];
    _mm256_storeu_ps(tb, best); _mm256_storeu_si256((__m256i *)ib, bi);
    for (int l = 0; l < cnt; ++l) { t_out[r0 + l] = tb[l]; tri_out[r0 + l] = ib[l]; }
}

LSO_API int lso_trace_bvh_packet(const lso_bvh *bvh, const float *orgs, int org_stride, const float *dirs, int64_t nR,
                                 float *t_out, int32_t *tri_out)
{
    if (g_watertight || !__builtin_cpu_supports("avx2")) return 0;
    if (bvh->n_tris <= 0) {
        for (int64_t r = 0; r < nR; ++r) { t_out[r] = INFINITY; tri_out[r] = -1; }
        return 1;
    }
    const int64_t nP = (nR + 7) / 8;
#pragma omp parallel for schedule(dynamic, 32)
    for (int64_t p = 0; p < nP; ++p) {
        const int64_t r0 = 8 * p;
        const int cnt = (int)(nR - r0 < 8 ? nR - r0 : 8);
        trace_packet8(bvh, orgs, org_stride, dirs, r0, cnt, t_out, tri_out);
    }
    return 1;
}
#else
LSO_API int lso_trace_bvh_packet(const lso_bvh *bvh, const float *orgs, int org_stride, const float *dirs, int64_t nR,
                                 float *t_out, int32_t *tri_out)
{
    (void)bvh; (void)orgs; (void)org_stride; (void)dirs; (void)nR; (void)t_out; (void)tri_out;
    return 0;
}
#endif

/* ------------------------------------------------------------------------------------------------
 * Float64 SECOND OPINION (not a restatement of anything in the reference): textbook Moeller-Trumbore 1997
 * in double precision on the ORIGINAL float32 vertices, closest hit = lexicographic min of (t, triangle id),
 * through the same CPU BVH (slabs evaluated in double on the padded boxes).  tests/ and bench.py run it over
 * the same rays as the fp32 oracle; the rays where the two disagree in hit mask or triangle id are the
 * "grazing" rays north_star asks to be counted and reported -- each is then shown (tests/grazing.py) to pass
 * within a few fp32 ulp of a triangle edge, of a plane seen edge-on, or of a tie between two triangles.
 * uvw_out (optional, nR x 3): barycentrics (u, v) and det of the accepted hit.
 * ---------------------------------------------------------------------------------------------- */
static inline int mt_f64(const float *a, const float *b, const float *c, const double *o, const double *d,
                         double *t_out, double *u_out, double *v_out, double *det_out)
{
    const double e1[3] = {(double)b[0] - a[0], (double)b[1] - a[1], (double)b[2] - a[2]};
    const double e2[3] = {(double)c[0] - a[0], (double)c[1] - a[1], (double)c[2] - a[2]};
    const double p[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
    const double det = e1[0] * p[0] + e1[1] * p[1] + e1[2] * p[2];
    *det_out = det;
    if (det == 0.0) return 0;
    const double inv = 1.0 / det;
    const double tv[3] = {o[0] - a[0], o[1] - a[1], o[2] - a[2]};
    const double u = (tv[0] * p[0] + tv[1] * p[1] + tv[2] * p[2]) * inv;
    const double q[3] = {tv[1] * e1[2] - tv[2] * e1[1], tv[2] * e1[0] - tv[0] * e1[2], tv[0] * e1[1] - tv[1] * e1[0]};
    const double v = (d[0] * q[0] + d[1] * q[1] + d[2] * q[2]) * inv;
    const double t = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * inv;
    *u_out = u; *v_out = v; *t_out = t;
    if (u < 0.0 || v < 0.0 || u + v > 1.0) return 0;
    if (!(t > 0.0) || !(t <= DBL_MAX)) return 0;
    return 1;
}

/* Barycentrics of ray r against triangle tri_ids[r] in double, hit or not (u, v, t, det per ray; NaN for id < 0). */
LSO_API void lso_bary_f64(const float *verts, const uint32_t *tris, const float *orgs, int org_stride, const float *dirs,
                          const int32_t *tri_ids, int64_t nR, double *uvtd_out)
{
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < nR; ++r) {
        double *out = uvtd_out + 4 * r;
        if (tri_ids[r] < 0) { out[0] = out[1] = out[2] = out[3] = NAN; continue; }
        const uint32_t *tv = tris + 3 * (size_t)tri_ids[r];
        const float *of = orgs + (size_t)org_stride * r, *df = dirs + 3 * r;
        const double o[3] = {of[0], of[1], of[2]}, d[3] = {df[0], df[1], df[2]};
        double t = NAN, u = NAN, v = NAN, det = 0.0;
        mt_f64(verts + 3 * (size_t)tv[0], verts + 3 * (size_t)tv[1], verts + 3 * (size_t)tv[2], o, d, &t, &u, &v, &det);
        out[0] = u; out[1] = v; out[2] = t; out[3] = det;
    }
}

static inline int slab_f64(const lso_node *n, const double *o, const double *inv, double tbest)
{
    double t0 = 0.0, t1 = tbest;
    for (int k = 0; k < 3; ++k) {
        double a = ((double)n->lo[k] - o[k]) * inv[k];
        double b = ((double)n->hi[k] - o[k]) * inv[k];
        double mn = a < b ? a : b, mx = a < b ? b : a;
        if (mn != mn) mn = -INFINITY;
        if (mx != mx) mx = INFINITY;
        if (mn > t0) t0 = mn;
        if (mx < t1) t1 = mx;
    }
    return t0 <= t1 * (1.0 + 1e-9);
}

LSO_API void lso_trace_bvh_f64(const lso_bvh *bvh, const float *verts, const uint32_t *tris, const float *orgs,
                               int org_stride, const float *dirs, int64_t nR, double *t_out, int32_t *tri_out)
{
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t r = 0; r < nR; ++r) {
        const float *of = orgs + (size_t)org_stride * r, *df = dirs + 3 * r;
        const double o[3] = {of[0], of[1], of[2]}, d[3] = {df[0], df[1], df[2]};
        const double inv[3] = {1.0 / d[0], 1.0 / d[1], 1.0 / d[2]};
        double best = INFINITY; int32_t bi = -1;
        if (bvh->n_tris > 0) {
            int32_t stack[128]; int sp = 0;
            stack[sp++] = 0;
            while (sp) {
                const lso_node *n = &bvh->nodes[stack[--sp]];
                if (!slab_f64(n, o, inv, best)) continue;
                if (n->count) {
                    for (int32_t i = n->left; i < n->left + n->count; ++i) {
                        const int32_t g = bvh->tri_ids[i];
                        const uint32_t *tv = tris + 3 * (size_t)g;
                        double t, u, v, det;
                        if (mt_f64(verts + 3 * (size_t)tv[0], verts + 3 * (size_t)tv[1], verts + 3 * (size_t)tv[2], o, d, &t, &u, &v, &det))
                            if (t < best || (t == best && g < bi)) { best = t; bi = g; }
                    }
                } else if (sp < 126) {
                    stack[sp++] = n->left; stack[sp++] = n->left + 1;
                }
            }
        }
        t_out[r] = best; tri_out[r] = bi;
    }
}

/* Brute-force form of the same (small scenes). */
LSO_API void lso_trace_brute_f64(const float *verts, const uint32_t *tris, int64_t nT, const float *orgs, int org_stride,
                                 const float *dirs, int64_t nR, double *t_out, int32_t *tri_out)
{
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t r = 0; r < nR; ++r) {
        const float *of = orgs + (size_t)org_stride * r, *df = dirs + 3 * r;
        const double o[3] = {of[0], of[1], of[2]}, d[3] = {df[0], df[1], df[2]};
        double best = INFINITY; int32_t bi = -1;
        for (int64_t i = 0; i < nT; ++i) {
            const uint32_t *tv = tris + 3 * (size_t)i;
            double t, u, v, det;
            if (mt_f64(verts + 3 * (size_t)tv[0], verts + 3 * (size_t)tv[1], verts + 3 * (size_t)tv[2], o, d, &t, &u, &v, &det))
                if (t < best) { best = t; bi = (int32_t)i; }
        }
        t_out[r] = best; tri_out[r] = bi;
    }
}

/* Reference-faithful threading (SRC/EmbreeTracer.cpp:304-313): fixed thread count. */
LSO_API void lso_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
LSO_API int lso_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------------------------------
 * Point packing (SRC/XYZIRBytes.cpp:24-40; SRC/EmbreeTracer.cpp:331-335): 32 bytes per hit,
 * x,y,z f32 | 4 zero | intensity f32 (64.0) | ring i32 | 8 zero.  Point = t * d_sensor.  Emitted in
 * ray-id order (the reference's order is mutex-nondeterministic; parity is on the set).
 * ring_mode 0: ring = channel (OptiX path, SRC/LidarDeviceKernels.cu:51); 1: ring = 0 (Embree path quirk).
 * Returns the number of points written.
 * ---------------------------------------------------------------------------------------------- */
LSO_API int64_t lso_pack_points(const float *dirs_s, const float *t, const int32_t *tri, int64_t nR, int H,
                                int ring_mode, uint8_t *out)
{
    /* ray-id order, threaded: per-chunk hit counts, exclusive prefix, then every chunk writes its own range */
    enum { MAXC = 256 };
    int64_t counts[MAXC + 1];
#ifdef _OPENMP
    int nc = omp_get_max_threads();
#else
    int nc = 1;
#endif
    if (nc > MAXC) nc = MAXC;
    if (nR < 65536) nc = 1;
    const int64_t per = (nR + nc - 1) / nc;
#pragma omp parallel for schedule(static, 1) if (nc > 1)
    for (int c = 0; c < nc; ++c) {
        const int64_t lo = c * per, hi = lo + per < nR ? lo + per : nR;
        int64_t k = 0;
        for (int64_t r = lo; r < hi; ++r) k += tri[r] >= 0;
        counts[c] = k;
    }
    int64_t n = 0;
    for (int c = 0; c < nc; ++c) { const int64_t k = counts[c]; counts[c] = n; n += k; }
#pragma omp parallel for schedule(static, 1) if (nc > 1)
    for (int c = 0; c < nc; ++c) {
        const int64_t lo = c * per, hi = lo + per < nR ? lo + per : nR;
        int64_t w = counts[c];
        for (int64_t r = lo; r < hi; ++r) {
            if (tri[r] < 0) continue;
            float rec[8];
            memset(rec, 0, sizeof(rec));
            rec[0] = t[r] * dirs_s[3 * r + 0];
            rec[1] = t[r] * dirs_s[3 * r + 1];
            rec[2] = t[r] * dirs_s[3 * r + 2];
            rec[4] = 64.0f;
            int32_t ring = ring_mode ? 0 : (int32_t)(r / H);
            memcpy(&rec[5], &ring, 4);
            memcpy(out + 32 * w, rec, 32);
            ++w;
        }
    }
    return n;
}
