// trace.cuh -- ray generation + closest-hit traversal + hit compaction.  Replaces EmbreeTracer::traceScene
// (SRC/EmbreeTracer.cpp:297-367): LidarDevice::nextRay16 (SRC/LidarDevice.cpp:294-342), rtcIntersect16
// (SRC/EmbreeTracer.cpp:472-480) and XYZIRBytes::addToCloud (SRC/XYZIRBytes.cpp:24-40).
//
// The intersection arithmetic is the oracle's (oracle/lsb_oracle.c: lso_intersect), written with explicit
// __f*_rn intrinsics so that ptxas can never contract it into FMA: per-ray t and triangle id are bit-exact.
// Box culling is allowed to be sloppy in one direction only (never reject a box whose triangle the exact
// test would accept): the ray origin is widened by eps on both sides and t_far is relaxed by 4 ulp.
#pragma once
#include "common.cuh"

namespace lsb {

constexpr int kTraceThreads = 512;     // 2 CTAs / SM at <= 64 registers
constexpr int kStackSize = 64;         // LBVH depth <= 30 Morton bits + index tie-break bits

struct TraceParams {
    const float4 *nodes;          // [top block SoA: 4 planes x top_cap][Karras nodes AoS]
    const float4 *tris;           // leaf-ordered 48-byte records
    const int32_t *leaf_ids;
    const uint32_t *bounds;       // scene bounds (ordered uint) -> box padding
    const int *n_top;             // nodes actually present in the top block
    const SensorDev *sensors;
    int n_sensors;
    int unit_lo, n_units;         // this launch covers the 32-ray work units [unit_lo, n_units) of the sensor table
    int top_cap, n_tris;
    int world_frame;
    uint32_t flags;
    uint32_t *work;               // [0] next unit, [1] finished CTAs
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra WAIT_%=;\n\t}"
        ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA bulk copy global -> shared (SASS: UBLKCP), completion counted in bytes on the mbarrier.
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct Ray {
    float ox, oy, oz, dx, dy, dz;
};

// Embree-ordered Moeller-Trumbore, bit-identical to lso_intersect().  q0..q2 = the 48-byte triangle record.
__device__ __forceinline__ bool intersect_tri(const float4 q0, const float4 q1, const float4 q2, const Ray &r, float &t_out)
{
    const float v0x = q0.x, v0y = q0.y, v0z = q0.z;
    const float e1x = q0.w, e1y = q1.x, e1z = q1.y;
    const float e2x = q1.z, e2y = q1.w, e2z = q2.x;
    const float ngx = q2.y, ngy = q2.z, ngz = q2.w;
    const float cx = __fsub_rn(v0x, r.ox), cy = __fsub_rn(v0y, r.oy), cz = __fsub_rn(v0z, r.oz);
    const float rx = __fsub_rn(__fmul_rn(cy, r.dz), __fmul_rn(cz, r.dy));
    const float ry = __fsub_rn(__fmul_rn(cz, r.dx), __fmul_rn(cx, r.dz));
    const float rz = __fsub_rn(__fmul_rn(cx, r.dy), __fmul_rn(cy, r.dx));
    const float den = __fadd_rn(__fadd_rn(__fmul_rn(ngx, r.dx), __fmul_rn(ngy, r.dy)), __fmul_rn(ngz, r.dz));
    const float absDen = fabsf(den);
    const bool neg = (__float_as_uint(den) >> 31) != 0u;
    float U = __fadd_rn(__fadd_rn(__fmul_rn(rx, e2x), __fmul_rn(ry, e2y)), __fmul_rn(rz, e2z));
    float V = __fadd_rn(__fadd_rn(__fmul_rn(rx, e1x), __fmul_rn(ry, e1y)), __fmul_rn(rz, e1z));
    if (neg) { U = -U; V = -V; }
    if (!(den != 0.0f)) return false;
    if (!(U >= 0.0f)) return false;
    if (!(V >= 0.0f)) return false;
    if (!(__fadd_rn(U, V) <= absDen)) return false;
    float T = __fadd_rn(__fadd_rn(__fmul_rn(ngx, cx), __fmul_rn(ngy, cy)), __fmul_rn(ngz, cz));
    if (neg) T = -T;
    if (!(__fmul_rn(absDen, 0.0f) < T)) return false;      // tnear = 0, strict
    const float t = __fdiv_rn(T, absDen);
    if (!(t <= 3.402823466e+38f)) return false;            // tfar = +inf; rejects inf / NaN
    t_out = t;
    return true;
}

__device__ __forceinline__ void test_leaf(const TraceParams &P, int ref, const Ray &r, float &best, int &bi)
{
    const int enc = ~ref;
    const int first = enc & ((1 << kLeafShift) - 1);
    const int count = (enc >> kLeafShift) + 1;
    for (int k = 0; k < count; ++k) {
        const float4 *rec = P.tris + 3 * (size_t)(first + k);
        const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1), q2 = __ldg(rec + 2);
        float t;
        if (intersect_tri(q0, q1, q2, r, t) && t <= best) {
            const int g = __ldg(P.leaf_ids + first + k);
            if (t < best || g < bi) { best = t; bi = g; }   // lexicographic minimum of (t, global triangle id)
        }
    }
}

// Closest hit of one ray.  TOP: node references below n_top are read from the shared-memory copy.
template <bool TOP>
__device__ __forceinline__ void trace_ray(const TraceParams &P, const float4 *s_top, int n_top, float eps, const Ray &r,
                                          float &best_out, int &bi_out)
{
    // slab set-up: t_lo = (lo - (o + eps)) * inv, t_hi = (hi - (o - eps)) * inv, in FMA form
    const float tiny = 1e-30f;
    const float ix = 1.0f / (fabsf(r.dx) < tiny ? copysignf(tiny, r.dx) : r.dx);
    const float iy = 1.0f / (fabsf(r.dy) < tiny ? copysignf(tiny, r.dy) : r.dy);
    const float iz = 1.0f / (fabsf(r.dz) < tiny ? copysignf(tiny, r.dz) : r.dz);
    const float ax = -(r.ox + eps) * ix, ay = -(r.oy + eps) * iy, az = -(r.oz + eps) * iz;
    const float bx = -(r.ox - eps) * ix, by = -(r.oy - eps) * iy, bz = -(r.oz - eps) * iz;
    const int cap = P.top_cap;

    float best = INFINITY;
    int bi = -1;
    int stack[kStackSize];
    int sp = 0;
    int cur = TOP ? 0 : cap;          // root: top slot 0, or Karras node 0
    if (TOP && n_top == 0) cur = cap;
    while (true) {
        if (cur >= 0) {
            float4 q0, q1, q2, q3;
            if (TOP && cur < cap) {
                q0 = s_top[cur]; q1 = s_top[cap + cur]; q2 = s_top[2 * cap + cur]; q3 = s_top[3 * cap + cur];
            } else {
                const float4 *n = P.nodes + 4 * (size_t)cur;
                q0 = __ldg(n); q1 = __ldg(n + 1); q2 = __ldg(n + 2); q3 = __ldg(n + 3);
            }
            const float far_lim = best * 1.0000005f;
            // child 0
            float l0 = fmaf(q0.x, ix, ax), h0 = fmaf(q0.y, ix, bx);
            float l1 = fmaf(q0.z, iy, ay), h1 = fmaf(q0.w, iy, by);
            float l2 = fmaf(q2.x, iz, az), h2 = fmaf(q2.y, iz, bz);
            const float n0 = fmaxf(fmaxf(fminf(l0, h0), fminf(l1, h1)), fmaxf(fminf(l2, h2), 0.0f));
            const float f0 = fminf(fminf(fmaxf(l0, h0), fmaxf(l1, h1)), fminf(fmaxf(l2, h2), far_lim));
            // child 1
            l0 = fmaf(q1.x, ix, ax); h0 = fmaf(q1.y, ix, bx);
            l1 = fmaf(q1.z, iy, ay); h1 = fmaf(q1.w, iy, by);
            l2 = fmaf(q2.z, iz, az); h2 = fmaf(q2.w, iz, bz);
            const float n1 = fmaxf(fmaxf(fminf(l0, h0), fminf(l1, h1)), fmaxf(fminf(l2, h2), 0.0f));
            const float f1 = fminf(fminf(fmaxf(l0, h0), fmaxf(l1, h1)), fminf(fmaxf(l2, h2), far_lim));
            const bool hit0 = n0 <= f0 * 1.0000005f, hit1 = n1 <= f1 * 1.0000005f;
            const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            if (hit0 && hit1) {
                const bool swap = n1 < n0;
                cur = swap ? c1 : c0;
                stack[sp++] = swap ? c0 : c1;
                continue;
            }
            if (hit0) { cur = c0; continue; }
            if (hit1) { cur = c1; continue; }
        } else {
            test_leaf(P, cur, r, best, bi);
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    best_out = best;
    bi_out = bi;
}

// Brute force over all leaf-ordered triangles (validation variant; also the tiny-scene path).
__device__ __forceinline__ void brute_ray(const TraceParams &P, const Ray &r, float &best, int &bi)
{
    best = INFINITY; bi = -1;
    for (int k = 0; k < P.n_tris; ++k) {
        const float4 *rec = P.tris + 3 * (size_t)k;
        float t;
        if (intersect_tri(__ldg(rec), __ldg(rec + 1), __ldg(rec + 2), r, t) && t <= best) {
            const int g = __ldg(P.leaf_ids + k);
            if (t < best || g < bi) { best = t; bi = g; }
        }
    }
}

__device__ __forceinline__ int find_sensor_by_unit(const SensorDev *s, int n, int unit)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (s[mid].unit_base <= unit) lo = mid; else hi = mid - 1; }
    return lo;
}

// Writes one hit.  32-byte record (SRC/XYZIRBytes.cpp:24-40) or 16-byte XYZI.
__device__ __forceinline__ void store_point(uint8_t *points, uint32_t slot, uint32_t flags, float x, float y, float z, int ring)
{
    if (flags & 4u /*LSB200_TRACE_XYZI16*/) {
        reinterpret_cast<float4 *>(points)[slot] = make_float4(x, y, z, 64.0f);
    } else {
        float4 *p = reinterpret_cast<float4 *>(points) + 2 * (size_t)slot;
        p[0] = make_float4(x, y, z, 0.0f);
        p[1] = make_float4(64.0f, __int_as_float(ring), 0.0f, 0.0f);
    }
}

__device__ __forceinline__ void hit_point(const SensorDev &S, uint32_t flags, float t, float dsx, float dsy, float dsz,
                                          float &x, float &y, float &z)
{
    x = __fmul_rn(t, dsx); y = __fmul_rn(t, dsy); z = __fmul_rn(t, dsz);      // t * d_sensor (SRC/EmbreeTracer.cpp:331-333)
    if (flags & 16u /*LSB200_TRACE_WORLD_POINTS*/) {                          // R p + T (SRC/LidarDevice.cpp:393-401)
        // Eigen 3-term dot order a0 + (a1 + a2), then the translation (LidarDevice::originToSensorInverse)
        const float wx = __fadd_rn(__fadd_rn(__fmul_rn(S.R[0], x), __fadd_rn(__fmul_rn(S.R[1], y), __fmul_rn(S.R[2], z))), S.T[0]);
        const float wy = __fadd_rn(__fadd_rn(__fmul_rn(S.R[3], x), __fadd_rn(__fmul_rn(S.R[4], y), __fmul_rn(S.R[5], z))), S.T[1]);
        const float wz = __fadd_rn(__fadd_rn(__fmul_rn(S.R[6], x), __fadd_rn(__fmul_rn(S.R[7], y), __fmul_rn(S.R[8], z))), S.T[2]);
        x = wx; y = wy; z = wz;
    }
}

// MODE 0: persistent CTAs, top BVH levels staged into shared memory by TMA bulk copy
// MODE 1: persistent CTAs, all nodes from global memory (L1/L2)
// MODE 2: one thread per ray, no work fetching
// MODE 3: brute force (persistent)
template <int MODE>
__global__ void __launch_bounds__(kTraceThreads, 2) k_trace(TraceParams P)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *s_top = reinterpret_cast<float4 *>(smem_raw);
    __shared__ __align__(8) uint64_t s_bar;
    int n_top = 0;
    if (MODE == 0) {
        n_top = min(*P.n_top, P.top_cap);
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
        __syncthreads();
        if (n_top > 0) {
            if (threadIdx.x == 0) {
                const uint32_t plane = (uint32_t)n_top * 16u;
                mbar_expect_tx(&s_bar, 4u * plane);
                for (int q = 0; q < 4; ++q)
                    tma_bulk_g2s(s_top + (size_t)q * P.top_cap, P.nodes + (size_t)q * P.top_cap, plane, &s_bar);
            }
            mbar_wait(&s_bar, 0);
        }
    }
    // box padding: see the header comment; scaled by the largest coordinate any (origin, vertex) pair can have
    float smax = 0.0f;
#pragma unroll
    for (int k = 0; k < 6; ++k) smax = fmaxf(smax, fabsf(ord2f(__ldg(P.bounds + k))));

    const unsigned lane = lane_id();
    const int warps_per_cta = kTraceThreads / 32;
    int unit = (MODE == 2) ? P.unit_lo + (int)(blockIdx.x * warps_per_cta + (threadIdx.x >> 5)) : 0;
    while (true) {
        if (MODE != 2) {
            if (lane == 0) unit = P.unit_lo + (int)atomicAdd(&P.work[0], 1u);
            unit = __shfl_sync(0xffffffffu, unit, 0);
        }
        if (unit >= P.n_units) break;
        const int si = find_sensor_by_unit(P.sensors, P.n_sensors, unit);
        const SensorDev &S = P.sensors[si];
        const int ray = (unit - S.unit_base) * 32 + (int)lane;
        const bool valid = ray < S.n_rays;
        float best = INFINITY; int bi = -1;
        float dsx = 0.f, dsy = 0.f, dsz = 1.f;
        int v = 0;
        if (valid) {
            v = ray / S.H;
            const int h = ray - v * S.H;
            // d_s = (sin t cos p, sin t sin p, cos t)  (SRC/LidarDevice.cpp:319-321); ray id = channel * H + azimuth
            const float st = __ldg(S.sin_t + v), ct = __ldg(S.cos_t + v), sp_ = __ldg(S.sin_p + h), cp = __ldg(S.cos_p + h);
            dsx = __fmul_rn(st, cp); dsy = __fmul_rn(st, sp_); dsz = ct;
            Ray r;
            float omax = 0.0f;
            if (P.world_frame) {
                r.ox = S.T[0]; r.oy = S.T[1]; r.oz = S.T[2];
                r.dx = __fadd_rn(__fadd_rn(__fmul_rn(S.R[0], dsx), __fmul_rn(S.R[1], dsy)), __fmul_rn(S.R[2], dsz));
                r.dy = __fadd_rn(__fadd_rn(__fmul_rn(S.R[3], dsx), __fmul_rn(S.R[4], dsy)), __fmul_rn(S.R[5], dsz));
                r.dz = __fadd_rn(__fadd_rn(__fmul_rn(S.R[6], dsx), __fmul_rn(S.R[7], dsy)), __fmul_rn(S.R[8], dsz));
                omax = fmaxf(fabsf(r.ox), fmaxf(fabsf(r.oy), fabsf(r.oz)));
            } else {
                r.ox = r.oy = r.oz = 0.0f; r.dx = dsx; r.dy = dsy; r.dz = dsz;
            }
            const float eps = 8e-6f * (smax + omax) + 1e-30f;
            if (MODE == 3) brute_ray(P, r, best, bi);
            else trace_ray<MODE == 0>(P, s_top, n_top, eps, r, best, bi);
            S.t_out[ray] = best;
            S.tri_out[ray] = bi;
        }
        if (!(P.flags & (1u | 8u))) {     // neither ORDERED nor NO_POINTS: warp-aggregated atomic append
            const bool hit = valid && bi >= 0;
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (m) {
                uint32_t base = 0;
                if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(S.count, (uint32_t)__popc(m));
                base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
                if (hit) {
                    float x, y, z;
                    hit_point(S, P.flags, best, dsx, dsy, dsz, x, y, z);
                    store_point(S.points, base + __popc(m & lanemask_lt()), P.flags, x, y, z, (P.flags & 2u) ? 0 : v);
                }
            }
        }
        if (MODE == 2) break;
    }
    if (MODE != 2) {
        // the last CTA to leave re-arms the work counters for the next launch
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(&P.work[1], 1u) == gridDim.x - 1) { P.work[0] = 0; P.work[1] = 0; __threadfence(); }
        }
    }
}

// ---- ordered compaction: count -> scan -> scatter (points in ray-id order) ---------------------------------
constexpr int kCompactThreads = 256;
constexpr int kCompactRays = 1024;      // rays per block

__global__ void __launch_bounds__(kCompactThreads) k_count_hits(const SensorDev *__restrict__ sensors, int sensor_base)
{
    const SensorDev &S = sensors[sensor_base + blockIdx.y];
    const int base = blockIdx.x * kCompactRays;
    if (base >= S.n_rays) return;
    int c = 0;
#pragma unroll
    for (int k = 0; k < kCompactRays / kCompactThreads; ++k) {
        const int r = base + k * kCompactThreads + threadIdx.x;
        c += (r < S.n_rays && S.tri_out[r] >= 0) ? 1 : 0;
    }
    c = __reduce_add_sync(0xffffffffu, c);
    __shared__ int s_c[kCompactThreads / 32];
    if ((threadIdx.x & 31) == 0) s_c[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < kCompactThreads / 32; ++w) t += s_c[w];
        S.block_counts[blockIdx.x] = (uint32_t)t;
    }
}

__global__ void __launch_bounds__(256) k_scan_counts(const SensorDev *__restrict__ sensors, int sensor_base)
{
    const SensorDev &S = sensors[sensor_base + blockIdx.x];
    const int nblk = (S.n_rays + kCompactRays - 1) / kCompactRays;
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < nblk; base += 256) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < nblk ? S.block_counts[i] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) s_w[warp] = incl;
        __syncthreads();
        uint32_t woff = 0;
        for (int w = 0; w < warp; ++w) woff += s_w[w];
        const uint32_t carry = s_carry;
        if (i < nblk) S.block_counts[i] = carry + woff + incl - v;
        __syncthreads();
        if (threadIdx.x == 255) s_carry = carry + woff + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) *S.count = s_carry;
}

__global__ void __launch_bounds__(kCompactThreads) k_scatter_hits(const SensorDev *__restrict__ sensors, int sensor_base,
                                                                  uint32_t flags)
{
    const SensorDev &S = sensors[sensor_base + blockIdx.y];
    const int base = blockIdx.x * kCompactRays;
    if (base >= S.n_rays) return;
    __shared__ uint32_t s_w[kCompactThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // thread owns 4 consecutive rays so that the block-local order is the ray order
    const int r0 = base + threadIdx.x * 4;
    int hit[4]; uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) { hit[k] = (r0 + k < S.n_rays) && (S.tri_out[r0 + k] >= 0); c += hit[k]; }
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; ++w) woff += s_w[w];
    uint32_t slot = S.block_counts[blockIdx.x] + woff + incl - c;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (!hit[k]) continue;
        const int ray = r0 + k;
        const int v = ray / S.H, h = ray - v * S.H;
        const float st = __ldg(S.sin_t + v), ct = __ldg(S.cos_t + v), sp_ = __ldg(S.sin_p + h), cp = __ldg(S.cos_p + h);
        float x, y, z;
        hit_point(S, flags, S.t_out[ray], __fmul_rn(st, cp), __fmul_rn(st, sp_), ct, x, y, z);
        store_point(S.points, slot++, flags, x, y, z, (flags & 2u) ? 0 : v);
    }
}

}  // namespace lsb
