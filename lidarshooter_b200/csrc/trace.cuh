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

#ifndef LSB_TRACE_THREADS
#define LSB_TRACE_THREADS 384
#define LSB_TRACE_CTAS 3
#endif
#ifndef LSB_SMEM_STACK
#define LSB_SMEM_STACK 8         // first N stack entries per ray in shared memory (variant 6; 0 = all in local memory)
#endif
#ifndef LSB_MIXED_TILE_LISTS
#define LSB_MIXED_TILE_LISTS 1   // sign-mixed tiles (class 8) get entry lists too: interval test with an unbounded 1/d on the flipping axis
#endif
#ifndef LSB_HEAVY_FIRST
#define LSB_HEAVY_FIRST 1         // 1: heavy units (grazing rays) queued at the front of their octant class (k_scatter_units)
#endif
#ifndef LSB_PHASE_ORDER
#define LSB_PHASE_ORDER 1          // 0: mixed, downward heavy fronts, upward, downward rest; 1: mixed, upward, downward heavy fronts, downward rest
#endif
#ifndef LSB_SPLIT_DOWNWARD
#define LSB_SPLIT_DOWNWARD 1     // 1: the heavy fronts (grazing rows, LSB_HEAVY_FIRST) of the downward octant queues are drained before the upward octants
#endif
#ifndef LSB_LAZY_LIST
#define LSB_LAZY_LIST 1          // variant 6: list entries are taken lazily instead of being pushed up front
#endif
#ifndef LSB_DEFER_APPEND
#define LSB_DEFER_APPEND 0       // 1: hit-append slots requested one unit ahead (measured slower: the parked state spills)
#endif
constexpr int kTraceThreads = LSB_TRACE_THREADS;     // x LSB_TRACE_CTAS CTAs / SM: 384 x 3 = 1152 threads at <= 56 registers (512 x 2 measured 2 % slower)
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
    int tile_w_log2;              // a work unit is a tile of 2^w azimuth steps x 32 / 2^w channels of one sensor
    uint32_t flags;
    uint32_t *work;               // [0] next unit, [1] finished CTAs, [2..10] next entry of each octant queue
    unsigned long long *counters; // STATS builds only: see TraceCounter
    const int4 *unit_list;        // MODE 0: unit descriptors (sensor, first ray, its channel, its azimuth step) grouped by octant class
    const uint32_t *oct_hist;     // MODE 0: [9] units per class (0..7 octants, 8 mixed); [16, 25): heavy units per class
    const float4 *chains;         // [n_sensors][9][kMaxChain][2] origin-chain lists per octant class (k_origin_chain)
    const int *chain_counts;      // [n_sensors][9] entries per list; nullptr = lists off (start at the root)
    float4 *tile_lists;           // [queue slot][tile_cap][2] per-TILE entry lists (k_tile_entries); nullptr = off
    uint8_t *tile_counts;         // [queue slot] entries of the tile's list; kTileFallback = use the sensor's origin chain
    int tile_cap;                 // entries per tile list (<= kMaxTileEntries)
    float4 *sup_lists;            // [supertile][kMaxTileEntries][2] entry lists of the supertiles (k_super_entries); nullptr = off
    uint8_t *sup_counts;          // [supertile] entries, or kTileFallback
    float4 *tile_iv;              // [queue slot][2] direction interval of the tile: (dlo.xyz, class), (dhi.xyz, max |origin|)
    float4 *sup_iv;               // [supertile][2] same for the supertiles.  Both depend on the sensor tables / poses only.
    int refill;                   // lane refill (run_phase_refill) per group of queues: bit 0 sign-mixed, 1 upward octants, 2 heavy fronts of
                                  // the downward octants, 3 their light remainders (LSB200_OPT_LANE_REFILL)
};
constexpr int kMaxTileEntries = 16;
constexpr int kTileFallback = 0xFF;

// Instrumented (STATS) launches accumulate these per-launch totals (lsb200_get_trace_counters).
enum TraceCounter { kCntRays = 0, kCntNodes, kCntLeaves, kCntTris, kCntUnitsUniform, kCntUnitsMixed, kCntMaxStack, kCntHits, kCntCulled, kCntChainTests,
                    kCntTileVisits, kCntTileEntries, kCntTileFallbacks, kCntTileEmpty, kCntSuperVisits,
                    // per-unit cost histogram (instrumented launches): bin = log2(SM cycles the unit took); units, cycles, node visits,
                    // list size per bin
                    kCntCycUnits, kCntCycSum = kCntCycUnits + 24, kCntCycNodes = kCntCycSum + 24, kCntCycList = kCntCycNodes + 24,
                    // heavy units (>= 2^17 cycles) by octant class [9] and by elevation band [32] (band = first channel * 32 / V)
                    kCntHeavyClass = kCntCycList + 24, kCntHeavyBand = kCntHeavyClass + 9,
                    kCntCount = kCntHeavyBand + 32 };

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// One lane's atomic add as the plain instruction.  atomicAdd() on a global address lets nvcc wrap the call in its
// warp-aggregation sequence (activemask, match, leader election, popc, shuffle: ~25 instructions) even when the call
// site already runs in exactly one lane -- the traversal's ticket and hit-append atomics are such sites (2.4 % of
// k_trace's instructions, profiles/ncu_r02 source view).
__device__ __forceinline__ uint32_t atom_add_u32(uint32_t *p, uint32_t v)
{
    uint32_t old;
    asm volatile("atom.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}

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

#ifndef LSB_NODE_EVICT_LAST
#define LSB_NODE_EVICT_LAST 1
#endif
#ifndef LSB_LEAF_PAIR
#define LSB_LEAF_PAIR 0          // 1: leaves of <= 2 triangles load both records before testing either
#endif
#ifndef LSB_BRANCHY_INTERSECT
#define LSB_BRANCHY_INTERSECT 0   // 1: Moeller-Trumbore with its five early exits (round 1); 0: one predicate
#endif
// L2 eviction priority for the BVH nodes (the results stream through L2 with evict-first stores)
__device__ __forceinline__ uint64_t l2_policy_evict_last()
{
    uint64_t p;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ float4 ldg_f4_hint(const float4 *a, uint64_t pol)
{
    float4 v;
    asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ float2 ldg_f2_hint(const float2 *a, uint64_t pol)
{
    float2 v;
    asm volatile("ld.global.nc.L2::cache_hint.v2.f32 {%0, %1}, [%2], %3;" : "=f"(v.x), "=f"(v.y) : "l"(a), "l"(pol));
    return v;
}

struct Ray {
    float ox, oy, oz, dx, dy, dz;
};
#ifndef LSB_EXACT_RCP
#define LSB_EXACT_RCP 1          // 1: IEEE 1 / d for the slab set-up; 0: MUFU.RCP (measured: no gain, 10 344 - 10 410 vs 10 312 - 10 380 Mrays/s)
#endif
__device__ __forceinline__ float rcp_approx(float x)
{
#if LSB_EXACT_RCP
    return 1.0f / x;
#else
    float y;
    asm("rcp.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
#endif
}
template <bool WT = false>
__device__ __forceinline__ void brute_ray(const TraceParams &P, const Ray &r, float &best, int &bi);   // depth-guard fallback
constexpr int kRayOverflow = -2;       // trace_ray_lz: "my stack overflowed" (never a triangle id)

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
#if LSB_BRANCHY_INTERSECT
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
#else
    // the same tests as one predicate: in a warp whose lanes test different triangles the five early exits save nothing
    // (some lane always goes on) and cost a divergence stack push / pop each -- 16 % of k_trace's instructions were
    // BRA / BSSY / BSYNC / BREAK.  Same arithmetic, same accept set: den = 0 gives t = NaN or inf, rejected by the last test.
    float T = __fadd_rn(__fadd_rn(__fmul_rn(ngx, cx), __fmul_rn(ngy, cy)), __fmul_rn(ngz, cz));
    if (neg) T = -T;
    const float t = __fdiv_rn(T, absDen);
    const bool ok = (den != 0.0f) & (U >= 0.0f) & (V >= 0.0f) & (__fadd_rn(U, V) <= absDen) & (__fmul_rn(absDen, 0.0f) < T) &
                    (t <= 3.402823466e+38f);
    t_out = t;
    return ok;
#endif
}

// Watertight ray / triangle test (Woop, Benthin, Wald 2013, "Watertight Ray/Triangle Intersection"), the optional
// LSB200_OPT_WATERTIGHT mode; bit-identical to oracle/lsb_oracle.c: lso_intersect_wt.  The triangle record of this
// mode holds the three ORIGINAL vertices (q0 = v0.xyz v1.x, q1 = v1.yz v2.xy, q2 = v2.z): two triangles that share
// an edge then evaluate that edge's function on bit-identical operands with opposite sign, so a ray cannot pass
// between them.  Ray-space set-up (dominant axis kz, shear Sx Sy, scale Sz) is recomputed from the direction (cheap,
// and it keeps six registers out of the traversal loop); the edge functions fall back to exact double products when
// one of them is 0 in fp32.  No back-face culling, tnear = 0 (strict), tfar = +inf, like the default test.
__device__ __forceinline__ float sel3(int k, float x, float y, float z) { return k == 0 ? x : (k == 1 ? y : z); }

__device__ __forceinline__ bool intersect_tri_wt(const float4 q0, const float4 q1, const float4 q2, const Ray &r, float &t_out)
{
    const float ax = fabsf(r.dx), ay = fabsf(r.dy), az = fabsf(r.dz);
    int kz = (ax > ay) ? (ax > az ? 0 : 2) : (ay > az ? 1 : 2);
    int kx = kz == 2 ? 0 : kz + 1, ky = kx == 2 ? 0 : kx + 1;
    const float dkz = sel3(kz, r.dx, r.dy, r.dz);
    if (dkz < 0.0f) { const int tmp = kx; kx = ky; ky = tmp; }
    const float Sx = __fdiv_rn(sel3(kx, r.dx, r.dy, r.dz), dkz), Sy = __fdiv_rn(sel3(ky, r.dx, r.dy, r.dz), dkz);
    const float Sz = __fdiv_rn(1.0f, dkz);
    const float Ax_ = __fsub_rn(q0.x, r.ox), Ay_ = __fsub_rn(q0.y, r.oy), Az_ = __fsub_rn(q0.z, r.oz);
    const float Bx_ = __fsub_rn(q0.w, r.ox), By_ = __fsub_rn(q1.x, r.oy), Bz_ = __fsub_rn(q1.y, r.oz);
    const float Cx_ = __fsub_rn(q1.z, r.ox), Cy_ = __fsub_rn(q1.w, r.oy), Cz_ = __fsub_rn(q2.x, r.oz);
    const float Akz = sel3(kz, Ax_, Ay_, Az_), Bkz = sel3(kz, Bx_, By_, Bz_), Ckz = sel3(kz, Cx_, Cy_, Cz_);
    const float Ax = __fsub_rn(sel3(kx, Ax_, Ay_, Az_), __fmul_rn(Sx, Akz)), Ay = __fsub_rn(sel3(ky, Ax_, Ay_, Az_), __fmul_rn(Sy, Akz));
    const float Bx = __fsub_rn(sel3(kx, Bx_, By_, Bz_), __fmul_rn(Sx, Bkz)), By = __fsub_rn(sel3(ky, Bx_, By_, Bz_), __fmul_rn(Sy, Bkz));
    const float Cx = __fsub_rn(sel3(kx, Cx_, Cy_, Cz_), __fmul_rn(Sx, Ckz)), Cy = __fsub_rn(sel3(ky, Cx_, Cy_, Cz_), __fmul_rn(Sy, Ckz));
    float U = __fsub_rn(__fmul_rn(Cx, By), __fmul_rn(Cy, Bx));
    float V = __fsub_rn(__fmul_rn(Ax, Cy), __fmul_rn(Ay, Cx));
    float W = __fsub_rn(__fmul_rn(Bx, Ay), __fmul_rn(By, Ax));
    if (U == 0.0f || V == 0.0f || W == 0.0f) {           // on (or within fp32 rounding of) an edge: exact products
        U = (float)__dsub_rn(__dmul_rn((double)Cx, (double)By), __dmul_rn((double)Cy, (double)Bx));
        V = (float)__dsub_rn(__dmul_rn((double)Ax, (double)Cy), __dmul_rn((double)Ay, (double)Cx));
        W = (float)__dsub_rn(__dmul_rn((double)Bx, (double)Ay), __dmul_rn((double)By, (double)Ax));
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    const float det = __fadd_rn(__fadd_rn(U, V), W);
    if (!(det != 0.0f)) return false;
    const float Az = __fmul_rn(Sz, Akz), Bz = __fmul_rn(Sz, Bkz), Cz = __fmul_rn(Sz, Ckz);
    const float T = __fadd_rn(__fadd_rn(__fmul_rn(U, Az), __fmul_rn(V, Bz)), __fmul_rn(W, Cz));
    const float t = __fdiv_rn(T, det);
    if (!(t > 0.0f)) return false;                         // tnear = 0, strict; also rejects NaN
    if (!(t <= 3.402823466e+38f)) return false;            // tfar = +inf
    t_out = t;
    return true;
}

template <bool WT = false>
__device__ __forceinline__ void test_leaf(const TraceParams &P, int ref, const Ray &r, float &best, int &bi)
{
    const int enc = ~ref;
    const int first = enc & ((1 << kLeafShift) - 1);
    const int count = (enc >> kLeafShift) + 1;
#if LSB_LEAF_PAIR == 2
    // the second record of a two-triangle leaf is requested (into L1, no registers) while the first is loaded and tested
    if (count >= 2) {
        const float4 *nx = P.tris + 3 * (size_t)first + 3;
        asm volatile("prefetch.global.L1 [%0];" ::"l"(nx));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(nx + 2));
    }
#endif
#if LSB_LEAF_PAIR == 1
    if (count <= 2) {
        // the default leaf holds one or two triangles: both records are requested before either is tested, so a leaf costs
        // ONE memory round trip instead of two, and there is no loop to diverge on
        const float4 *rec = P.tris + 3 * (size_t)first;
        const bool two = count == 2;
        const float4 a0 = __ldg(rec), a1 = __ldg(rec + 1), a2 = __ldg(rec + 2);
        const int ga = __ldg(P.leaf_ids + first);
        float4 b0 = a0, b1 = a1, b2 = a2;
        int gb = ga;
        if (two) { b0 = __ldg(rec + 3); b1 = __ldg(rec + 4); b2 = __ldg(rec + 5); gb = __ldg(P.leaf_ids + first + 1); }
        float t;
        if ((WT ? intersect_tri_wt(a0, a1, a2, r, t) : intersect_tri(a0, a1, a2, r, t)) && t <= best) {
            if (t < best || ga < bi) { best = t; bi = ga; }
        }
        if (two && (WT ? intersect_tri_wt(b0, b1, b2, r, t) : intersect_tri(b0, b1, b2, r, t)) && t <= best) {
            if (t < best || gb < bi) { best = t; bi = gb; }
        }
        return;
    }
#endif
    for (int k = 0; k < count; ++k) {
        const float4 *rec = P.tris + 3 * (size_t)(first + k);
#if LSB_NODE_EVICT_LAST >= 2
        const uint64_t pol = l2_policy_evict_last();
        const float4 q0 = ldg_f4_hint(rec, pol), q1 = ldg_f4_hint(rec + 1, pol), q2 = ldg_f4_hint(rec + 2, pol);
#else
        const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1), q2 = __ldg(rec + 2);
#endif
        const int g = __ldg(P.leaf_ids + first + k);        // with the record: no dependent round trip after a hit
        float t;
        if ((WT ? intersect_tri_wt(q0, q1, q2, r, t) : intersect_tri(q0, q1, q2, r, t)) && t <= best) {
            if (t < best || g < bi) { best = t; bi = g; }   // lexicographic minimum of (t, global triangle id)
        }
    }
}

struct RayStats { unsigned nodes, leaves, tris, max_sp, culled, chain_tests; };

// Entry / exit distance of one child box from the six slab products.  OCT 0..7: the signs of (dx, dy, dz) are
// known (bit k set = component k negative), so the near and far plane of every slab are picked by name instead
// of by min/max -- 6 FMNMX fewer per child.  OCT 8: generic min/max form (also relaxes the exit by 4 ulp).
template <int OCT>
__device__ __forceinline__ void slab_nf(float l0, float h0, float l1, float h1, float l2, float h2, float far_lim,
                                        float &n, float &f)
{
    if (OCT < 8) {
        n = fmaxf(fmaxf((OCT & 1) ? h0 : l0, (OCT & 2) ? h1 : l1), fmaxf((OCT & 4) ? h2 : l2, 0.0f));
        f = fminf(fminf((OCT & 1) ? l0 : h0, (OCT & 2) ? l1 : h1), fminf((OCT & 4) ? l2 : h2, far_lim));
    } else {
        n = fmaxf(fmaxf(fminf(l0, h0), fminf(l1, h1)), fmaxf(fminf(l2, h2), 0.0f));
        f = fminf(fminf(fmaxf(l0, h0), fmaxf(l1, h1)), fminf(fmaxf(l2, h2), far_lim)) * 1.0000005f;
    }
}

// Origin chain of one sensor (k_origin_chain): every ray of a LiDAR starts at the same point, so the root-to-leaf
// chain of nodes whose box contains that point is entered by all of its rays.  Instead of walking it ray by ray
// (one 56-byte node visit per level, the chain child trivially hit), the chain is cut out once per sensor and
// frame: what remains is the list of sibling subtrees that hang off it, shallow (far) to deep (near), plus the
// subtree or leaf the chain ends in.  A ray tests the boxes of that list (32 bytes each, no divergence) and
// seeds its stack with the ones it enters.  The union of the listed subtrees is the whole scene and closest hit
// = min (t, triangle id) does not depend on the visiting order, so results are unchanged.
// kMaxChain (entries per list, 2 x float4 each) and the walk itself live in build.cuh: origin_chain_walk.

#ifdef LSB_LAB   // evidence variants (modes 0, 1, 2, 4, 5): built into liblsb200_lab.so only
// Closest hit of one ray.  TOP: node references below n_top are read from the shared-memory copy.
// OCT 0..7: sign-specialised slab test (the caller guarantees the ray's direction signs); 8: generic.
// chain / n_chain: the sensor's origin-chain list (n_chain == 0: start at the root).
// Every form only ever rejects a box that the eps-widened exact slab test rejects, so the per-ray results do
// not depend on the form.
template <bool TOP, int OCT, bool STATS>
__device__ __forceinline__ void trace_ray(const TraceParams &P, const float4 *s_top, int n_top, float eps, const Ray &r,
                                          const float4 *__restrict__ chain, int n_chain,
                                          float &best_out, int &bi_out, RayStats &rs)
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
    float far_lim = INFINITY;          // best * (1 + 4 ulp), refreshed when best changes
    int bi = -1;
    // stack entries carry the entry distance of the deferred child: when the entry is popped after a closer
    // hit has been found it is dropped without loading the node
    int2 stack[kStackSize];
    int sp = 0;
    bool ovf = false;
    if (n_chain > 0) {
        for (int k = 0; k < n_chain; ++k) {
            const float4 e0 = __ldg(chain + 2 * k), e1 = __ldg(chain + 2 * k + 1);
            float n, f;
            slab_nf<OCT>(fmaf(e0.x, ix, ax), fmaf(e0.y, ix, bx), fmaf(e0.z, iy, ay), fmaf(e0.w, iy, by),
                         fmaf(e1.x, iz, az), fmaf(e1.y, iz, bz), far_lim, n, f);
            if (STATS) ++rs.chain_tests;
            if (n <= f) { if (sp < kStackSize) stack[sp++] = make_int2(__float_as_int(e1.z), __float_as_int(n)); else ovf = true; }
        }
        if (STATS) rs.max_sp = max(rs.max_sp, (unsigned)sp);
    } else {
        { if (sp < kStackSize) stack[sp++] = make_int2((TOP && n_top > 0) ? 0 : cap, 0); else ovf = true; }      // root: top slot 0, or Karras node 0
    }
    while (true) {
        int cur = 0;
        bool found = false;
        while (sp > 0) {
            const int2 e = stack[--sp];
            if (__int_as_float(e.y) <= far_lim * 1.0000005f) { cur = e.x; found = true; break; }
            if (STATS) ++rs.culled;
        }
        if (!found) break;
        while (true) {                                               // descend until a leaf or a dead end
            if (cur < 0) {
                if (STATS) { ++rs.leaves; rs.tris += (unsigned)(((~cur) >> kLeafShift) + 1); }
                test_leaf(P, cur, r, best, bi);
                far_lim = best * 1.0000005f;
                break;
            }
            float4 q0, q1, q2, q3;
            if (TOP && cur < cap) {
                q0 = s_top[cur]; q1 = s_top[cap + cur]; q2 = s_top[2 * cap + cur]; q3 = s_top[3 * cap + cur];
            } else {
                const float4 *n = P.nodes + 4 * (size_t)cur;
                q0 = __ldg(n); q1 = __ldg(n + 1); q2 = __ldg(n + 2); q3 = __ldg(n + 3);
            }
            if (STATS) ++rs.nodes;
            float n0, f0, n1, f1;
            slab_nf<OCT>(fmaf(q0.x, ix, ax), fmaf(q0.y, ix, bx), fmaf(q0.z, iy, ay), fmaf(q0.w, iy, by),
                         fmaf(q2.x, iz, az), fmaf(q2.y, iz, bz), far_lim, n0, f0);
            slab_nf<OCT>(fmaf(q1.x, ix, ax), fmaf(q1.y, ix, bx), fmaf(q1.z, iy, ay), fmaf(q1.w, iy, by),
                         fmaf(q2.z, iz, az), fmaf(q2.w, iz, bz), far_lim, n1, f1);
            const bool hit0 = n0 <= f0, hit1 = n1 <= f1;
            const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            if (hit0 && hit1) {
                const bool swap = n1 < n0;
                cur = swap ? c1 : c0;
                { if (sp < kStackSize) stack[sp++] = make_int2(swap ? c0 : c1, __float_as_int(swap ? n0 : n1)); else ovf = true; }
                if (STATS) rs.max_sp = max(rs.max_sp, (unsigned)sp);
            } else if (hit0) {
                cur = c0;
            } else if (hit1) {
                cur = c1;
            } else {
                break;
            }
        }
    }
    if (ovf) brute_ray(P, r, best, bi);          // depth guard tripped (see trace_ray_lz)
    best_out = best;
    bi_out = bi;
}

// "while-while" form of trace_ray (Aila & Laine 2009): an inner loop that only visits nodes, left by a lane when
// it reaches a leaf (or runs out of work); the warp reconverges behind it, so the (long) triangle tests of the
// lanes run together instead of being interleaved with other lanes' node visits.  Nodes from global memory.
template <int OCT, bool STATS, bool SMEM>
__device__ __forceinline__ void trace_ray_ww(const TraceParams &P, float eps, const Ray &r,
                                             const float4 *__restrict__ chain, int n_chain, bool reversed,
                                             float &best_out, int &bi_out, RayStats &rs)
{
    const float tiny = 1e-30f;
    const float ix = 1.0f / (fabsf(r.dx) < tiny ? copysignf(tiny, r.dx) : r.dx);
    const float iy = 1.0f / (fabsf(r.dy) < tiny ? copysignf(tiny, r.dy) : r.dy);
    const float iz = 1.0f / (fabsf(r.dz) < tiny ? copysignf(tiny, r.dz) : r.dz);
    const float ax = -(r.ox + eps) * ix, ay = -(r.oy + eps) * iy, az = -(r.oz + eps) * iz;
    const float bx = -(r.ox - eps) * ix, by = -(r.oy - eps) * iy, bz = -(r.oz - eps) * iz;
    constexpr int kDone = 0x7fffffff;       // never a node index (n_tris < 2^28)
    float best = INFINITY, far_lim = INFINITY;
    int bi = -1;
    int2 stack[kStackSize];
    int sp = 0;
    bool ovf = false;
    int cur = P.top_cap;
    if (n_chain > 0) {
        // sensor chains are listed far -> near, tile lists (k_tile_entries) near -> far: the nearest entry ends on top
        const float4 *e = reversed ? chain + 2 * (n_chain - 1) : chain;
        const int step = reversed ? -2 : 2;
        for (int k = 0; k < n_chain; ++k, e += step) {
            // SMEM: the warp staged the list in shared memory with one coalesced load (process_unit)
            const float4 e0 = SMEM ? e[0] : __ldg(e), e1 = SMEM ? e[1] : __ldg(e + 1);
            float n, f;
            slab_nf<OCT>(fmaf(e0.x, ix, ax), fmaf(e0.y, ix, bx), fmaf(e0.z, iy, ay), fmaf(e0.w, iy, by),
                         fmaf(e1.x, iz, az), fmaf(e1.y, iz, bz), far_lim, n, f);
            if (STATS) ++rs.chain_tests;
            if (n <= f) { if (sp < kStackSize) stack[sp++] = make_int2(__float_as_int(e1.z), __float_as_int(n)); else ovf = true; }
        }
        if (STATS) rs.max_sp = max(rs.max_sp, (unsigned)sp);
        cur = sp > 0 ? stack[--sp].x : kDone;
    }
    while (cur != kDone) {
        while ((unsigned)cur < (unsigned)kDone) {           // cur >= 0 and not kDone: an inner node
            const float4 *n = P.nodes + 4 * (size_t)cur;
            const float4 q0 = __ldg(n), q1 = __ldg(n + 1), q2 = __ldg(n + 2);
            const float2 q3 = __ldg(reinterpret_cast<const float2 *>(n + 3));
            if (STATS) ++rs.nodes;
            float n0, f0, n1, f1;
            slab_nf<OCT>(fmaf(q0.x, ix, ax), fmaf(q0.y, ix, bx), fmaf(q0.z, iy, ay), fmaf(q0.w, iy, by),
                         fmaf(q2.x, iz, az), fmaf(q2.y, iz, bz), far_lim, n0, f0);
            slab_nf<OCT>(fmaf(q1.x, ix, ax), fmaf(q1.y, ix, bx), fmaf(q1.z, iy, ay), fmaf(q1.w, iy, by),
                         fmaf(q2.z, iz, az), fmaf(q2.w, iz, bz), far_lim, n1, f1);
            const bool hit0 = n0 <= f0, hit1 = n1 <= f1;
            const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            if (hit0 && hit1) {
                const bool swap = n1 < n0;
                cur = swap ? c1 : c0;
                { if (sp < kStackSize) stack[sp++] = make_int2(swap ? c0 : c1, __float_as_int(swap ? n0 : n1)); else ovf = true; }
                if (STATS) rs.max_sp = max(rs.max_sp, (unsigned)sp);
            } else if (hit0) {
                cur = c0;
            } else if (hit1) {
                cur = c1;
            } else {
                cur = kDone;
                while (sp > 0) {
                    const int2 e = stack[--sp];
                    if (__int_as_float(e.y) <= far_lim * 1.0000005f) { cur = e.x; break; }
                    if (STATS) ++rs.culled;
                }
            }
        }
        if (cur < 0) {
            if (STATS) { ++rs.leaves; rs.tris += (unsigned)(((~cur) >> kLeafShift) + 1); }
            test_leaf(P, cur, r, best, bi);
            far_lim = best * 1.0000005f;
            cur = kDone;
            while (sp > 0) {
                const int2 e = stack[--sp];
                if (__int_as_float(e.y) <= far_lim * 1.0000005f) { cur = e.x; break; }
                if (STATS) ++rs.culled;
            }
        }
    }
    if (ovf) brute_ray(P, r, best, bi);          // depth guard tripped (see trace_ray_lz)
    best_out = best;
    bi_out = bi;
}

#endif  // LSB_LAB


// Lazy-list form of trace_ray_ww (variant 6 with LSB_LAZY_LIST): instead of testing every list entry up front and
// pushing the ones the ray enters (one 8-byte local-memory store per entry and lane -- written through to L2: 2.6 GB
// of the 4.3 GB the SMs sent to L2 per launch -- and one load per pop), the list itself is the bottom of the stack:
// entries are taken in near -> far order only when the ray has nothing else left, and tested against the hit
// distance known by then, so whatever lies behind the first hit fails its own box test.  The stack holds deferred
// children only.  Same set of subtrees entered or rejected by conservative tests: same results.
template <int OCT, bool STATS, bool SMEM, bool WT = false>
__device__ __forceinline__ void trace_ray_lz(const TraceParams &P, float eps, const Ray &r,
                                             const float4 *__restrict__ chain, int n_chain, bool reversed,
                                             float &best_out, int &bi_out, RayStats &rs, int2 *s_stack = nullptr)
{
    // 1 / d only feeds the BOX tests (rcp_approx: IEEE by default; the MUFU.RCP form would be safe -- <= 1.2e-7 relative
    // against a box padding of 8e-6 x the scene's extent -- but measured no faster, so the exact quotient stays)
    const float tiny = 1e-30f;
    const float ix = rcp_approx(fabsf(r.dx) < tiny ? copysignf(tiny, r.dx) : r.dx);
    const float iy = rcp_approx(fabsf(r.dy) < tiny ? copysignf(tiny, r.dy) : r.dy);
    const float iz = rcp_approx(fabsf(r.dz) < tiny ? copysignf(tiny, r.dz) : r.dz);
    const float ax = -(r.ox + eps) * ix, ay = -(r.oy + eps) * iy, az = -(r.oz + eps) * iz;
    const float bx = -(r.ox - eps) * ix, by = -(r.oy - eps) * iy, bz = -(r.oz - eps) * iz;
    constexpr int kDone = 0x7fffffff;       // never a node index (n_tris < 2^28)
#if LSB_NODE_EVICT_LAST
    const uint64_t pol = l2_policy_evict_last();
#endif
    float best = INFINITY, far_lim = INFINITY;
    int bi = -1;
    int2 stack[kStackSize];
    int sp = 0;
    // Depth guard: a Karras tree over n coincident centroids is ~32 + log2(n) deep, so kStackSize deferred children are
    // enough for anything but an adversarial scene.  A ray that would overflow drops the push, remembers it (sp gets
    // the sentinel bit kOvf, popped entries are unaffected) and is re-done by brute force below: never an
    // out-of-bounds write, never a wrong result.
    constexpr int kOvf = 1 << 20;
#if LSB_SMEM_STACK
    // the first LSB_SMEM_STACK entries of the stack live in shared memory ([slot][thread]: no bank conflicts), deeper
    // ones in the local array
    auto push = [&](int2 v) {
        const int d = sp & (kOvf - 1);
        if (d < LSB_SMEM_STACK) s_stack[d * kTraceThreads] = v;
        else if (d < kStackSize) stack[d] = v;
        else { sp |= kOvf; return; }
        ++sp;
    };
    auto pop = [&]() -> int2 { --sp; const int d = sp & (kOvf - 1); return d < LSB_SMEM_STACK ? s_stack[d * kTraceThreads] : stack[d]; };
#else
    auto push = [&](int2 v) { const int d = sp & (kOvf - 1); if (d < kStackSize) { stack[d] = v; ++sp; } else sp |= kOvf; };
    auto pop = [&]() -> int2 { --sp; return stack[sp & (kOvf - 1)]; };
#endif
    auto depth = [&]() -> int { return sp & (kOvf - 1); };
    // sensor chains are listed far -> near (walked backwards), tile lists near -> far
    const float4 *e = reversed ? chain : chain + 2 * (n_chain - 1);
    const int step = reversed ? 2 : -2;
    int left = n_chain;                     // list entries not looked at yet
    int cur = n_chain > 0 ? kDone : P.top_cap;
    while (true) {
        if (cur == kDone) {                 // nothing in hand: next list entry the ray enters
            while (left > 0) {
                const float4 e0 = SMEM ? e[0] : __ldg(e), e1 = SMEM ? e[1] : __ldg(e + 1);
                e += step; --left;
                float n, f;
                slab_nf<OCT>(fmaf(e0.x, ix, ax), fmaf(e0.y, ix, bx), fmaf(e0.z, iy, ay), fmaf(e0.w, iy, by),
                             fmaf(e1.x, iz, az), fmaf(e1.y, iz, bz), far_lim, n, f);
                if (STATS) ++rs.chain_tests;
                if (n <= f) { cur = __float_as_int(e1.z); break; }
            }
            if (cur == kDone) break;
        }
        while ((unsigned)cur < (unsigned)kDone) {           // cur >= 0 and not kDone: an inner node
            const float4 *n = P.nodes + 4 * (size_t)cur;
#if LSB_NODE_EVICT_LAST
            const float4 q0 = ldg_f4_hint(n, pol), q1 = ldg_f4_hint(n + 1, pol), q2 = ldg_f4_hint(n + 2, pol);
            const float2 q3 = ldg_f2_hint(reinterpret_cast<const float2 *>(n + 3), pol);
#else
            const float4 q0 = __ldg(n), q1 = __ldg(n + 1), q2 = __ldg(n + 2);
            const float2 q3 = __ldg(reinterpret_cast<const float2 *>(n + 3));
#endif
            if (STATS) ++rs.nodes;
            float n0, f0, n1, f1;
            slab_nf<OCT>(fmaf(q0.x, ix, ax), fmaf(q0.y, ix, bx), fmaf(q0.z, iy, ay), fmaf(q0.w, iy, by),
                         fmaf(q2.x, iz, az), fmaf(q2.y, iz, bz), far_lim, n0, f0);
            slab_nf<OCT>(fmaf(q1.x, ix, ax), fmaf(q1.y, ix, bx), fmaf(q1.z, iy, ay), fmaf(q1.w, iy, by),
                         fmaf(q2.z, iz, az), fmaf(q2.w, iz, bz), far_lim, n1, f1);
            const bool hit0 = n0 <= f0, hit1 = n1 <= f1;
            const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            if (hit0 && hit1) {
                const bool swap = n1 < n0;
                cur = swap ? c1 : c0;
                push(make_int2(swap ? c0 : c1, __float_as_int(swap ? n0 : n1)));
                if (STATS) rs.max_sp = max(rs.max_sp, (unsigned)depth());
            } else if (hit0) {
                cur = c0;
            } else if (hit1) {
                cur = c1;
            } else {
                cur = kDone;
                while (depth() > 0) {
                    const int2 t = pop();
                    if (__int_as_float(t.y) <= far_lim * 1.0000005f) { cur = t.x; break; }
                    if (STATS) ++rs.culled;
                }
            }
        }
        if (cur < 0) {
            if (STATS) { ++rs.leaves; rs.tris += (unsigned)(((~cur) >> kLeafShift) + 1); }
            test_leaf<WT>(P, cur, r, best, bi);
            far_lim = best * 1.0000005f;
            cur = kDone;
            while (depth() > 0) {
                const int2 t = pop();
                if (__int_as_float(t.y) <= far_lim * 1.0000005f) { cur = t.x; break; }
                if (STATS) ++rs.culled;
            }
        }
    }
    if (sp & kOvf) bi = kRayOverflow;                 // depth guard tripped: process_unit re-does this ray exhaustively
    best_out = best;
    bi_out = bi;
}

#ifdef LSB_LAB
// Single-loop form for the tile-list regime (variant 7): a step of a lane is whatever its next item is -- a node visit
// or a leaf test -- so a warp needs max-over-lanes(items) dependent memory round trips per unit instead of the
// while-while form's (node rounds + leaf rounds).  More instructions issued (both bodies per mixed step), fewer
// round trips: the right trade once the kernel is latency-bound rather than issue-bound.
template <int OCT, bool STATS, bool SMEM>
__device__ __forceinline__ void trace_ray_il(const TraceParams &P, float eps, const Ray &r,
                                             const float4 *__restrict__ chain, int n_chain, bool reversed,
                                             float &best_out, int &bi_out, RayStats &rs)
{
    const float tiny = 1e-30f;
    const float ix = 1.0f / (fabsf(r.dx) < tiny ? copysignf(tiny, r.dx) : r.dx);
    const float iy = 1.0f / (fabsf(r.dy) < tiny ? copysignf(tiny, r.dy) : r.dy);
    const float iz = 1.0f / (fabsf(r.dz) < tiny ? copysignf(tiny, r.dz) : r.dz);
    const float ax = -(r.ox + eps) * ix, ay = -(r.oy + eps) * iy, az = -(r.oz + eps) * iz;
    const float bx = -(r.ox - eps) * ix, by = -(r.oy - eps) * iy, bz = -(r.oz - eps) * iz;
    constexpr int kDone = 0x7fffffff;
    float best = INFINITY, far_lim = INFINITY;
    int bi = -1;
    int2 stack[kStackSize];
    int sp = 0;
    bool ovf = false;
    int cur = P.top_cap;
    if (n_chain > 0) {
        const float4 *e = reversed ? chain + 2 * (n_chain - 1) : chain;
        const int step = reversed ? -2 : 2;
        for (int k = 0; k < n_chain; ++k, e += step) {
            const float4 e0 = SMEM ? e[0] : __ldg(e), e1 = SMEM ? e[1] : __ldg(e + 1);
            float n, f;
            slab_nf<OCT>(fmaf(e0.x, ix, ax), fmaf(e0.y, ix, bx), fmaf(e0.z, iy, ay), fmaf(e0.w, iy, by),
                         fmaf(e1.x, iz, az), fmaf(e1.y, iz, bz), far_lim, n, f);
            if (STATS) ++rs.chain_tests;
            if (n <= f) { if (sp < kStackSize) stack[sp++] = make_int2(__float_as_int(e1.z), __float_as_int(n)); else ovf = true; }
        }
        if (STATS) rs.max_sp = max(rs.max_sp, (unsigned)sp);
        cur = sp > 0 ? stack[--sp].x : kDone;
    }
    while (cur != kDone) {
        bool pop = false;
        if (cur >= 0) {
            const float4 *n = P.nodes + 4 * (size_t)cur;
            const float4 q0 = __ldg(n), q1 = __ldg(n + 1), q2 = __ldg(n + 2);
            const float2 q3 = __ldg(reinterpret_cast<const float2 *>(n + 3));
            if (STATS) ++rs.nodes;
            float n0, f0, n1, f1;
            slab_nf<OCT>(fmaf(q0.x, ix, ax), fmaf(q0.y, ix, bx), fmaf(q0.z, iy, ay), fmaf(q0.w, iy, by),
                         fmaf(q2.x, iz, az), fmaf(q2.y, iz, bz), far_lim, n0, f0);
            slab_nf<OCT>(fmaf(q1.x, ix, ax), fmaf(q1.y, ix, bx), fmaf(q1.z, iy, ay), fmaf(q1.w, iy, by),
                         fmaf(q2.z, iz, az), fmaf(q2.w, iz, bz), far_lim, n1, f1);
            const bool hit0 = n0 <= f0, hit1 = n1 <= f1;
            const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            if (hit0 && hit1) {
                const bool swap = n1 < n0;
                cur = swap ? c1 : c0;
                { if (sp < kStackSize) stack[sp++] = make_int2(swap ? c0 : c1, __float_as_int(swap ? n0 : n1)); else ovf = true; }
                if (STATS) rs.max_sp = max(rs.max_sp, (unsigned)sp);
            } else if (hit0) {
                cur = c0;
            } else if (hit1) {
                cur = c1;
            } else {
                pop = true;
            }
        } else {
            if (STATS) { ++rs.leaves; rs.tris += (unsigned)(((~cur) >> kLeafShift) + 1); }
            test_leaf(P, cur, r, best, bi);
            far_lim = best * 1.0000005f;
            pop = true;
        }
        if (pop) {
            cur = kDone;
            while (sp > 0) {
                const int2 e = stack[--sp];
                if (__int_as_float(e.y) <= far_lim * 1.0000005f) { cur = e.x; break; }
                if (STATS) ++rs.culled;
            }
        }
    }
    if (ovf) brute_ray(P, r, best, bi);          // depth guard tripped (see trace_ray_lz)
    best_out = best;
    bi_out = bi;
}

#endif  // LSB_LAB

// Brute force over all leaf-ordered triangles (validation variant; also the tiny-scene path).
template <bool WT>
__device__ __forceinline__ void brute_ray(const TraceParams &P, const Ray &r, float &best, int &bi)
{
    best = INFINITY; bi = -1;
    for (int k = 0; k < P.n_tris; ++k) {
        const float4 *rec = P.tris + 3 * (size_t)k;
        float t;
        const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1), q2 = __ldg(rec + 2);
        if ((WT ? intersect_tri_wt(q0, q1, q2, r, t) : intersect_tri(q0, q1, q2, r, t)) && t <= best) {
            const int g = __ldg(P.leaf_ids + k);
            if (t < best || g < bi) { best = t; bi = g; }
        }
    }
}

// Out-of-line copy for the depth-guard fallback (adversarial scenes only): kept out of the traversal's register budget.
template <bool WT>
__device__ __noinline__ void brute_ray_cold(const float4 *tris, const int32_t *leaf_ids, int n_tris, const Ray &r, float &best, int &bi)
{
    TraceParams P{};
    P.tris = tris; P.leaf_ids = leaf_ids; P.n_tris = n_tris;
    brute_ray<WT>(P, r, best, bi);
}

__device__ __forceinline__ int find_sensor_by_unit(const SensorDev *s, int n, int unit)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (s[mid].unit_base <= unit) lo = mid; else hi = mid - 1; }
    return lo;
}

// Writes one hit.  32-byte record (SRC/XYZIRBytes.cpp:24-40) or 16-byte XYZI.
__device__ __forceinline__ void store_point(uint8_t *points, uint32_t slot, uint32_t flags, float x, float y, float z, int ring,
                                            float intensity = 64.0f /* the reference's constant, SRC/EmbreeTracer.cpp:333 */)
{
    // streaming stores: the 1.2 GB of results per frame should not evict the BVH from L2
    if (flags & 4u /*LSB200_TRACE_XYZI16*/) {
        __stcs(reinterpret_cast<float4 *>(points) + slot, make_float4(x, y, z, intensity));
    } else {
        float4 *p = reinterpret_cast<float4 *>(points) + 2 * (size_t)slot;
        __stcs(p, make_float4(x, y, z, 0.0f));
        __stcs(p + 1, make_float4(intensity, __int_as_float(ring), 0.0f, 0.0f));
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

// Ray `ray` of sensor S: d_s = (sin t cos p, sin t sin p, cos t) (SRC/LidarDevice.cpp:319-321), ray id = channel * H
// + azimuth; world-frame mode moves the ray instead of the scene (o = T, d = R d_s).  Used by the traversal AND by
// the octant classification, so both see bit-identical direction signs.
__device__ __forceinline__ void ray_from_tables(const SensorDev &S, int world_frame, float st, float ct, float sp_, float cp, Ray &r,
                                                float &dsx, float &dsy, float &dsz, float &omax)
{
    dsx = __fmul_rn(st, cp); dsy = __fmul_rn(st, sp_); dsz = ct;
    if (world_frame) {
        r.ox = S.T[0]; r.oy = S.T[1]; r.oz = S.T[2];
        r.dx = __fadd_rn(__fadd_rn(__fmul_rn(S.R[0], dsx), __fmul_rn(S.R[1], dsy)), __fmul_rn(S.R[2], dsz));
        r.dy = __fadd_rn(__fadd_rn(__fmul_rn(S.R[3], dsx), __fmul_rn(S.R[4], dsy)), __fmul_rn(S.R[5], dsz));
        r.dz = __fadd_rn(__fadd_rn(__fmul_rn(S.R[6], dsx), __fmul_rn(S.R[7], dsy)), __fmul_rn(S.R[8], dsz));
        omax = fmaxf(fabsf(r.ox), fmaxf(fabsf(r.oy), fabsf(r.oz)));
    } else {
        r.ox = r.oy = r.oz = 0.0f; r.dx = dsx; r.dy = dsy; r.dz = dsz;
        omax = 0.0f;
    }
}
__device__ __forceinline__ void make_ray(const SensorDev &S, int world_frame, int v, int h, Ray &r, float &dsx, float &dsy,
                                         float &dsz, float &omax)
{
    const float st = __ldg(S.sin_t + v), ct = __ldg(S.cos_t + v), sp_ = __ldg(S.sin_p + h), cp = __ldg(S.cos_p + h);
    ray_from_tables(S, world_frame, st, ct, sp_, cp, r, dsx, dsy, dsz, omax);
}

__device__ __forceinline__ int ray_octant(const Ray &r)
{
    return (int)((__float_as_uint(r.dx) >> 31) | ((__float_as_uint(r.dy) >> 31) << 1) | ((__float_as_uint(r.dz) >> 31) << 2));
}

// Position of work unit `unit`: (sensor, first channel, first azimuth step) of its tile.
__device__ __forceinline__ int4 unit_descriptor(const SensorDev *sensors, int n_sensors, int unit, int tile_w_log2)
{
    const int si = find_sensor_by_unit(sensors, n_sensors, unit);
    const int local = unit - sensors[si].unit_base;
    const int tv = local / sensors[si].tiles_h;
    return make_int4(si, tv << (5 - tile_w_log2), (local - tv * sensors[si].tiles_h) << tile_w_log2, 0);
}

// Lane -> (channel, azimuth step) inside the tile of descriptor ud; false when the lane falls off the table.
__device__ __forceinline__ bool unit_lane_ray(const SensorDev &S, int4 ud, int tile_w_log2, unsigned lane, int &v, int &h)
{
    v = ud.y + (int)(lane >> tile_w_log2);
    h = ud.z + (int)(lane & ((1u << tile_w_log2) - 1u));
    return v < S.V && h < S.H;
}

// A warp's hit append in flight: slots requested (raw = the atomic's return value in the leader lane), points parked in
// the warp's shared-memory row until the next unit has been traced.
struct PendingAppend { uint32_t raw; unsigned mask; int sensor; };
__device__ __forceinline__ void flush_append(const TraceParams &P, PendingAppend &pd, const float4 *s_pts, unsigned lane)
{
    if (pd.mask == 0u) return;
    const uint32_t base = __shfl_sync(0xffffffffu, pd.raw, __ffs(pd.mask) - 1);
    if ((pd.mask >> lane) & 1u) {
        const float4 q = s_pts[lane];
        store_point(P.sensors[pd.sensor].points, base + __popc(pd.mask & lanemask_lt()), P.flags, q.x, q.y, q.z, __float_as_int(q.w));
    }
    pd.mask = 0u;
}

// One 32-ray work unit: generate the rays, trace them, store t / triangle id, append the hits.
// MODE: see k_trace.  OCT: slab form for every ray of the unit (0..7 only when classified uniform).
template <int MODE, int OCT, bool STATS, bool WT = false>
__device__ __forceinline__ void process_unit(const TraceParams &P, const float4 *s_top, int n_top, float smax, int4 ud,
                                             unsigned lane, int tile_n = -1, const float4 *s_list = nullptr,
                                             PendingAppend *pend = nullptr, float4 *s_pts = nullptr, int2 *s_stack = nullptr)
{
#ifdef LSB_LAB
    constexpr bool TOP = (MODE == 4);
#endif
    const int si = ud.x;
    const SensorDev &S = P.sensors[si];
    int v, h;
    const bool valid = unit_lane_ray(S, ud, P.tile_w_log2, lane, v, h);
    const int ray = v * S.H + h;                          // ray id = channel * H + azimuth step
    float best = INFINITY; int bi = -1;
    float dsx = 0.f, dsy = 0.f, dsz = 1.f;
    RayStats rs{0u, 0u, 0u, 0u, 0u, 0u};
    const long long unit_t0 = STATS ? clock64() : 0ll;
    // MODE 6: a unit whose tile list (k_tile_entries, near -> far) was prefetched into shared memory by run_phase
    // starts from it (tile_n entries, 0 = the tile's pyramid reaches nothing); tile_n < 0: no tile list -- the rays
    // read their sensor's origin chain (far -> near) from global memory.
    const bool w_tile = (MODE == 6 || MODE == 7) && tile_n >= 0;
    const bool fired = valid && (S.az_mask == nullptr || __ldg(S.az_mask + h) != 0);
    if (valid && !fired) { S.t_out[ray] = INFINITY; S.tri_out[ray] = -1; }      // outside the azimuth window: no return
    if (fired) {
        Ray r;
        float omax;
        make_ray(S, P.world_frame, v, h, r, dsx, dsy, dsz, omax);
        const float eps = 8e-6f * (smax + omax) + 1e-30f;
        if (MODE == 3) brute_ray<WT>(P, r, best, bi);
        else if (w_tile && tile_n == 0) { /* nothing to hit */ }
        else {
            const int list = si * 9 + OCT;
            const int n_chain = P.chain_counts ? __ldg(P.chain_counts + list) : 0;
#ifdef LSB_LAB
            if (MODE == 7) {
                if (w_tile) trace_ray_il<OCT, STATS, true>(P, eps, r, s_list, tile_n, true, best, bi, rs);
                else trace_ray_il<OCT, STATS, false>(P, eps, r, P.chains + (size_t)list * (2 * kMaxChain), n_chain, false, best, bi, rs);
            } else
            if (MODE == 6 && LSB_LAZY_LIST) {
                if (w_tile) trace_ray_lz<OCT, STATS, true, WT>(P, eps, r, s_list, tile_n, true, best, bi, rs, s_stack);
                else trace_ray_lz<OCT, STATS, false, WT>(P, eps, r, P.chains + (size_t)list * (2 * kMaxChain), n_chain, false, best, bi, rs, s_stack);
            } else
            if (w_tile) trace_ray_ww<OCT, STATS, true>(P, eps, r, s_list, tile_n, true, best, bi, rs);
            else if (MODE == 5 || MODE == 6) trace_ray_ww<OCT, STATS, false>(P, eps, r, P.chains + (size_t)list * (2 * kMaxChain), n_chain, false, best, bi, rs);
            else
            trace_ray<TOP, OCT, STATS>(P, s_top, n_top, eps, r, P.chains + (size_t)list * (2 * kMaxChain), n_chain, best, bi, rs);
#else
            // product build: the one shipped traversal (variant 6, lazy tile / chain list)
            if (w_tile) trace_ray_lz<OCT, STATS, true, WT>(P, eps, r, s_list, tile_n, true, best, bi, rs, s_stack);
            else trace_ray_lz<OCT, STATS, false, WT>(P, eps, r, P.chains + (size_t)list * (2 * kMaxChain), n_chain, false, best, bi, rs, s_stack);
#endif
        }
        if (bi == kRayOverflow) brute_ray_cold<WT>(P.tris, P.leaf_ids, P.n_tris, r, best, bi);
        if (!(best <= S.max_range)) { best = INFINITY; bi = -1; }                // detection radius (closest hit beyond it: no return)
        __stcs(S.t_out + ray, best);
        __stcs(S.tri_out + ray, bi);
    }
    if (STATS) {
        unsigned a = __reduce_add_sync(0xffffffffu, rs.nodes), b = __reduce_add_sync(0xffffffffu, rs.leaves);
        unsigned c = __reduce_add_sync(0xffffffffu, rs.tris), d = __reduce_max_sync(0xffffffffu, rs.max_sp);
        unsigned cu = __reduce_add_sync(0xffffffffu, rs.culled), ct = __reduce_add_sync(0xffffffffu, rs.chain_tests);
        unsigned nv = __reduce_add_sync(0xffffffffu, valid ? 1u : 0u), nh = __reduce_add_sync(0xffffffffu, (valid && bi >= 0) ? 1u : 0u);
        if (lane == 0) {
            const long long cyc = clock64() - unit_t0;
            const int bin = min(23, max(0, 63 - __clzll(cyc > 0 ? cyc : 1)));
            atomicAdd(P.counters + kCntCycUnits + bin, 1ull); atomicAdd(P.counters + kCntCycSum + bin, (unsigned long long)cyc);
            atomicAdd(P.counters + kCntCycNodes + bin, (unsigned long long)a);
            atomicAdd(P.counters + kCntCycList + bin, (unsigned long long)(tile_n < 0 ? 99 : tile_n));
            if (bin >= 17) {
                atomicAdd(P.counters + kCntHeavyClass + OCT, (unsigned long long)cyc);
                atomicAdd(P.counters + kCntHeavyBand + min(31, ud.y * 32 / max(S.V, 1)), (unsigned long long)cyc);
            }
            atomicAdd(P.counters + kCntRays, (unsigned long long)nv); atomicAdd(P.counters + kCntNodes, (unsigned long long)a);
            atomicAdd(P.counters + kCntLeaves, (unsigned long long)b); atomicAdd(P.counters + kCntTris, (unsigned long long)c);
            atomicAdd(P.counters + kCntHits, (unsigned long long)nh); atomicMax(P.counters + kCntMaxStack, (unsigned long long)d);
            atomicAdd(P.counters + kCntCulled, (unsigned long long)cu);
            atomicAdd(P.counters + kCntChainTests, (unsigned long long)ct);
            if (MODE == 0 || MODE == 6 || MODE == 7) atomicAdd(P.counters + (OCT < 8 ? kCntUnitsUniform : kCntUnitsMixed), 1ull);
        }
    }
    if (!(P.flags & (1u | 8u))) {     // neither ORDERED nor NO_POINTS: warp-aggregated atomic append
        const bool hit = valid && bi >= 0;
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if ((MODE == 6 || MODE == 7) && pend != nullptr) {
            // deferred form: the slot reservation of THIS unit is requested now and used after the next unit has been
            // traced (flush_append), so nobody waits for the atomic's round trip; the points wait in shared memory
            flush_append(P, *pend, s_pts, lane);
            if (m) {
                if (hit) {
                    float x, y, z;
                    hit_point(S, P.flags, best, dsx, dsy, dsz, x, y, z);
                    s_pts[lane] = make_float4(x, y, z, __int_as_float((P.flags & 2u) ? 0 : v));
                }
                if (lane == (unsigned)(__ffs(m) - 1)) pend->raw = atom_add_u32(S.count, (uint32_t)__popc(m));
                pend->mask = m; pend->sensor = si;
            }
        } else if (m) {
            uint32_t base = 0;
            if (lane == (unsigned)(__ffs(m) - 1)) base = atom_add_u32(S.count, (uint32_t)__popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            if (hit) {
                float x, y, z;
                hit_point(S, P.flags, best, dsx, dsy, dsz, x, y, z);
                store_point(S.points, base + __popc(m & lanemask_lt()), P.flags, x, y, z, (P.flags & 2u) ? 0 : v);
            }
        }
    }
}

// MODE 0 / 6: drains the queue of octant class OCT (units whose 32 rays all have the direction signs OCT; 8 = mixed).
// An empty class costs no atomic.
// The loop is software-pipelined, because with the tile lists the kernel is bound by the chain of dependent memory
// round trips at the head of every unit (ticket -> descriptor / list size -> list -> ...), not by the SM's load path:
// the ticket is requested two units ahead, descriptor and list size one unit ahead, and the unit's tile list is
// copied into the warp's other shared-memory buffer with cp.async (one 16-byte piece per lane, no registers) while
// the current unit is traced.
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int MODE, int OCT, bool STATS, bool WT = false>
__device__ __forceinline__ void run_phase(const TraceParams &P, float smax, unsigned lane, float4 *s_list, float4 *s_pts, int2 *s_stack,
                                          int part = 0)
{
    int begin = 0;
#pragma unroll
    for (int o = 0; o < OCT; ++o) begin += (int)__ldg(P.oct_hist + o);
    const int count = (int)__ldg(P.oct_hist + OCT);
    if (count == 0) return;
    if (MODE != 6 && MODE != 7) {                                  // variant 0 keeps the plain loop it was measured with
        while (true) {
            int k = 0;
            if (lane == 0) k = (int)atomicAdd(&P.work[2 + OCT], 1u);
            k = __shfl_sync(0xffffffffu, k, 0);
            if (k >= count) break;
            process_unit<MODE, OCT, STATS, WT>(P, nullptr, 0, smax, __ldg(P.unit_list + begin + k), lane);
        }
        return;
    }
    const bool lists = (MODE == 6 || MODE == 7) && (OCT < 8 || LSB_MIXED_TILE_LISTS) && P.tile_counts != nullptr;
    constexpr int kBuf = 2 * kMaxTileEntries;                     // float4 per buffer
    // part 0: the whole queue; 1 / 2: its heavy front (k_scatter_units' front cursor = number of heavy units of the class) /
    // the rest (own ticket counters) -- see the phase order in k_trace
    const int n_front = LSB_HEAVY_FIRST ? min(count, (int)__ldg(P.oct_hist + 16 + OCT)) : (int)(((long long)count * 3) / 8);
    const int lo = part == 2 ? n_front : 0, hi = part == 1 ? n_front : count;
    if (lo >= hi) return;
    uint32_t *ctr = &P.work[(part == 2 ? 11 : 2) + OCT];
    auto ticket = [&]() -> int { return lane == 0 ? lo + (int)atom_add_u32(ctr, 1u) : 0; };
    auto fetch = [&](int k, int4 &ud, int &tn, float4 *buf) {     // descriptor, list size and list of queue entry k
        ud = __ldg(P.unit_list + begin + k);
        tn = kTileFallback;                                        // raw list size; looked at when the unit's turn comes
        if (lists) {
            tn = (int)__ldg(P.tile_counts + begin + k);
            if ((int)lane < 2 * P.tile_cap) cp_async16(buf + lane, P.tile_lists + (size_t)(begin + k) * (2 * P.tile_cap) + lane);
        }
        cp_async_commit();
    };
    int k_cur = __shfl_sync(0xffffffffu, ticket(), 0);
    if (k_cur >= hi) return;
    int raw_next = ticket();                                       // consumed at the top of the loop
    int4 ud; int tn; int b = 0;
    PendingAppend pend{0u, 0u, 0};
    fetch(k_cur, ud, tn, s_list);
    while (true) {
        const int k_next = __shfl_sync(0xffffffffu, raw_next, 0);
        const bool more = k_next < hi;
        int4 ud_n = make_int4(0, 0, 0, 0); int tn_n = kTileFallback;
        if (more) { fetch(k_next, ud_n, tn_n, s_list + (b ^ 1) * kBuf); raw_next = ticket(); }
        else cp_async_commit();
        cp_async_wait<1>();                                        // everything but the newest group: this unit's list is in
        __syncwarp();
        process_unit<MODE, OCT, STATS, WT>(P, nullptr, 0, smax, ud, lane, tn == kTileFallback ? -1 : tn, s_list + b * kBuf, LSB_DEFER_APPEND ? &pend : nullptr, s_pts, s_stack);
        if (!more) break;
        __syncwarp();                                              // all lanes are done with buffer b before it is refilled
        ud = ud_n; tn = tn_n; b ^= 1;
    }
    cp_async_wait<0>();
    flush_append(P, pend, s_pts, lane);
}

// ---- lane refill (ray compaction) ---------------------------------------------------------------------------
// run_phase hands a warp one 32-ray unit at a time: a lane whose ray ends early waits for the slowest ray of its
// tile, and the heavy units (grazing rays: 10 - 300 node visits, next to rays that stop at the first wall) are where
// the warp runs with a handful of live lanes (ncu: 22.9 of 32 threads per instruction over the launch).  Here the
// RAY is the unit of work inside a warp: the traversal of a lane is a resumable state machine that runs for a
// quantum of steps; after every quantum the warp reconverges, free lanes are counted with a ballot and the next
// rays of the feeding unit are dealt to them by rank (popc of the free mask below the lane) -- the reference's
// analogue is the per-packet loop of SRC/EmbreeTracer.cpp:297-367, which has the same tail per 16-ray packet.
// The rays in flight may belong to up to kRing different units: a unit's tile list stays in its slot of the warp's
// shared-memory ring until the last of its rays has ended (ballot over the slot tags), the next unit's descriptor
// and list are prefetched into a free slot with cp.async while the current one is being dealt out.  Hits are
// appended per finishing group (match.any on the sensor, one atomic per sensor and round).  Per-ray arithmetic,
// culling rules and result stores are those of trace_ray_lz / process_unit: same t, same triangle id.
#ifndef LSB_REFILL
#ifdef LSB_LAB
#define LSB_REFILL 1             // measured slower (design log 18): compiled into liblsb200_lab.so only, like the other rejected variants
#else
#define LSB_REFILL 0
#endif
#endif
#ifndef LSB_REFILL_RING
#define LSB_REFILL_RING 4
#endif
#ifndef LSB_REFILL_QUANTUM
#define LSB_REFILL_QUANTUM 32
#endif
constexpr int kRing = LSB_REFILL_RING;              // unit slots per warp (<= 4: the slot travels in two bits of the lane's tag)
constexpr int kRefillQuantum = LSB_REFILL_QUANTUM;  // traversal steps (list tests + node visits + leaf tests) between two refills

template <int OCT, bool WT>
__device__ __forceinline__ void run_phase_refill(const TraceParams &P, float smax, unsigned lane, float4 *s_ring, int4 *s_ud,
                                                 int2 *s_stack, int part)
{
    int begin = 0;
#pragma unroll
    for (int o = 0; o < OCT; ++o) begin += (int)__ldg(P.oct_hist + o);
    const int count = (int)__ldg(P.oct_hist + OCT);
    if (count == 0) return;
    constexpr int kBuf = 2 * kMaxTileEntries;                     // float4 per ring slot
    const int n_front = LSB_HEAVY_FIRST ? min(count, (int)__ldg(P.oct_hist + 16 + OCT)) : (int)(((long long)count * 3) / 8);
    const int lo = part == 2 ? n_front : 0, hi = part == 1 ? n_front : count;
    if (lo >= hi) return;
    uint32_t *ctr = &P.work[(part == 2 ? 11 : 2) + OCT];
    auto ticket = [&]() -> int { return lane == 0 ? lo + (int)atom_add_u32(ctr, 1u) : 0; };
    // descriptor (+ list size in .w) and list of queue entry k into ring slot `slot`
    auto fetch = [&](int k, int slot) {
        if (lane == 0) {
            int4 ud = __ldg(P.unit_list + begin + k);
            ud.w = (int)__ldg(P.tile_counts + begin + k);
            s_ud[slot] = ud;
        }
        if ((int)lane < 2 * P.tile_cap) cp_async16(s_ring + slot * kBuf + lane, P.tile_lists + (size_t)(begin + k) * (2 * P.tile_cap) + lane);
        cp_async_commit();
    };
    constexpr int kDone = 0x7fffffff;        // lane state in `cur`: nothing in hand (take the next list entry)
    constexpr int kFinished = 0x7ffffffe;    //                      ray ended, results not yet stored
    constexpr int kIdle = 0x7ffffffd;        //                      no ray
    constexpr unsigned kNodeLimit = 0x7ffffff0u;   // cur < kNodeLimit (unsigned): inner node; negative: leaf reference
    constexpr int kOvf = 1 << 20;            // depth-guard sentinel in sp (see trace_ray_lz)
#if LSB_NODE_EVICT_LAST
    const uint64_t pol = l2_policy_evict_last();
#endif
    // per-lane ray state
    Ray r{0.f, 0.f, 0.f, 0.f, 0.f, 1.f};
    float ix = 0.f, iy = 0.f, iz = 0.f, ax = 0.f, ay = 0.f, az = 0.f, bx = 0.f, by = 0.f, bz = 0.f;
    float best = INFINITY, far_lim = INFINITY;
    int bi = -1, cur = kIdle, sp = 0, left = 0, tag = 0;     // tag = ring slot | ray of the unit << 2
    int eo = 0;                               // next list entry: float4 index into the warp's ring
    int2 stack[kStackSize];
    auto push = [&](int2 v) {
        const int d = sp & (kOvf - 1);
        if (d < LSB_SMEM_STACK) s_stack[d * kTraceThreads] = v;
        else if (d < kStackSize) stack[d] = v;
        else { sp |= kOvf; return; }
        ++sp;
    };
    auto pop = [&]() -> int2 { --sp; const int d = sp & (kOvf - 1); return d < LSB_SMEM_STACK ? s_stack[d * kTraceThreads] : stack[d]; };
    auto pop_next = [&]() {                  // next deferred child that can still hold a closer hit, else kDone
        cur = kDone;
        while ((sp & (kOvf - 1)) > 0) {
            const int2 t = pop();
            if (__int_as_float(t.y) <= far_lim * 1.0000005f) { cur = t.x; break; }
        }
    };

    // warp-uniform queue / ring state
    int feed = -1, feed_next = 32;           // slot being dealt out, rays already dealt
    int load = -1;                           // slot being prefetched
    bool more = true;                        // the queue may hold further units
    int k0 = __shfl_sync(0xffffffffu, ticket(), 0);
    if (k0 >= hi) return;
    int raw_next = ticket();
    fetch(k0, 0);
    load = 0;
    while (true) {
        unsigned idle = __ballot_sync(0xffffffffu, cur == kIdle);
        // prefetch the next unit into a free slot (not the feeding one, no ray of it in flight)
        if (load < 0 && more) {
            int slot = -1;
#pragma unroll
            for (int k = 0; k < kRing; ++k) {
                const unsigned inflight = __ballot_sync(0xffffffffu, cur != kIdle && (tag & 3) == k);
                if (slot < 0 && (k != feed || feed_next >= 32) && inflight == 0u) slot = k;
            }
            if (slot >= 0) {
                const int k_next = __shfl_sync(0xffffffffu, raw_next, 0);
                if (k_next < hi) { __syncwarp(); fetch(k_next, slot); load = slot; raw_next = ticket(); }
                else more = false;
            }
        }
        // deal rays to the free lanes
        while (idle != 0u) {
            if (feed_next >= 32) {
                if (load < 0) break;                               // nothing prefetched: ring full or queue exhausted
                cp_async_wait<0>();
                __syncwarp();
                feed = load; load = -1; feed_next = 0;
                if (s_ud[feed].w == kTileFallback) {
                    // no tile list for this unit (its pyramid could not be bounded): drained warp, classic path from the origin chain
                    if (idle != 0xffffffffu) { load = feed; feed = -1; feed_next = 32; break; }
                    process_unit<6, OCT, false, WT>(P, nullptr, 0, smax, s_ud[feed], lane, -1, nullptr, nullptr, nullptr, s_stack);
                    feed_next = 32; feed = -1;
                    continue;
                }
            }
            const int4 ud = s_ud[feed];
            const unsigned rank = __popc(idle & lanemask_lt());
            const int n_avail = 32 - feed_next;
            if (((idle >> lane) & 1u) && (int)rank < n_avail) {
                const int j = feed_next + (int)rank;
                const SensorDev &S = P.sensors[ud.x];
                int v, h;
                const bool valid = unit_lane_ray(S, ud, P.tile_w_log2, (unsigned)j, v, h);
                const bool fired = valid && (S.az_mask == nullptr || __ldg(S.az_mask + h) != 0);
                if (valid && (!fired || ud.w == 0)) {             // outside the azimuth window, or the tile's pyramid reaches nothing
                    __stcs(S.t_out + v * S.H + h, INFINITY); __stcs(S.tri_out + v * S.H + h, -1);
                } else if (fired) {
                    float dsx, dsy, dsz, omax;
                    make_ray(S, P.world_frame, v, h, r, dsx, dsy, dsz, omax);
                    const float eps = 8e-6f * (smax + omax) + 1e-30f;
                    const float tiny = 1e-30f;
                    ix = 1.0f / (fabsf(r.dx) < tiny ? copysignf(tiny, r.dx) : r.dx);
                    iy = 1.0f / (fabsf(r.dy) < tiny ? copysignf(tiny, r.dy) : r.dy);
                    iz = 1.0f / (fabsf(r.dz) < tiny ? copysignf(tiny, r.dz) : r.dz);
                    ax = -(r.ox + eps) * ix; ay = -(r.oy + eps) * iy; az = -(r.oz + eps) * iz;
                    bx = -(r.ox - eps) * ix; by = -(r.oy - eps) * iy; bz = -(r.oz - eps) * iz;
                    best = INFINITY; far_lim = INFINITY; bi = -1; sp = 0;
                    left = ud.w; eo = feed * kBuf;
                    tag = feed | (j << 2);
                    cur = kDone;
                }
            }
            feed_next += min(n_avail, (int)__popc(idle));
            idle = __ballot_sync(0xffffffffu, cur == kIdle);      // lanes whose ray needed no traversal are served again
        }
        if (idle == 0xffffffffu && feed_next >= 32 && load < 0 && !more) break;

        // one quantum of traversal per lane (trace_ray_lz, resumable)
        if (cur != kIdle) {
            int budget = kRefillQuantum;
            while (budget > 0) {
                if (cur == kDone) {                  // nothing in hand: next list entry the ray enters
                    while (left > 0) {
                        const float4 e0 = s_ring[eo], e1 = s_ring[eo + 1];
                        eo += 2; --left; --budget;
                        float n, f;
                        slab_nf<OCT>(fmaf(e0.x, ix, ax), fmaf(e0.y, ix, bx), fmaf(e0.z, iy, ay), fmaf(e0.w, iy, by),
                                     fmaf(e1.x, iz, az), fmaf(e1.y, iz, bz), far_lim, n, f);
                        if (n <= f) { cur = __float_as_int(e1.z); break; }
                    }
                    if (cur == kDone) { cur = kFinished; break; }
                }
                while ((unsigned)cur < kNodeLimit && budget > 0) {
                    const float4 *nd = P.nodes + 4 * (size_t)cur;
#if LSB_NODE_EVICT_LAST
                    const float4 q0 = ldg_f4_hint(nd, pol), q1 = ldg_f4_hint(nd + 1, pol), q2 = ldg_f4_hint(nd + 2, pol);
                    const float2 q3 = ldg_f2_hint(reinterpret_cast<const float2 *>(nd + 3), pol);
#else
                    const float4 q0 = __ldg(nd), q1 = __ldg(nd + 1), q2 = __ldg(nd + 2);
                    const float2 q3 = __ldg(reinterpret_cast<const float2 *>(nd + 3));
#endif
                    --budget;
                    float n0, f0, n1, f1;
                    slab_nf<OCT>(fmaf(q0.x, ix, ax), fmaf(q0.y, ix, bx), fmaf(q0.z, iy, ay), fmaf(q0.w, iy, by),
                                 fmaf(q2.x, iz, az), fmaf(q2.y, iz, bz), far_lim, n0, f0);
                    slab_nf<OCT>(fmaf(q1.x, ix, ax), fmaf(q1.y, ix, bx), fmaf(q1.z, iy, ay), fmaf(q1.w, iy, by),
                                 fmaf(q2.z, iz, az), fmaf(q2.w, iz, bz), far_lim, n1, f1);
                    const bool hit0 = n0 <= f0, hit1 = n1 <= f1;
                    const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
                    if (hit0 && hit1) {
                        const bool swap = n1 < n0;
                        cur = swap ? c1 : c0;
                        push(make_int2(swap ? c0 : c1, __float_as_int(swap ? n0 : n1)));
                    } else if (hit0) {
                        cur = c0;
                    } else if (hit1) {
                        cur = c1;
                    } else {
                        pop_next();
                    }
                }
                if (cur < 0) {
                    test_leaf<WT>(P, cur, r, best, bi);
                    far_lim = best * 1.0000005f;
                    --budget;
                    pop_next();
                }
            }
        }
        __syncwarp();
        // rays that ended in this quantum: results, hit append, lane free again
        const bool fin = cur == kFinished;
        int v = 0, h = 0, si = 0;
        if (fin) {
            const int4 ud = s_ud[tag & 3];
            si = ud.x;
            const SensorDev &S = P.sensors[si];
            unit_lane_ray(S, ud, P.tile_w_log2, (unsigned)(tag >> 2), v, h);
            if (sp & kOvf) brute_ray_cold<WT>(P.tris, P.leaf_ids, P.n_tris, r, best, bi);     // depth guard tripped
            if (!(best <= S.max_range)) { best = INFINITY; bi = -1; }                          // detection radius
            __stcs(S.t_out + v * S.H + h, best);
            __stcs(S.tri_out + v * S.H + h, bi);
        }
        if (!(P.flags & (1u | 8u))) {                // neither ORDERED nor NO_POINTS: warp-aggregated atomic append, per sensor
            const bool hit = fin && bi >= 0;
            const unsigned hm = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const SensorDev &S = P.sensors[si];
                const unsigned m = __match_any_sync(hm, si);
                const int leader = __ffs(m) - 1;
                uint32_t base = 0;
                if ((int)lane == leader) base = atom_add_u32(S.count, (uint32_t)__popc(m));
                base = __shfl_sync(m, base, leader);
                Ray r2; float dsx, dsy, dsz, omax, x, y, z;
                make_ray(S, P.world_frame, v, h, r2, dsx, dsy, dsz, omax);
                hit_point(S, P.flags, best, dsx, dsy, dsz, x, y, z);
                store_point(S.points, base + __popc(m & lanemask_lt()), P.flags, x, y, z, (P.flags & 2u) ? 0 : v);
            }
        }
        if (fin) cur = kIdle;
    }
    cp_async_wait<0>();
}

// MODE 0: persistent CTAs, nodes through L1/L2, units pre-sorted by ray octant (k_classify_units / k_scatter_units)
//         and drained octant by octant with a sign-specialised slab test -- every CTA runs the same small loop at
//         a time, so the eight specialisations do not compete for the instruction cache
// MODE 1: persistent CTAs, nodes through L1/L2, generic min/max slab test, units in ray order
// MODE 2: one thread per ray, no work fetching (generic slab test)
// MODE 3: brute force (persistent)
// MODE 4: MODE 1 + top BVH levels staged into shared memory by TMA bulk copy
// MODE 5: MODE 1 in while-while form (node loop and leaf tests separated)
// MODE 6: MODE 0 in while-while form
// STATS: per-ray node / leaf / triangle counters summed into P.counters (diagnostic launches only)
// WT: watertight intersector (LSB200_OPT_WATERTIGHT; variants 6 and 3 only -- the triangle records then hold raw vertices)
template <int MODE, bool STATS, bool WT = false>
__global__ void __launch_bounds__(kTraceThreads, LSB_TRACE_CTAS) k_trace(TraceParams P)
{
    constexpr bool TOP = (MODE == 4);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *s_top = reinterpret_cast<float4 *>(smem_raw);
    __shared__ __align__(8) uint64_t s_bar;
    int n_top = 0;
    if (TOP) {
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
    __shared__ float4 s_lists[(MODE >= 6) ? kTraceThreads / 32 : 1][(MODE >= 6) ? (LSB_REFILL ? kRing : 2) * 2 * kMaxTileEntries : 1];      // per warp: 2 list buffers (run_phase) / a ring of kRing (run_phase_refill)
    float4 *s_list = s_lists[(MODE >= 6) ? (threadIdx.x >> 5) : 0];
    __shared__ int4 s_uds[(MODE == 6) ? kTraceThreads / 32 : 1][kRing];
    int4 *s_ud = s_uds[(MODE == 6) ? (threadIdx.x >> 5) : 0];
    __shared__ float4 s_points[(MODE >= 6 && LSB_DEFER_APPEND) ? kTraceThreads : 1];                  // parked hit points (deferred append)
    float4 *s_pts = s_points + ((MODE >= 6 && LSB_DEFER_APPEND) ? (threadIdx.x & ~31u) : 0u);
    int2 *s_stack = (MODE == 6 && LSB_SMEM_STACK) ? reinterpret_cast<int2 *>(smem_raw) + threadIdx.x : nullptr;
    // a queue (or its heavy front / light remainder) is drained unit by unit (run_phase) or ray by ray with lane refill
    // (run_phase_refill: MODE 6 with tile lists, per group of queues as P.refill says)
#define LSB_PHASE(OCTV, PART) do { \
        bool refilled_ = false; \
        if constexpr (LSB_REFILL != 0 && MODE == 6 && !STATS && LSB_LAZY_LIST != 0 && ((OCTV) < 8 || LSB_MIXED_TILE_LISTS != 0)) { \
            const int group_ = (OCTV) == 8 ? 0 : ((OCTV) < 4 ? 1 : ((PART) == 2 ? 3 : 2)); \
            if (P.tile_counts != nullptr && ((P.refill >> group_) & 1)) { \
                run_phase_refill<OCTV, WT>(P, smax, lane, s_list, s_ud, s_stack, PART); \
                refilled_ = true; \
            } \
        } \
        if (!refilled_) run_phase<MODE, OCTV, STATS, WT>(P, smax, lane, s_list, s_pts, s_stack, PART); \
    } while (0)
    if (MODE == 0 || MODE == 6 || MODE == 7) {
        // Order of the queues.  The kernel ends when its last unit ends, and a few units run ~100 x the median (long grazing
        // rays, design log section 15), so the launch must END with light units only.  k_scatter_units puts the heavy units
        // (a ray within ~4 degrees of the horizon or above it) at the front of every queue; the heavy fronts of the downward
        // octants (4..7) are drained early, together with the sign-mixed and the upward queues, and the light remainders
        // (steep rays that hit the ground nearby, ~10 us a unit) make the tail: 1.67 -> 1.53 ms for the stage.
        if (LSB_PHASE_ORDER != 2) LSB_PHASE(8, 0);
        if (MODE == 6 && LSB_SPLIT_DOWNWARD && LSB_PHASE_ORDER == 0) {
            LSB_PHASE(4, 1); LSB_PHASE(5, 1);
            LSB_PHASE(6, 1); LSB_PHASE(7, 1);
        }
        LSB_PHASE(0, 0); LSB_PHASE(1, 0);
        LSB_PHASE(2, 0); LSB_PHASE(3, 0);
        if (LSB_PHASE_ORDER == 2) LSB_PHASE(8, 0);
        if (MODE == 6 && LSB_SPLIT_DOWNWARD && LSB_PHASE_ORDER >= 1) {
            LSB_PHASE(4, 1); LSB_PHASE(5, 1);
            LSB_PHASE(6, 1); LSB_PHASE(7, 1);
        }
        const int rest = (MODE == 6 && LSB_SPLIT_DOWNWARD) ? 2 : 0;
        LSB_PHASE(4, rest); LSB_PHASE(5, rest);
        LSB_PHASE(6, rest); LSB_PHASE(7, rest);
    } else if (MODE == 2) {
        const int unit = P.unit_lo + (int)(blockIdx.x * (kTraceThreads / 32) + (threadIdx.x >> 5));
        if (unit < P.n_units) process_unit<MODE, 8, STATS, WT>(P, s_top, n_top, smax, unit_descriptor(P.sensors, P.n_sensors, unit, P.tile_w_log2), lane);
    } else {
        while (true) {
            int unit = 0;
            if (lane == 0) unit = P.unit_lo + (int)atomicAdd(&P.work[0], 1u);
            unit = __shfl_sync(0xffffffffu, unit, 0);
            if (unit >= P.n_units) break;
            process_unit<MODE, 8, STATS, WT>(P, s_top, n_top, smax, unit_descriptor(P.sensors, P.n_sensors, unit, P.tile_w_log2), lane);
        }
    }
    if (MODE != 2) {
        // the last CTA to leave re-arms the work counters for the next launch
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(&P.work[1], 1u) == gridDim.x - 1) {
                P.work[0] = 0; P.work[1] = 0;
                if (MODE == 0 || MODE == 6 || MODE == 7) for (int o = 0; o < 18; ++o) P.work[2 + o] = 0;
                __threadfence();
            }
        }
    }
}

#undef LSB_PHASE

// Heavy units first (LSB_HEAVY_FIRST): a unit with a ray at or above ~4 degrees below the horizon is "heavy" -- grazing rays
// fly hundreds of metres along streets and over roofs and cost up to 100 x the median unit (design log, section 15).
// The queues are filled sensor by sensor, so without this the grazing rows of the LAST sensors sit a few microseconds
// before the end of a queue and the launch ends with a tail behind them.  Heavy units are queued from the front of
// their class, the others from the back.  Order of work only: results do not depend on it.
#ifndef LSB_HEAVY_DZ
#define LSB_HEAVY_DZ -0.07f
#endif
constexpr float kHeavyDz = LSB_HEAVY_DZ;
constexpr int kUnitHeavy = 0x80;

// Rays (channels [v0, v0 + nv) x azimuth steps [h0, h0 + nh)) of sensor S, generated with the traversal's own
// arithmetic: their direction interval and origin.  Returns the common sign class (bit k = component k negative) or
// 8 when the signs differ.  The tables are read once per row / column.
__device__ __forceinline__ int block_interval(const SensorDev &S, int world_frame, int v0, int nv, int h0, int nh,
                                              float dlo[3], float dhi[3], float o3[3], float &omax)
{
    dlo[0] = dlo[1] = dlo[2] = INFINITY; dhi[0] = dhi[1] = dhi[2] = -INFINITY;
    o3[0] = o3[1] = o3[2] = 0.0f; omax = 0.0f;
    unsigned sign_or = 0u, sign_and = 7u;
    const int v1 = min(v0 + nv, S.V), h1 = min(h0 + nh, S.H);
    auto one = [&](float st, float ct, float sp_, float cp) {
        Ray r; float x, y, z;
        ray_from_tables(S, world_frame, st, ct, sp_, cp, r, x, y, z, omax);
        o3[0] = r.ox; o3[1] = r.oy; o3[2] = r.oz;
        dlo[0] = fminf(dlo[0], r.dx); dhi[0] = fmaxf(dhi[0], r.dx);
        dlo[1] = fminf(dlo[1], r.dy); dhi[1] = fmaxf(dhi[1], r.dy);
        dlo[2] = fminf(dlo[2], r.dz); dhi[2] = fmaxf(dhi[2], r.dz);
        const unsigned o = (unsigned)ray_octant(r);
        sign_or |= o; sign_and &= o;
    };
    if (nv <= 4) {
        float stv[4], ctv[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int v = min(v0 + a, S.V - 1);
            stv[a] = __ldg(S.sin_t + v); ctv[a] = __ldg(S.cos_t + v);
        }
        for (int h = h0; h < h1; ++h) {
            const float sp_ = __ldg(S.sin_p + h), cp = __ldg(S.cos_p + h);
#pragma unroll
            for (int a = 0; a < 4; ++a) if (a < nv && v0 + a < S.V) one(stv[a], ctv[a], sp_, cp);
        }
    } else {
        for (int v = v0; v < v1; ++v) {
            const float st = __ldg(S.sin_t + v), ct = __ldg(S.cos_t + v);
            for (int h = h0; h < h1; ++h) one(st, ct, __ldg(S.sin_p + h), __ldg(S.cos_p + h));
        }
    }
    return sign_or == sign_and ? (int)sign_or : 8;
}

// ---- octant classification of the 32-ray units (MODE 0) ----------------------------------------------------
// unit_oct[u - unit_lo] = common octant of the unit's rays (bit k = sign bit of direction component k) or 8 when
// they disagree; hist[o] = number of units per class.  Only depends on the sensor table: cached by the host.
__global__ void __launch_bounds__(256) k_classify_units(const SensorDev *__restrict__ sensors, int n_sensors, int unit_lo,
                                                        int unit_hi, int world_frame, int tile_w_log2,
                                                        uint8_t *__restrict__ unit_oct, uint32_t *__restrict__ hist,
                                                        float4 *__restrict__ unit_iv /* [unit][2] direction interval, or nullptr */)
{
    const int u = unit_lo + blockIdx.x * blockDim.x + threadIdx.x;
    int oct = -1;
    if (u < unit_hi) {
        // one pass over the unit's rays gives its octant class, its heavy flag AND its direction interval (what
        // k_tile_intervals computes per queue slot): a moving sensor pays this every frame, so it is done once
        const int4 ud = unit_descriptor(sensors, n_sensors, u, tile_w_log2);
        const SensorDev &S = sensors[ud.x];
        float dlo[3], dhi[3], o3[3], omax;
        oct = block_interval(S, world_frame, ud.y, 32 >> tile_w_log2, ud.z, 1 << tile_w_log2, dlo, dhi, o3, omax);
        const bool heavy = LSB_HEAVY_FIRST && dhi[2] > kHeavyDz;          // some ray of the unit has d.z > kHeavyDz
        unit_oct[u - unit_lo] = (uint8_t)(oct | (heavy ? kUnitHeavy : 0));
        if (unit_iv != nullptr) {
            unit_iv[2 * (size_t)(u - unit_lo)] = make_float4(dlo[0], dlo[1], dlo[2], __int_as_float(oct));
            unit_iv[2 * (size_t)(u - unit_lo) + 1] = make_float4(dhi[0], dhi[1], dhi[2], omax);
        }
    }
#pragma unroll
    for (int o = 0; o < 9; ++o) {
        const unsigned m = __ballot_sync(0xffffffffu, oct == o);
        if (m && lane_id() == 0) atomicAdd(&hist[o], (uint32_t)__popc(m));
    }
}

// unit_list = unit descriptors grouped by class (class o occupies [sum hist[<o], +hist[o])), in chunks of up to 32
// consecutive units (one atomic per class per warp), so neighbouring units stay neighbours in the queues.
__global__ void __launch_bounds__(256) k_scatter_units(const SensorDev *__restrict__ sensors, int n_sensors, int unit_lo,
                                                       int unit_hi, int tile_w_log2, const uint8_t *__restrict__ unit_oct,
                                                       const uint32_t *__restrict__ hist, uint32_t *__restrict__ cursor,
                                                       int4 *__restrict__ unit_list, const float4 *__restrict__ unit_iv = nullptr,
                                                       float4 *__restrict__ tile_iv = nullptr /* [queue slot][2] */)
{
    const int u = unit_lo + blockIdx.x * blockDim.x + threadIdx.x;
    const int raw = (u < unit_hi) ? (int)unit_oct[u - unit_lo] : -1;
    const int oct = raw < 0 ? -1 : (raw & 15);
    const bool light = raw >= 0 && LSB_HEAVY_FIRST && !(raw & kUnitHeavy);      // queued from the back of the class
    uint32_t base = 0;
#pragma unroll
    for (int o = 0; o < 9; ++o) {
#pragma unroll
        for (int back = 0; back < (LSB_HEAVY_FIRST ? 2 : 1); ++back) {
            const bool mine = oct == o && light == (back != 0);
            const unsigned m = __ballot_sync(0xffffffffu, mine);
            if (m == 0) continue;
            const uint32_t k = (uint32_t)__popc(m);
            uint32_t off = 0;
            if (lane_id() == (unsigned)(__ffs(m) - 1)) off = atomicAdd(&cursor[o + 16 * back], k);
            off = __shfl_sync(0xffffffffu, off, __ffs(m) - 1);
            if (back) off = hist[o] - off - k;
            if (mine) {
                const uint32_t slot = base + off + __popc(m & lanemask_lt());
                unit_list[slot] = unit_descriptor(sensors, n_sensors, u, tile_w_log2);
                if (tile_iv != nullptr) {
                    tile_iv[2 * (size_t)slot] = unit_iv[2 * (size_t)(u - unit_lo)];
                    tile_iv[2 * (size_t)slot + 1] = unit_iv[2 * (size_t)(u - unit_lo) + 1];
                }
            }
        }
        base += hist[o];
    }
}

// ---- origin chains -----------------------------------------------------------------------------------------
// One thread per (sensor, octant class): walks from the root along the children whose box contains the sensor
// origin and lists the sibling of every step (box + reference), then the node or leaf the walk ends in.
// Entry layout:   e0 = (lo.x, hi.x, lo.y, hi.y)   e1 = (lo.z, hi.z, ref, 0)
// The walk stops when neither child box contains the origin, at a leaf, or when the list is nearly full.
// Class 8 keeps every entry.  Class o < 8 (direction signs known: bit k set = component k negative) drops the
// entries that lie entirely behind the origin along some axis -- hi < o - eps on a non-negative axis, lo > o + eps
// on a negative one: no ray of that octant can reach them (eps = the traversal's own box padding), so the rays
// of a sign-uniform unit test about half of the list.
__global__ void k_origin_chain(const float4 *__restrict__ nodes, int top_cap, const SensorDev *__restrict__ sensors,
                               int n_sensors, int world_frame, const uint32_t *__restrict__ bounds,
                               float4 *__restrict__ chains, int *__restrict__ counts)
{
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_sensors * 9) return;
    const int si = id / 9, cls = id - si * 9;
    const SensorDev &S = sensors[si];
    const float ox = world_frame ? S.T[0] : 0.0f, oy = world_frame ? S.T[1] : 0.0f, oz = world_frame ? S.T[2] : 0.0f;
    float smax = 0.0f;
    for (int k = 0; k < 6; ++k) smax = fmaxf(smax, fabsf(ord2f(bounds[k])));
    origin_chain_walk(nodes, top_cap, ox, oy, oz, smax, cls, chains + (size_t)id * (2 * kMaxChain), counts + id);
}

// ---- tile entry lists --------------------------------------------------------------------------------------
// Every ray of a LiDAR starts at the same point, so the 32 rays of a work unit (a tile of 8 azimuth steps x 4
// channels, about 0.8 x 1.2 degrees) form a thin pyramid with its apex at the sensor: they walk almost the same
// root-to-leaf path, 32 times.  k_tile_entries walks that shared part ONCE per tile (one THREAD per tile): a
// near-first depth-first descent in which a box is entered when the tile's direction interval can reach it, until
// `tile_cap` subtrees / leaves are listed.  The rays of the tile then test the <= tile_cap listed boxes (32 bytes
// each, no divergence) instead of descending ~17 levels each (trace_ray_lz) -- near part of the pyramid resolved
// down to leaves, far part left coarse (it fails its box test once a closer hit is known).
//
// Two levels: the walk of a tile does not start from its sensor's origin chain but from the list of its SUPERTILE
// (4 x 4 tiles = 32 azimuth steps x 16 channels, k_super_entries, one thread per supertile, same walk with a wider
// interval and 16 entries): the ~15 levels between the chain's big sibling subtrees and the neighbourhood of the
// pyramid are descended once per 16 tiles instead of once per tile.
//
// Tile test = interval form of the ray's own slab test.  For a sign-uniform tile |1/d| of axis k lies in
// [imin_k, imax_k]; entry distances are bounded from below with imin, exit distances from above with imax:
//     n_tile = max_k (near plane_k - o_k -+ eps) * imin_k  <=  n_ray        f_tile = min_k (far plane_k ...) * imax_k  >=  f_ray
// so n_ray <= f_ray implies n_tile <= f_tile: a tile never drops a box one of its rays would enter (eps is 1.5 x the
// rays' padding and the interval is widened by 1e-6, far above the rounding noise of either form).  The union of
// the listed subtrees therefore covers everything any ray of the tile can hit, and closest hit = min (t, triangle
// id) does not depend on how the tree was cut: per-ray results are unchanged (tests: every variant bit-identical).
// Sign-mixed tiles (class 8, ~1 % of a 128 x 3601 table) and axes whose interval touches 0 use the both-signs form of the
// test on that axis (1/d unbounded: entry bound from the gap to the slab, no exit bound); tiles of a sign-mixed
// supertile start from the chain.
#ifndef LSB_TILE_BLOCK
#define LSB_TILE_BLOCK 128
#endif
#ifndef LSB_TILE_SMEM_STACK
#define LSB_TILE_SMEM_STACK 1     // k_tile_entries: deferred entries of the walk in shared memory ([slot][thread]) instead of a local array
#endif
constexpr int kTileBlock = LSB_TILE_BLOCK;   // threads per CTA of k_tile_entries (one thread per tile)
static_assert(!LSB_TILE_SMEM_STACK || 16 * LSB_TILE_BLOCK * 16 <= 48 * 1024, "k_tile_entries: [kMaxTileEntries][kTileBlock] int4 walk stack must fit the default 48 KB of dynamic shared memory");
constexpr int kSuperTiles = 4;          // a supertile is kSuperTiles x kSuperTiles tiles

// Direction interval of a block of rays -> the constants of the interval slab test.
struct TileFrustum {
    int mode[3];                 // per axis: 0 = every ray goes up the axis, 1 = down, 2 = both signs occur (or |d| touches 0)
    bool any_mixed;              // some axis is in mode 2
    float in[3], an[3], ifr[3], af[3];
    float max_range;
    // one axis: lower bound of the entry distance, upper bound of the exit distance.  Mode 2: 1/d is unbounded, so a
    // ray can stay inside the slab for ever (no exit bound) and enters it no earlier than gap / max |d|.
    __device__ __forceinline__ void axis(int k, float lo, float hi, float &n, float &f) const
    {
        if (mode[k] == 2) {
            n = fmaxf(fmaf(lo, in[k], an[k]), fmaf(hi, -in[k], af[k]));
            f = INFINITY;
        } else {
            n = fmaf(mode[k] ? hi : lo, in[k], an[k]);
            f = fmaf(mode[k] ? lo : hi, ifr[k], af[k]);
        }
    }
    // box = (lo.x, hi.x, lo.y, hi.y), (lo.z, hi.z): can a ray of the block enter it, and not before which distance
    __device__ __forceinline__ bool test(const float4 bxy, float bzlo, float bzhi, float &n) const
    {
        float nx, fx, ny, fy, nz, fz;
        if (!any_mixed) {            // the usual case, without the per-axis mode branches
            nx = fmaf(mode[0] ? bxy.y : bxy.x, in[0], an[0]); fx = fmaf(mode[0] ? bxy.x : bxy.y, ifr[0], af[0]);
            ny = fmaf(mode[1] ? bxy.w : bxy.z, in[1], an[1]); fy = fmaf(mode[1] ? bxy.z : bxy.w, ifr[1], af[1]);
            nz = fmaf(mode[2] ? bzhi : bzlo, in[2], an[2]); fz = fmaf(mode[2] ? bzlo : bzhi, ifr[2], af[2]);
        } else {
            axis(0, bxy.x, bxy.y, nx, fx); axis(1, bxy.z, bxy.w, ny, fy); axis(2, bzlo, bzhi, nz, fz);
        }
        n = fmaxf(fmaxf(nx, ny), fmaxf(nz, 0.0f));
        const float f = fminf(fminf(fx, fy), fz);
        return n <= f && n <= max_range;
    }
};

// The interval itself tells the mode of every axis (for a sign-uniform tile they agree with its octant class).
// false: the block holds no ray.
__device__ __forceinline__ bool make_frustum(const float dlo[3], const float dhi[3], const float o3[3], float eps, float max_range,
                                             TileFrustum &F)
{
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (!(dlo[k] <= dhi[k])) ok = false;
        const int mode = dlo[k] > 1e-6f ? 0 : (dhi[k] < -1e-6f ? 1 : 2);
        F.mode[k] = mode;
        if (mode == 2) {
            const float amax = fmaxf(fmaxf(fabsf(dlo[k]), fabsf(dhi[k])), 1e-30f);
            const float in_abs = (1.0f / amax) * 0.999999f;
            F.in[k] = in_abs;
            F.an[k] = -(o3[k] + eps) * in_abs;          // (lo - o - eps) * in
            F.af[k] = (o3[k] - eps) * in_abs;           // (o - eps - hi) * in   (used with -in)
            F.ifr[k] = 0.0f;
        } else {
            const float amin = mode ? -dhi[k] : dlo[k], amax = mode ? -dlo[k] : dhi[k];     // |d| interval
            const float in_abs = (1.0f / amax) * 0.999999f, if_abs = (1.0f / amin) * 1.000001f;
            F.in[k] = mode ? -in_abs : in_abs;
            F.ifr[k] = mode ? -if_abs : if_abs;
            F.an[k] = -o3[k] * F.in[k] - eps * in_abs;
            F.af[k] = -o3[k] * F.ifr[k] + eps * if_abs;
        }
    }
    F.max_range = max_range;
    F.any_mixed = F.mode[0] == 2 || F.mode[1] == 2 || F.mode[2] == 2;
    return ok;
}

__device__ __forceinline__ int find_sensor_by_super(const SensorDev *s, int lo, int hi, int sup)
{
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (s[mid].sup_base <= sup) lo = mid; else hi = mid - 1; }
    return lo;
}

// Direction intervals depend on the sensor tables and poses only (not on the scene): computed when those change and
// kept.  Tiles: one thread per queue entry.  Supertiles (512 rays): one WARP per supertile, a lane per azimuth
// column, intervals merged with shuffles.
__global__ void __launch_bounds__(128) k_tile_intervals(TraceParams P)
{
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    int total = 0;
#pragma unroll
    for (int o = 0; o < 9; ++o) total += (int)__ldg(P.oct_hist + o);
    if (slot >= total) return;
    const int4 ud = __ldg(P.unit_list + slot);
    const SensorDev &S = P.sensors[ud.x];
    const int tw = 1 << P.tile_w_log2, th = 32 >> P.tile_w_log2;
    float dlo[3], dhi[3], o3[3], omax;
    const int cls = block_interval(S, P.world_frame, ud.y, th, ud.z, tw, dlo, dhi, o3, omax);
    P.tile_iv[2 * (size_t)slot] = make_float4(dlo[0], dlo[1], dlo[2], __int_as_float(cls));
    P.tile_iv[2 * (size_t)slot + 1] = make_float4(dhi[0], dhi[1], dhi[2], omax);
}

#ifdef LSB_LAB
__global__ void __launch_bounds__(128) k_super_intervals(TraceParams P, int s_first, int s_count, int sup_lo, int sup_hi)
{
    const int sup = sup_lo + (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (sup >= sup_hi) return;                                   // whole warps leave together
    const unsigned lane = lane_id();
    const int si = find_sensor_by_super(P.sensors, s_first, s_first + s_count - 1, sup);
    const SensorDev &S = P.sensors[si];
    const int local = sup - S.sup_base;
    const int sv = local / S.sup_w, sh = local - sv * S.sup_w;
    const int tw = 1 << P.tile_w_log2, th = 32 >> P.tile_w_log2;
    const int v0 = sv * kSuperTiles * th, nv = kSuperTiles * th, h0 = sh * kSuperTiles * tw, nh = kSuperTiles * tw;
    float dlo[3] = {INFINITY, INFINITY, INFINITY}, dhi[3] = {-INFINITY, -INFINITY, -INFINITY}, o3[3] = {0.f, 0.f, 0.f}, omax = 0.0f;
    unsigned sign_or = 0u, sign_and = 7u;
    for (int h = h0 + (int)lane; h < min(h0 + nh, S.H); h += 32) {
        float lo[3], hi[3], om;
        const int c = block_interval(S, P.world_frame, v0, nv, h, 1, lo, hi, o3, om);
        omax = om;
#pragma unroll
        for (int k = 0; k < 3; ++k) { dlo[k] = fminf(dlo[k], lo[k]); dhi[k] = fmaxf(dhi[k], hi[k]); }
        if (c == 8) { sign_or = 7u; sign_and = 0u; }
        else { sign_or |= (unsigned)c; sign_and &= (unsigned)c; }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            dlo[k] = fminf(dlo[k], __shfl_xor_sync(0xffffffffu, dlo[k], off));
            dhi[k] = fmaxf(dhi[k], __shfl_xor_sync(0xffffffffu, dhi[k], off));
        }
        omax = fmaxf(omax, __shfl_xor_sync(0xffffffffu, omax, off));
        sign_or |= __shfl_xor_sync(0xffffffffu, sign_or, off);
        sign_and &= __shfl_xor_sync(0xffffffffu, sign_and, off);
    }
    if (lane == 0) {
        const int cls = sign_or == sign_and ? (int)sign_or : 8;
        P.sup_iv[2 * (size_t)sup] = make_float4(dlo[0], dlo[1], dlo[2], __int_as_float(cls));
        P.sup_iv[2 * (size_t)sup + 1] = make_float4(dhi[0], dhi[1], dhi[2], omax);
    }
}

#endif  // LSB_LAB

// The walk: seeds (a list of (box, reference) entries -- the sensor's origin chain, far -> near, or the supertile's
// list, near -> far) are filtered with the block's interval test, then refined near-first until `cap` entries are
// listed.  Returns the number of entries written to `out` (near -> far), or -1 when more than `cap` seeds pass.
// Deferred entries are 16 bytes: (reference, entry distance, where the box is stored); side 0 / 1 = child slot of
// node `from`, side 2 = seed `from`.  The box itself (24 bytes) is only re-read when a popped entry is listed
// without having been split.  One flat loop (a step = list the entry in hand, or split it): the lanes of a warp --
// neighbouring blocks -- stay in step instead of serialising an outer pop loop against an inner descent loop.
// The deferred entries live at stk[d * STRIDE]: a local array (STRIDE 1; k_super_entries) or a [slot][thread] block of
// shared memory (STRIDE = threads per CTA; k_tile_entries).  Lanes of a warp sit at different depths: in local memory
// every distinct depth of a 16-byte access is its own four wavefronts (a third of the kernel's L1 wavefronts were
// local loads / stores), in shared memory the bank is the thread's column whatever the depth -- always four.
template <int STRIDE>
__device__ __forceinline__ int frustum_walk(const float4 *__restrict__ nodes, const TileFrustum &F, const float4 *__restrict__ seeds,
                                            int n_seeds, bool seeds_near_first, int cap, int max_visits, float4 *__restrict__ out,
                                            int &visits, int4 *__restrict__ stk)
{
    int sp = 0;
    for (int q = 0; q < n_seeds; ++q) {                          // far -> near: the nearest seed ends on top
        const int k = seeds_near_first ? n_seeds - 1 - q : q;
        const float4 e0 = __ldg(seeds + 2 * k), e1 = __ldg(seeds + 2 * k + 1);
        float n;
        if (!F.test(e0, e1.x, e1.y, n)) continue;
        if (sp == cap) return -1;
        stk[(sp++) * STRIDE] = make_int4(__float_as_int(e1.z), __float_as_int(n), k, 2);
    }
    int n_out = 0;
    int4 cur = make_int4(0, 0, 0, 0);
    float4 cxy = make_float4(0.f, 0.f, 0.f, 0.f);
    float2 cz = make_float2(0.f, 0.f);
    bool have = sp > 0, have_box = false;
    if (have) cur = stk[(--sp) * STRIDE];
    while (have) {
        bool pop = false;
        // a split turns the entry in hand into two: only while the list has room for it
        if (cur.x < 0 || n_out + sp + 2 > cap || visits >= max_visits) {
            if (!have_box) {
                if (cur.w == 2) {
                    const float4 e1 = __ldg(seeds + 2 * cur.z + 1);
                    cxy = __ldg(seeds + 2 * cur.z); cz = make_float2(e1.x, e1.y);
                } else {
                    const float4 *nd = nodes + 4 * (size_t)cur.z;
                    cxy = __ldg(nd + cur.w);
                    cz = __ldg(reinterpret_cast<const float2 *>(nd + 2) + cur.w);
                }
            }
            __stcs(out + 2 * n_out, cxy);                        // read once, by one warp, a kernel later: streaming
            __stcs(out + 2 * n_out + 1, make_float4(cz.x, cz.y, __int_as_float(cur.x), __int_as_float(cur.y)));
            ++n_out;
            pop = true;
        } else {
            const int node = cur.x;
            const float4 *nd = nodes + 4 * (size_t)node;
            const float4 q0 = __ldg(nd), q1 = __ldg(nd + 1), q2 = __ldg(nd + 2);
            const float2 q3 = __ldg(reinterpret_cast<const float2 *>(nd + 3));
            ++visits;
            float n0, n1;
            const bool h0 = F.test(q0, q2.x, q2.y, n0), h1 = F.test(q1, q2.z, q2.w, n1);
            const bool first1 = h1 && (!h0 || n1 < n0);              // continue with child 1 (the nearer, or the only one)
            if (h0 && h1)                                            // defer the farther child
                stk[(sp++) * STRIDE] = first1 ? make_int4(__float_as_int(q3.x), __float_as_int(n0), node, 0)
                                   : make_int4(__float_as_int(q3.y), __float_as_int(n1), node, 1);
            cur = first1 ? make_int4(__float_as_int(q3.y), __float_as_int(n1), node, 1) : make_int4(__float_as_int(q3.x), __float_as_int(n0), node, 0);
            cxy = first1 ? q1 : q0;
            cz = first1 ? make_float2(q2.z, q2.w) : make_float2(q2.x, q2.y);
            have_box = true;
            pop = !h0 && !h1;
        }
        if (pop) {
            have = sp > 0;
            have_box = false;
            if (have) cur = stk[(--sp) * STRIDE];
        }
    }
    return n_out;
}

__device__ __forceinline__ float scene_smax(const uint32_t *__restrict__ bounds)
{
    float smax = 0.0f;
#pragma unroll
    for (int k = 0; k < 6; ++k) smax = fmaxf(smax, fabsf(ord2f(__ldg(bounds + k))));
    return smax;
}

#ifdef LSB_LAB
// One thread per supertile of the traced sensors [s_first, s_first + s_count): list of <= kMaxTileEntries entries from
// the sensor's origin chain.  sup_counts: entries, or kTileFallback (sign-mixed block / interval touching 0 / too
// many seeds).
template <bool STATS>
__global__ void __launch_bounds__(128) k_super_entries(TraceParams P, int s_first, int s_count, int sup_lo, int sup_hi, int max_visits)
{
    const int sup = sup_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (sup >= sup_hi) return;
    const int si = find_sensor_by_super(P.sensors, s_first, s_first + s_count - 1, sup);
    const SensorDev &S = P.sensors[si];
    const float4 iv0 = __ldg(P.sup_iv + 2 * (size_t)sup), iv1 = __ldg(P.sup_iv + 2 * (size_t)sup + 1);
    const float dlo[3] = {iv0.x, iv0.y, iv0.z}, dhi[3] = {iv1.x, iv1.y, iv1.z};
    const int cls = __float_as_int(iv0.w);
    const float omax = iv1.w;
    const float o3[3] = {P.world_frame ? S.T[0] : 0.0f, P.world_frame ? S.T[1] : 0.0f, P.world_frame ? S.T[2] : 0.0f};
    TileFrustum F;
    const float eps = 1.5f * (8e-6f * (scene_smax(P.bounds) + omax) + 1e-30f);
    int n_out = -1, visits = 0;
    if (cls < 8 && make_frustum(dlo, dhi, o3, eps, S.max_range, F)) {
        const int list = si * 9 + cls;
        int4 stk[kMaxTileEntries];
        n_out = frustum_walk<1>(P.nodes, F, P.chains + (size_t)list * (2 * kMaxChain), __ldg(P.chain_counts + list), false,
                                kMaxTileEntries, max_visits, P.sup_lists + (size_t)sup * (2 * kMaxTileEntries), visits, stk);
    }
    P.sup_counts[sup] = (uint8_t)(n_out < 0 ? kTileFallback : n_out);
    if (STATS) atomicAdd(P.counters + kCntSuperVisits, (unsigned long long)visits);
}

#endif  // LSB_LAB

// One thread per queue entry (32-ray tile).  tile_counts: entries, or kTileFallback.
template <bool STATS>
__global__ void __launch_bounds__(kTileBlock) k_tile_entries(TraceParams P, int max_visits)
{
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    int begin8 = 0;                                    // first slot of class 8 = number of sign-uniform units
#pragma unroll
    for (int o = 0; o < 8; ++o) begin8 += (int)__ldg(P.oct_hist + o);
    const int total = begin8 + (int)__ldg(P.oct_hist + 8);
    if (slot >= total) return;
    if (slot >= begin8 && !LSB_MIXED_TILE_LISTS) { P.tile_counts[slot] = (uint8_t)kTileFallback; return; }
    int cls = 0, acc = 0;                              // 0..7: the octant queue the slot is in; 8: sign-mixed
#pragma unroll
    for (int o = 0; o < 8; ++o) { acc += (int)__ldg(P.oct_hist + o); if (slot >= acc) cls = o + 1; }

    const int4 ud = __ldg(P.unit_list + slot);
    const SensorDev &S = P.sensors[ud.x];
    const int tw = 1 << P.tile_w_log2, th = 32 >> P.tile_w_log2;
    const float4 iv0 = __ldg(P.tile_iv + 2 * (size_t)slot), iv1 = __ldg(P.tile_iv + 2 * (size_t)slot + 1);
    const float dlo[3] = {iv0.x, iv0.y, iv0.z}, dhi[3] = {iv1.x, iv1.y, iv1.z};
    const float omax = iv1.w;
    const float o3[3] = {P.world_frame ? S.T[0] : 0.0f, P.world_frame ? S.T[1] : 0.0f, P.world_frame ? S.T[2] : 0.0f};
    TileFrustum F;
    const float eps = 1.5f * (8e-6f * (scene_smax(P.bounds) + omax) + 1e-30f);
    if (!make_frustum(dlo, dhi, o3, eps, S.max_range, F)) { P.tile_counts[slot] = (uint8_t)kTileFallback; return; }

    // seeds: the supertile's list when it has one, else the sensor's origin chain
    const int list = ud.x * 9 + cls;
    const float4 *seeds = P.chains + (size_t)list * (2 * kMaxChain);
    int n_seeds = __ldg(P.chain_counts + list);
    bool near_first = false;
    if (P.sup_counts != nullptr && cls < 8) {
        const int sup = S.sup_base + ((ud.y / th) / kSuperTiles) * S.sup_w + (ud.z / tw) / kSuperTiles;
        const int sc = (int)__ldg(P.sup_counts + sup);
        if (sc != kTileFallback) { seeds = P.sup_lists + (size_t)sup * (2 * kMaxTileEntries); n_seeds = sc; near_first = true; }
    }
    int visits = 0;
#if LSB_TILE_SMEM_STACK
    extern __shared__ int4 s_tile_stk[];                // [tile_cap][kTileBlock]
    int4 *stk = s_tile_stk + threadIdx.x;
    constexpr int kStride = kTileBlock;
#else
    int4 stk[kMaxTileEntries];
    constexpr int kStride = 1;
#endif
    int n_out = frustum_walk<kStride>(P.nodes, F, seeds, n_seeds, near_first, P.tile_cap, max_visits,
                                      P.tile_lists + (size_t)slot * (2 * P.tile_cap), visits, stk);
    if (n_out < 0 && near_first)            // more supertile entries pass than the list holds: start from the chain instead
        n_out = frustum_walk<kStride>(P.nodes, F, P.chains + (size_t)list * (2 * kMaxChain), __ldg(P.chain_counts + list), false, P.tile_cap,
                                      max_visits, P.tile_lists + (size_t)slot * (2 * P.tile_cap), visits, stk);
    P.tile_counts[slot] = (uint8_t)(n_out < 0 ? kTileFallback : n_out);
    if (STATS) {
        atomicAdd(P.counters + kCntTileVisits, (unsigned long long)visits);
        if (n_out < 0) atomicAdd(P.counters + kCntTileFallbacks, 1ull);
        else {
            atomicAdd(P.counters + kCntTileEntries, (unsigned long long)n_out);
            if (n_out == 0) atomicAdd(P.counters + kCntTileEmpty, 1ull);
        }
    }
}

// ---- small launches: one thread per ray ----------------------------------------------------------------------
// A single XT-32 scan is 4 800 rays = 150 work units: the persistent kernel's machinery (444 CTAs that mostly find no
// work, ticket / descriptor / list fetches, origin chains read ray by ray) is pure latency at that size (50 - 55 us per
// launch, design log 6).  Launches of fewer than LSB200_OPT_SMALL_TRACE_UNITS units take this kernel instead: one thread
// per ray, 128-thread CTAs spread over the SMs, the shipped lazy-list traversal (trace_ray_lz) started from the sensor's
// origin chain, generic slab test.  Same conservative culling, same intersector, same stores and the same
// warp-aggregated append as process_unit: identical per-ray results.  grid.y = sensor of the batch.
template <bool WT>
__global__ void __launch_bounds__(128) k_trace_small(TraceParams P, int sensor_base, int spread_log2)
{
    // trace_ray_lz keeps the first LSB_SMEM_STACK stack entries of a thread at s_stack[d * kTraceThreads]
    __shared__ int2 s_stack_buf[(LSB_SMEM_STACK ? LSB_SMEM_STACK : 1) * kTraceThreads];
    const int si = sensor_base + blockIdx.y;
    const SensorDev &S = P.sensors[si];
    // Only every 2^spread-th lane carries a ray.  An XT-32's neighbouring rays are 2.4 degrees apart: 32 of them in one warp
    // walk 32 different paths one after the other (3 000 warp instructions at ~29 cycles each = 45 us, with the machine
    // otherwise empty); spread over 32 warps the same rays run side by side.
    const int gthread = blockIdx.x * blockDim.x + threadIdx.x;
    const int ray = gthread >> spread_log2;
    const bool valid = ray < S.n_rays && (gthread & ((1 << spread_log2) - 1)) == 0;
    const int v = valid ? ray / S.H : 0, h = valid ? ray - v * S.H : 0;
    const bool fired = valid && (S.az_mask == nullptr || __ldg(S.az_mask + h) != 0);
    float best = INFINITY; int bi = -1;
    float dsx = 0.f, dsy = 0.f, dsz = 1.f;
    if (fired) {
        float smax = 0.0f;
#pragma unroll
        for (int k = 0; k < 6; ++k) smax = fmaxf(smax, fabsf(ord2f(__ldg(P.bounds + k))));
        Ray r; float omax;
        make_ray(S, P.world_frame, v, h, r, dsx, dsy, dsz, omax);
        const float eps = 8e-6f * (smax + omax) + 1e-30f;
        // the sensor's origin chain for the ray's own octant (the entries no ray of these signs can reach are already
        // dropped) replaces the ~17-level descent from the root by a handful of box tests: at this size a ray is one
        // dependent instruction stream with nothing to hide its latency behind, so fewer instructions per ray is what counts
        const int list = si * 9 + ray_octant(r);
        const int n_chain = P.chain_counts ? __ldg(P.chain_counts + list) : 0;
        RayStats rs{0u, 0u, 0u, 0u, 0u, 0u};
        trace_ray_lz<8, false, false, WT>(P, eps, r, P.chains + (size_t)list * (2 * kMaxChain), n_chain, false, best, bi, rs,
                                          s_stack_buf + threadIdx.x);
        if (bi == kRayOverflow) brute_ray_cold<WT>(P.tris, P.leaf_ids, P.n_tris, r, best, bi);      // depth guard tripped
        if (!(best <= S.max_range)) { best = INFINITY; bi = -1; }
    }
    if (valid) { S.t_out[ray] = best; S.tri_out[ray] = bi; }
    if (!(P.flags & (1u | 8u))) {        // neither ORDERED nor NO_POINTS: warp-aggregated atomic append
        const bool hit = valid && bi >= 0;
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (m) {
            const unsigned lane = lane_id();
            uint32_t base = 0;
            if (lane == (unsigned)(__ffs(m) - 1)) base = atom_add_u32(S.count, (uint32_t)__popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            if (hit) {
                float x, y, z;
                hit_point(S, P.flags, best, dsx, dsy, dsz, x, y, z);
                store_point(S.points, base + __popc(m & lanemask_lt()), P.flags, x, y, z, (P.flags & 2u) ? 0 : v);
            }
        }
    }
}

// ---- second return (multi-return) ---------------------------------------------------------------------------
// The reference leaves multi-return to "an anyhit program which will produce multiple returns" (SRC/LidarShooter.hpp.in:
// 87-93, never written).  LSB200_OPT_RETURNS = 2 reports, next to the closest hit (t1, id1), the NEXT surface the ray
// would reach behind it: the lexicographically smallest (t, triangle id) > (t1, id1) -- a triangle at exactly the same
// distance counts (two sheets in contact), the first-hit triangle itself does not.  One thread per ray, plain
// depth-first walk from the root with the traversal's own conservative slab test (only the upper bound prunes; the
// lower bound is applied to accepted triangles, never to boxes).  Bit-identical to oracle/lsb_oracle.c:
// lso_trace_bvh_second.  An optional feature off the benchmarked path: no tile lists, no queues.
template <bool WT>
__global__ void __launch_bounds__(256) k_second_return(TraceParams P, int sensor_base)
{
    const SensorDev &S = P.sensors[sensor_base + blockIdx.y];
    const int ray = blockIdx.x * blockDim.x + threadIdx.x;
    if (ray >= S.n_rays) return;
    const float t1 = S.t_out[ray];
    const int id1 = S.tri_out[ray];
    float best = INFINITY; int bi = -1;
    if (id1 >= 0) {
        float smax = 0.0f;
#pragma unroll
        for (int k = 0; k < 6; ++k) smax = fmaxf(smax, fabsf(ord2f(__ldg(P.bounds + k))));
        const int v = ray / S.H, h = ray - v * S.H;
        Ray r; float dsx, dsy, dsz, omax;
        make_ray(S, P.world_frame, v, h, r, dsx, dsy, dsz, omax);
        const float eps = 8e-6f * (smax + omax) + 1e-30f;
        const float tiny = 1e-30f;
        const float ix = 1.0f / (fabsf(r.dx) < tiny ? copysignf(tiny, r.dx) : r.dx);
        const float iy = 1.0f / (fabsf(r.dy) < tiny ? copysignf(tiny, r.dy) : r.dy);
        const float iz = 1.0f / (fabsf(r.dz) < tiny ? copysignf(tiny, r.dz) : r.dz);
        const float ax = -(r.ox + eps) * ix, ay = -(r.oy + eps) * iy, az = -(r.oz + eps) * iz;
        const float bx = -(r.ox - eps) * ix, by = -(r.oy - eps) * iy, bz = -(r.oz - eps) * iz;
        auto accept = [&](float t, int g) {          // (t1, id1) < (t, g) < (best, bi), lexicographically
            if (!(t > t1 || (t == t1 && g > id1))) return;
            if (t < best || (t == best && (bi < 0 || g < bi))) { best = t; bi = g; }
        };
        auto leaf = [&](int ref) {
            const int enc = ~ref;
            const int first = enc & ((1 << kLeafShift) - 1), count = (enc >> kLeafShift) + 1;
            for (int k = 0; k < count; ++k) {
                const float4 *rec = P.tris + 3 * (size_t)(first + k);
                const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1), q2 = __ldg(rec + 2);
                float t;
                if (WT ? intersect_tri_wt(q0, q1, q2, r, t) : intersect_tri(q0, q1, q2, r, t)) accept(t, __ldg(P.leaf_ids + first + k));
            }
        };
        int stack[kStackSize];
        int sp = 0;
        bool overflow = false;
        int cur = P.top_cap;                         // Karras node 0 (a single-triangle scene has a root too: leaf + empty box)
        while (cur != 0x7fffffff) {
            if (cur < 0) { leaf(cur); cur = sp > 0 ? stack[--sp] : 0x7fffffff; continue; }
            const float4 *n = P.nodes + 4 * (size_t)cur;
            const float4 q0 = __ldg(n), q1 = __ldg(n + 1), q2 = __ldg(n + 2);
            const float2 q3 = __ldg(reinterpret_cast<const float2 *>(n + 3));
            const float far_lim = best * 1.0000005f;
            float n0, f0, n1, f1;
            slab_nf<8>(fmaf(q0.x, ix, ax), fmaf(q0.y, ix, bx), fmaf(q0.z, iy, ay), fmaf(q0.w, iy, by),
                       fmaf(q2.x, iz, az), fmaf(q2.y, iz, bz), far_lim, n0, f0);
            slab_nf<8>(fmaf(q1.x, ix, ax), fmaf(q1.y, ix, bx), fmaf(q1.z, iy, ay), fmaf(q1.w, iy, by),
                       fmaf(q2.z, iz, az), fmaf(q2.w, iz, bz), far_lim, n1, f1);
            const bool hit0 = n0 <= f0, hit1 = n1 <= f1;
            const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            if (hit0 && hit1) {
                const bool swap = n1 < n0;
                if (sp < kStackSize) stack[sp++] = swap ? c0 : c1; else overflow = true;
                cur = swap ? c1 : c0;
            } else if (hit0) cur = c0;
            else if (hit1) cur = c1;
            else cur = sp > 0 ? stack[--sp] : 0x7fffffff;
        }
        if (overflow) {                              // depth guard tripped (adversarial scene): exhaustive
            best = INFINITY; bi = -1;
            for (int k = 0; k < P.n_tris; ++k) {
                const float4 *rec = P.tris + 3 * (size_t)k;
                const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1), q2 = __ldg(rec + 2);
                float t;
                if (WT ? intersect_tri_wt(q0, q1, q2, r, t) : intersect_tri(q0, q1, q2, r, t)) accept(t, __ldg(P.leaf_ids + k));
            }
        }
        if (!(best <= S.max_range)) { best = INFINITY; bi = -1; }
    }
    S.t2_out[ray] = best;
    S.tri2_out[ray] = bi;
}

// 32-bit fill (empty-scene results: t = +inf)
__global__ void __launch_bounds__(256) k_fill_u32(uint32_t *__restrict__ p, uint32_t v, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ---- intensity model + intensity filter --------------------------------------------------------
// The reference writes the constant 64.0 (SRC/EmbreeTracer.cpp:333) and carries `use_intensity_filtering` /
// `minimum_intensity` in its device config without applying them (config/hesai-pandar-XT-32-lidar_0000.json:19-20).
// Optional post-pass over the per-ray results, off the traversal's hot path: model 1 = Lambertian cos-incidence
// 255 * |Ng . d| / (|Ng| |d|) from the hit triangle (Ng = cross(e2, e1) of the transformed vertices, the same
// normal the intersector uses); a return below the sensor's minimum intensity becomes a miss.  Fixed fp32 operation
// order == oracle/lsb_oracle.c: lso_intensity.
constexpr uint32_t kFlagIntensityBuffer = 1u << 30;      // internal: k_scatter_hits takes the intensity from S.intensity
__global__ void __launch_bounds__(256) k_intensity(SceneDev Sc, const SensorDev *__restrict__ sensors, int sensor_base, int world_frame, int model)
{
    const SensorDev &S = sensors[sensor_base + blockIdx.y];
    const int ray = blockIdx.x * blockDim.x + threadIdx.x;
    if (ray >= S.n_rays) return;
    const int g = S.tri_out[ray];
    float I = 0.0f;
    if (g >= 0) {
        I = 64.0f;
        if (model == 1) {
            int a, b, c;
            tri_vertex_ids(Sc, g, a, b, c);
            const float4 A = Sc.xverts[a], B = Sc.xverts[b], C = Sc.xverts[c];
            const float e1x = __fsub_rn(A.x, B.x), e1y = __fsub_rn(A.y, B.y), e1z = __fsub_rn(A.z, B.z);
            const float e2x = __fsub_rn(C.x, A.x), e2y = __fsub_rn(C.y, A.y), e2z = __fsub_rn(C.z, A.z);
            const float ngx = __fsub_rn(__fmul_rn(e2y, e1z), __fmul_rn(e2z, e1y));
            const float ngy = __fsub_rn(__fmul_rn(e2z, e1x), __fmul_rn(e2x, e1z));
            const float ngz = __fsub_rn(__fmul_rn(e2x, e1y), __fmul_rn(e2y, e1x));
            const int v = ray / S.H, h = ray - v * S.H;
            Ray r; float dsx, dsy, dsz, om;
            make_ray(S, world_frame, v, h, r, dsx, dsy, dsz, om);
            const float den = __fadd_rn(__fadd_rn(__fmul_rn(ngx, r.dx), __fmul_rn(ngy, r.dy)), __fmul_rn(ngz, r.dz));
            const float nn = __fadd_rn(__fadd_rn(__fmul_rn(ngx, ngx), __fmul_rn(ngy, ngy)), __fmul_rn(ngz, ngz));
            const float dd = __fadd_rn(__fadd_rn(__fmul_rn(r.dx, r.dx), __fmul_rn(r.dy, r.dy)), __fmul_rn(r.dz, r.dz));
            const float len = __fsqrt_rn(__fmul_rn(nn, dd));
            I = len > 0.0f ? __fdiv_rn(__fmul_rn(255.0f, fabsf(den)), len) : 0.0f;
        }
        if (I < S.min_intensity) { S.t_out[ray] = INFINITY; S.tri_out[ray] = -1; I = 0.0f; }
    }
    S.intensity[ray] = I;
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
        store_point(S.points, slot++, flags, x, y, z, (flags & 2u) ? 0 : v, (flags & kFlagIntensityBuffer) ? S.intensity[ray] : 64.0f);
    }
}

}  // namespace lsb
