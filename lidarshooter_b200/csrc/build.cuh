// build.cuh -- per-frame scene kernels: vertex transform, Morton codes, Karras hierarchy, bottom-up AABB fit,
// top-level extraction.  Replaces MeshTransformer::transformIntoBuffer (SRC/MeshTransformer.cpp:142-205) and
// rtcCommitScene (SRC/EmbreeTracer.cpp:290-295).  All arithmetic that feeds the per-ray results is plain
// IEEE fp32 with explicit __f*_rn intrinsics (never contracted into FMA) in the order of oracle/lsb_oracle.c.
#pragma once
#include "common.cuh"

namespace lsb {

struct SceneDev {
    const GeomDesc *geoms; int n_geoms;
    int n_verts, n_tris;
    float4 *xverts;          // [n_verts] transformed vertices (w unused)
    uint32_t *bounds;        // [6] ordered-uint scene bounds: lo xyz (atomicMin), hi xyz (atomicMax)
    uint32_t *keys, *vals;   // [n_tris] Morton code / global triangle id
    float4 *tris;            // [n_tris * 3] leaf-ordered triangle records
    int32_t *leaf_ids;       // [n_tris] leaf order -> global triangle id
    float4 *nodes;           // [top_cap + n_tris] * 4  (top-level copy first, then Karras nodes)
    int32_t *node_parent;    // [n_tris] internal node -> (parent << 2 | block-local << 1 | slot), see k_hierarchy
    int32_t *leaf_parent;    // [n_tris]
    uint32_t *flags;         // [n_tris] arrival counters of the bottom-up pass (parity: never re-zeroed)
    int top_cap;             // nodes reserved for the top-level copy
    int leaf_max;
    int morton_uniform;      // 1: one scale (the largest extent) for all three axes -> cubic Morton cells
    int morton_bxy, morton_bz;   // explicit Morton layout (bits of x and y / of z) when morton_bxy > 0, see morton_plan()
    float morton_alpha;          // otherwise: bit allocation by extent^alpha (0 = round robin, 1 = cubic cells)
    struct MortonPlan *plan;     // device copy of the plan of this commit (k_morton_plan)
    int watertight;              // 1: leaf records hold the three raw vertices (LSB200_OPT_WATERTIGHT, trace.cuh: intersect_tri_wt)
};

// Eigen-ordered 3-term dot: a0*b0 + (a1*b1 + a2*b2)  (oracle/lsb_oracle.c: dot3_eigen)
__device__ __forceinline__ float dot3_eigen(float a0, float b0, float a1, float b1, float a2, float b2)
{
    return __fadd_rn(__fmul_rn(a0, b0), __fadd_rn(__fmul_rn(a1, b1), __fmul_rn(a2, b2)));
}

// ---- (1) vertex transform: p1 = t + L p ; sensor frame: p2 = Rinv (p1 - T) ------------------------------
// SRC/MeshTransformer.cpp:187-188, SRC/LidarDevice.cpp:383-391.  One thread per vertex of the whole scene;
// block-reduced scene bounds (needed to normalise the Morton codes) go to `bounds` with ordered-uint atomics.
__device__ __forceinline__ void transform_vertex(const SceneDev &S, int sensor_frame, const float *__restrict__ RinvT, int j,
                                                 float &x, float &y, float &z, int geom_hint = -1)
{
    const int gi = geom_hint >= 0 ? geom_hint : find_geom_by_vert(S.geoms, S.n_geoms, j);
    const GeomDesc &G = S.geoms[gi];
    const float *src = reinterpret_cast<const float *>(G.verts + (size_t)(j - G.vert_off) * (size_t)G.stride);
    const float px = __ldg(src), py = __ldg(src + 1), pz = __ldg(src + 2);
    x = __fadd_rn(G.A[3], dot3_eigen(G.A[0], px, G.A[1], py, G.A[2], pz));
    y = __fadd_rn(G.A[7], dot3_eigen(G.A[4], px, G.A[5], py, G.A[6], pz));
    z = __fadd_rn(G.A[11], dot3_eigen(G.A[8], px, G.A[9], py, G.A[10], pz));
    if (sensor_frame) {
        const float dx = __fsub_rn(x, RinvT[9]), dy = __fsub_rn(y, RinvT[10]), dz = __fsub_rn(z, RinvT[11]);
        x = dot3_eigen(RinvT[0], dx, RinvT[1], dy, RinvT[2], dz);
        y = dot3_eigen(RinvT[3], dx, RinvT[4], dy, RinvT[5], dz);
        z = dot3_eigen(RinvT[6], dx, RinvT[7], dy, RinvT[8], dz);
    }
    S.xverts[j] = make_float4(x, y, z, 0.0f);
}

#ifndef LSB_XFORM_PER_THREAD
#define LSB_XFORM_PER_THREAD 4
#endif
constexpr int kXformPerThread = LSB_XFORM_PER_THREAD;     // vertices per thread of k_transform
constexpr int kXformBlock = 256 * kXformPerThread;        // vertices per CTA

// One thread transforms kXformPerThread vertices (block-strided: coalesced) and keeps their bounds in registers; the warp
// reduction is ONE redux.sync per bound on the ordered-uint encoding (the shuffle tree of the one-vertex-per-thread form
// was 30 SHFL per 32 vertices: two thirds of the kernel's L1 data-pipe wavefronts, which it was bound by at 74 %).
__global__ void __launch_bounds__(256) k_transform(SceneDev S, int sensor_frame, const float *__restrict__ RinvT /*9+3*/)
{
    const int j0 = blockIdx.x * kXformBlock;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    // vertices are numbered geometry by geometry: almost every block lies inside ONE geometry, found once per block
    __shared__ int s_geom;
    if (threadIdx.x == 0) {
        const int j1 = min(j0 + kXformBlock, S.n_verts) - 1;
        const int g0 = find_geom_by_vert(S.geoms, S.n_geoms, j0);
        s_geom = (g0 == find_geom_by_vert(S.geoms, S.n_geoms, j1)) ? g0 : -1;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kXformPerThread; ++k) {
        const int j = j0 + k * 256 + (int)threadIdx.x;
        if (j < S.n_verts) {
            float x, y, z;
            transform_vertex(S, sensor_frame, RinvT, j, x, y, z, s_geom);
            lo[0] = fminf(lo[0], x); hi[0] = fmaxf(hi[0], x);       // a NaN coordinate leaves the bounds alone, as before
            lo[1] = fminf(lo[1], y); hi[1] = fmaxf(hi[1], y);
            lo[2] = fminf(lo[2], z); hi[2] = fmaxf(hi[2], z);
        }
    }
    __shared__ uint32_t s_lo[8][3], s_hi[8][3];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const uint32_t l = __reduce_min_sync(0xffffffffu, f2ord(lo[k])), h = __reduce_max_sync(0xffffffffu, f2ord(hi[k]));
        if (lane == 0) { s_lo[warp][k] = l; s_hi[warp][k] = h; }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int k = threadIdx.x;
        uint32_t l = s_lo[0][k], h = s_hi[0][k];
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { l = min(l, s_lo[w][k]); h = max(h, s_hi[w][k]); }
        if (l <= h) { atomicMin(&S.bounds[k], l); atomicMax(&S.bounds[3 + k], h); }
    }
}

// Vertex ids (into the concatenated vertex array) of global triangle g.  Quads (v0 v1 v2 v3) are split like
// Embree does: (v0 v1 v3) and (v2 v3 v1).
__device__ __forceinline__ void tri_vertex_ids(const SceneDev &S, int g, int &a, int &b, int &c, int geom_hint = -1)
{
    const int gi = geom_hint >= 0 ? geom_hint : find_geom_by_tri(S.geoms, S.n_geoms, g);
    const GeomDesc &G = S.geoms[gi];
    const int k = g - G.tri_off;
    if (G.verts_per_elem == 3) {
        const uint32_t *e = G.elems + 3 * (size_t)k;
        a = (int)__ldg(e) + G.vert_off; b = (int)__ldg(e + 1) + G.vert_off; c = (int)__ldg(e + 2) + G.vert_off;
    } else {
        const uint32_t *e = G.elems + 4 * (size_t)(k >> 1);
        const int v0 = (int)__ldg(e), v1 = (int)__ldg(e + 1), v2 = (int)__ldg(e + 2), v3 = (int)__ldg(e + 3);
        if ((k & 1) == 0) { a = v0; b = v1; c = v3; } else { a = v2; b = v3; c = v1; }
        a += G.vert_off; b += G.vert_off; c += G.vert_off;
    }
}

__device__ __forceinline__ uint32_t expand10(uint32_t v)
{
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// ---- (2) Morton code of each triangle's AABB centre, normalised to the scene bounds ------------------------
// Which key bit belongs to which axis is decided per commit from the scene's extent (MortonPlan): walking down from
// bit 31, the next bit goes to the axis whose current cell size, raised to the power alpha, is largest, and that
// cell size is halved.  alpha = 0 is the textbook round-robin x, y, z (every axis normalised to its own extent);
// alpha = 1 splits the longest side first (cubic cells); the default in between gives a flat scene (a city:
// 2 km x 2 km x 80 m) a few x / y bits ahead of z without starving z -- measured in profiles/design_choices_r01.md.
// An explicit layout (bxy bits for x and y, bz for z; the extra x / y bits on top) can be forced for experiments.
struct MortonPlan {
    uint32_t mask[3];        // key bits of x, y, z
    float scale[3];          // 2^bits of each axis
};

__device__ __forceinline__ void morton_plan(const SceneDev &S, const float slo[3], const float shi[3], MortonPlan &P)
{
    P.mask[0] = P.mask[1] = P.mask[2] = 0u;
    int bits[3] = {0, 0, 0};
    if (S.morton_bxy > 0) {
        int L = 2 * S.morton_bxy + S.morton_bz - 1;
        for (int k = 0; k < S.morton_bxy - S.morton_bz; ++k) { P.mask[0] |= 1u << L--; P.mask[1] |= 1u << L--; }
        for (int k = 0; k < S.morton_bz; ++k) { P.mask[0] |= 1u << L--; P.mask[1] |= 1u << L--; P.mask[2] |= 1u << L--; }
        bits[0] = bits[1] = S.morton_bxy; bits[2] = S.morton_bz;
    } else {
        float cur[3];
        float emax = 0.0f;
        for (int k = 0; k < 3; ++k) emax = fmaxf(emax, shi[k] - slo[k]);
        for (int k = 0; k < 3; ++k) {
            const float e = S.morton_uniform ? emax : (shi[k] - slo[k]);
            cur[k] = (e > 0.0f && emax > 0.0f) ? __powf(e / emax, S.morton_alpha) : (S.morton_alpha > 0.0f ? 0.0f : 1.0f);
        }
        for (int L = 31; L >= 0; --L) {
            int a = 0;
            if (cur[1] > cur[a]) a = 1;
            if (cur[2] > cur[a]) a = 2;
            if (bits[a] >= 16) {                         // an axis never takes more than 16 bits
                a = (a + 1) % 3;
                if (bits[a] >= 16 || cur[(a + 1) % 3] > cur[a]) a = bits[(a + 1) % 3] < 16 ? (a + 1) % 3 : a;
            }
            P.mask[a] |= 1u << L;
            ++bits[a];
            cur[a] *= 0.5f;
        }
    }
    for (int k = 0; k < 3; ++k) P.scale[k] = (float)(1u << bits[k]);
}

// bit j of v -> the j-th lowest set bit of mask
__device__ __forceinline__ uint32_t deposit_bits(uint32_t v, uint32_t mask)
{
    uint32_t key = 0;
    while (mask) {
        const uint32_t low = mask & (0u - mask);
        if (v & 1u) key |= low;
        v >>= 1;
        mask ^= low;
    }
    return key;
}

// Per-CTA lookup tables of the plan: tab[axis][nibble][value] = the key bits of that 4-bit slice of the coordinate.
struct MortonTables {
    uint32_t tab[3][4][16];
    float scale[3];
};
__device__ __forceinline__ void morton_tables_fill(const MortonPlan &P, MortonTables &T, int tid)      // tid in [0, 192)
{
    if (tid < 192) {
        const int a = tid >> 6, nib = (tid >> 4) & 3, v = tid & 15;
        T.tab[a][nib][v] = deposit_bits((uint32_t)v << (4 * nib), P.mask[a]);
    }
    if (tid < 3) T.scale[tid] = P.scale[tid];
}
__device__ __forceinline__ uint32_t morton_lookup(const MortonTables &T, int a, uint32_t i)
{
    return T.tab[a][0][i & 15u] | T.tab[a][1][(i >> 4) & 15u] | T.tab[a][2][(i >> 8) & 15u] | T.tab[a][3][(i >> 12) & 15u];
}

__device__ __forceinline__ uint32_t morton_key(const SceneDev &S, int g, const float slo[3], const float shi[3], const MortonTables &P,
                                               int geom_hint = -1)
{
    int a, b, c;
    tri_vertex_ids(S, g, a, b, c, geom_hint);
    const float4 A = S.xverts[a], B = S.xverts[b], C = S.xverts[c];
    const float cx = 0.5f * (fminf(A.x, fminf(B.x, C.x)) + fmaxf(A.x, fmaxf(B.x, C.x)));
    const float cy = 0.5f * (fminf(A.y, fminf(B.y, C.y)) + fmaxf(A.y, fmaxf(B.y, C.y)));
    const float cz = 0.5f * (fminf(A.z, fminf(B.z, C.z)) + fmaxf(A.z, fmaxf(B.z, C.z)));
    float ex = shi[0] - slo[0], ey = shi[1] - slo[1], ez = shi[2] - slo[2];
    if (S.morton_uniform) ex = ey = ez = fmaxf(ex, fmaxf(ey, ez));
    const float nx = ex > 0.0f ? (cx - slo[0]) / ex : 0.0f;
    const float ny = ey > 0.0f ? (cy - slo[1]) / ey : 0.0f;
    const float nz = ez > 0.0f ? (cz - slo[2]) / ez : 0.0f;
    const uint32_t ix = (uint32_t)fminf(fmaxf(nx * P.scale[0], 0.0f), P.scale[0] - 1.0f);
    const uint32_t iy = (uint32_t)fminf(fmaxf(ny * P.scale[1], 0.0f), P.scale[1] - 1.0f);
    const uint32_t iz = (uint32_t)fminf(fmaxf(nz * P.scale[2], 0.0f), P.scale[2] - 1.0f);
    return morton_lookup(P, 0, ix) | morton_lookup(P, 1, iy) | morton_lookup(P, 2, iz);
}

// one thread: the plan of this commit (after k_transform has reduced the scene bounds)
__global__ void k_morton_plan(SceneDev S)
{
    const float slo[3] = {ord2f(S.bounds[0]), ord2f(S.bounds[1]), ord2f(S.bounds[2])};
    const float shi[3] = {ord2f(S.bounds[3]), ord2f(S.bounds[4]), ord2f(S.bounds[5])};
    MortonPlan P;
    morton_plan(S, slo, shi, P);
    *S.plan = P;
}

#ifndef LSB_MORTON_PER_THREAD
#define LSB_MORTON_PER_THREAD 4
#endif
constexpr int kMortonPerThread = LSB_MORTON_PER_THREAD;   // triangles per thread of k_morton (block-strided)
constexpr int kMortonBlock = 256 * kMortonPerThread;      // triangles per CTA: the plan's tables, the bounds and the geometry are set up once for all of them

__global__ void __launch_bounds__(256) k_morton(SceneDev S)
{
    __shared__ MortonTables s_plan;
    __shared__ int s_geom;                   // the one geometry this block's triangles belong to, or -1 (see k_transform)
    morton_tables_fill(*S.plan, s_plan, threadIdx.x);
    const int g_base = blockIdx.x * kMortonBlock;
    if (threadIdx.x == 255) {
        const int g0 = g_base, g1 = min(g0 + kMortonBlock, S.n_tris) - 1;
        const int q0 = find_geom_by_tri(S.geoms, S.n_geoms, g0);
        s_geom = (q0 == find_geom_by_tri(S.geoms, S.n_geoms, g1)) ? q0 : -1;
    }
    __syncthreads();
    const float slo[3] = {ord2f(S.bounds[0]), ord2f(S.bounds[1]), ord2f(S.bounds[2])};
    const float shi[3] = {ord2f(S.bounds[3]), ord2f(S.bounds[4]), ord2f(S.bounds[5])};
#pragma unroll
    for (int k = 0; k < kMortonPerThread; ++k) {
        const int g = g_base + k * 256 + (int)threadIdx.x;
        if (g < S.n_tris) {
            S.keys[g] = morton_key(S, g, slo, shi, s_plan, s_geom);
            S.vals[g] = (uint32_t)g;
        }
    }
}

// ---- (3) Karras 2012 hierarchy over the sorted codes ------------------------------------------------------
// Sorted Morton codes as the hierarchy reads them: from global memory through the read-only path (multi-kernel
// pipeline) or from the shared-memory sort buffer of the fused small build (code in the high word of each entry).
struct KeysGlobal {
    const uint32_t *p;
    __device__ __forceinline__ uint32_t operator()(int j) const { return __ldg(p + j); }
};
struct KeysShared64 {
    const unsigned long long *p;
    __device__ __forceinline__ uint32_t operator()(int j) const { return (uint32_t)(p[j] >> 32); }
};

template <class Keys>
__device__ __forceinline__ int delta(const Keys &keys, int n, int i, uint32_t ki, int j)
{
    if (j < 0 || j >= n) return -1;
    const uint32_t kj = keys(j);
    return (ki == kj) ? 32 + __clz((uint32_t)i ^ (uint32_t)j) : __clz(ki ^ kj);
}

#ifndef LSB_FIT_BLOCK
#define LSB_FIT_BLOCK 256
#endif
constexpr int kFitBlock = LSB_FIT_BLOCK;         // leaves per k_fit CTA

__device__ __forceinline__ int leaf_ref(int first, int count) { return ~(first | ((count - 1) << kLeafShift)); }

// One thread per internal node i in [0, n-1).  Node i covers sorted leaves [first, last]; its left child covers
// [first, split], its right child [split+1, last].  A child whose range holds <= leaf_max triangles is
// referenced as a leaf (the internal nodes below it stay in the bottom-up pass but are never traversed).
template <class Keys>
__device__ __forceinline__ void hierarchy_node(const SceneDev &S, const Keys &keys, int i)
{
    const int n = S.n_tris;
    const uint32_t ki = keys(i);
    const int d = (delta(keys, n, i, ki, i + 1) - delta(keys, n, i, ki, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, ki, i - d);
    int lmax = 2;
    while (delta(keys, n, i, ki, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, ki, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, ki, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, ki, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int split = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const int lcount = split - first + 1, rcount = last - split;
    int c0, c1;
    // parent links: (parent << 2) | (parent's leaf range lies inside one kFitBlock-aligned block) << 1 | child slot.
    // k_fit handles such block-local parents with shared-memory arrival flags instead of global atomics.
    const int link = (i << 2) | ((first / kFitBlock == last / kFitBlock) ? 2 : 0);
    if (lcount == 1) S.leaf_parent[split] = link; else S.node_parent[split] = link;
    if (rcount == 1) S.leaf_parent[split + 1] = link | 1; else S.node_parent[split + 1] = link | 1;
    // internal children are stored in the traversal kernel's unified index space: Karras node k -> top_cap + k
    c0 = (lcount <= S.leaf_max) ? leaf_ref(first, lcount) : S.top_cap + split;
    c1 = (rcount <= S.leaf_max) ? leaf_ref(split + 1, rcount) : S.top_cap + split + 1;
    if (i == 0) S.node_parent[0] = -1;
    float4 *node = S.nodes + 4 * ((size_t)S.top_cap + i);
    node[3] = make_float4(__int_as_float(c0), __int_as_float(c1), __int_as_float(first), __int_as_float(last));
}

__global__ void __launch_bounds__(256) k_hierarchy(SceneDev S, const uint32_t *__restrict__ keys)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S.n_tris - 1) return;
    hierarchy_node(S, KeysGlobal{keys}, i);
}

// ---- (4) leaf set-up + bottom-up AABB fit (rebuild and refit) --------------------------------------------
// One thread per leaf slot i: gathers the triangle's transformed vertices, writes the 48-byte intersection
// record, then climbs: it stores its box into its parent's child slot, and the second thread to arrive at a
// node merges both slots and continues.  A Karras node whose leaf range lies inside this CTA's 256 leaves (its
// index then lies in the range too) is met by two threads of THIS CTA: its arrival flag lives in shared memory
// and a block-scope fence orders the box store -- ~99 % of the nodes.  Only nodes that straddle CTAs use the
// global arrival counters, which are never reset: each receives exactly two arrivals per fit, so the first
// arrival always sees an even count.
// Leaf slot i: triangle record + climb.  LOCAL 1: arrival flags of block-local parents (link bit 1) live in s_flags
// (indexed from flag_base), the others in the global parity counters.  LOCAL 2 (fused small build, one CTA): every
// arrival flag lives in s_flags.
template <int LOCAL>
__device__ __forceinline__ void fit_leaf(const SceneDev &S, int g, int i, uint32_t *s_flags, int flag_base)
{
    int a, b, c;
    tri_vertex_ids(S, g, a, b, c);
    const float4 A = S.xverts[a], B = S.xverts[b], C = S.xverts[c];
    // e1 = v0 - v1, e2 = v2 - v0, Ng = cross(e2, e1)   (oracle/lsb_oracle.c: tri_setup, cross3)
    const float e1x = __fsub_rn(A.x, B.x), e1y = __fsub_rn(A.y, B.y), e1z = __fsub_rn(A.z, B.z);
    const float e2x = __fsub_rn(C.x, A.x), e2y = __fsub_rn(C.y, A.y), e2z = __fsub_rn(C.z, A.z);
    const float ngx = __fsub_rn(__fmul_rn(e2y, e1z), __fmul_rn(e2z, e1y));
    const float ngy = __fsub_rn(__fmul_rn(e2z, e1x), __fmul_rn(e2x, e1z));
    const float ngz = __fsub_rn(__fmul_rn(e2x, e1y), __fmul_rn(e2y, e1x));
    float4 *rec = S.tris + 3 * (size_t)i;
    if (S.watertight) {                       // raw vertices: shared edges are evaluated on bit-identical operands
        rec[0] = make_float4(A.x, A.y, A.z, B.x);
        rec[1] = make_float4(B.y, B.z, C.x, C.y);
        rec[2] = make_float4(C.z, 0.0f, 0.0f, 0.0f);
    } else {
        rec[0] = make_float4(A.x, A.y, A.z, e1x);
        rec[1] = make_float4(e1y, e1z, e2x, e2y);
        rec[2] = make_float4(e2z, ngx, ngy, ngz);
    }
    S.leaf_ids[i] = g;

    float lox = fminf(A.x, fminf(B.x, C.x)), hix = fmaxf(A.x, fmaxf(B.x, C.x));
    float loy = fminf(A.y, fminf(B.y, C.y)), hiy = fmaxf(A.y, fmaxf(B.y, C.y));
    float loz = fminf(A.z, fminf(B.z, C.z)), hiz = fmaxf(A.z, fmaxf(B.z, C.z));
    int up = S.leaf_parent[i];
    while (up >= 0) {
        const int p = up >> 2, slot = up & 1;
        float4 *node = S.nodes + 4 * ((size_t)S.top_cap + p);
        node[slot] = make_float4(lox, hix, loy, hiy);
        reinterpret_cast<float2 *>(node + 2)[slot] = make_float2(loz, hiz);
        uint32_t old;
        if (LOCAL == 2 || (LOCAL == 1 && (up & 2))) {               // both children arrive from this CTA
            __threadfence_block();
            old = atomicAdd(&s_flags[p - flag_base], 1u);
        } else {
            __threadfence();
            old = atomicAdd(&S.flags[p], 1u);
        }
        if ((old & 1u) == 0u) return;                           // first to arrive: the sibling finishes the node
        const float4 sxy = __ldcg(node + (slot ^ 1));
        const float2 sz = __ldcg(reinterpret_cast<const float2 *>(node + 2) + (slot ^ 1));
        lox = fminf(lox, sxy.x); hix = fmaxf(hix, sxy.y);
        loy = fminf(loy, sxy.z); hiy = fmaxf(hiy, sxy.w);
        loz = fminf(loz, sz.x); hiz = fmaxf(hiz, sz.y);
        up = S.node_parent[p];
    }
}

__global__ void __launch_bounds__(kFitBlock) k_fit(SceneDev S, const uint32_t *__restrict__ sorted_ids)
{
    __shared__ uint32_t s_flags[kFitBlock];
    s_flags[threadIdx.x] = 0u;
    __syncthreads();
    const int i = blockIdx.x * kFitBlock + threadIdx.x;
    if (i >= S.n_tris) return;
    fit_leaf<1>(S, (int)__ldg(sorted_ids + i), i, s_flags, blockIdx.x * kFitBlock);
}

struct Box6 { float lox, hix, loy, hiy, loz, hiz; };
__device__ __forceinline__ void box_merge(Box6 &a, const Box6 &b)
{
    a.lox = fminf(a.lox, b.lox); a.hix = fmaxf(a.hix, b.hix);
    a.loy = fminf(a.loy, b.loy); a.hiy = fmaxf(a.hiy, b.hiy);
    a.loz = fminf(a.loz, b.loz); a.hiz = fmaxf(a.hiz, b.hiz);
}

// Writes leaf record i (triangle g) and returns its box.
__device__ __forceinline__ Box6 leaf_record(const SceneDev &S, int g, int i)
{
    int a, b, c;
    tri_vertex_ids(S, g, a, b, c);
    const float4 A = S.xverts[a], B = S.xverts[b], C = S.xverts[c];
    const float e1x = __fsub_rn(A.x, B.x), e1y = __fsub_rn(A.y, B.y), e1z = __fsub_rn(A.z, B.z);
    const float e2x = __fsub_rn(C.x, A.x), e2y = __fsub_rn(C.y, A.y), e2z = __fsub_rn(C.z, A.z);
    const float ngx = __fsub_rn(__fmul_rn(e2y, e1z), __fmul_rn(e2z, e1y));
    const float ngy = __fsub_rn(__fmul_rn(e2z, e1x), __fmul_rn(e2x, e1z));
    const float ngz = __fsub_rn(__fmul_rn(e2x, e1y), __fmul_rn(e2y, e1x));
    float4 *rec = S.tris + 3 * (size_t)i;
    if (S.watertight) {                       // raw vertices: shared edges are evaluated on bit-identical operands
        rec[0] = make_float4(A.x, A.y, A.z, B.x);
        rec[1] = make_float4(B.y, B.z, C.x, C.y);
        rec[2] = make_float4(C.z, 0.0f, 0.0f, 0.0f);
    } else {
        rec[0] = make_float4(A.x, A.y, A.z, e1x);
        rec[1] = make_float4(e1y, e1z, e2x, e2y);
        rec[2] = make_float4(e2z, ngx, ngy, ngz);
    }
    S.leaf_ids[i] = g;
    Box6 bx;
    bx.lox = fminf(A.x, fminf(B.x, C.x)); bx.hix = fmaxf(A.x, fmaxf(B.x, C.x));
    bx.loy = fminf(A.y, fminf(B.y, C.y)); bx.hiy = fmaxf(A.y, fmaxf(B.y, C.y));
    bx.loz = fminf(A.z, fminf(B.z, C.z)); bx.hiz = fmaxf(A.z, fmaxf(B.z, C.z));
    return bx;
}

// ---- (3 + 4) fused: hierarchy and boxes in ONE bottom-up pass over the sorted codes (k_tree) ---------------
// Replaces k_hierarchy + k_fit (0.23 + 0.53 ms at 10 M triangles) for scenes beyond the small fused build.  The
// binary radix tree over the (code, index) keys is unique, so it can be discovered from the leaves upwards
// (Apetrei 2014, "Fast and Simple Agglomerative LBVH Construction"): a node that covers the sorted leaves [F, L]
// merges across the boundary with the LONGER common prefix -- delta(L) > delta(F - 1): its parent splits at L
// and the node is the left child; otherwise the parent splits at F - 1 and the node is the right child.  No
// binary searches, no parent links in memory.  What this kernel adds to that scheme:
//   * Karras numbering is kept (left child of a split s is node s, right child s + 1, root 0), so the node array is
//     bit-identical to k_hierarchy + k_fit (tests/test_gpu_round2.py: test_fused_tree_equals_pipeline) and every
//     consumer (traversal, origin chains, tile lists) is untouched.  A node learns its own index when it picks its
//     parent; its two children meet at a rendezvous slot indexed by the SPLIT, which both know without it.
//   * each child leaves (box, far end of its range, delta of its outer boundary) in its half of the slot; the
//     second to arrive merges and writes the finished 64-byte node in one go.  The deltas travel with the boxes: a merged
//     node inherits its left delta from its left child and its right delta from its right child -- no key loads
//     above the leaves.
//   * rendezvous slots of splits whose node lies inside one kTreeBlock-leaf block live in SHARED memory (~99 % of the
//     nodes: both children are then completed by threads of this CTA); whether a split is block-local is decided
//     exactly, from a prefix / suffix minimum of the block's deltas.  Only nodes that straddle blocks meet in
//     global memory (rendezvous array + the parity arrival counters of k_fit).
// The same kernel refits: the sorted codes and ids of the last rebuild stay where the sort left them.
#ifndef LSB_TREE_BLOCK
#define LSB_TREE_BLOCK 128
#endif
#ifndef LSB_TREE_SKIP_LEAF_NODES
#define LSB_TREE_SKIP_LEAF_NODES 1
#endif
constexpr int kTreeBlock = LSB_TREE_BLOCK;        // leaves per k_tree CTA

// common-prefix length of the (code, index) keys of the sorted leaves j and j + 1; -1 outside the scene
__device__ __forceinline__ int boundary_delta(uint32_t ka, uint32_t kb, int j, int n)
{
    if (j < 0 || j >= n - 1) return -1;
    return (ka == kb) ? 32 + __clz((uint32_t)j ^ (uint32_t)(j + 1)) : __clz(ka ^ kb);
}

// One merge step, shared by both kernels: this child (range end / outer delta / box in registers) and its sibling's
// half of the rendezvous slot (sxy, sz) form the node that splits at s.  Writes the finished 64-byte node under its Karras
// index and returns the merged state; true when the node was the root.
struct Climber { float lox, hix, loy, hiy, loz, hiz; int F, L, dL, dR; };
__device__ __forceinline__ bool tree_merge(const SceneDev &S, Climber &c, bool left, int s, const float4 sxy, const float4 sz)
{
    const int sib_end = __float_as_int(sz.z), sib_outer = __float_as_int(sz.w);
    const int nF = left ? c.F : sib_end, nL = left ? sib_end : c.L;
    const int ndL = left ? c.dL : sib_outer, ndR = left ? sib_outer : c.dR;
    const int lcount = s - nF + 1, rcount = nL - s;
    const int c0 = (lcount <= S.leaf_max) ? leaf_ref(nF, lcount) : S.top_cap + s;
    const int c1 = (rcount <= S.leaf_max) ? leaf_ref(s + 1, rcount) : S.top_cap + s + 1;
    const bool root = ndL < 0 && ndR < 0;
    const int idx = root ? 0 : (ndR > ndL ? nL : nF);          // Karras index: left children sit at their last leaf, right at their first
    // A node whose whole range fits a leaf is referenced by its parent as a leaf (leaf_ref above), never by index:
    // nobody reads its record, so it is not written (a third of the nodes at leaf size 2: 64 bytes each).
    if (!LSB_TREE_SKIP_LEAF_NODES || root || nL - nF >= S.leaf_max) {
        float4 *node = S.nodes + 4 * ((size_t)S.top_cap + idx);
        const float4 mxy = make_float4(c.lox, c.hix, c.loy, c.hiy);
        node[0] = left ? mxy : sxy;
        node[1] = left ? sxy : mxy;
        node[2] = left ? make_float4(c.loz, c.hiz, sz.x, sz.y) : make_float4(sz.x, sz.y, c.loz, c.hiz);
        node[3] = make_float4(__int_as_float(c0), __int_as_float(c1), __int_as_float(nF), __int_as_float(nL));
    }
    c.lox = fminf(c.lox, sxy.x); c.hix = fmaxf(c.hix, sxy.y);
    c.loy = fminf(c.loy, sxy.z); c.hiy = fmaxf(c.hiy, sxy.w);
    c.loz = fminf(c.loz, sz.x); c.hiz = fmaxf(c.hiz, sz.y);
    c.F = nF; c.L = nL; c.dL = ndL; c.dR = ndR;
    return root;
}

// Phase 1: leaf records + every node whose leaf range lies inside one block, rendezvous in shared memory.  A child of a
// node that straddles blocks (the root of a maximal block-local subtree, or a leaf) leaves its half of the global
// rendezvous slot and an entry in the work list of phase 2 -- it does NOT wait or climb on, so the CTA retires as soon as
// its local subtrees are done (threads that lingered in the global climb held whole warps: 29 % of the stall samples
// of the one-kernel form, profiles/design_choices_r02.md).
// tree_block<B>: one block of B leaves [base, base + B) by B threads (all of them call it; it synchronises the CTA).
// Used by k_tree (B = kTreeBlock, one block per CTA) and by the fused small build (B = 1024, a CTA's round).
template <int B>
struct TreeSmem {
    uint32_t key[B + 2];              // codes of the leaves base - 1 .. base + B
    int8_t delta[B + 1];              // delta of the boundaries base - 1 .. base + B - 1
    int8_t sufmin[B + 2];
    int8_t wpre[B / 32], wsuf[B / 32];
    uint8_t local[B];                 // split base + t: its node lies inside this block
    uint32_t flag[B];
    uint32_t work[B];
    uint32_t nwork, wbase;
    float4 rxy[2 * B], rz[2 * B];     // per split: the left child's half and the right child's half, at rslot()
};
// Where the children of block-local split ls leave their halves.  Left halves at [ls], right halves at [B + (ls + 1) % B]:
// at the leaf level lane t deposits at split t (left child) or t - 1 (right child) -- either way its 16-byte slot index is
// t modulo 8, so the eight lanes of a quarter warp hit eight different bank groups (the interleaved [2 ls + side] layout
// put them at a 32-byte stride: two-way conflicts, twice the conflict-free wavefronts in ncu).
#ifndef LSB_TREE_SPLIT_SLOTS
#define LSB_TREE_SPLIT_SLOTS 1
#endif
template <int B>
__device__ __forceinline__ int rslot(int ls, int side)
{
#if LSB_TREE_SPLIT_SLOTS
    return side ? B + ((ls + 1) & (B - 1)) : ls;
#else
    return 2 * ls + side;
#endif
}

template <int B>
__device__ __forceinline__ void tree_block(const SceneDev &S, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ sorted_ids,
                                           float4 *__restrict__ rendezvous, uint32_t *__restrict__ work, uint32_t *__restrict__ work_count,
                                           int base, int t, TreeSmem<B> &M, bool ids_uncached)
{
    const int lane = t & 31, warp = t >> 5;
    const int n = S.n_tris;
    const int i = base + t;
    M.flag[t] = 0u;
    if (t == 0) M.nwork = 0u;
    for (int k = t; k < B + 2; k += B) {
        const int j = base - 1 + k;
        M.key[k] = (j >= 0 && j < n) ? __ldcg(keys + j) : 0u;
    }
    __syncthreads();
    // delta of boundary b = t (leaves i - 1 | i); thread 0 also owns the block's last boundary
    const int d_lo = boundary_delta(M.key[t], M.key[t + 1], i - 1, n);
    M.delta[t] = (int8_t)d_lo;
    if (t == 0) M.delta[B] = (int8_t)boundary_delta(M.key[B], M.key[B + 1], base + B - 1, n);
    // inclusive prefix / suffix minima of the deltas over the block's boundaries
    int pm = d_lo, sm = d_lo;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, pm, o), w = __shfl_down_sync(0xffffffffu, sm, o);
        if (lane >= o) pm = min(pm, u);
        if (lane + o < 32) sm = min(sm, w);
    }
    if (lane == 31) M.wpre[warp] = (int8_t)pm;
    if (lane == 0) M.wsuf[warp] = (int8_t)sm;
    __syncthreads();
    for (int w = 0; w < warp; ++w) pm = min(pm, (int)M.wpre[w]);
    for (int w = warp + 1; w < B / 32; ++w) sm = min(sm, (int)M.wsuf[w]);
    sm = min(sm, (int)M.delta[B]);
    M.sufmin[t] = (int8_t)sm;
    if (t == 0) { M.sufmin[B] = M.delta[B]; M.sufmin[B + 1] = 127; }
    __syncthreads();
    {
        // split s = i separates the leaves i | i + 1 (boundary t + 1): its node [F, L] stays inside the block iff a
        // boundary with a shorter prefix stops it on either side before the block ends
        const int d_s = M.delta[t + 1];
        M.local[t] = (t + 1 < B && pm < d_s && (int)M.sufmin[t + 2] < d_s) ? 1 : 0;
    }
    __syncthreads();
    if (i < n) {
        const Box6 leaf = leaf_record(S, (int)(ids_uncached ? __ldcg(sorted_ids + i) : __ldg(sorted_ids + i)), i);
        Climber c{leaf.lox, leaf.hix, leaf.loy, leaf.hiy, leaf.loz, leaf.hiz, i, i, d_lo, (int)M.delta[t + 1]};
        for (int level = 0; level < 96; ++level) {       // a chain is at most the tree's depth (< 64) long; the bound only guards against a hang
            const bool left = c.dR > c.dL;               // this node is the left child of the split at L
            const int s = left ? c.L : c.F - 1;
            const int ls = s - base;
            const int mine = left ? 0 : 1;
            const float4 pxy = make_float4(c.lox, c.hix, c.loy, c.hiy);
            const float4 pz = make_float4(c.loz, c.hiz, __int_as_float(left ? c.F : c.L), __int_as_float(left ? c.dL : c.dR));
            if (!(ls >= 0 && ls < B && M.local[ls])) {
                // the parent straddles blocks: park this subtree for phase 2
                float4 *slot = rendezvous + 4 * (size_t)s + 2 * mine;
                __stcg(slot, pxy);
                __stcg(slot + 1, pz);
                M.work[atomicAdd(&M.nwork, 1u)] = ((uint32_t)s << 1) | (uint32_t)mine;
                break;
            }
            M.rxy[rslot<B>(ls, mine)] = pxy;
            M.rz[rslot<B>(ls, mine)] = pz;
            __threadfence_block();
            if (atomicAdd(&M.flag[ls], 1u) == 0u) break;           // first to arrive: the sibling finishes the node
            __threadfence_block();
            if (tree_merge(S, c, left, s, M.rxy[rslot<B>(ls, mine ^ 1)], M.rz[rslot<B>(ls, mine ^ 1)])) break;
        }
    }
    __syncthreads();
    const uint32_t nw = M.nwork;
    if (t == 0 && nw) M.wbase = atomicAdd(work_count, nw);
    __syncthreads();
    if ((uint32_t)t < nw) work[M.wbase + t] = M.work[t];
}

#ifndef LSB_TREE_ROUNDS
#define LSB_TREE_ROUNDS 0
#endif
#if LSB_TREE_ROUNDS
// Round-synchronous form of tree_block (measured and rejected: -DLSB_TREE_ROUNDS=1 builds it for the A/B, design log 22;
// 0.648 vs 0.631 ms -- k_tree is bound by the L1 data pipe, gathers + scattered stores + the shared-memory rendezvous,
// 68 % of its peak, and those wavefronts are the same in both forms; the rounds only add barriers): same block set-up, same rendezvous slots, same
// parked subtrees and the same node records -- but the climb is not a chain per thread.  In tree_block a thread follows its
// own subtree upwards for as long as it is the second child to arrive, so every level halves the live lanes of a warp
// (14.7 of 32 threads per instruction over the kernel, ~100 instructions per level and warp whatever the lane count).  Here
// a node whose second child has arrived is put on a shared-memory queue; after a block barrier the queue of the round is
// dealt to threads 0 .. n - 1, so the merges of a round sit in full warps (the other warps only pass the barrier), and
// the barrier orders the children's shared-memory stores: no fences.  Rounds = height of the block's tallest subtree.
template <int B>
struct TreeRounds {
    uint16_t q[2][B / 2];             // splits (block-local) whose two children have arrived, by round parity
    uint32_t cnt[3];                  // entries of the queue of round r: cnt[r % 3]
};

template <int B>
__device__ __forceinline__ void tree_block_rounds(const SceneDev &S, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ sorted_ids,
                                                  float4 *__restrict__ rendezvous, uint32_t *__restrict__ work, uint32_t *__restrict__ work_count,
                                                  int base, int t, TreeSmem<B> &M, TreeRounds<B> &Q, bool ids_uncached)
{
    const int lane = t & 31, warp = t >> 5;
    const int n = S.n_tris;
    const int i = base + t;
    M.flag[t] = 0u;
    if (t == 0) { M.nwork = 0u; Q.cnt[0] = 0u; Q.cnt[1] = 0u; Q.cnt[2] = 0u; }
    for (int k = t; k < B + 2; k += B) {
        const int j = base - 1 + k;
        M.key[k] = (j >= 0 && j < n) ? __ldcg(keys + j) : 0u;
    }
    __syncthreads();
    const int d_lo = boundary_delta(M.key[t], M.key[t + 1], i - 1, n);
    M.delta[t] = (int8_t)d_lo;
    if (t == 0) M.delta[B] = (int8_t)boundary_delta(M.key[B], M.key[B + 1], base + B - 1, n);
    int pm = d_lo, sm = d_lo;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, pm, o), w = __shfl_down_sync(0xffffffffu, sm, o);
        if (lane >= o) pm = min(pm, u);
        if (lane + o < 32) sm = min(sm, w);
    }
    if (lane == 31) M.wpre[warp] = (int8_t)pm;
    if (lane == 0) M.wsuf[warp] = (int8_t)sm;
    __syncthreads();
    for (int w = 0; w < warp; ++w) pm = min(pm, (int)M.wpre[w]);
    for (int w = warp + 1; w < B / 32; ++w) sm = min(sm, (int)M.wsuf[w]);
    sm = min(sm, (int)M.delta[B]);
    M.sufmin[t] = (int8_t)sm;
    if (t == 0) { M.sufmin[B] = M.delta[B]; M.sufmin[B + 1] = 127; }
    __syncthreads();
    {
        const int d_s = M.delta[t + 1];
        M.local[t] = (t + 1 < B && pm < d_s && (int)M.sufmin[t + 2] < d_s) ? 1 : 0;
    }
    __syncthreads();
    // a finished subtree [F, L] with outer deltas (dL, dR) goes to its parent's slot: queue of the NEXT round when it is
    // the second child there, the global rendezvous + work list when the parent straddles blocks
    auto deposit = [&](const float4 pxy, float loz, float hiz, int F, int L, int dL, int dR, uint32_t *next_cnt, uint16_t *next_q) {
        const bool left = dR > dL;
        const int s = left ? L : F - 1;
        const int ls = s - base;
        const int mine = left ? 0 : 1;
        const float4 pz = make_float4(loz, hiz, __int_as_float(left ? F : L), __int_as_float(left ? dL : dR));
        if (!(ls >= 0 && ls < B && M.local[ls])) {
            float4 *slot = rendezvous + 4 * (size_t)s + 2 * mine;
            __stcg(slot, pxy);
            __stcg(slot + 1, pz);
            M.work[atomicAdd(&M.nwork, 1u)] = ((uint32_t)s << 1) | (uint32_t)mine;
            return;
        }
        M.rxy[rslot<B>(ls, mine)] = pxy;
        M.rz[rslot<B>(ls, mine)] = pz;
        if (atomicAdd(&M.flag[ls], 1u) != 0u) next_q[atomicAdd(next_cnt, 1u)] = (uint16_t)ls;
    };
    if (i < n) {
        const Box6 leaf = leaf_record(S, (int)(ids_uncached ? __ldcg(sorted_ids + i) : __ldg(sorted_ids + i)), i);
        deposit(make_float4(leaf.lox, leaf.hix, leaf.loy, leaf.hiy), leaf.loz, leaf.hiz, i, i, d_lo, (int)M.delta[t + 1], &Q.cnt[0], Q.q[0]);
    }
    for (int r = 0; r < B; ++r) {                          // r < B only guards against a hang
        __syncthreads();
        const int nq = (int)Q.cnt[r % 3];
        if (nq == 0) break;
        if (t == 0) Q.cnt[(r + 2) % 3] = 0u;               // last read before this round's barrier; next filled in round r + 1
        if (t < nq) {
            const int ls = Q.q[r & 1][t];
            const int s = base + ls;
            const float4 lxy = M.rxy[rslot<B>(ls, 0)], lz = M.rz[rslot<B>(ls, 0)], rxy = M.rxy[rslot<B>(ls, 1)], rz = M.rz[rslot<B>(ls, 1)];
            const int nF = __float_as_int(lz.z), ndL = __float_as_int(lz.w), nL = __float_as_int(rz.z), ndR = __float_as_int(rz.w);
            const int lcount = s - nF + 1, rcount = nL - s;
            const bool root = ndL < 0 && ndR < 0;
            if (!LSB_TREE_SKIP_LEAF_NODES || root || nL - nF >= S.leaf_max) {
                const int c0 = (lcount <= S.leaf_max) ? leaf_ref(nF, lcount) : S.top_cap + s;
                const int c1 = (rcount <= S.leaf_max) ? leaf_ref(s + 1, rcount) : S.top_cap + s + 1;
                const int idx = root ? 0 : (ndR > ndL ? nL : nF);
                float4 *node = S.nodes + 4 * ((size_t)S.top_cap + idx);
                node[0] = lxy;
                node[1] = rxy;
                node[2] = make_float4(lz.x, lz.y, rz.x, rz.y);
                node[3] = make_float4(__int_as_float(c0), __int_as_float(c1), __int_as_float(nF), __int_as_float(nL));
            }
            if (!root)
                deposit(make_float4(fminf(lxy.x, rxy.x), fmaxf(lxy.y, rxy.y), fminf(lxy.z, rxy.z), fmaxf(lxy.w, rxy.w)),
                        fminf(lz.x, rz.x), fmaxf(lz.y, rz.y), nF, nL, ndL, ndR, &Q.cnt[(r + 1) % 3], Q.q[(r + 1) & 1]);
        }
    }
    __syncthreads();
    const uint32_t nw = M.nwork;
    if (t == 0 && nw) M.wbase = atomicAdd(work_count, nw);
    __syncthreads();
    if ((uint32_t)t < nw) work[M.wbase + t] = M.work[t];
}

#endif  // LSB_TREE_ROUNDS
__global__ void __launch_bounds__(kTreeBlock) k_tree(SceneDev S, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ sorted_ids,
                                                     float4 *__restrict__ rendezvous /* [4 * n_tris] */,
                                                     uint32_t *__restrict__ work /* [n_tris]: split << 1 | side */, uint32_t *__restrict__ work_count)
{
    __shared__ TreeSmem<kTreeBlock> M;
#if LSB_TREE_ROUNDS
    __shared__ TreeRounds<kTreeBlock> Q;
    tree_block_rounds<kTreeBlock>(S, keys, sorted_ids, rendezvous, work, work_count, blockIdx.x * kTreeBlock, threadIdx.x, M, Q, false);
#else
    tree_block<kTreeBlock>(S, keys, sorted_ids, rendezvous, work, work_count, blockIdx.x * kTreeBlock, threadIdx.x, M, false);
#endif
}

// Phase 2: the nodes that straddle blocks (~1 % of the tree, its top).  One thread per parked subtree: arrival at the
// global rendezvous slot (parity counters that are never re-zeroed, see k_fit), the second to arrive merges, writes the
// node and carries on to the next split.
__device__ __forceinline__ void tree_top_entry(const SceneDev &S, float4 *__restrict__ rendezvous, uint32_t w)
{
    int s = (int)(w >> 1);
    bool left = (w & 1u) == 0u;
    Climber c;
    {
        const float4 *slot = rendezvous + 4 * (size_t)s + 2 * (left ? 0 : 1);
        const float4 pxy = __ldcg(slot), pz = __ldcg(slot + 1);
        c.lox = pxy.x; c.hix = pxy.y; c.loy = pxy.z; c.hiy = pxy.w; c.loz = pz.x; c.hiz = pz.y;
        // only the fields the merge reads on this side are known -- and needed
        c.F = left ? __float_as_int(pz.z) : s + 1; c.L = left ? s : __float_as_int(pz.z);
        c.dL = left ? __float_as_int(pz.w) : 0; c.dR = left ? 0 : __float_as_int(pz.w);
    }
    bool parked = true;                               // the payload of the first level is already in the slot
    for (int level = 0; level < 96; ++level) {
        float4 *slot = rendezvous + 4 * (size_t)s;
        const int mine = left ? 0 : 1;
        if (!parked) {
            __stcg(slot + 2 * mine, make_float4(c.lox, c.hix, c.loy, c.hiy));
            __stcg(slot + 2 * mine + 1, make_float4(c.loz, c.hiz, __int_as_float(left ? c.F : c.L), __int_as_float(left ? c.dL : c.dR)));
        }
        // arrival: ONE acquire-release atomic (releases the payload stores above, acquires the sibling's) instead of
        // fence + relaxed atomic + fence -- the chain of straddling nodes is a dozen dependent levels deep
        uint32_t old;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(S.flags + s) : "memory");
        if ((old & 1u) == 0u) break;                             // first to arrive: the sibling finishes the node
        if (tree_merge(S, c, left, s, __ldcg(slot + 2 * (mine ^ 1)), __ldcg(slot + 2 * (mine ^ 1) + 1))) break;
        left = c.dR > c.dL;
        s = left ? c.L : c.F - 1;
        parked = false;
    }
}

// The top of a SMALL tree (fused small build) finished by ONE CTA.  The nodes that straddle the 1024-leaf rounds form, per
// round boundary, a chain up to the root -- a dozen dependent levels; through global memory each level costs two
// device-scope fences and two L2 round trips (~4 k cycles: the old climb's 58 k-cycle critical path).  Here the children
// meet in a shared-memory hash table keyed by the split (CTA-scope fences, shared-memory atomics: a few hundred cycles a
// level).  At most (rounds - 1) x depth < 16 x 60 distinct splits exist, so 1024 slots never fill.
constexpr int kTopSlots = 1024;
struct TopSmem {
    int key[kTopSlots];
    uint32_t flag[kTopSlots];
    float4 rxy[2 * kTopSlots], rz[2 * kTopSlots];
};

__device__ __forceinline__ void tree_top_local(const SceneDev &S, const float4 *__restrict__ rendezvous, const uint32_t *__restrict__ work,
                                               uint32_t nw, int tid, int n_threads, TopSmem &T)
{
    for (int k = tid; k < kTopSlots; k += n_threads) { T.key[k] = -1; T.flag[k] = 0u; }
    __syncthreads();
    for (uint32_t e = (uint32_t)tid; e < nw; e += (uint32_t)n_threads) {
        const uint32_t w = __ldcg(work + e);
        int s = (int)(w >> 1);
        bool left = (w & 1u) == 0u;
        Climber c;
        {
            const float4 *slot = rendezvous + 4 * (size_t)s + 2 * (left ? 0 : 1);
            const float4 pxy = __ldcg(slot), pz = __ldcg(slot + 1);
            c.lox = pxy.x; c.hix = pxy.y; c.loy = pxy.z; c.hiy = pxy.w; c.loz = pz.x; c.hiz = pz.y;
            c.F = left ? __float_as_int(pz.z) : s + 1; c.L = left ? s : __float_as_int(pz.z);
            c.dL = left ? __float_as_int(pz.w) : 0; c.dR = left ? 0 : __float_as_int(pz.w);
        }
        for (int level = 0; level < 96; ++level) {
            // slot of split s: linear probing, claimed with a compare-and-swap on the key
            int hslot = (int)(((uint32_t)s * 2654435761u) >> 22) & (kTopSlots - 1);
            for (int probe = 0; probe < kTopSlots; ++probe) {
                const int prev = atomicCAS(&T.key[hslot], -1, s);
                if (prev == -1 || prev == s) break;
                hslot = (hslot + 1) & (kTopSlots - 1);
            }
            const int mine = left ? 0 : 1;
            T.rxy[2 * hslot + mine] = make_float4(c.lox, c.hix, c.loy, c.hiy);
            T.rz[2 * hslot + mine] = make_float4(c.loz, c.hiz, __int_as_float(left ? c.F : c.L), __int_as_float(left ? c.dL : c.dR));
            __threadfence_block();
            if (atomicAdd(&T.flag[hslot], 1u) == 0u) break;          // first to arrive: the sibling finishes the node
            __threadfence_block();
            if (tree_merge(S, c, left, s, T.rxy[2 * hslot + (mine ^ 1)], T.rz[2 * hslot + (mine ^ 1)])) break;
            left = c.dR > c.dL;
            s = left ? c.L : c.F - 1;
        }
    }
}

__global__ void __launch_bounds__(128) k_tree_top(SceneDev S, float4 *__restrict__ rendezvous, const uint32_t *__restrict__ work,
                                                  const uint32_t *__restrict__ work_count)
{
    const uint32_t nw = *work_count;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < nw; e += gridDim.x * blockDim.x)
        tree_top_entry(S, rendezvous, __ldg(work + e));
}

// ---- (4b) fit without the climb: range queries over the block's leaf boxes --------------------------------
// k_fit's climb is a chain of dependent L2 round trips (ncu: 29 % issue, 46 % DRAM).  A Karras node whose leaf range
// lies inside one 256-leaf block (99 % of the nodes; their index lies in the block too) does not need its children
// to "arrive": its child boxes are min / max over contiguous runs of the block's leaf boxes.  Each CTA therefore
//   1. writes its 256 leaf records and puts the leaf boxes into shared memory,
//   2. builds a sparse table over them (windows of 1, 2, 4 .. 64 leaves: 7 levels x 256 x 24 B = 42 KB),
//   3. lets thread t handle internal node (block base + t): two range queries (2-4 table reads each, min / max
//      tolerate overlapping windows) give both child boxes, written as three coalesced 16-byte stores,
//   4. and only the roots of the maximal block-local subtrees (and leaves hanging directly off a straddling node)
//      continue upwards through the global arrival counters, as k_fit does.
// Same boxes bit for bit (min / max are exact and order-free).
constexpr int kRmqLevels = 7;

// Continues upwards from a child of a node that straddles 256-leaf blocks (every ancestor straddles too).
__device__ __forceinline__ void climb_global(const SceneDev &S, int up, Box6 bx)
{
    while (up >= 0) {
        const int p = up >> 2, slot = up & 1;
        float4 *node = S.nodes + 4 * ((size_t)S.top_cap + p);
        node[slot] = make_float4(bx.lox, bx.hix, bx.loy, bx.hiy);
        reinterpret_cast<float2 *>(node + 2)[slot] = make_float2(bx.loz, bx.hiz);
        __threadfence();
        if ((atomicAdd(&S.flags[p], 1u) & 1u) == 0u) return;          // first to arrive: the sibling finishes the node
        const float4 sxy = __ldcg(node + (slot ^ 1));
        const float2 sz = __ldcg(reinterpret_cast<const float2 *>(node + 2) + (slot ^ 1));
        bx.lox = fminf(bx.lox, sxy.x); bx.hix = fmaxf(bx.hix, sxy.y);
        bx.loy = fminf(bx.loy, sxy.z); bx.hiy = fmaxf(bx.hiy, sxy.w);
        bx.loz = fminf(bx.loz, sz.x); bx.hiz = fmaxf(bx.hiz, sz.y);
        up = S.node_parent[p];
    }
}

struct RmqTable { float v[6][kRmqLevels][kFitBlock]; };

__device__ __forceinline__ Box6 rmq_get(const RmqTable &T, int k, int x)
{
    Box6 b;
    b.lox = T.v[0][k][x]; b.hix = T.v[1][k][x]; b.loy = T.v[2][k][x]; b.hiy = T.v[3][k][x]; b.loz = T.v[4][k][x]; b.hiz = T.v[5][k][x];
    return b;
}
// box of the block-relative leaf run [f, l]
__device__ __forceinline__ Box6 rmq_query(const RmqTable &T, int f, int l)
{
    const int len = l - f + 1;
    const int k = min(kRmqLevels - 1, 31 - __clz(len));
    const int w = 1 << k;
    Box6 b = rmq_get(T, k, f);
    for (int x = f + w; x + w <= l + 1; x += w) box_merge(b, rmq_get(T, k, x));
    box_merge(b, rmq_get(T, k, l - w + 1));
    return b;
}

__global__ void __launch_bounds__(kFitBlock) k_fit_rmq(SceneDev S, const uint32_t *__restrict__ sorted_ids)
{
    extern __shared__ __align__(16) unsigned char s_raw[];
    RmqTable &T = *reinterpret_cast<RmqTable *>(s_raw);
    const int t = threadIdx.x;
    const int base = blockIdx.x * kFitBlock;
    const int i = base + t;
    const int n = S.n_tris;
    Box6 leaf;
    leaf.lox = leaf.loy = leaf.loz = INFINITY; leaf.hix = leaf.hiy = leaf.hiz = -INFINITY;
    if (i < n) leaf = leaf_record(S, (int)__ldg(sorted_ids + i), i);
    T.v[0][0][t] = leaf.lox; T.v[1][0][t] = leaf.hix; T.v[2][0][t] = leaf.loy; T.v[3][0][t] = leaf.hiy; T.v[4][0][t] = leaf.loz; T.v[5][0][t] = leaf.hiz;
    __syncthreads();
#pragma unroll
    for (int k = 1; k < kRmqLevels; ++k) {
        const int h = 1 << (k - 1);
        if (t + 2 * h <= kFitBlock) {
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                const float a = T.v[c][k - 1][t], b = T.v[c][k - 1][t + h];
                T.v[c][k][t] = (c & 1) ? fmaxf(a, b) : fminf(a, b);
            }
        }
        __syncthreads();
    }
    // leaves hanging directly off a straddling node
    if (i < n) {
        const int up = S.leaf_parent[i];
        if (!(up & 2)) climb_global(S, up, leaf);
    }
    // internal node i (exists for i < n - 1): block-local when its leaf range lies in this block
    if (i < n - 1) {
        float4 *node = S.nodes + 4 * ((size_t)S.top_cap + i);
        const float4 q3 = node[3];
        const int c0 = __float_as_int(q3.x), first = __float_as_int(q3.z), last = __float_as_int(q3.w);
        if (first >= base && last < base + kFitBlock) {
            int split;
            if (c0 >= 0) split = c0 - S.top_cap;
            else { const int enc = ~c0; split = (enc & ((1 << kLeafShift) - 1)) + (enc >> kLeafShift); }
            const Box6 b0 = rmq_query(T, first - base, split - base);
            Box6 b1 = rmq_query(T, split + 1 - base, last - base);
            node[0] = make_float4(b0.lox, b0.hix, b0.loy, b0.hiy);
            node[1] = make_float4(b1.lox, b1.hix, b1.loy, b1.hiy);
            node[2] = make_float4(b0.loz, b0.hiz, b1.loz, b1.hiz);
            const int up = S.node_parent[i];
            if (up >= 0 && !(up & 2)) { box_merge(b1, b0); climb_global(S, up, b1); }
        }
    }
}

// ---- origin chains (see trace.cuh: trace_ray) --------------------------------------------------------------
// Walks from the root along the children whose box contains the sensor origin and lists the sibling of every step
// (box + reference), then the node or leaf the walk ends in.  Entry: e0 = (lo.x, hi.x, lo.y, hi.y), e1 = (lo.z,
// hi.z, ref, 0).  The walk stops when neither child box contains the origin, at a leaf, or when the list is nearly
// full.  Class 8 keeps every entry.  Class o < 8 (direction signs known: bit k set = component k negative) drops
// the entries that lie entirely behind the origin along some axis -- hi < o - eps on a non-negative axis, lo > o +
// eps on a negative one: no ray of that octant can reach them (eps = the traversal's own box padding), so the
// rays of a sign-uniform unit test about half of the list.  Nodes are read past L1 (they may have just been
// written by other CTAs of the fused small build).
constexpr int kMaxChain = 48;          // entries per list (2 x float4 each)

__device__ __forceinline__ void origin_chain_walk(const float4 *nodes, int top_cap, float ox, float oy, float oz, float smax,
                                                  int cls, float4 *out, int *count)
{
    const float eps = 8e-6f * (smax + fmaxf(fabsf(ox), fmaxf(fabsf(oy), fabsf(oz)))) + 1e-30f;
    auto reachable = [&](const float4 bxy, const float2 bz) -> bool {
        if (cls == 8) return true;
        const bool x_ok = (cls & 1) ? (bxy.x <= ox + eps) : (bxy.y >= ox - eps);
        const bool y_ok = (cls & 2) ? (bxy.z <= oy + eps) : (bxy.w >= oy - eps);
        const bool z_ok = (cls & 4) ? (bz.x <= oz + eps) : (bz.y >= oz - eps);
        return x_ok && y_ok && z_ok;
    };
    int n = 0;
    int cur = top_cap;                                   // Karras node 0
    float4 cb0 = make_float4(-INFINITY, INFINITY, -INFINITY, INFINITY);      // box of `cur`
    float2 cbz = make_float2(-INFINITY, INFINITY);
    int steps = 0;
    while (cur >= 0 && steps < kMaxChain - 2) {
        const float4 *nd = nodes + 4 * (size_t)cur;
        const float4 q0 = __ldcg(nd), q1 = __ldcg(nd + 1), q2 = __ldcg(nd + 2), q3 = __ldcg(nd + 3);
        const bool in0 = ox >= q0.x && ox <= q0.y && oy >= q0.z && oy <= q0.w && oz >= q2.x && oz <= q2.y;
        const bool in1 = ox >= q1.x && ox <= q1.y && oy >= q1.z && oy <= q1.w && oz >= q2.z && oz <= q2.w;
        if (!in0 && !in1) break;
        const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
        if (in0) {
            if (reachable(q1, make_float2(q2.z, q2.w))) {
                out[2 * n] = q1; out[2 * n + 1] = make_float4(q2.z, q2.w, __int_as_float(c1), 0.0f); ++n;
            }
            cur = c0; cb0 = q0; cbz = make_float2(q2.x, q2.y);
        } else {
            if (reachable(q0, make_float2(q2.x, q2.y))) {
                out[2 * n] = q0; out[2 * n + 1] = make_float4(q2.x, q2.y, __int_as_float(c0), 0.0f); ++n;
            }
            cur = c1; cb0 = q1; cbz = make_float2(q2.z, q2.w);
        }
        ++steps;
    }
    if (reachable(cb0, cbz)) { out[2 * n] = cb0; out[2 * n + 1] = make_float4(cbz.x, cbz.y, __int_as_float(cur), 0.0f); ++n; }
    *count = n;
}

// ---- fused build of a small scene: ONE thread-block cluster ------------------------------------------------
// Scenes of a few thousand triangles (the reference's own: XT-32 + ground + ben = 5 651) are launch-latency bound:
// eleven tiny dependent launches cost more than the work.  One cluster of 8 CTAs x 1024 threads (8 SMs of one GPC)
// does the whole commit instead, with hardware cluster barriers between the phases:
//   transform + bounds (per-CTA partial bounds, read by every CTA through distributed shared memory)
//   Morton codes
//   LSD radix sort, 4 x 8 bits: per-warp digit counts with MATCH.ANY, per-CTA digit totals exchanged through
//   distributed shared memory, stable scatter (global ping-pong buffers, L2-resident)
//   Karras hierarchy with the sorted codes staged in each CTA's shared memory
//   leaf records + bottom-up boxes (same arrival-flag scheme as k_fit)
//   origin chains of up to 8 sensors (saves the k_origin_chain launch of the next trace)
// Same tree, same records, same results as the multi-kernel pipeline (tests/test_gpu_parity.py).
constexpr int kClusterCtas = 8;
constexpr int kSmallThreads = 1024;
constexpr int kSmallMaxTris = 16384;      // sorted codes of the whole scene fit in each CTA's shared memory (64 KB)
constexpr int kSmallMaxGeoms = 16;        // descriptors travel as kernel parameters: no H2D copy per commit
struct SmallArgs {
    GeomDesc geoms[kSmallMaxGeoms];
    float RinvT[12];                      // sensor-frame mode: Rinv (9) and T (3) of sensor 0
    uint32_t *keys_tmp, *vals_tmp;        // ping-pong partner of S.keys / S.vals
    float origins[8][3];                  // ray origins of the first n_origins sensors (sensor frame: one origin, 0)
    int n_origins;                        // 0 = the trace call computes the origin chains itself
    float4 *chains; int *chain_counts;    // [n_origins][9] lists (trace.cuh: TraceParams)
    float4 *rendezvous;                   // tree phase: global rendezvous slots of the nodes that straddle CTA rounds (k_tree)
    uint32_t *work, *work_count;          // tree phase: parked subtrees
};

struct KeysShared {
    const uint32_t *p;
    __device__ __forceinline__ uint32_t operator()(int j) const { return p[j]; }
};

__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `p` (a shared-memory variable of this CTA) in the shared memory of CTA `rank` of the cluster
__device__ __forceinline__ uint32_t dsmem_ld_u32(const uint32_t *p, int rank)
{
    uint32_t remote, v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(rank));
    asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(remote) : "memory");
    return v;
}

__global__ void __cluster_dims__(kClusterCtas, 1, 1) __launch_bounds__(kSmallThreads, 1)
k_build_small(SceneDev S, const __grid_constant__ SmallArgs args, int sensor_frame, int rebuild)
{
    extern __shared__ __align__(16) uint32_t s_codes[];                  // the tree phase's TreeSmem<kSmallThreads>
    __shared__ uint32_t s_hist[kSmallThreads / 32][256];                 // per-warp digit counts -> running offsets
    __shared__ uint32_t s_ctot[256], s_base[256], s_wsum[8];
    __shared__ uint32_t s_bounds[6], s_gbounds[6];
    __shared__ uint32_t s_flags[kSmallThreads];
    S.geoms = args.geoms;
    const float *RinvT = args.RinvT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rank = blockIdx.x;                                          // one cluster: rank in cluster == blockIdx.x
    constexpr int kAll = kClusterCtas * kSmallThreads;
    const int gtid = rank * kSmallThreads + tid;
    const int n = S.n_tris;
#ifdef LSB_SMALL_CLK
    long long clk[16]; int nclk = 0;
#define LSB_CLK() do { if (gtid == 0) clk[nclk++] = clock64(); } while (0)
#else
#define LSB_CLK() do { } while (0)
#endif
    LSB_CLK();
    if (tid < 3) s_bounds[tid] = 0xFFFFFFFFu; else if (tid < 6) s_bounds[tid] = 0u;
    if (gtid == 0) *args.work_count = 0u;      // parked-subtree counter of the tree phase (ordered by the cluster barriers in between)
    __syncthreads();
    // (1) transform + bounds
    {
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int j = gtid; j < S.n_verts; j += kAll) {
            float x, y, z;
            transform_vertex(S, sensor_frame, RinvT, j, x, y, z);
            lo[0] = fminf(lo[0], x); hi[0] = fmaxf(hi[0], x);
            lo[1] = fminf(lo[1], y); hi[1] = fmaxf(hi[1], y);
            lo[2] = fminf(lo[2], z); hi[2] = fmaxf(hi[2], z);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
                hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
            }
            if (lane == 0 && lo[k] <= hi[k]) { atomicMin(&s_bounds[k], f2ord(lo[k])); atomicMax(&s_bounds[3 + k], f2ord(hi[k])); }
        }
    }
    cluster_sync_all();
    LSB_CLK();
    if (tid < 6) {
        uint32_t v = dsmem_ld_u32(&s_bounds[tid], 0);
        for (int r = 1; r < kClusterCtas; ++r) {
            const uint32_t w = dsmem_ld_u32(&s_bounds[tid], r);
            v = tid < 3 ? min(v, w) : max(v, w);
        }
        s_gbounds[tid] = v;
        if (rank == 0) S.bounds[tid] = v;
    }
    __syncthreads();
    if (rebuild) {
        // (2) Morton codes
        const float slo[3] = {ord2f(s_gbounds[0]), ord2f(s_gbounds[1]), ord2f(s_gbounds[2])};
        const float shi[3] = {ord2f(s_gbounds[3]), ord2f(s_gbounds[4]), ord2f(s_gbounds[5])};
        __shared__ MortonPlan s_plan0;
        __shared__ MortonTables s_plan;
        if (tid == 0) morton_plan(S, slo, shi, s_plan0);
        __syncthreads();
        morton_tables_fill(s_plan0, s_plan, tid);
        __syncthreads();
        for (int g = gtid; g < n; g += kAll) { S.keys[g] = morton_key(S, g, slo, shi, s_plan); S.vals[g] = (uint32_t)g; }
        cluster_sync_all();
        LSB_CLK();
        // (3) stable LSD radix sort; warp w of CTA r owns the contiguous chunk [(r * 32 + w) * per_warp, +per_warp)
        const int per_warp = (((n + kAll / 32 - 1) / (kAll / 32)) + 31) & ~31;
        const int chunk_lo = (rank * (kSmallThreads / 32) + warp) * per_warp;
        const int chunk_hi = min(n, chunk_lo + per_warp);
        uint32_t *src_k = S.keys, *src_v = S.vals, *dst_k = args.keys_tmp, *dst_v = args.vals_tmp;
        for (int shift = 0; shift < 32; shift += 8) {
            for (int k = tid; k < (kSmallThreads / 32) * 256; k += kSmallThreads) (&s_hist[0][0])[k] = 0u;
            __syncthreads();
            for (int base = chunk_lo; base < chunk_hi; base += 32) {
                const int i = base + lane;
                const uint32_t d = i < chunk_hi ? ((__ldcg(src_k + i) >> shift) & 255u) : (256u + (uint32_t)lane);
                const unsigned m = __match_any_sync(0xffffffffu, d);
                if (d < 256u && lane == __ffs(m) - 1) s_hist[warp][d] += (uint32_t)__popc(m);
                __syncwarp();
            }
            __syncthreads();
            if (tid < 256) {                     // counts -> exclusive prefix over the warps of this CTA, and the CTA total
                uint32_t run = 0;
                for (int w = 0; w < kSmallThreads / 32; ++w) { const uint32_t c = s_hist[w][tid]; s_hist[w][tid] = run; run += c; }
                s_ctot[tid] = run;
            }
            cluster_sync_all();
            if (tid < 256) {                     // digit total over the cluster, and what earlier CTAs hold of this digit
                uint32_t total = 0, before = 0;
                for (int r = 0; r < kClusterCtas; ++r) {
                    const uint32_t v = dsmem_ld_u32(&s_ctot[tid], r);
                    before += r < rank ? v : 0u;
                    total += v;
                }
                uint32_t incl = total;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
                if (lane == 31) s_wsum[warp] = incl;
                s_base[tid] = incl - total + before;      // + sum of the earlier warps' totals, added below
            }
            __syncthreads();
            if (tid < 256) { uint32_t add = 0; for (int w = 0; w < warp; ++w) add += s_wsum[w]; s_base[tid] += add; }
            __syncthreads();
            for (int base = chunk_lo; base < chunk_hi; base += 32) {
                const int i = base + lane;
                const bool ok = i < chunk_hi;
                const uint32_t key = ok ? __ldcg(src_k + i) : 0u, val = ok ? __ldcg(src_v + i) : 0u;
                const uint32_t d = ok ? ((key >> shift) & 255u) : (256u + (uint32_t)lane);
                const unsigned m = __match_any_sync(0xffffffffu, d);
                if (ok) {
                    const uint32_t pos = s_base[d] + s_hist[warp][d] + (uint32_t)__popc(m & lanemask_lt());
                    dst_k[pos] = key; dst_v[pos] = val;
                }
                __syncwarp();
                if (ok && lane == __ffs(m) - 1) s_hist[warp][d] += (uint32_t)__popc(m);
                __syncwarp();
            }
            cluster_sync_all();
            LSB_CLK();
            uint32_t *t = src_k; src_k = dst_k; dst_k = t;
            t = src_v; src_v = dst_v; dst_v = t;
        }
        // (4) single-triangle scene: the root is (leaf 0, empty box); larger scenes get their hierarchy from the tree phase below
        if (n == 1 && gtid == 0) {
            float4 *node = S.nodes + 4 * (size_t)S.top_cap;
            node[1] = make_float4(INFINITY, -INFINITY, INFINITY, -INFINITY);
            reinterpret_cast<float2 *>(node + 2)[1] = make_float2(INFINITY, -INFINITY);
            node[3] = make_float4(__int_as_float(leaf_ref(0, 1)), __int_as_float(leaf_ref(0, 1)), __int_as_float(0), __int_as_float(0));
            S.leaf_parent[0] = 0; S.node_parent[0] = -1; S.flags[0] |= 1u;
        }
        if (n == 1) cluster_sync_all();
        LSB_CLK();
    }
    if (n == 1) {
        // (5') the lone leaf: record + its box into slot 0 of the root (the arrival protocol of k_fit, parity counter)
        if (rank == 0) {
            s_flags[tid] = 0u;
            __syncthreads();
            if (tid == 0) fit_leaf<1>(S, (int)__ldcg(S.vals), 0, s_flags, 0);
        }
    } else {
        // (5) hierarchy AND boxes in one bottom-up pass (tree_block, the scheme of k_tree): CTA r, round k owns the leaves
        // [k * 8192 + r * 1024, +1024); nodes inside a round meet in shared memory, the few that straddle rounds are
        // parked and finished after a cluster barrier by tree_top_entry.  The same pass refits (sorted codes persist).
        TreeSmem<kSmallThreads> &M = *reinterpret_cast<TreeSmem<kSmallThreads> *>(s_codes);
        for (int base = rank * kSmallThreads; base < n; base += kAll) {
            tree_block<kSmallThreads>(S, S.keys, S.vals, args.rendezvous, args.work, args.work_count, base, tid, M, true);
            __syncthreads();
        }
        cluster_sync_all();
        // the nodes that straddle rounds: one CTA, rendezvous in its shared memory (tree_top_local)
        if (rank == 0) tree_top_local(S, args.rendezvous, args.work, __ldcg(args.work_count), tid, kSmallThreads, *reinterpret_cast<TopSmem *>(s_codes));
    }
    cluster_sync_all();                       // every box is final; no CTA leaves while its shared memory may still be read
    // (6) origin chains
    if (rank == 0 && tid < args.n_origins * 9) {
        const int si = tid / 9, cls = tid - si * 9;
        float smax = 0.0f;
        for (int k = 0; k < 6; ++k) smax = fmaxf(smax, fabsf(ord2f(s_gbounds[k])));
        origin_chain_walk(S.nodes, S.top_cap, args.origins[si][0], args.origins[si][1], args.origins[si][2], smax, cls,
                          args.chains + (size_t)tid * (2 * kMaxChain), args.chain_counts + tid);
    }
    LSB_CLK();
#ifdef LSB_SMALL_CLK
    if (gtid == 0) { printf("small build clocks:"); for (int k = 1; k < nclk; ++k) printf(" %lld", clk[k] - clk[k - 1]); printf("\n"); }
#endif
}

// Single-triangle scene: node 0 = (leaf 0, empty box).
__global__ void k_single_leaf_root(SceneDev S)
{
    float4 *node = S.nodes + 4 * (size_t)S.top_cap;
    node[1] = make_float4(INFINITY, -INFINITY, INFINITY, -INFINITY);
    reinterpret_cast<float2 *>(node + 2)[1] = make_float2(INFINITY, -INFINITY);
    node[3] = make_float4(__int_as_float(leaf_ref(0, 1)), __int_as_float(leaf_ref(0, 1)), __int_as_float(0), __int_as_float(0));
    S.leaf_parent[0] = 0;            // parent 0, slot 0, global arrival counter
    S.node_parent[0] = -1;
    S.flags[0] |= 1u;    // the lone leaf arrives as the "second" child, so the counter is even again after the fit
}

// ---- (5) top-level extraction -----------------------------------------------------------------------------
// Breadth-first copy of the first `top_cap` internal nodes below the root into the top block, stored as four
// planes of float4 (plane q of slot s at nodes[q * top_cap + s]) so that shared-memory reads of different
// nodes by the lanes of a warp spread over the banks.  Child references use the unified index space:
//     ref >= 0 && ref < top_cap : slot in the top block (staged into shared memory by TMA bulk copy)
//     ref >= top_cap            : Karras node (ref - top_cap), AoS at nodes + 4 * ref
//     ref < 0                   : leaf reference
// One block; each sweep processes one level in parallel.  The order inside a level is whatever the shared
// atomic hands out -- the block is a cache, not part of the results.
__global__ void __launch_bounds__(1024) k_top_level(SceneDev S, int *__restrict__ n_top_out)
{
    extern __shared__ int s_queue[];          // [top_cap] Karras index of the node in each top slot
    __shared__ int s_count, s_level_begin, s_level_end;
    const int cap = S.top_cap;
    if (threadIdx.x == 0) { s_queue[0] = 0; s_count = 1; s_level_begin = 0; s_level_end = 1; }
    __syncthreads();
    while (true) {
        const int lb = s_level_begin, le = s_level_end;
        if (lb >= le) break;
        for (int slot = lb + threadIdx.x; slot < le; slot += blockDim.x) {
            const int k = s_queue[slot];
            const float4 *src = S.nodes + 4 * ((size_t)cap + k);
            float4 q0 = src[0], q1 = src[1], q2 = src[2], q3 = src[3];
            int c[2] = {__float_as_int(q3.x), __float_as_int(q3.y)};
#pragma unroll
            for (int s = 0; s < 2; ++s) {
                if (c[s] >= 0) {
                    const int pos = atomicAdd(&s_count, 1);
                    if (pos < cap) { s_queue[pos] = c[s] - cap; c[s] = pos; }
                }
            }
            S.nodes[slot] = q0; S.nodes[(size_t)cap + slot] = q1; S.nodes[2 * (size_t)cap + slot] = q2;
            S.nodes[3 * (size_t)cap + slot] = make_float4(__int_as_float(c[0]), __int_as_float(c[1]), q3.z, q3.w);
        }
        __syncthreads();
        if (threadIdx.x == 0) { s_level_begin = le; s_level_end = min(s_count, cap); s_count = min(s_count, cap); }
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_top_out = s_level_end;
}

}  // namespace lsb
