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
};

// Eigen-ordered 3-term dot: a0*b0 + (a1*b1 + a2*b2)  (oracle/lsb_oracle.c: dot3_eigen)
__device__ __forceinline__ float dot3_eigen(float a0, float b0, float a1, float b1, float a2, float b2)
{
    return __fadd_rn(__fmul_rn(a0, b0), __fadd_rn(__fmul_rn(a1, b1), __fmul_rn(a2, b2)));
}

// ---- (1) vertex transform: p1 = t + L p ; sensor frame: p2 = Rinv (p1 - T) ------------------------------
// SRC/MeshTransformer.cpp:187-188, SRC/LidarDevice.cpp:383-391.  One thread per vertex of the whole scene;
// block-reduced scene bounds (needed to normalise the Morton codes) go to `bounds` with ordered-uint atomics.
__global__ void __launch_bounds__(256) k_transform(SceneDev S, int sensor_frame, const float *__restrict__ RinvT /*9+3*/)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (j < S.n_verts) {
        const int gi = find_geom_by_vert(S.geoms, S.n_geoms, j);
        const GeomDesc &G = S.geoms[gi];
        const float *src = reinterpret_cast<const float *>(G.verts + (size_t)(j - G.vert_off) * (size_t)G.stride);
        const float px = __ldg(src), py = __ldg(src + 1), pz = __ldg(src + 2);
        float x = __fadd_rn(G.A[3], dot3_eigen(G.A[0], px, G.A[1], py, G.A[2], pz));
        float y = __fadd_rn(G.A[7], dot3_eigen(G.A[4], px, G.A[5], py, G.A[6], pz));
        float z = __fadd_rn(G.A[11], dot3_eigen(G.A[8], px, G.A[9], py, G.A[10], pz));
        if (sensor_frame) {
            const float dx = __fsub_rn(x, RinvT[9]), dy = __fsub_rn(y, RinvT[10]), dz = __fsub_rn(z, RinvT[11]);
            x = dot3_eigen(RinvT[0], dx, RinvT[1], dy, RinvT[2], dz);
            y = dot3_eigen(RinvT[3], dx, RinvT[4], dy, RinvT[5], dz);
            z = dot3_eigen(RinvT[6], dx, RinvT[7], dy, RinvT[8], dz);
        }
        S.xverts[j] = make_float4(x, y, z, 0.0f);
        lo[0] = hi[0] = x; lo[1] = hi[1] = y; lo[2] = hi[2] = z;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    __shared__ float s_lo[8][3], s_hi[8][3];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) for (int k = 0; k < 3; ++k) { s_lo[warp][k] = lo[k]; s_hi[warp][k] = hi[k]; }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int k = threadIdx.x;
        float l = s_lo[0][k], h = s_hi[0][k];
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { l = fminf(l, s_lo[w][k]); h = fmaxf(h, s_hi[w][k]); }
        if (l <= h) { atomicMin(&S.bounds[k], f2ord(l)); atomicMax(&S.bounds[3 + k], f2ord(h)); }
    }
}

// Vertex ids (into the concatenated vertex array) of global triangle g.  Quads (v0 v1 v2 v3) are split like
// Embree does: (v0 v1 v3) and (v2 v3 v1).
__device__ __forceinline__ void tri_vertex_ids(const SceneDev &S, int g, int &a, int &b, int &c)
{
    const int gi = find_geom_by_tri(S.geoms, S.n_geoms, g);
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

// ---- (2) 30-bit Morton code of each triangle's AABB centre, normalised to the scene bounds -----------------
__global__ void __launch_bounds__(256) k_morton(SceneDev S)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= S.n_tris) return;
    int a, b, c;
    tri_vertex_ids(S, g, a, b, c);
    const float4 A = S.xverts[a], B = S.xverts[b], C = S.xverts[c];
    const float slo[3] = {ord2f(S.bounds[0]), ord2f(S.bounds[1]), ord2f(S.bounds[2])};
    const float shi[3] = {ord2f(S.bounds[3]), ord2f(S.bounds[4]), ord2f(S.bounds[5])};
    const float cx = 0.5f * (fminf(A.x, fminf(B.x, C.x)) + fmaxf(A.x, fmaxf(B.x, C.x)));
    const float cy = 0.5f * (fminf(A.y, fminf(B.y, C.y)) + fmaxf(A.y, fmaxf(B.y, C.y)));
    const float cz = 0.5f * (fminf(A.z, fminf(B.z, C.z)) + fmaxf(A.z, fmaxf(B.z, C.z)));
    float ex = shi[0] - slo[0], ey = shi[1] - slo[1], ez = shi[2] - slo[2];
    if (S.morton_uniform) ex = ey = ez = fmaxf(ex, fmaxf(ey, ez));
    const float nx = ex > 0.0f ? (cx - slo[0]) / ex : 0.0f;
    const float ny = ey > 0.0f ? (cy - slo[1]) / ey : 0.0f;
    const float nz = ez > 0.0f ? (cz - slo[2]) / ez : 0.0f;
    const uint32_t ix = (uint32_t)fminf(fmaxf(nx * 1024.0f, 0.0f), 1023.0f);
    const uint32_t iy = (uint32_t)fminf(fmaxf(ny * 1024.0f, 0.0f), 1023.0f);
    const uint32_t iz = (uint32_t)fminf(fmaxf(nz * 1024.0f, 0.0f), 1023.0f);
    S.keys[g] = (expand10(ix) << 2) | (expand10(iy) << 1) | expand10(iz);
    S.vals[g] = (uint32_t)g;
}

// ---- (3) Karras 2012 hierarchy over the sorted codes ------------------------------------------------------
__device__ __forceinline__ int delta(const uint32_t *__restrict__ keys, int n, int i, uint32_t ki, int j)
{
    if (j < 0 || j >= n) return -1;
    const uint32_t kj = __ldg(keys + j);
    return (ki == kj) ? 32 + __clz((uint32_t)i ^ (uint32_t)j) : __clz(ki ^ kj);
}

constexpr int kFitBlock = 256;         // leaves per k_fit CTA

__device__ __forceinline__ int leaf_ref(int first, int count) { return ~(first | ((count - 1) << kLeafShift)); }

// One thread per internal node i in [0, n-1).  Node i covers sorted leaves [first, last]; its left child covers
// [first, split], its right child [split+1, last].  A child whose range holds <= leaf_max triangles is
// referenced as a leaf (the internal nodes below it stay in the bottom-up pass but are never traversed).
__global__ void __launch_bounds__(256) k_hierarchy(SceneDev S, const uint32_t *__restrict__ keys)
{
    const int n = S.n_tris;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const uint32_t ki = __ldg(keys + i);
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

// ---- (4) leaf set-up + bottom-up AABB fit (rebuild and refit) --------------------------------------------
// One thread per leaf slot i: gathers the triangle's transformed vertices, writes the 48-byte intersection
// record, then climbs: it stores its box into its parent's child slot, and the second thread to arrive at a
// node merges both slots and continues.  A Karras node whose leaf range lies inside this CTA's 256 leaves (its
// index then lies in the range too) is met by two threads of THIS CTA: its arrival flag lives in shared memory
// and a block-scope fence orders the box store -- ~99 % of the nodes.  Only nodes that straddle CTAs use the
// global arrival counters, which are never reset: each receives exactly two arrivals per fit, so the first
// arrival always sees an even count.
__global__ void __launch_bounds__(kFitBlock) k_fit(SceneDev S, const uint32_t *__restrict__ sorted_ids)
{
    __shared__ uint32_t s_flags[kFitBlock];
    s_flags[threadIdx.x] = 0u;
    __syncthreads();
    const int i = blockIdx.x * kFitBlock + threadIdx.x;
    if (i >= S.n_tris) return;
    const int g = (int)__ldg(sorted_ids + i);
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
    rec[0] = make_float4(A.x, A.y, A.z, e1x);
    rec[1] = make_float4(e1y, e1z, e2x, e2y);
    rec[2] = make_float4(e2z, ngx, ngy, ngz);
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
        if (up & 2) {                                               // both children arrive from this CTA
            __threadfence_block();
            old = atomicAdd(&s_flags[p - blockIdx.x * kFitBlock], 1u);
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
