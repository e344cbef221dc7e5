// radix_sort.cuh -- device LSD radix sort of (uint32 key, uint32 value) pairs, hand-written for sm_100a.
//
// Onesweep organisation (Adinets & Merrill 2022): ONE histogram kernel reads the keys once and produces the
// digit histograms of all four 8-bit passes; each pass is then ONE kernel that reads every pair once, ranks
// it inside an 8192-pair tile (warp-level MATCH.ANY multi-split, stable), obtains the tile's global digit
// offsets by a decoupled look-back over the preceding tiles' published counts, reorders the tile in shared
// memory and writes it out in digit runs.  Algorithmic traffic per pair: 4 (histogram) + 4 x (8 read + 8
// write) = 68 B.  Tile ids are handed out by an atomic ticket so that every tile a block waits on has already
// been started (forward progress of the look-back).
//
// Replaces the sort that is implicit in rtcCommitScene's BVH build (SRC/EmbreeTracer.cpp:290-295) -- the
// reference has no device sort; this one orders the 30-bit Morton codes for the Karras hierarchy.
#pragma once
#include "common.cuh"

namespace lsb {
namespace rsort {

constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;            // 256
constexpr int kPasses = 4;
// tile shape: measured in profiles/design_choices_r01.md (512 x 16, two CTAs per SM at 64 registers)
#ifndef LSB_SORT_THREADS
#define LSB_SORT_THREADS 512
#endif
#ifndef LSB_SORT_LOOKBACK
#define LSB_SORT_LOOKBACK 4
#endif
#ifndef LSB_SORT_ITEMS
#define LSB_SORT_ITEMS 16
#define LSB_SORT_MINB 2
#endif
constexpr int kThreads = LSB_SORT_THREADS;          // >= kRadix (thread d < 256 owns digit d of the tile)
constexpr int kWarps = kThreads / 32;
constexpr int kItems = LSB_SORT_ITEMS;              // pairs per thread
constexpr int kTile = kThreads * kItems;            // 8192 pairs per tile
constexpr uint32_t kFlagAgg = 1u << 30, kFlagIncl = 2u << 30, kFlagMask = 3u << 30, kValMask = ~kFlagMask;

inline int64_t num_tiles(int64_t n) { return (n + kTile - 1) / kTile; }
#ifdef LSB_SORT_PROF
// A/B instrumentation only: cycles thread 0 of every CTA spends in each phase of k_onesweep (summed over CTAs and passes)
__device__ unsigned long long g_sort_prof[8];
#define LSB_SORT_STAMP(k) do { if (threadIdx.x == 0) { const long long now__ = clock64(); atomicAdd(&g_sort_prof[k], (unsigned long long)(now__ - t_prof)); t_prof = now__; } } while (0)
#else
#define LSB_SORT_STAMP(k) do { } while (0)
#endif
// scratch layout (uint32): [kPasses*kRadix] global histograms -> exclusive digit bases,
//                          [kPasses] tile tickets, pad to 16, then [kPasses][tiles][kRadix] status words
inline size_t scratch_words(int64_t n) { return (size_t)kPasses * kRadix + 16 + (size_t)kPasses * num_tiles(n) * kRadix; }

// ---- kernel 1: all four digit histograms in one read of the keys ----------------------------------------
__global__ void __launch_bounds__(kThreads) k_histogram(const uint32_t *__restrict__ keys, int64_t n,
                                                        uint32_t *__restrict__ ghist)
{
    __shared__ uint32_t sh[kPasses * kRadix];
    for (int i = threadIdx.x; i < kPasses * kRadix; i += kThreads) sh[i] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * kThreads * 4;
    for (int64_t base = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4; base < n; base += stride) {
        uint32_t k[4];
        if (base + 3 < n) {
            const uint4 v = *reinterpret_cast<const uint4 *>(keys + base);
            k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
        } else {
            for (int j = 0; j < 4; ++j) k[j] = (base + j < n) ? keys[base + j] : 0u;
        }
        const int cnt = (int)min((int64_t)4, n - base);
        for (int j = 0; j < cnt; ++j) {
#pragma unroll
            for (int p = 0; p < kPasses; ++p) atomicAdd(&sh[p * kRadix + ((k[j] >> (p * kRadixBits)) & (kRadix - 1))], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kPasses * kRadix; i += kThreads)
        if (sh[i]) atomicAdd(&ghist[i], sh[i]);
}

// ---- kernel 2: exclusive scan of each pass's 256-bin histogram (one block, 4 warps = 4 passes) -----------
__global__ void __launch_bounds__(kPasses * 32) k_scan_histograms(uint32_t *__restrict__ ghist)
{
    const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t v[8], sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) { v[j] = ghist[p * kRadix + lane * 8 + j]; sum += v[j]; }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    uint32_t run = incl - sum;
#pragma unroll
    for (int j = 0; j < 8; ++j) { ghist[p * kRadix + lane * 8 + j] = run; run += v[j]; }
}

// Lanes of the warp whose 8-bit digit equals this lane's: the MATCH.ANY instruction (eight ballots give the same mask
// and were measured slower: 0.351 vs 0.327 ms per 10 M-pair sort; LSB_SORT_BALLOT selects them for the A/B).
__device__ __forceinline__ unsigned digit_peers(uint32_t d)
{
#ifndef LSB_SORT_BALLOT
    return __match_any_sync(0xffffffffu, d);
#else
    unsigned peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < kRadixBits; ++b) {
        const unsigned vote = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        peers &= ((d >> b) & 1u) ? vote : ~vote;
    }
    return peers;
#endif
}

// ---- kernel 3: one onesweep pass -------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, LSB_SORT_MINB) k_onesweep(const uint32_t *__restrict__ keys_in,
                                                       const uint32_t *__restrict__ vals_in,
                                                       uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                                                       int64_t n, int shift, const uint32_t *__restrict__ digit_base,
                                                       uint32_t *__restrict__ ticket, volatile uint32_t *status)
{
    extern __shared__ __align__(16) uint32_t s_dyn[];          // [2][kTile] reordered keys and values
    uint32_t *s_keys = s_dyn, *s_vals = s_dyn + kTile;
    __shared__ uint32_t s_whist[kWarps][kRadix];   // per-warp digit counts -> per-warp digit bases within the tile
    __shared__ uint32_t s_local[kRadix];           // tile-local exclusive digit offsets
    __shared__ uint32_t s_global[kRadix];          // global position of the tile's first pair of each digit, minus its tile-local offset
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef LSB_SORT_PROF
    long long t_prof = clock64();
#endif
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < kWarps * kRadix; i += kThreads) (&s_whist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const int64_t tile_base = (int64_t)tile * kTile;
    const int n_valid = (int)min((int64_t)kTile, n - tile_base);

    // warp-striped load: item i of lane l is element warp*512 + i*32 + l of the tile
    uint32_t key[kItems], val[kItems];
    uint16_t rank[kItems];
    {
        const int64_t wbase = tile_base + warp * (32 * kItems);
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            const int64_t idx = wbase + i * 32 + lane;
            const bool ok = idx < n;
            key[i] = ok ? keys_in[idx] : 0xffffffffu;
            val[i] = ok ? vals_in[idx] : 0u;
        }
    }
    LSB_SORT_STAMP(0);       // ticket + zeroing + barrier + load issue (the loads are only waited for below)
    // stable multi-split inside the warp
    const unsigned lt = lanemask_lt();
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const uint32_t d = (key[i] >> shift) & (kRadix - 1);
        const unsigned peers = digit_peers(d);
        const uint32_t prev = s_whist[warp][d];
        __syncwarp();
        if ((peers & lt) == 0) s_whist[warp][d] = prev + __popc(peers);
        __syncwarp();
        rank[i] = (uint16_t)(prev + __popc(peers & lt));
    }
    LSB_SORT_STAMP(1);       // load latency + ranking of warp 0
    __syncthreads();
    LSB_SORT_STAMP(2);       // barrier: the slowest warp's ranking

    // thread d < 256 owns digit d: scan over warps, publish, look back
    {
        const int d = tid;
        uint32_t run = 0;
        if (d < kRadix) {
#pragma unroll
            for (int w = 0; w < kWarps; ++w) { const uint32_t c = s_whist[w][d]; s_whist[w][d] = run; run += c; }
            volatile uint32_t *my = status + (size_t)tile * kRadix + d;
            uint32_t excl = 0;
            if (tile == 0) {
                *my = run | kFlagIncl;
            } else {
                *my = run | kFlagAgg;
                // look-back, kLook predecessors per round: their status words are independent loads, so the round trips
                // overlap instead of queueing one behind the other (the serial walk was 18 % of the pass's stall samples)
                constexpr int kLook = LSB_SORT_LOOKBACK;
                int64_t t = (int64_t)tile - 1;
                bool done = false;
                while (!done) {
                    uint32_t sv[kLook];
#pragma unroll
                    for (int j = 0; j < kLook; ++j) { sv[j] = 2u << 30; /* before tile 0: inclusive 0 */ if (t - j >= 0) sv[j] = status[(size_t)(t - j) * kRadix + d]; }
#pragma unroll
                    for (int j = 0; j < kLook; ++j) {
                        if (done) break;
                        const uint32_t f = sv[j] & kFlagMask;
                        if (f == 0) break;                      // not published yet: poll again from this tile
                        excl += sv[j] & kValMask;
                        --t;
                        if (f == kFlagIncl) done = true;
                    }
                }
                *my = (excl + run) | kFlagIncl;
            }
            s_global[d] = digit_base[d] + excl;
        }
        LSB_SORT_STAMP(3);   // scan over warps + look-back (thread 0 = digit 0)
        // tile-local exclusive scan over digits (block scan of `run`; threads beyond the radix carry 0)
        uint32_t incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t t2 = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t2; }
        __shared__ uint32_t s_wsum[kWarps];
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        uint32_t woff = 0;
#pragma unroll
        for (int w = 0; w < kRadix / 32; ++w) if (w < warp) woff += s_wsum[w];
        if (d < kRadix) {
            s_local[d] = woff + incl - run;
            s_global[d] -= woff + incl - run;      // global position of the tile's pair j of digit d = s_global[d] + j (one load per pair at write-out)
        }
    }
    __syncthreads();
    LSB_SORT_STAMP(4);       // block scan + barrier: the slowest digit's look-back

    // reorder the tile in shared memory
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const uint32_t d = (key[i] >> shift) & (kRadix - 1);
        const uint32_t pos = s_local[d] + s_whist[warp][d] + rank[i];
        s_keys[pos] = key[i];
        s_vals[pos] = val[i];
    }
    LSB_SORT_STAMP(5);       // reorder of thread 0
    __syncthreads();
    LSB_SORT_STAMP(6);       // barrier
    // write out: consecutive threads write consecutive positions of a digit run
    for (int j = tid; j < n_valid; j += kThreads) {
        const uint32_t k = s_keys[j];
        const uint32_t d = (k >> shift) & (kRadix - 1);
        const uint32_t g = s_global[d] + (uint32_t)j;
        keys_out[g] = k;
        vals_out[g] = s_vals[j];
    }
    LSB_SORT_STAMP(7);       // write-out issue
}

// Function attributes are per DEVICE: every context calls this once for its own device (lsb200_create), so that a
// process may hold contexts on several GPUs.
inline cudaError_t prepare_device()
{
    return cudaFuncSetAttribute(k_onesweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * kTile * sizeof(uint32_t)));
}

// Sorts n pairs.  keys/vals: input and (after the call) output; keys_tmp/vals_tmp: ping-pong buffers of n.
// scratch: scratch_words(n) uint32.  `passes` 8-bit passes starting at bit 0 (4 covers 32 bits; an even
// number leaves the result in keys/vals).  Returns the number of kernel launches issued.
inline int sort_pairs(uint32_t *keys, uint32_t *vals, uint32_t *keys_tmp, uint32_t *vals_tmp, int64_t n,
                      uint32_t *scratch, cudaStream_t stream)
{
    if (n <= 0) return 0;
    const int64_t tiles = num_tiles(n);
    uint32_t *ghist = scratch;
    uint32_t *tickets = scratch + kPasses * kRadix;
    uint32_t *status = scratch + kPasses * kRadix + 16;
    cudaMemsetAsync(scratch, 0, scratch_words(n) * sizeof(uint32_t), stream);
    int64_t hb = (n + kThreads * 4 - 1) / (kThreads * 4);
    const int hist_blocks = (int)(hb < kSmCount * 8 ? hb : kSmCount * 8);
    k_histogram<<<hist_blocks, kThreads, 0, stream>>>(keys, n, ghist);
    k_scan_histograms<<<1, kPasses * 32, 0, stream>>>(ghist);
    uint32_t *ki = keys, *vi = vals, *ko = keys_tmp, *vo = vals_tmp;
    for (int p = 0; p < kPasses; ++p) {
        k_onesweep<<<(unsigned)tiles, kThreads, 2 * kTile * sizeof(uint32_t), stream>>>(ki, vi, ko, vo, n, p * kRadixBits, ghist + p * kRadix,
                                                             tickets + p, status + (size_t)p * tiles * kRadix);
        uint32_t *t = ki; ki = ko; ko = t;
        t = vi; vi = vo; vo = t;
    }
    return 2 + kPasses;
}

}  // namespace rsort
}  // namespace lsb
