// primitives.cuh — hand-written device-wide building blocks: exclusive scan, ordered stream
// compaction and an LSD radix sort of (u64 key, u32 value) pairs.  No CUB/Thrust.
#pragma once
#include "common.cuh"

// =====================================================================================
// exclusive scan of uint32 (recursive: 2048 items per block)
// =====================================================================================
#define SCAN_THREADS 256
#define SCAN_ITEMS 8
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

__device__ __forceinline__ uint32_t block_exclusive_scan_u32(uint32_t v, uint32_t *total, uint32_t *sh /*>=33*/)
{
    // inclusive warp scan
    int lane = lane_id(), w = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(FULL, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) sh[w] = inc;
    __syncthreads();
    if (w == 0) {
        int nw = blockDim.x >> 5;
        uint32_t s = lane < nw ? sh[lane] : 0u, si = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(FULL, si, d);
            if (lane >= d) si += t;
        }
        if (lane < nw) sh[lane] = si - s;  // exclusive warp base
        if (lane == 31) sh[32] = si;       // block total
    }
    __syncthreads();
    uint32_t r = sh[w] + inc - v;
    *total = sh[32];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, int64_t n,
                  uint32_t *__restrict__ tile_sums)
{
    __shared__ uint32_t sh[34];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        v[i] = (base + i < n) ? in[base + i] : 0u;
        s += v[i];
    }
    uint32_t total;
    uint32_t ex = block_exclusive_scan_u32(s, &total, sh);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        if (base + i < n) out[base + i] = ex;
        ex += v[i];
    }
    if (threadIdx.x == 0 && tile_sums) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_add_kernel(uint32_t *__restrict__ out, int64_t n, const uint32_t *__restrict__ tile_offsets)
{
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t off = tile_offsets[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i)
        if (base + i < n) out[base + i] += off;
}

static int scan_rec(o3dpp_ctx *c, const uint32_t *in, uint32_t *out, int64_t n, uint32_t *scratch)
{
    int64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (tiles <= 1) {
        LAUNCH(c, scan_tiles_kernel, 1, SCAN_THREADS, 0, in, out, n, (uint32_t *)nullptr);
        return 0;
    }
    uint32_t *sums = scratch;
    LAUNCH(c, scan_tiles_kernel, (int)tiles, SCAN_THREADS, 0, in, out, n, sums);
    O3_TRY(scan_rec(c, sums, sums, tiles, scratch + tiles + 64));  // in place: tiles are disjoint
    LAUNCH(c, scan_add_kernel, (int)tiles, SCAN_THREADS, 0, out, n, (const uint32_t *)sums);
    return 0;
}

// exclusive prefix sum; in == out is allowed
static int exclusive_scan_u32(o3dpp_ctx *c, const uint32_t *in, uint32_t *out, int64_t n)
{
    if (n <= 0) return 0;
    size_t scratch = 0;
    for (int64_t m = (n + SCAN_TILE - 1) / SCAN_TILE; m > 1; m = (m + SCAN_TILE - 1) / SCAN_TILE)
        scratch += (size_t)m + 64;
    O3_TRY(c->s_scan.ensure((scratch + 64) * sizeof(uint32_t)));
    O3_TRY(scan_rec(c, in, out, n, c->s_scan.as<uint32_t>()));
    return check_launch("scan");
}

// =====================================================================================
// ordered compaction: out[k] = i for the k-th i in [0,n) with pred(i), ascending
// =====================================================================================
#define CMP_THREADS 256
#define CMP_ITEMS 8
#define CMP_TILE (CMP_THREADS * CMP_ITEMS)

template <class Pred>
__global__ void __launch_bounds__(CMP_THREADS)
compact_count_kernel(Pred pred, int64_t n, uint32_t *__restrict__ tile_counts)
{
    __shared__ uint32_t sh[8];
    int64_t base = (int64_t)blockIdx.x * CMP_TILE;
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < CMP_ITEMS; ++i) {
        int64_t j = base + (int64_t)i * CMP_THREADS + threadIdx.x;
        bool f = j < n && pred(j);
        c += __popc(__ballot_sync(FULL, f));  // identical on all lanes of the warp
    }
    if (lane_id() == 0) sh[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t s = 0;
        for (int w = 0; w < CMP_THREADS / 32; ++w) s += sh[w];
        tile_counts[blockIdx.x] = s;
    }
}

template <class Pred>
__global__ void __launch_bounds__(CMP_THREADS)
compact_write_kernel(Pred pred, int64_t n, const uint32_t *__restrict__ tile_offsets,
                     int32_t *__restrict__ out)
{
    // tile layout: item (i, thread) = base + i*256 + thread ; ascending order = (i, warp, lane)
    __shared__ uint32_t wc[CMP_ITEMS][CMP_THREADS / 32];
    int64_t base = (int64_t)blockIdx.x * CMP_TILE;
    int lane = lane_id(), w = threadIdx.x >> 5;
    uint32_t ball[CMP_ITEMS];
#pragma unroll
    for (int i = 0; i < CMP_ITEMS; ++i) {
        int64_t j = base + (int64_t)i * CMP_THREADS + threadIdx.x;
        bool f = j < n && pred(j);
        ball[i] = __ballot_sync(FULL, f);
        if (lane == 0) wc[i][w] = __popc(ball[i]);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t s = tile_offsets[blockIdx.x];
        for (int i = 0; i < CMP_ITEMS; ++i)
            for (int ww = 0; ww < CMP_THREADS / 32; ++ww) {
                uint32_t t = wc[i][ww];
                wc[i][ww] = s;
                s += t;
            }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < CMP_ITEMS; ++i) {
        if (ball[i] & (1u << lane)) {
            int64_t j = base + (int64_t)i * CMP_THREADS + threadIdx.x;
            uint32_t pos = wc[i][w] + __popc(ball[i] & ((1u << lane) - 1u));
            out[pos] = (int32_t)j;
        }
    }
}

__global__ void read_total_kernel(const uint32_t *offs, const uint32_t *counts, int64_t tiles, uint32_t *total)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) *total = tiles > 0 ? offs[tiles - 1] + counts[tiles - 1] : 0u;
}

// returns the number of selected items in *m (synchronises the stream)
template <class Pred>
static int compact_indices(o3dpp_ctx *c, Pred pred, int64_t n, int32_t *out, int64_t *m)
{
    *m = 0;
    if (n <= 0) return 0;
    int64_t tiles = (n + CMP_TILE - 1) / CMP_TILE;
    O3_TRY(c->s_flags.ensure((size_t)tiles * 2 * sizeof(uint32_t) + 256));
    uint32_t *counts = c->s_flags.as<uint32_t>();
    uint32_t *offs = counts + tiles;
    LAUNCH(c, (compact_count_kernel<Pred>), (int)tiles, CMP_THREADS, 0, pred, n, counts);
    O3_TRY(exclusive_scan_u32(c, counts, offs, tiles));
    uint32_t *d_total = c->d_small.as<uint32_t>() + 12288;  // byte 48 KiB of d_small; bounds live at the front
    LAUNCH(c, read_total_kernel, 1, 32, 0, offs, counts, tiles, d_total);
    LAUNCH(c, (compact_write_kernel<Pred>), (int)tiles, CMP_THREADS, 0, pred, n, offs, out);
    O3_TRY(check_launch("compact"));
    CU_TRY(cudaMemcpyAsync(c->h_small, d_total, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(cudaStreamSynchronize(c->stream));
    *m = *(uint32_t *)c->h_small;
    return 0;
}

// =====================================================================================
// LSD radix sort, 8-bit digits, (u64 key, u32 value) pairs, stable
// =====================================================================================
#define RS_THREADS 256
#define RS_WARPS 8
#ifndef RS_ITEMS
#define RS_ITEMS 6      // 1 536 keys per tile (4..12 measured at 10 M: within 3 %, 6 with six resident blocks best)
#endif
#define RS_TILE (RS_THREADS * RS_ITEMS)
#ifndef RS_MINB
#define RS_MINB 6
#endif

__global__ void __launch_bounds__(RS_THREADS)
rs_hist_kernel(const uint64_t *__restrict__ keys, int64_t n, int shift, uint32_t dmask,
               uint32_t *__restrict__ thist, int64_t T)
{
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    int64_t base = (int64_t)blockIdx.x * RS_TILE;
    uint64_t k[RS_ITEMS];                        // all loads first: one at a time left the kernel latency bound at 42 % of the copy rate
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        int64_t j = base + (int64_t)i * RS_THREADS + threadIdx.x;
        k[i] = j < n ? keys[j] : 0ull;
    }
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        int64_t j = base + (int64_t)i * RS_THREADS + threadIdx.x;
        if (j < n) atomicAdd(&h[(uint32_t)(k[i] >> shift) & dmask], 1u);
    }
    __syncthreads();
    thist[(int64_t)threadIdx.x * T + blockIdx.x] = h[threadIdx.x];
}

template <bool VALS>        // false: keys only (the payload rides in the key bits below the sorted range)
__global__ void __launch_bounds__(RS_THREADS, RS_MINB)
rs_scatter_kernel(const uint64_t *__restrict__ kin, const uint32_t *__restrict__ vin,
                  uint64_t *__restrict__ kout, uint32_t *__restrict__ vout, int64_t n, int shift,
                  uint32_t dmask, const uint32_t *__restrict__ toff, int64_t T)
{
    __shared__ uint32_t wcnt[RS_WARPS][256];
    __shared__ uint32_t dbase[256];
    __shared__ uint32_t sh[34];
    __shared__ uint64_t skey[RS_TILE];
    __shared__ uint32_t sval[VALS ? RS_TILE : 1];
    const int lane = lane_id(), w = threadIdx.x >> 5;
    const int64_t tbase = (int64_t)blockIdx.x * RS_TILE;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&wcnt[0][0])[i] = 0;
    __syncthreads();

    uint64_t key[RS_ITEMS];
    uint32_t val[RS_ITEMS], rk[RS_ITEMS];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        int64_t j = tbase + w * (RS_ITEMS * 32) + r * 32 + lane;
        bool valid = j < n;
        key[r] = valid ? kin[j] : 0ull;
        val[r] = !VALS ? 0u : valid ? (vin ? vin[j] : (uint32_t)j) : 0u;
        uint32_t d = (uint32_t)(key[r] >> shift) & dmask;
        uint32_t deff = valid ? d : (256u + lane);
        uint32_t peers = __match_any_sync(FULL, deff);
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader && valid) {
            old = wcnt[w][d];
            wcnt[w][d] = old + __popc(peers);
        }
        old = __shfl_sync(FULL, old, leader);
        rk[r] = old + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();
    // per digit: exclusive prefix over warps, tile total
    uint32_t tot = 0;
    {
        int d = threadIdx.x;
#pragma unroll
        for (int ww = 0; ww < RS_WARPS; ++ww) {
            uint32_t t = wcnt[ww][d];
            wcnt[ww][d] = tot;
            tot += t;
        }
    }
    uint32_t tile_total;
    uint32_t ex = block_exclusive_scan_u32(tot, &tile_total, sh);
    dbase[threadIdx.x] = ex;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        int64_t j = tbase + w * (RS_ITEMS * 32) + r * 32 + lane;
        if (j < n) {
            uint32_t d = (uint32_t)(key[r] >> shift) & dmask;
            uint32_t pos = dbase[d] + wcnt[w][d] + rk[r];
            skey[pos] = key[r];
            if (VALS) sval[pos] = val[r];
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < tile_total; i += RS_THREADS) {
        uint64_t k = skey[i];
        uint32_t d = (uint32_t)(k >> shift) & dmask;
        int64_t g = (int64_t)toff[(int64_t)d * T + blockIdx.x] + (i - dbase[d]);
        kout[g] = k;
        if (VALS) vout[g] = sval[i];
    }
}

// Sorts by key bits [0, nbits).  keys0 holds the input keys; vals are implicit iota on input
// unless vals0_initialised says vals0 already carries the payload.
// On return *keys_sorted / *vals_sorted point at whichever ping-pong buffer holds the result.
static int radix_sort_pairs(o3dpp_ctx *c, uint64_t *keys0, uint64_t *keys1, uint32_t *vals0,
                            uint32_t *vals1, int64_t n, int nbits, uint64_t **keys_sorted,
                            uint32_t **vals_sorted, bool vals0_initialised = false)
{
    int64_t T = (n + RS_TILE - 1) / RS_TILE;
    if (T < 1) T = 1;
    O3_TRY(c->s_hist.ensure((size_t)T * 256 * sizeof(uint32_t) + 1024));
    uint32_t *thist = c->s_hist.as<uint32_t>();
    uint64_t *kin = keys0, *kout = keys1;
    uint32_t *vin = vals0_initialised ? vals0 : nullptr, *vout = vals1, *vother = vals0;
    if (nbits < 1) nbits = 1;  // at least one pass so that vals are materialised
    for (int shift = 0; shift < nbits; shift += 8) {
        int bits = nbits - shift < 8 ? nbits - shift : 8;
        uint32_t dmask = (1u << bits) - 1u;
        LAUNCH(c, rs_hist_kernel, (int)T, RS_THREADS, 0, kin, n, shift, dmask, thist, T);
        O3_TRY(exclusive_scan_u32(c, thist, thist, T * 256));
        LAUNCH(c, rs_scatter_kernel<true>, (int)T, RS_THREADS, 0, kin, vin, kout, vout, n, shift, dmask, thist, T);
        O3_TRY(check_launch("radix pass"));
        uint64_t *tk = kin; kin = kout; kout = tk;
        vin = vout;
        uint32_t *tv = vout; vout = vother; vother = tv;
    }
    *keys_sorted = kin;
    *vals_sorted = vin;
    return 0;
}

// Keys only, sorted by the bits [shift0, shift0 + nbits): the caller packs its payload (the original index) into the bits
// below shift0, so a pass moves 8 bytes per element instead of 12 and the result is the stable order of the pair sort.
// RS_PACK_BITS = 24 index bits: clouds of up to 16.7 M points whose key needs at most 40 bits.
#define RS_PACK_BITS 24
static int radix_sort_packed(o3dpp_ctx *c, uint64_t *keys0, uint64_t *keys1, int64_t n, int shift0, int nbits, uint64_t **keys_sorted)
{
    int64_t T = (n + RS_TILE - 1) / RS_TILE;
    if (T < 1) T = 1;
    O3_TRY(c->s_hist.ensure((size_t)T * 256 * sizeof(uint32_t) + 1024));
    uint32_t *thist = c->s_hist.as<uint32_t>();
    uint64_t *kin = keys0, *kout = keys1;
    for (int shift = 0; shift < nbits; shift += 8) {
        int bits = nbits - shift < 8 ? nbits - shift : 8;
        uint32_t dmask = (1u << bits) - 1u;
        LAUNCH(c, rs_hist_kernel, (int)T, RS_THREADS, 0, kin, n, shift0 + shift, dmask, thist, T);
        O3_TRY(exclusive_scan_u32(c, thist, thist, T * 256));
        LAUNCH(c, rs_scatter_kernel<false>, (int)T, RS_THREADS, 0, kin, nullptr, kout, nullptr, n, shift0 + shift, dmask, thist, T);
        O3_TRY(check_launch("radix pass (packed)"));
        uint64_t *tk = kin; kin = kout; kout = tk;
    }
    *keys_sorted = kin;
    return 0;
}

__global__ void unpack_index_kernel(const uint64_t *__restrict__ packed, int64_t n, uint32_t *__restrict__ idx)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        idx[i] = (uint32_t)packed[i] & ((1u << RS_PACK_BITS) - 1u);
}
