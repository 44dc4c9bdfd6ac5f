// common.cuh — context, error plumbing and small device helpers shared by every kernel file.
// Everything in csrc/ is compiled as ONE translation unit (o3dpp.cu) with -fmad=false so that
// a*b+c is never contracted: the FP64 expressions below must round exactly like the CPU path
// they replace (Open3D 0.19 built without -ffast-math).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <string>
#include <vector>
#include "../../include/o3dpp.h"

#define O3_SM_COUNT_FALLBACK 148
#define O3_NEV 10

// tuning knobs of the neighbour search (o3dpp_set_option; the O3DPP_KNN_* environment variables are read ONCE, in
// o3dpp_create, so that sweeps need no rebuild and no call pays a getenv).  0 / negative = the built-in default.
struct KnnTune {
    float rho = 0.f;        // staging radius / home-box diagonal
    int gsplit = 0;         // cell groups to scan beyond which a staging attempt is abandoned
    int max_shrink = -1;    // shrink retries before a part is given up (12 = never halve)
    float shrink = 0.f;     // radius factor per retry
    int bits = 0;           // curve cells per axis = 2^bits
    bool force_list = false;   // estimate_normals through the FP64 list kernel (A/B timing, parity cross-checks)
    bool keep = false;      // record the neighbour sets of every tier (o3dpp_fetch_neighbors)
    int64_t query_first = 0;   // > 0: estimate_normals answers only the first query_first points (all points stay candidates)
};

static thread_local std::string g_last_error;

static inline int set_err(int code, const std::string &msg)
{
    g_last_error = msg;
    return code;
}

#define CU_TRY(expr)                                                                         \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            char _b[512];                                                                    \
            snprintf(_b, sizeof _b, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),  \
                     __FILE__, __LINE__);                                                    \
            return set_err(O3DPP_E_CUDA, _b);                                                \
        }                                                                                    \
    } while (0)

#define O3_TRY(expr)                  \
    do {                              \
        int _r = (expr);              \
        if (_r != 0) return _r;       \
    } while (0)

// grow-only device buffer
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes)
    {
        if (bytes <= cap) return 0;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + (bytes >> 3) + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            p = nullptr;
            char b[256];
            snprintf(b, sizeof b, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
            return set_err(O3DPP_E_CUDA, b);
        }
        cap = want;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return (T *)p; }
};

struct Attr3 {           // one (n,3) attribute stored as three planes (SoA), FP64
    DevBuf buf;          // 3 * stride doubles
    int64_t stride = 0;  // elements per plane (>= n, multiple of 32 so planes stay 256B aligned)
    bool present = false;
    double *x() const { return buf.as<double>(); }
    double *y() const { return buf.as<double>() + stride; }
    double *z() const { return buf.as<double>() + 2 * stride; }
};

static inline int64_t plane_stride(int64_t n) { return ((n + 31) / 32) * 32 + 32; }

struct o3dpp_ctx {
    int device = 0;
    int sm_count = O3_SM_COUNT_FALLBACK;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // o3dpp_download_async: device->host copies that overlap the next kernels
    cudaEvent_t copy_ev = nullptr;
    DevBuf stage2;                        // interleaved rows of an asynchronous download
    cudaEvent_t ev[O3_NEV] = {};
    uint64_t launches = 0;
    double timings[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool want_stats = false;              // O3DPP_STATS=1: search kernels count their work
    KnnTune tune;
    unsigned long long stats[64] = {0};

    int64_t n = 0;
    Attr3 pts, nrm, col;

    // staging for AoS <-> SoA conversion and small host round trips
    DevBuf stage;
    void *h_small = nullptr;  // pinned, 64 KiB
    DevBuf d_small;           // 64 KiB device scratch for reductions / counters

    // generic scratch (sort ping-pong, scan partials, ...)
    DevBuf s_keys0, s_keys1, s_vals0, s_vals1, s_hist, s_scan, s_scan2, s_flags;

    // voxel results (device resident until the next voxel()/upload)
    bool vox_valid = false;
    bool kth_valid = false;               // g_ncnt holds the per-query k-th distance bounds of the last estimate_normals
    int64_t vox_V = 0, vox_n = 0;
    DevBuf v_order, v_seg, v_oct, v_rep, v_rep_asc, v_count, v_centroid, v_normal, v_color, v_cubic, v_flags;
    DevBuf sel_flags, sel_idx;            // select_by_index scratch (kept: cudaMalloc/cudaFree per call cost more than the kernels)

    // KNN grid (Morton-sorted cells, 32-point cell groups and their box hierarchy)
    DevBuf g_sx, g_sy, g_sz, g_sp4, g_sidx, g_box, g_boxf, g_nbr, g_nd2, g_ncnt;
    DevBuf g_p4;                          // the points as (x, y, z, w) 32-byte records in original order: gather source of grid build / voxel
    DevBuf g_glist, g_gcnt;               // tier 1: cell-group lists of knn_groups_kernel
    DevBuf g_park;                        // tier 2: cumulants of the answered queries, [11][failed queries]
    DevBuf g_nbrpos, g_tier, g_inv;       // tier chain: neighbour positions [K][pitch], answering tier, original index -> position
    bool nbr_valid = false;               // g_nbrpos / g_tier hold the sets of the last estimate_normals (keep mode)
    int nbr_K = 0, nbr_mode = 0;
    int64_t nbr_pitch = 0;
    double nbr_radius = 0.0;

    // RANSAC
    DevBuf r_samples, r_planes, r_cnt, r_err, r_inl, r_part;
    int64_t r_ninl = -1;
    int r_H = 0, r_rn = 0;
    std::vector<int64_t> r_samples_host;
};

#define LAUNCH(ctx, kernel, grid, block, smem, ...)                          \
    do {                                                                     \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);     \
        (ctx)->launches++;                                                   \
    } while (0)

static inline int check_launch(const char *what)
{
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        char b[256];
        snprintf(b, sizeof b, "kernel launch %s failed: %s", what, cudaGetErrorString(e));
        return set_err(O3DPP_E_CUDA, b);
    }
    return 0;
}

static inline int grid_for(const o3dpp_ctx *c, int64_t work_items, int per_block, int waves_cap)
{
    int64_t g = (work_items + per_block - 1) / per_block;
    int64_t cap = (int64_t)c->sm_count * waves_cap;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

// ---- device helpers ------------------------------------------------------------------
#define FULL 0xffffffffu
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ double shfl_d(double v, int src)
{
    return __shfl_sync(FULL, v, src);
}
__device__ __forceinline__ double shfl_xor_d(double v, int m) { return __shfl_xor_sync(FULL, v, m); }
__device__ __forceinline__ double shfl_up_d(double v, int d) { return __shfl_up_sync(FULL, v, d); }
__device__ __forceinline__ double shfl_down_d(double v, int d) { return __shfl_down_sync(FULL, v, d); }

// a 32-byte point record in ONE load (LDG.E.256, sm_100): half the memory instructions and L1 wavefronts of the two 16-byte
// loads the compiler emits for a double4; read-only path (the records are never written by the kernel that gathers them).
// Measured at 10 M: finish 1.31 -> 1.26 ms, tier 2 1.94 -> 1.86 ms; no change in the sorted gather; voxel_reduce 0.2 ms SLOWER.
#ifndef O3_LD256
#define O3_LD256 1
#endif
__device__ __forceinline__ double4 ld_rec(const double4 *p)
{
#if O3_LD256
    double4 r;
    asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
    return r;
#else
    return *p;
#endif
}
__device__ __forceinline__ double pos_inf() { return __longlong_as_double(0x7ff0000000000000LL); }

// (dx*dx + dy*dy) + dz*dz, unfused: the nanoflann L2_Simple_Adaptor expression
__device__ __forceinline__ double dist2_rn(double ax, double ay, double az, double bx, double by, double bz)
{
    double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by), dz = __dsub_rn(az, bz);
    return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}
