// o3dpp.cu — the C ABI declared in include/o3dpp.h.  Single translation unit:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false
//        -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=hidden -shared -o libo3dpp_b200.so o3dpp.cu
// There is no CPU path in this library: every entry point needs a live CUDA context.
#include "extras.cuh"
#include <stdlib.h>
#include <mutex>

#define CTX_CHECK(c)                                                      \
    do {                                                                  \
        if (!(c)) return set_err(O3DPP_E_ARG, "null context");            \
        cudaError_t _e = cudaSetDevice((c)->device);                      \
        if (_e != cudaSuccess) return set_err(O3DPP_E_CUDA, cudaGetErrorString(_e)); \
    } while (0)

extern "C" {

O3DPP_API const char *o3dpp_last_error(void) { return g_last_error.c_str(); }
O3DPP_API const char *o3dpp_version(void) { return "o3dpp_b200 0.1 (sm_100a)"; }

O3DPP_API int o3dpp_device_count(int *count)
{
    if (!count) return set_err(O3DPP_E_ARG, "null count");
    *count = 0;
    CU_TRY(cudaGetDeviceCount(count));
    return 0;
}

O3DPP_API int o3dpp_create(int device, o3dpp_ctx **out)
{
    if (!out) return set_err(O3DPP_E_ARG, "null out");
    *out = nullptr;
    int cnt = 0;
    CU_TRY(cudaGetDeviceCount(&cnt));
    if (cnt <= 0) return set_err(O3DPP_E_CUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= cnt) return set_err(O3DPP_E_ARG, "device index out of range");
    CU_TRY(cudaSetDevice(device));
    o3dpp_ctx *c = new o3dpp_ctx();
    c->device = device;
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount > 0 ? prop.multiProcessorCount : O3_SM_COUNT_FALLBACK;
    CU_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (int i = 0; i < O3_NEV; ++i) CU_TRY(cudaEventCreate(&c->ev[i]));
    CU_TRY(cudaHostAlloc(&c->h_small, 64 * 1024, cudaHostAllocDefault));
    O3_TRY(c->d_small.ensure(64 * 1024));
    // tuning defaults from the environment, read once per context (sweeps without a rebuild); o3dpp_set_option overrides
    const char *e = getenv("O3DPP_STATS");
    c->want_stats = e && e[0] == '1';
    if ((e = getenv("O3DPP_KNN_RHO"))) c->tune.rho = (float)atof(e);
    if ((e = getenv("O3DPP_KNN_GSPLIT"))) c->tune.gsplit = atoi(e);
    if ((e = getenv("O3DPP_KNN_MAXSHRINK"))) c->tune.max_shrink = atoi(e);
    if ((e = getenv("O3DPP_KNN_SHRINK"))) c->tune.shrink = (float)atof(e);
    if ((e = getenv("O3DPP_KNN_BITS"))) c->tune.bits = atoi(e);
    if (getenv("O3DPP_KNN_LIST")) c->tune.force_list = true;
    *out = c;
    return 0;
}

O3DPP_API int o3dpp_destroy(o3dpp_ctx *c)
{
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); cudaEventDestroy(c->copy_ev); }
    c->stage2.release(); c->v_flags.release(); c->sel_flags.release(); c->sel_idx.release();
    DevBuf *bufs[] = {&c->pts.buf, &c->nrm.buf, &c->col.buf, &c->stage, &c->d_small, &c->s_keys0, &c->s_keys1,
                      &c->s_vals0, &c->s_vals1, &c->s_hist, &c->s_scan, &c->s_scan2, &c->s_flags, &c->v_order,
                      &c->v_seg, &c->v_oct, &c->v_rep, &c->v_rep_asc, &c->v_count, &c->v_centroid, &c->v_normal,
                      &c->v_color, &c->v_cubic, &c->g_sx, &c->g_sy, &c->g_sz, &c->g_sidx, &c->g_box, &c->g_nbr,
                      &c->g_nd2, &c->g_ncnt, &c->g_sp4, &c->g_boxf, &c->g_nbrpos, &c->g_tier, &c->g_inv, &c->g_glist, &c->g_gcnt, &c->g_p4, &c->g_park, &c->r_samples, &c->r_planes, &c->r_cnt, &c->r_err, &c->r_inl, &c->r_part};
    for (DevBuf *b : bufs) b->release();
    if (c->h_small) cudaFreeHost(c->h_small);
    for (int i = 0; i < O3_NEV; ++i) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return 0;
}

O3DPP_API int o3dpp_synchronize(o3dpp_ctx *c)
{
    CTX_CHECK(c);
    CU_TRY(cudaStreamSynchronize(c->stream));
    if (c->copy_stream) CU_TRY(cudaStreamSynchronize(c->copy_stream));
    return 0;
}

O3DPP_API int o3dpp_stream(o3dpp_ctx *c, uintptr_t *stream)
{
    if (!c || !stream) return set_err(O3DPP_E_ARG, "null argument");
    *stream = (uintptr_t)c->stream;
    return 0;
}

O3DPP_API int o3dpp_host_alloc(void **ptr, uint64_t bytes)
{
    if (!ptr) return set_err(O3DPP_E_ARG, "null ptr");
    *ptr = nullptr;
    CU_TRY(cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocDefault));
    return 0;
}

O3DPP_API int o3dpp_host_free(void *ptr)
{
    if (ptr) CU_TRY(cudaFreeHost(ptr));
    return 0;
}

O3DPP_API int o3dpp_host_register(void *ptr, uint64_t bytes)
{
    if (!ptr || !bytes) return set_err(O3DPP_E_ARG, "null range");
    CU_TRY(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return 0;
}

O3DPP_API int o3dpp_host_unregister(void *ptr)
{
    if (ptr) CU_TRY(cudaHostUnregister(ptr));
    return 0;
}

O3DPP_API int o3dpp_set_option(o3dpp_ctx *c, int option, double value)
{
    if (!c) return set_err(O3DPP_E_ARG, "null context");
    switch (option) {
    case O3DPP_OPT_KEEP_NEIGHBORS: c->tune.keep = value != 0.0; if (!c->tune.keep) c->nbr_valid = false; break;
    case O3DPP_OPT_STATS: c->want_stats = value != 0.0; break;
    case O3DPP_OPT_KNN_RHO: c->tune.rho = (float)value; break;
    case O3DPP_OPT_KNN_GSPLIT: c->tune.gsplit = (int)value; break;
    case O3DPP_OPT_KNN_MAXSHRINK: c->tune.max_shrink = (int)value; break;
    case O3DPP_OPT_KNN_SHRINK: c->tune.shrink = (float)value; break;
    case O3DPP_OPT_KNN_BITS: c->tune.bits = (int)value; break;
    case O3DPP_OPT_KNN_FORCE_LIST: c->tune.force_list = value != 0.0; break;
    case O3DPP_OPT_QUERY_FIRST: c->tune.query_first = value > 0.0 ? (int64_t)value : 0; break;
    default: return set_err(O3DPP_E_ARG, "unknown option");
    }
    return 0;
}

O3DPP_API int o3dpp_launch_count(o3dpp_ctx *c, uint64_t *count)
{
    if (!c || !count) return set_err(O3DPP_E_ARG, "null argument");
    *count = c->launches;
    return 0;
}

O3DPP_API int o3dpp_last_timings(o3dpp_ctx *c, double *out8)
{
    if (!c || !out8) return set_err(O3DPP_E_ARG, "null argument");
    memcpy(out8, c->timings, sizeof c->timings);
    return 0;
}

O3DPP_API int o3dpp_debug_counters(o3dpp_ctx *c, uint64_t *out8)
{
    if (!c || !out8) return set_err(O3DPP_E_ARG, "null argument");
    for (int i = 0; i < 8; ++i) out8[i] = c->stats[i];
    return 0;
}

O3DPP_API int o3dpp_upload(o3dpp_ctx *c, int attr, const void *host, int64_t n, int dtype)
{
    CTX_CHECK(c);
    O3_TRY(upload_impl(c, attr, host, false, n, dtype));
    // the staging buffer is reused by the next call: finish the copy before returning
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

O3DPP_API int o3dpp_upload_dev(o3dpp_ctx *c, int attr, uintptr_t dev, int64_t n, int dtype)
{
    CTX_CHECK(c);
    O3_TRY(upload_impl(c, attr, (const void *)dev, true, n, dtype));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

O3DPP_API int o3dpp_download(o3dpp_ctx *c, int attr, double *host)
{
    CTX_CHECK(c);
    return download_impl(c, attr, host, false, O3DPP_F64);
}

O3DPP_API int o3dpp_download_as(o3dpp_ctx *c, int attr, void *host, int dtype)
{
    CTX_CHECK(c);
    if (dtype != O3DPP_F32 && dtype != O3DPP_F64) return set_err(O3DPP_E_ARG, "dtype must be O3DPP_F32 or O3DPP_F64");
    return download_impl(c, attr, host, false, dtype);
}

O3DPP_API int o3dpp_download_rows6(o3dpp_ctx *c, void *host, int dtype)
{
    CTX_CHECK(c);
    return download_rows6_impl(c, host, dtype);
}

O3DPP_API int o3dpp_download_async(o3dpp_ctx *c, int attr, double *host)
{
    CTX_CHECK(c);
    return download_async_impl(c, attr, host);
}

O3DPP_API int o3dpp_download_dev(o3dpp_ctx *c, int attr, uintptr_t dev, int dtype)
{
    CTX_CHECK(c);
    if (dtype != O3DPP_F32 && dtype != O3DPP_F64) return set_err(O3DPP_E_ARG, "dtype must be O3DPP_F32 or O3DPP_F64");
    O3_TRY(download_impl(c, attr, (void *)dev, true, dtype));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

O3DPP_API int o3dpp_size(o3dpp_ctx *c, int64_t *n)
{
    if (!c || !n) return set_err(O3DPP_E_ARG, "null argument");
    *n = c->n;
    return 0;
}

O3DPP_API int o3dpp_has(o3dpp_ctx *c, int attr, int *flag)
{
    if (!c || !flag) return set_err(O3DPP_E_ARG, "null argument");
    Attr3 *a = attr_of(c, attr);
    if (!a) return set_err(O3DPP_E_ARG, "unknown attribute");
    *flag = a->present ? 1 : 0;
    return 0;
}

O3DPP_API int o3dpp_drop(o3dpp_ctx *c, int attr)
{
    if (!c) return set_err(O3DPP_E_ARG, "null context");
    if (attr == O3DPP_ATTR_NORMALS) c->nrm.present = false;
    else if (attr == O3DPP_ATTR_COLORS) c->col.present = false;
    else if (attr == O3DPP_ATTR_POINTS) { c->n = 0; c->pts.present = false; c->nrm.present = false; c->col.present = false; invalidate_derived(c); }
    else return set_err(O3DPP_E_ARG, "unknown attribute");
    return 0;
}

O3DPP_API int o3dpp_bounds(o3dpp_ctx *c, double *min3, double *max3)
{
    CTX_CHECK(c);
    double mn[3], mx[3];
    O3_TRY(bounds_impl(c, mn, mx));
    if (min3) memcpy(min3, mn, sizeof mn);
    if (max3) memcpy(max3, mx, sizeof mx);
    return 0;
}

O3DPP_API int o3dpp_transform(o3dpp_ctx *c, const double *T16)
{
    CTX_CHECK(c);
    O3_TRY(transform_impl(c, T16));
    CU_TRY(cudaStreamSynchronize(c->stream));
    if (c->n > 0) {
        float ms = 0;
        cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]);
        c->timings[0] = ms;
    }
    return 0;
}

O3DPP_API int o3dpp_transform_points(o3dpp_ctx *c, const double *T16)
{
    CTX_CHECK(c);
    O3_TRY(transform_impl(c, T16, 1));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

O3DPP_API int o3dpp_voxel(o3dpp_ctx *c, double voxel_size, int64_t *V)
{
    CTX_CHECK(c);
    return voxel_impl(c, voxel_size, nullptr, nullptr, V);
}

O3DPP_API int o3dpp_voxel_bounded(o3dpp_ctx *c, double voxel_size, const double *min_bound3, const double *max_bound3, int64_t *V)
{
    CTX_CHECK(c);
    return voxel_impl(c, voxel_size, min_bound3, max_bound3, V);
}

O3DPP_API int o3dpp_voxel_fetch(o3dpp_ctx *c, int32_t *rep_asc, int32_t *rep, int32_t *count, double *centroid,
                                double *normal, double *color, int32_t *cubic_id, int32_t *order, int64_t *seg)
{
    CTX_CHECK(c);
    return voxel_fetch_impl(c, rep_asc, rep, count, centroid, normal, color, cubic_id, order, seg);
}

O3DPP_API int o3dpp_voxel_select(o3dpp_ctx *c)
{
    CTX_CHECK(c);
    return voxel_select_impl(c);
}

O3DPP_API int o3dpp_estimate_normals(o3dpp_ctx *c, int mode, int knn, double radius)
{
    CTX_CHECK(c);
    return search_impl(c, mode, knn, radius, true, nullptr, nullptr, nullptr);
}

O3DPP_API int o3dpp_knn(o3dpp_ctx *c, int mode, int knn, double radius, int32_t *idx, double *d2, int32_t *cnt)
{
    CTX_CHECK(c);
    if (!idx) return set_err(O3DPP_E_ARG, "idx is required");
    return search_impl(c, mode, knn, radius, false, idx, d2, cnt);
}

O3DPP_API int o3dpp_segment_plane(o3dpp_ctx *c, double distance_threshold, int ransac_n, int num_iterations,
                                  double probability, int sampler, uint64_t seed, const int64_t *samples,
                                  double *plane4, int64_t *n_inliers, double *best_plane4, int32_t *best_iter,
                                  int64_t *counts, double *errs)
{
    CTX_CHECK(c);
    return segment_plane_impl(c, distance_threshold, ransac_n, num_iterations, probability, sampler, seed, samples,
                              plane4, n_inliers, best_plane4, best_iter, counts, errs);
}

O3DPP_API int o3dpp_fetch_inliers(o3dpp_ctx *c, int32_t *out)
{
    CTX_CHECK(c);
    if (c->r_ninl < 0) return set_err(O3DPP_E_STATE, "o3dpp_segment_plane has not been run on the current cloud");
    if (!out && c->r_ninl > 0) return set_err(O3DPP_E_ARG, "null output buffer");
    O3_TRY(d2h(c, out, c->r_inl.p, (size_t)c->r_ninl * 4));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

O3DPP_API int o3dpp_fetch_samples(o3dpp_ctx *c, int64_t *out)
{
    if (!c || !out) return set_err(O3DPP_E_ARG, "null argument");
    if (c->r_samples_host.empty() && c->r_H > 0) return set_err(O3DPP_E_STATE, "no samples");
    memcpy(out, c->r_samples_host.data(), c->r_samples_host.size() * sizeof(int64_t));
    return 0;
}

O3DPP_API int o3dpp_select(o3dpp_ctx *c, const int32_t *idx, int64_t m, int invert)
{
    CTX_CHECK(c);
    if (m < 0 || (m > 0 && !idx)) return set_err(O3DPP_E_ARG, "bad index list");
    DevBuf d;
    O3_TRY(d.ensure((size_t)(m + 16) * 4));
    int rc = 0;
    if (m > 0 && cudaMemcpyAsync(d.p, idx, (size_t)m * 4, cudaMemcpyHostToDevice, c->stream) != cudaSuccess)
        rc = set_err(O3DPP_E_CUDA, "copy failed");
    if (!rc) rc = select_impl(c, d.as<int32_t>(), m, invert);
    cudaStreamSynchronize(c->stream);
    d.release();
    return rc;
}

O3DPP_API int o3dpp_select_inliers(o3dpp_ctx *c, int invert)
{
    CTX_CHECK(c);
    if (c->r_ninl < 0) return set_err(O3DPP_E_STATE, "o3dpp_segment_plane has not been run on the current cloud");
    DevBuf d;
    int64_t m = c->r_ninl;
    O3_TRY(d.ensure((size_t)(m + 16) * 4));
    int rc = 0;
    if (m > 0 && cudaMemcpyAsync(d.p, c->r_inl.p, (size_t)m * 4, cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess)
        rc = set_err(O3DPP_E_CUDA, "copy failed");
    if (!rc) rc = select_impl(c, d.as<int32_t>(), m, invert);
    cudaStreamSynchronize(c->stream);
    d.release();
    return rc;
}

}  // extern "C"

/* ---- slab-sharded scans: the pieces of segment_plane and the k-th distance bounds ------------------- */
O3DPP_API int o3dpp_sample_philox(o3dpp_ctx *c, uint64_t seed, int64_t n, int ransac_n, int num_iterations, int64_t *out)
{
    CTX_CHECK(c);
    if (!out) return set_err(O3DPP_E_ARG, "null argument");
    return sample_philox_impl(c, seed, n, ransac_n, num_iterations, out);
}
O3DPP_API int o3dpp_planes_from_points(o3dpp_ctx *c, const double *pts, int ransac_n, int num_iterations, double *planes)
{
    CTX_CHECK(c);
    if (!pts || !planes) return set_err(O3DPP_E_ARG, "null argument");
    return planes_from_points_impl(c, pts, ransac_n, num_iterations, planes);
}
O3DPP_API int o3dpp_score_planes(o3dpp_ctx *c, const double *planes, int num_planes, double distance_threshold, int64_t n_use,
                                 int64_t *counts, double *errs)
{
    CTX_CHECK(c);
    if (!planes || !counts || !errs) return set_err(O3DPP_E_ARG, "null argument");
    return score_planes_impl(c, planes, num_planes, distance_threshold, n_use, counts, errs);
}
O3DPP_API int o3dpp_plane_inliers(o3dpp_ctx *c, const double *plane4, double distance_threshold, int64_t n_use, int64_t *n_inliers,
                                  double *sum3)
{
    CTX_CHECK(c);
    if (!plane4) return set_err(O3DPP_E_ARG, "null argument");
    return plane_inliers_impl(c, plane4, distance_threshold, n_use, n_inliers, sum3);
}
O3DPP_API int o3dpp_inlier_moments(o3dpp_ctx *c, const double *centroid3, double *mom6)
{
    CTX_CHECK(c);
    if (!centroid3 || !mom6) return set_err(O3DPP_E_ARG, "null argument");
    return inlier_moments_impl(c, centroid3, mom6);
}
O3DPP_API int o3dpp_fetch_neighbors(o3dpp_ctx *c, const int64_t *queries, int64_t m, int32_t *idx, int32_t *cnt, int8_t *tier)
{
    CTX_CHECK(c);
    return fetch_neighbors_impl(c, queries, m, idx, cnt, tier);
}
O3DPP_API int o3dpp_fetch_kth_bound(o3dpp_ctx *c, float *out)
{
    CTX_CHECK(c);
    if (!out) return set_err(O3DPP_E_ARG, "null argument");
    return fetch_kth_impl(c, out);
}
O3DPP_API int o3dpp_fetch_kth_bound_dev(o3dpp_ctx *c, uintptr_t dev_out)
{
    CTX_CHECK(c);
    if (!dev_out) return set_err(O3DPP_E_ARG, "null argument");
    return fetch_kth_impl(c, nullptr, (float *)dev_out);
}

/* ---- callers either side of the path ------------------------------------------------ */
O3DPP_API int o3dpp_knn_mean_distance(o3dpp_ctx *c, int knn, double *out)
{
    CTX_CHECK(c);
    if (!out) return set_err(O3DPP_E_ARG, "null argument");
    return knn_mean_distance_impl(c, knn, out);
}
O3DPP_API int o3dpp_fp64_peak(o3dpp_ctx *c, double *tflops)
{
    CTX_CHECK(c);
    if (!tflops) return set_err(O3DPP_E_ARG, "null argument");
    return fp64_peak_impl(c, tflops);
}
O3DPP_API int o3dpp_zsplat(o3dpp_ctx *c, int grid_size, double x_min, double x_max, double y_min, double y_max, float *z_grid,
                           int32_t *last_index)
{
    CTX_CHECK(c);
    if (!z_grid) return set_err(O3DPP_E_ARG, "null argument");
    return zsplat_impl(c, grid_size, x_min, x_max, y_min, y_max, z_grid, last_index);
}

/* ---- o3dpp_bounds_static: min/max bound with NO allocation ------------------------------------------------------------
 * No counterpart in the reference; added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops).  Same result as o3dpp_bounds, but context-free: input and output are the
 * caller's device buffers, the per-block partial bounds live in a file-scope __device__ array of fixed size that every
 * GPU loading the library holds until exit.  Fixed tiles of SB_TILE rows per block (no grid-stride loop), so the scratch
 * a call needs is ceil(n / SB_TILE) x 6 doubles and n is bounded by O3DPP_BOUNDS_STATIC_MAX_POINTS.
 * Nothing in the scratch has to start at zero: the final kernel reads exactly the slots the partial kernel of the same
 * call has written.  Every call on a device shares the array: two calls that overlap on the GPU (different streams) race. */
#define SB_THREADS 256
#define SB_ROWS 16                                   /* rows per thread */
#define SB_TILE (SB_THREADS * SB_ROWS)               /* 4096 rows = 96 KiB of input per block */
#define SB_MAX_BLOCKS (int)(O3DPP_BOUNDS_STATIC_MAX_POINTS / SB_TILE)
static_assert(O3DPP_BOUNDS_STATIC_MAX_POINTS % SB_TILE == 0, "whole tiles");
#define SB_MAX_DEVICES 64

static __device__ __align__(256) double g_static_bounds_part[(size_t)SB_MAX_BLOCKS * 6];      /* 768 KiB */

// rows are interleaved (x y z x y z ...): the tile is read as 3 * SB_TILE consecutive doubles, thread t takes elements
// t, t + 256, ...; 256 = 1 (mod 3), so element j of a thread belongs to coordinate (t + j) mod 3: three accumulator
// pairs indexed by j mod 3 at compile time, rotated to (x, y, z) once at the end
__global__ void __launch_bounds__(SB_THREADS)
bounds_static_partial_kernel(const double *__restrict__ rows, int64_t n, double *__restrict__ part)
{
    const int64_t e0 = (int64_t)blockIdx.x * (3 * SB_TILE), e1 = 3 * n;     // flat element range of this tile
    const double *p = rows + e0;
    int64_t left = e1 - e0;                                                 // > 0: the grid is ceil(n / SB_TILE)
    int t = threadIdx.x, t3 = t % 3;
    double lo[3], hi[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {                                           // start from a real element of the same coordinate
        int64_t e = t + SB_THREADS * r;
        if (e >= left) e = (t3 + r) % 3;                                    // element (t3 + r) % 3 of row 0 of the tile
        lo[r] = hi[r] = p[e];
    }
#pragma unroll
    for (int j = 0; j < 3 * SB_ROWS; ++j) {
        int64_t e = t + SB_THREADS * j;
        if (e < left) {
            double v = p[e];
            if (v < lo[j % 3]) lo[j % 3] = v;
            if (v > hi[j % 3]) hi[j % 3] = v;
        }
    }
    double mn[3], mx[3];                                                    // accumulator r holds coordinate (t3 + r) % 3
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        int r = (k - t3 + 3) % 3;
        mn[k] = r == 0 ? lo[0] : r == 1 ? lo[1] : lo[2];
        mx[k] = r == 0 ? hi[0] : r == 1 ? hi[1] : hi[2];
    }
    __shared__ double sh[SB_THREADS / 32][6];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            double o = shfl_xor_d(mn[k], d);
            if (o < mn[k]) mn[k] = o;
            o = shfl_xor_d(mx[k], d);
            if (o > mx[k]) mx[k] = o;
        }
    }
    if (lane_id() == 0) {
        for (int k = 0; k < 3; ++k) { sh[t >> 5][k] = mn[k]; sh[t >> 5][3 + k] = mx[k]; }
    }
    __syncthreads();
    if (t < 6) {
        double v = sh[0][t];
        for (int w = 1; w < SB_THREADS / 32; ++w) {
            double o = sh[w][t];
            if (t < 3 ? o < v : o > v) v = o;
        }
        part[(int64_t)blockIdx.x * 6 + t] = v;
    }
}

// one block of 6 warps: warp k reduces column k of the `blocks` partial rows
__global__ void __launch_bounds__(192)
bounds_static_final_kernel(const double *__restrict__ part, int blocks, double *__restrict__ out6)
{
    int k = threadIdx.x >> 5, lane = lane_id();
    double v = part[k];
    for (int b = lane; b < blocks; b += 32) {
        double o = part[(int64_t)b * 6 + k];
        if (k < 3 ? o < v : o > v) v = o;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        double o = shfl_xor_d(v, d);
        if (k < 3 ? o < v : o > v) v = o;
    }
    if (lane == 0) out6[k] = v;
}

// the array's name is a host-side stand-in in host code: the device address is resolved once per device, with that
// device current, and handed to the kernels as an ordinary pointer
static double *g_static_bounds_addr[SB_MAX_DEVICES];
static std::mutex g_static_bounds_mutex;

static int static_bounds_scratch(int device, double **out)
{
    std::lock_guard<std::mutex> lk(g_static_bounds_mutex);
    if (!g_static_bounds_addr[device]) {
        double *ptr = nullptr;
        CU_TRY(cudaGetSymbolAddress((void **)&ptr, g_static_bounds_part));
        if (!ptr) return set_err(O3DPP_E_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_static_bounds_addr[device] = ptr;
    }
    *out = g_static_bounds_addr[device];
    return 0;
}

O3DPP_API int o3dpp_bounds_static(int device, uintptr_t stream, uintptr_t dev_rows, int64_t n, uintptr_t dev_out6)
{
    if (device < 0 || device >= SB_MAX_DEVICES) return set_err(O3DPP_E_ARG, "o3dpp_bounds_static: device index out of range");
    if (n < 0) return set_err(O3DPP_E_ARG, "o3dpp_bounds_static: negative point count");
    if (n > O3DPP_BOUNDS_STATIC_MAX_POINTS) {
        char b[256];
        snprintf(b, sizeof b, "o3dpp_bounds_static: %lld points exceed the static scratch (O3DPP_BOUNDS_STATIC_MAX_POINTS = %lld); use o3dpp_bounds",
                 (long long)n, (long long)O3DPP_BOUNDS_STATIC_MAX_POINTS);
        return set_err(O3DPP_E_ARG, b);
    }
    if (!dev_out6 || (n > 0 && !dev_rows)) return set_err(O3DPP_E_ARG, "o3dpp_bounds_static: null device pointer");
    CU_TRY(cudaSetDevice(device));
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) {                                     // o3dpp_bounds answers zeros for an empty cloud
        CU_TRY(cudaMemsetAsync((void *)dev_out6, 0, 6 * sizeof(double), s));
        return 0;
    }
    double *part = nullptr;
    O3_TRY(static_bounds_scratch(device, &part));
    int blocks = (int)((n + SB_TILE - 1) / SB_TILE);
    bounds_static_partial_kernel<<<blocks, SB_THREADS, 0, s>>>((const double *)dev_rows, n, part);
    bounds_static_final_kernel<<<1, 192, 0, s>>>(part, blocks, (double *)dev_out6);
    return check_launch("bounds_static");
}
