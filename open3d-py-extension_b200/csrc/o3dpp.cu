// o3dpp.cu — the C ABI declared in include/o3dpp.h.  Single translation unit:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false
//        -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=hidden -shared -o libo3dpp_b200.so o3dpp.cu
// There is no CPU path in this library: every entry point needs a live CUDA context.
#include "extras.cuh"
#include <stdlib.h>

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
                      &c->g_nd2, &c->g_ncnt, &c->g_sp4, &c->g_boxf, &c->g_nbrpos, &c->g_tier, &c->g_inv, &c->g_glist, &c->g_gcnt, &c->g_p4, &c->r_samples, &c->r_planes, &c->r_cnt, &c->r_err, &c->r_inl, &c->r_part};
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
