/*
 * o3dpp.h — C ABI of libo3dpp_b200.so, the B200 (sm_100a) implementation of the
 * open3dpypro point-cloud hot path.
 *
 * The reference (qinhy/Open3D-py-extension) has no FFI of its own: its hot path is
 * four one-line delegations from open3dpypro/PointCloud.py into the open3d==0.19.0
 * wheel.  Each entry point below names the reference call it replaces; the
 * reference-side binding a maintainer would add is a ctypes stub.
 *
 * Conventions
 *   - every function returns 0 on success, a negative O3DPP_E_* code on failure;
 *     o3dpp_last_error() returns the message of the last failure on this thread
 *   - no exceptions and no torch/pybind types cross this boundary: plain pointers
 *     and sizes only; the caller owns every buffer it passes
 *   - host arrays are C-contiguous (n,3) row-major, float64 unless a dtype says
 *     otherwise; *_dev variants take a CUDA device address (uintptr_t, e.g.
 *     torch.Tensor.data_ptr()) instead of a host pointer
 *   - one o3dpp_ctx belongs to one process and one GPU; it owns one CUDA stream
 *     and all device memory of one cloud.  Calls on one ctx must not overlap.
 *   - there is no CPU fallback: with no usable CUDA device o3dpp_create fails
 */
#ifndef O3DPP_H
#define O3DPP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define O3DPP_API __attribute__((visibility("default")))
#else
#define O3DPP_API
#endif

typedef struct o3dpp_ctx o3dpp_ctx;

enum {
    O3DPP_OK = 0,
    O3DPP_E_CUDA = -1,        /* CUDA runtime error (message has the detail)            */
    O3DPP_E_ARG = -2,         /* bad argument                                           */
    O3DPP_E_VOXEL_SIZE = -3,  /* Open3D: "voxel_size <= 0."                             */
    O3DPP_E_VOXEL_SMALL = -4, /* Open3D: "voxel_size is too small."                     */
    O3DPP_E_RANSAC_N = -5,    /* Open3D: "ransac_n should be set to higher than or equal to 3." */
    O3DPP_E_FEW_POINTS = -6,  /* Open3D: "There must be at least 'ransac_n' points."    */
    O3DPP_E_PROB = -7,        /* Open3D: "Probability must be > 0 and <= 1.0"            */
    O3DPP_E_STATE = -8,       /* call order (e.g. fetch before compute)                 */
    O3DPP_E_UNSUPPORTED = -9  /* e.g. knn > O3DPP_MAX_KNN                               */
};

enum { O3DPP_F32 = 0, O3DPP_F64 = 1 };
enum { O3DPP_SEARCH_KNN = 0, O3DPP_SEARCH_RADIUS = 1, O3DPP_SEARCH_HYBRID = 2 };
enum { O3DPP_SAMPLER_PHILOX = 0, O3DPP_SAMPLER_MT19937 = 1, O3DPP_SAMPLER_GIVEN = 2 };
enum { O3DPP_ATTR_POINTS = 0, O3DPP_ATTR_NORMALS = 1, O3DPP_ATTR_COLORS = 2 };

/* o3dpp_set_option: KEEP_NEIGHBORS = every tier of o3dpp_estimate_normals records the neighbour set it selected
 * (o3dpp_fetch_neighbors); STATS = the search kernels count their work; KNN_* = tuning of the neighbour search
 * (0 = built-in default; the same knobs are read once per context from the O3DPP_KNN_* environment variables). */
enum {
    O3DPP_OPT_KEEP_NEIGHBORS = 1, O3DPP_OPT_STATS = 2, O3DPP_OPT_KNN_RHO = 3, O3DPP_OPT_KNN_GSPLIT = 4,
    O3DPP_OPT_KNN_MAXSHRINK = 5, O3DPP_OPT_KNN_SHRINK = 6, O3DPP_OPT_KNN_BITS = 7, O3DPP_OPT_KNN_FORCE_LIST = 8,
    /* value = m > 0: o3dpp_estimate_normals (knn <= 32, KNN or hybrid) answers only the FIRST m points of the cloud; every point stays
     * a neighbour candidate, the normals of the others are left as they are.  0 = all points.  The second pass of a slab-sharded
     * pure-KNN scan queries only the points whose k-th distance reaches across a cut (open3dpypro/sharding.py). */
    O3DPP_OPT_QUERY_FIRST = 9
};

#define O3DPP_MAX_KNN 128

/* ---- library / context ---------------------------------------------------------- */
O3DPP_API const char *o3dpp_last_error(void);
O3DPP_API const char *o3dpp_version(void);
O3DPP_API int o3dpp_device_count(int *count);
O3DPP_API int o3dpp_create(int device, o3dpp_ctx **out);
O3DPP_API int o3dpp_destroy(o3dpp_ctx *ctx);
O3DPP_API int o3dpp_synchronize(o3dpp_ctx *ctx);
/* CUDA stream of the context as an integer (cudaStream_t), for callers that time with
 * events or enqueue their own work (torch.cuda.ExternalStream). */
O3DPP_API int o3dpp_stream(o3dpp_ctx *ctx, uintptr_t *stream);
/* Pinned host memory for the staging side of open3dpypro's Vector3dVector copies
 * (PointCloud.py:94-107) and shmIO frames (shmIO.py:229-240). */
O3DPP_API int o3dpp_host_alloc(void **ptr, uint64_t bytes);
O3DPP_API int o3dpp_host_free(void *ptr);
/* Page-lock memory the caller already owns (a named shared-memory frame segment of shmIO.py:39-44, :281-296), so
 * that uploads from it are direct DMA instead of a staged pageable copy.  Unregister before unmapping it. */
O3DPP_API int o3dpp_host_register(void *ptr, uint64_t bytes);
O3DPP_API int o3dpp_host_unregister(void *ptr);
/* per-context options (O3DPP_OPT_*); the reference has no counterpart: Open3D's search parameters are fixed at build time */
O3DPP_API int o3dpp_set_option(o3dpp_ctx *ctx, int option, double value);
/* number of kernels this library has launched on the context since creation */
O3DPP_API int o3dpp_launch_count(o3dpp_ctx *ctx, uint64_t *count);
/* ms spent in the last call's stages, measured with CUDA events on the ctx stream.
 * out[0..7]: operation specific, see each call. */
O3DPP_API int o3dpp_last_timings(o3dpp_ctx *ctx, double *out8);
/* work counters of the last search when the process runs with O3DPP_STATS=1 (else zeros):
 * [0] range-table visits [1] cell groups loaded [2] list insertions [3] candidates offered */
O3DPP_API int o3dpp_debug_counters(o3dpp_ctx *ctx, uint64_t *out8);

/* ---- cloud state: replaces o3d.utility.Vector3dVector (PointCloud.py:94-107) ----- */
/* attr = O3DPP_ATTR_*; uploading POINTS with a different n drops normals and colours */
O3DPP_API int o3dpp_upload(o3dpp_ctx *ctx, int attr, const void *host, int64_t n, int dtype);
O3DPP_API int o3dpp_upload_dev(o3dpp_ctx *ctx, int attr, uintptr_t dev, int64_t n, int dtype);
/* replaces np.asarray(pcd.points / normals / colors) (PointCloud.py:119-126): float64 (n,3) */
O3DPP_API int o3dpp_download(o3dpp_ctx *ctx, int attr, double *host);
O3DPP_API int o3dpp_download_dev(o3dpp_ctx *ctx, int attr, uintptr_t dev, int dtype);
/* o3dpp_download with the element type of the destination: float32 block mats (PointCloudMat.py:50-51) take
 * (n,3) float32 straight from the device (PlaneNormalize, processors.py:746-759, returns the mat's dtype) */
O3DPP_API int o3dpp_download_as(o3dpp_ctx *ctx, int attr, void *host, int dtype);
/* (n,6) rows [x y z nx ny nz]: np.hstack([pcd[:, :3], normals]) of the CPUNormals block (processors.py:243-249)
 * assembled on the device and copied once */
O3DPP_API int o3dpp_download_rows6(o3dpp_ctx *ctx, void *host, int dtype);
/* as o3dpp_download, but returns at once: the copy runs on a second stream after the work already queued and overlaps
 * the calls that follow (e.g. the normals travel to the host while o3dpp_segment_plane runs; np.asarray(pcd.normals)
 * of PointCloud.py:125 is then free).  `host` (page-locked for a real overlap) is complete, and the attribute may be
 * modified again, after o3dpp_synchronize(). */
O3DPP_API int o3dpp_download_async(o3dpp_ctx *ctx, int attr, double *host);
O3DPP_API int o3dpp_size(o3dpp_ctx *ctx, int64_t *n);
O3DPP_API int o3dpp_has(o3dpp_ctx *ctx, int attr, int *flag);
O3DPP_API int o3dpp_drop(o3dpp_ctx *ctx, int attr);
/* pcd.get_min_bound() / get_max_bound() (PointCloud.py:340): 3 + 3 float64 */
O3DPP_API int o3dpp_bounds(o3dpp_ctx *ctx, double *min3, double *max3);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same min / max bound as o3dpp_bounds
 * (bit-equal: min and max do not depend on the reduction order), but it allocates nothing and needs no context:
 *   dev_rows  n x 3 float64, row-major, DEVICE memory on `device` (the layout o3dpp_upload_dev takes);
 *   dev_out6  6 float64 of DEVICE memory: min x y z, max x y z (zeros for n = 0, as o3dpp_bounds answers);
 *   stream    a cudaStream_t of `device` (0 = the default stream); the call only enqueues, it does not synchronise.
 * The per-block partial bounds live in a file-scope __device__ array of 768 KiB that every GPU loading the library holds
 * until exit; n > O3DPP_BOUNDS_STATIC_MAX_POINTS is refused with O3DPP_E_ARG (use o3dpp_bounds).  No cudaMalloc, cudaFree
 * or allocating library routine is called.  All calls on one device share the array, so calls that overlap on the GPU
 * (two streams, two threads) race: keep them on one stream or order them with events.  The first call per device resolves
 * the array's address (cudaGetSymbolAddress): make it once before capturing a graph. */
#define O3DPP_BOUNDS_STATIC_MAX_POINTS 67108864LL    /* 2^26 points = 16 384 tiles of 4 096 rows x 48 B of scratch */
O3DPP_API int o3dpp_bounds_static(int device, uintptr_t stream, uintptr_t dev_rows, int64_t n, uintptr_t dev_out6);

/* ---- transform: replaces pcd.transform(T) (PointCloud.py:60-62) ------------------ */
/* T = 16 float64 row-major.  points: (T[p,1]).xyz / w ; normals (if present): (T[n,0]).xyz */
O3DPP_API int o3dpp_transform(o3dpp_ctx *ctx, const double *T16);
/* points only, normals left as they are: PlaneNormalize.rotate_to_plane (processors.py:725-744) transforms pcd[:, :3]
 * and passes every other column of the mat through unchanged, normal columns included */
O3DPP_API int o3dpp_transform_points(o3dpp_ctx *ctx, const double *T16);

/* ---- voxel_down_sample_and_trace: replaces pcd.voxel_down_sample_and_trace(
 *      voxel_size, min_bound, max_bound, approximate_class=False) (PointCloud.py:338-341)
 * Rows are in ascending (ix,iy,iz).  After o3dpp_voxel() the results stay on the device
 * until the next o3dpp_voxel()/upload; fetch what is needed (NULL skips a field):
 *   rep_asc   V  int32  idxmat.max(1) sorted ascending = PointCloud.py:341's mask order
 *   rep       V  int32  largest original index per voxel, row order
 *   count     V  int32  points per voxel
 *   centroid  V*3 f64   mean position; normal V*3 f64 normalised normal sum (if normals);
 *   color     V*3 f64   mean colour (if colours)
 *   cubic_id  V*8 int32 largest original index per octant, -1 if empty
 *   order     N  int32  original indices grouped by voxel row, ascending inside a row
 *   seg       V+1 int64 row offsets into `order`
 */
O3DPP_API int o3dpp_voxel(o3dpp_ctx *ctx, double voxel_size, int64_t *V);
/* the same with the explicit min_bound / max_bound of the call it replaces (PointCloud.py:339-340): min_bound3 is the
 * grid origin (Open3D's trace variant does not shift it), max_bound3 only enters the "voxel_size is too small" check;
 * NULL = the cloud's own bound (o3dpp_voxel).  Points may lie below min_bound3 (negative voxel indices, rows still
 * ascending).  A scan sharded by spatial slab passes the GLOBAL bounds on every rank, so every rank keys its voxels
 * exactly as the whole scan would. */
O3DPP_API int o3dpp_voxel_bounded(o3dpp_ctx *ctx, double voxel_size, const double *min_bound3, const double *max_bound3,
                                  int64_t *V);
O3DPP_API int o3dpp_voxel_fetch(o3dpp_ctx *ctx, int32_t *rep_asc, int32_t *rep, int32_t *count,
                                double *centroid, double *normal, double *color,
                                int32_t *cubic_id, int32_t *order, int64_t *seg);
/* Replace the context's cloud by the selected original points (PointCloud.py:185-204 with
 * idxmat.max(1)): points, normals, colours gathered in ascending original index. */
O3DPP_API int o3dpp_voxel_select(o3dpp_ctx *ctx);

/* ---- estimate_normals: replaces pcd.estimate_normals(param) (PointCloud.py:68-73),
 *      fast_normal_computation = True.  mode = O3DPP_SEARCH_*; knn <= O3DPP_MAX_KNN.
 * Normals are written into the context (ATTR_NORMALS); if the cloud already had normals
 * the new ones are flipped to agree with them, as Open3D does.
 * timings: [0] grid build (keys+sort+gather+cell boxes) [1] search+covariance+eigen; knn <= 32, KNN or hybrid (the
 * tier chain): [2] tier-1 selection kernel [7] finish kernel (cumulants + eigenvector) [3] compaction + tier 2
 * [6] compaction + tier 3, [4] / [5] = queries handed to tiers 2 / 3 */
O3DPP_API int o3dpp_estimate_normals(o3dpp_ctx *ctx, int mode, int knn, double radius);
/* The neighbour sets behind the last o3dpp_estimate_normals (knn <= 32, KNN or hybrid; O3DPP_OPT_KEEP_NEIGHBORS set before
 * it): what KDTreeFlann::Search* returned inside pcd.estimate_normals (PointCloud.py:72), as SETS.  queries = m original
 * point indices (NULL: all n points in order, m ignored).  idx: m*knn int32, each row the ORIGINAL indices of the selected
 * neighbours in ascending order, -1 padded (hybrid: only those with d2 < radius^2); cnt (nullable): m int32; tier (nullable):
 * m int8, which kernel of the chain answered the query (1 certified FP32 selection, 2 FP64 range query, 3 exact list search). */
O3DPP_API int o3dpp_fetch_neighbors(o3dpp_ctx *ctx, const int64_t *queries, int64_t m, int32_t *idx, int32_t *cnt, int8_t *tier);
/* KDTreeFlann search of every point against the cloud (PointCloud.py:148-163 batched):
 * idx (n*knn int32, -1 padded), d2 (n*knn f64, -1 padded, nullable), cnt (n int32, nullable);
 * rows sorted by (d2, index). Host buffers. */
O3DPP_API int o3dpp_knn(o3dpp_ctx *ctx, int mode, int knn, double radius,
                        int32_t *idx, double *d2, int32_t *cnt);

/* ---- segment_plane: replaces pcd.segment_plane(distance_threshold, ransac_n,
 *      num_iterations, probability) (PointCloud.py:75-77)
 * sampler PHILOX: counter-based Philox4x32-10 keyed by `seed`, generated on the device;
 *         MT19937: std::mt19937(seed) replay of Open3D's RandomSampler;
 *         GIVEN:   `samples` = num_iterations*ransac_n int64 host indices.
 * plane4: refit plane of the final inliers.  n_inliers: size of the inlier list, fetched
 * with o3dpp_fetch_inliers (ascending int32).  best_plane4 / best_iter (nullable): the
 * winning pre-refit hypothesis.  counts/errs (nullable, num_iterations each): per
 * hypothesis inlier count (-1 skipped by early exit, -2 degenerate) and distance sum.
 * timings: [0] sampling+models [1] scoring [2] final pass+refit ; out8[7] = hypotheses scored */
O3DPP_API int o3dpp_segment_plane(o3dpp_ctx *ctx, double distance_threshold, int ransac_n,
                                  int num_iterations, double probability, int sampler,
                                  uint64_t seed, const int64_t *samples, double *plane4,
                                  int64_t *n_inliers, double *best_plane4, int32_t *best_iter,
                                  int64_t *counts, double *errs);
O3DPP_API int o3dpp_fetch_inliers(o3dpp_ctx *ctx, int32_t *out);
/* the hypothesis sequence the last o3dpp_segment_plane used (H*ransac_n int64) */
O3DPP_API int o3dpp_fetch_samples(o3dpp_ctx *ctx, int64_t *out);

/* ---- selection: replaces PointCloudSelections._select_by_idx (PointCloud.py:185-204) */
/* keep (invert=0) or drop (invert=1) the given original indices, ascending-index order */
O3DPP_API int o3dpp_select(o3dpp_ctx *ctx, const int32_t *idx, int64_t m, int invert);
/* same with the last segment_plane's inlier list, without a host round trip */
O3DPP_API int o3dpp_select_inliers(o3dpp_ctx *ctx, int invert);

/* ---- a scan sharded across GPUs by spatial slab (the reference has no counterpart: it runs
 *      pcd.segment_plane / pcd.estimate_normals, PointCloud.py:68-77, on one host).  These are the steps of
 *      o3dpp_segment_plane as separate calls, so that the host side can all-reduce between them:
 *      sample -> (owners publish the sampled points) -> planes -> score on every rank -> all-reduce ->
 *      host replay of the selection -> inliers + coordinate sums per rank -> all-reduce -> moments -> all-reduce.
 *      n_use > 0 restricts a call to the first n_use points of the cloud (owned points first, halo after). */
O3DPP_API int o3dpp_sample_philox(o3dpp_ctx *ctx, uint64_t seed, int64_t n, int ransac_n, int num_iterations,
                                  int64_t *out);
O3DPP_API int o3dpp_planes_from_points(o3dpp_ctx *ctx, const double *pts, int ransac_n, int num_iterations,
                                       double *planes);
O3DPP_API int o3dpp_score_planes(o3dpp_ctx *ctx, const double *planes, int num_planes, double distance_threshold,
                                 int64_t n_use, int64_t *counts, double *errs);
O3DPP_API int o3dpp_plane_inliers(o3dpp_ctx *ctx, const double *plane4, double distance_threshold, int64_t n_use,
                                  int64_t *n_inliers, double *sum3);
O3DPP_API int o3dpp_inlier_moments(o3dpp_ctx *ctx, const double *centroid3, double *mom6);
/* per point, an upper bound of the squared distance to its k-th neighbour left by the last
 * o3dpp_estimate_normals (knn <= 32, KNN or hybrid): tells which queries near a slab cut need a wider halo */
O3DPP_API int o3dpp_fetch_kth_bound(o3dpp_ctx *ctx, float *out);
/* the same bounds written to n float32 of DEVICE memory (the sharded scan keeps its bookkeeping on the device: a 100 M-point
 * slab would otherwise move 0.4 GB to the host and back) */
O3DPP_API int o3dpp_fetch_kth_bound_dev(o3dpp_ctx *ctx, uintptr_t dev_out);

/* measured FP64 FMA throughput of the device in TFLOP/s (2 flop per FMA): the roofline the scoring pass of
 * o3dpp_segment_plane (replaces the evaluation loop behind PointCloud.py:75-77) is reported against by bench.py */
O3DPP_API int o3dpp_fp64_peak(o3dpp_ctx *ctx, double *tflops);

/* ---- callers either side of the path */
/* remove_statistical_outlier (PointCloud.py:370-372 -> Open3D RemoveStatisticalOutliers): out[i] = mean distance of
 * point i to its `knn` nearest points (itself included; summed in (d2, index) order), -1 if none */
O3DPP_API int o3dpp_knn_mean_distance(o3dpp_ctx *ctx, int knn, double *out);
/* ZDepthViewer splat (processors.py:859-915): z_grid[yi*grid+xi] = z of the LAST point (largest index) falling into the
 * cell, NaN where empty; xi = clip(int32((x - x_min) / (x_max - x_min) * (grid - 1))) in float32 as numpy computes it.
 * last_index (nullable): that point's index per cell, -1 where empty. */
O3DPP_API int o3dpp_zsplat(o3dpp_ctx *ctx, int grid_size, double x_min, double x_max, double y_min, double y_max,
                           float *z_grid, int32_t *last_index);

#ifdef __cplusplus
}
#endif
#endif /* O3DPP_H */
