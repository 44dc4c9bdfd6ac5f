// cloud.cuh — device-resident cloud state: AoS<->SoA staging, bounds, 4x4 transform, selection.
// Replaces o3d.utility.Vector3dVector / np.asarray(pcd.points) (reference PointCloud.py:94-126),
// pcd.get_min_bound/get_max_bound (:340), pcd.transform (:60-62) and _select_by_idx (:185-204).
#pragma once
#include "primitives.cuh"

// ---- (n,3) row-major <-> three planes ---------------------------------------------------
// One block moves 256 points: the 768 interleaved values are read/written fully coalesced and
// transposed through shared memory.
template <class TIn>
__global__ void __launch_bounds__(256)
deinterleave_kernel(const TIn *__restrict__ aos, int64_t n, double *__restrict__ x,
                    double *__restrict__ y, double *__restrict__ z)
{
    __shared__ double sh[768];
    for (int64_t tile = blockIdx.x; tile * 256 < n; tile += gridDim.x) {
        int64_t p0 = tile * 256;
        int64_t cnt = n - p0 < 256 ? n - p0 : 256;
        const TIn *src = aos + p0 * 3;
        for (int i = threadIdx.x; i < cnt * 3; i += 256) sh[i] = (double)src[i];
        __syncthreads();
        if (threadIdx.x < cnt) {
            x[p0 + threadIdx.x] = sh[3 * threadIdx.x];
            y[p0 + threadIdx.x] = sh[3 * threadIdx.x + 1];
            z[p0 + threadIdx.x] = sh[3 * threadIdx.x + 2];
        }
        __syncthreads();
    }
}

template <class TOut>
__global__ void __launch_bounds__(256)
interleave_kernel(const double *__restrict__ x, const double *__restrict__ y,
                  const double *__restrict__ z, int64_t n, TOut *__restrict__ aos)
{
    __shared__ double sh[768];
    for (int64_t tile = blockIdx.x; tile * 256 < n; tile += gridDim.x) {
        int64_t p0 = tile * 256;
        int64_t cnt = n - p0 < 256 ? n - p0 : 256;
        if (threadIdx.x < cnt) {
            sh[3 * threadIdx.x] = x[p0 + threadIdx.x];
            sh[3 * threadIdx.x + 1] = y[p0 + threadIdx.x];
            sh[3 * threadIdx.x + 2] = z[p0 + threadIdx.x];
        }
        __syncthreads();
        TOut *dst = aos + p0 * 3;
        for (int i = threadIdx.x; i < cnt * 3; i += 256) dst[i] = (TOut)sh[i];
        __syncthreads();
    }
}

// rows [x y z nx ny nz] (np.hstack([pcd[:, :3], normals]) of the CPUNormals block, reference processors.py:243-249) in one
// pass: a block moves 128 points, 768 values transposed through shared memory, coalesced both sides
template <class TOut>
__global__ void __launch_bounds__(256)
interleave6_kernel(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                   const double *__restrict__ nx, const double *__restrict__ ny, const double *__restrict__ nz,
                   int64_t n, TOut *__restrict__ rows)
{
    __shared__ double sh[768];
    for (int64_t tile = blockIdx.x; tile * 128 < n; tile += gridDim.x) {
        int64_t p0 = tile * 128;
        int64_t cnt = n - p0 < 128 ? n - p0 : 128;
        const int t = threadIdx.x & 127;
        if (t < cnt) {
            if (threadIdx.x < 128) { sh[6 * t] = x[p0 + t]; sh[6 * t + 1] = y[p0 + t]; sh[6 * t + 2] = z[p0 + t]; }
            else { sh[6 * t + 3] = nx[p0 + t]; sh[6 * t + 4] = ny[p0 + t]; sh[6 * t + 5] = nz[p0 + t]; }
        }
        __syncthreads();
        TOut *dst = rows + p0 * 6;
        for (int i = threadIdx.x; i < cnt * 6; i += 256) dst[i] = (TOut)sh[i];
        __syncthreads();
    }
}

static Attr3 *attr_of(o3dpp_ctx *c, int attr)
{
    switch (attr) {
    case O3DPP_ATTR_POINTS: return &c->pts;
    case O3DPP_ATTR_NORMALS: return &c->nrm;
    case O3DPP_ATTR_COLORS: return &c->col;
    }
    return nullptr;
}

static int attr_alloc(Attr3 *a, int64_t n)
{
    int64_t st = plane_stride(n);
    O3_TRY(a->buf.ensure((size_t)st * 3 * sizeof(double)));
    a->stride = st;
    return 0;
}

static void invalidate_derived(o3dpp_ctx *c)
{
    c->vox_valid = false;
    c->kth_valid = false;
    c->nbr_valid = false;
    c->r_ninl = -1;
}

static int upload_impl(o3dpp_ctx *c, int attr, const void *src, bool src_is_device, int64_t n, int dtype)
{
    Attr3 *a = attr_of(c, attr);
    if (!a) return set_err(O3DPP_E_ARG, "unknown attribute");
    if (n < 0 || n > 2147483647LL - 64) return set_err(O3DPP_E_ARG, "n out of range (int32 indices)");
    if (dtype != O3DPP_F32 && dtype != O3DPP_F64) return set_err(O3DPP_E_ARG, "dtype must be O3DPP_F32 or O3DPP_F64");
    if (n > 0 && !src) return set_err(O3DPP_E_ARG, "null source");
    if (attr == O3DPP_ATTR_POINTS) {
        if (n != c->n) { c->nrm.present = false; c->col.present = false; }
        c->n = n;
        invalidate_derived(c);
    } else if (n != c->n) {
        return set_err(O3DPP_E_ARG, "attribute length differs from the number of points");
    }
    O3_TRY(attr_alloc(a, n));
    a->present = n > 0 || attr == O3DPP_ATTR_POINTS;
    if (n == 0) return 0;
    size_t esz = dtype == O3DPP_F32 ? 4 : 8;
    const void *dsrc = src;
    if (!src_is_device) {
        O3_TRY(c->stage.ensure((size_t)n * 3 * esz));
        CU_TRY(cudaMemcpyAsync(c->stage.p, src, (size_t)n * 3 * esz, cudaMemcpyHostToDevice, c->stream));
        dsrc = c->stage.p;
    }
    int grid = grid_for(c, (n + 255) / 256, 1, 16);
    if (dtype == O3DPP_F32)
        LAUNCH(c, deinterleave_kernel<float>, grid, 256, 0, (const float *)dsrc, n, a->x(), a->y(), a->z());
    else
        LAUNCH(c, deinterleave_kernel<double>, grid, 256, 0, (const double *)dsrc, n, a->x(), a->y(), a->z());
    return check_launch("deinterleave");
}

static int download_impl(o3dpp_ctx *c, int attr, void *dst, bool dst_is_device, int dtype)
{
    Attr3 *a = attr_of(c, attr);
    if (!a) return set_err(O3DPP_E_ARG, "unknown attribute");
    if (!a->present) return set_err(O3DPP_E_STATE, "attribute not present");
    int64_t n = c->n;
    if (n == 0) return 0;
    if (!dst) return set_err(O3DPP_E_ARG, "null destination");
    size_t esz = dtype == O3DPP_F32 ? 4 : 8;
    void *ddst = dst;
    if (!dst_is_device) {
        O3_TRY(c->stage.ensure((size_t)n * 3 * esz));
        ddst = c->stage.p;
    }
    int grid = grid_for(c, (n + 255) / 256, 1, 16);
    if (dtype == O3DPP_F32)
        LAUNCH(c, interleave_kernel<float>, grid, 256, 0, a->x(), a->y(), a->z(), n, (float *)ddst);
    else
        LAUNCH(c, interleave_kernel<double>, grid, 256, 0, a->x(), a->y(), a->z(), n, (double *)ddst);
    O3_TRY(check_launch("interleave"));
    if (!dst_is_device) {
        CU_TRY(cudaMemcpyAsync(dst, ddst, (size_t)n * 3 * esz, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

// (n,6) rows [points | normals] to the host in one copy
static int download_rows6_impl(o3dpp_ctx *c, void *dst, int dtype)
{
    if (!c->pts.present || !c->nrm.present) return set_err(O3DPP_E_STATE, "points and normals must be present");
    if (dtype != O3DPP_F32 && dtype != O3DPP_F64) return set_err(O3DPP_E_ARG, "dtype must be O3DPP_F32 or O3DPP_F64");
    int64_t n = c->n;
    if (n == 0) return 0;
    if (!dst) return set_err(O3DPP_E_ARG, "null destination");
    size_t esz = dtype == O3DPP_F32 ? 4 : 8;
    O3_TRY(c->stage.ensure((size_t)n * 6 * esz));
    int grid = grid_for(c, (n + 127) / 128, 1, 16);
    if (dtype == O3DPP_F32)
        LAUNCH(c, interleave6_kernel<float>, grid, 256, 0, c->pts.x(), c->pts.y(), c->pts.z(), c->nrm.x(), c->nrm.y(), c->nrm.z(), n, c->stage.as<float>());
    else
        LAUNCH(c, interleave6_kernel<double>, grid, 256, 0, c->pts.x(), c->pts.y(), c->pts.z(), c->nrm.x(), c->nrm.y(), c->nrm.z(), n, c->stage.as<double>());
    O3_TRY(check_launch("interleave6"));
    CU_TRY(cudaMemcpyAsync(dst, c->stage.p, (size_t)n * 6 * esz, cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

// device -> host copy of one attribute that returns at once: the rows are interleaved and copied on a second stream,
// ordered after everything already queued on the main stream.  The host buffer (page-locked for a real overlap) is
// complete, and the attribute may be modified again, after o3dpp_synchronize().
static int download_async_impl(o3dpp_ctx *c, int attr, double *dst)
{
    Attr3 *a = attr_of(c, attr);
    if (!a) return set_err(O3DPP_E_ARG, "unknown attribute");
    if (!a->present) return set_err(O3DPP_E_STATE, "attribute not present");
    int64_t n = c->n;
    if (n == 0) return 0;
    if (!dst) return set_err(O3DPP_E_ARG, "null destination");
    if (!c->copy_stream) {
        int lo = 0, hi = 0;                                  // the interleave pass must not queue behind another scan's
        CU_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi)); // 300 k-block neighbour search: highest priority
        CU_TRY(cudaStreamCreateWithPriority(&c->copy_stream, cudaStreamNonBlocking, hi));
        CU_TRY(cudaEventCreateWithFlags(&c->copy_ev, cudaEventDisableTiming));
    }
    CU_TRY(cudaStreamSynchronize(c->copy_stream));          // stage2 may still feed an earlier copy
    O3_TRY(c->stage2.ensure((size_t)n * 24));
    CU_TRY(cudaEventRecord(c->copy_ev, c->stream));
    CU_TRY(cudaStreamWaitEvent(c->copy_stream, c->copy_ev, 0));
    int grid = grid_for(c, (n + 255) / 256, 1, 16);
    interleave_kernel<double><<<grid, 256, 0, c->copy_stream>>>(a->x(), a->y(), a->z(), n, c->stage2.as<double>());
    c->launches++;
    O3_TRY(check_launch("interleave"));
    CU_TRY(cudaMemcpyAsync(dst, c->stage2.p, (size_t)n * 24, cudaMemcpyDeviceToHost, c->copy_stream));
    return 0;
}

// ---- bounds ------------------------------------------------------------------------------
// min/max are order independent, so any reduction tree equals Open3D's serial loop.
#define BND_THREADS 256
__global__ void __launch_bounds__(BND_THREADS)
bounds_partial_kernel(const double *__restrict__ x, const double *__restrict__ y,
                      const double *__restrict__ z, int64_t n, double *__restrict__ part /*grid*6*/)
{
    double mn[3], mx[3];
    const double *a[3] = {x, y, z};
    int64_t i0 = (int64_t)blockIdx.x * BND_THREADS + threadIdx.x;
    int64_t first = i0 < n ? i0 : 0;  // every thread starts from a real element
#pragma unroll
    for (int k = 0; k < 3; ++k) mn[k] = mx[k] = a[k][first];
    for (int64_t i = i0; i < n; i += (int64_t)gridDim.x * BND_THREADS) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            double v = a[k][i];
            if (v < mn[k]) mn[k] = v;
            if (v > mx[k]) mx[k] = v;
        }
    }
    __shared__ double sh[BND_THREADS / 32][6];
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
        for (int k = 0; k < 3; ++k) { sh[threadIdx.x >> 5][k] = mn[k]; sh[threadIdx.x >> 5][3 + k] = mx[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        int k = threadIdx.x;
        double v = sh[0][k];
        for (int w = 1; w < BND_THREADS / 32; ++w) {
            double o = sh[w][k];
            if (k < 3 ? o < v : o > v) v = o;
        }
        part[blockIdx.x * 6 + k] = v;
    }
}

// one block of 6 warps: warp k reduces column k of the partial rows (a single thread walking them cost 42 us at 592 rows)
__global__ void __launch_bounds__(192)
bounds_final_kernel(const double *__restrict__ part, int blocks, double *__restrict__ out6)
{
    const int k = threadIdx.x >> 5, lane = lane_id();
    double v = part[k];
    for (int b = lane; b < blocks; b += 32) {
        double o = part[b * 6 + k];
        if (k < 3 ? o < v : o > v) v = o;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        double o = shfl_xor_d(v, d);
        if (k < 3 ? o < v : o > v) v = o;
    }
    if (lane == 0) out6[k] = v;
}

// leaves min/max (6 doubles) at d_small[0..5] on the device, and a host copy in mn/mx if non-null
static int bounds_impl(o3dpp_ctx *c, double *mn, double *mx)
{
    int64_t n = c->n;
    if (n == 0) {
        if (mn) mn[0] = mn[1] = mn[2] = 0.0;
        if (mx) mx[0] = mx[1] = mx[2] = 0.0;
        CU_TRY(cudaMemsetAsync(c->d_small.p, 0, 6 * sizeof(double), c->stream));
        return 0;
    }
    int grid = grid_for(c, n, BND_THREADS * 4, 4);
    double *part = c->d_small.as<double>() + 16;  // <= 148*4*6 doubles = 28 KiB
    LAUNCH(c, bounds_partial_kernel, grid, BND_THREADS, 0, c->pts.x(), c->pts.y(), c->pts.z(), n, part);
    LAUNCH(c, bounds_final_kernel, 1, 192, 0, part, grid, c->d_small.as<double>());
    O3_TRY(check_launch("bounds"));
    if (mn || mx) {
        CU_TRY(cudaMemcpyAsync(c->h_small, c->d_small.p, 6 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(cudaStreamSynchronize(c->stream));
        const double *h = (const double *)c->h_small;
        if (mn) { mn[0] = h[0]; mn[1] = h[1]; mn[2] = h[2]; }
        if (mx) { mx[0] = h[3]; mx[1] = h[4]; mx[2] = h[5]; }
    }
    return 0;
}

// ---- transform ---------------------------------------------------------------------------
// Geometry3D::TransformPoints / TransformNormals: r = ((T0*x + T1*y) + T2*z) + T3*w, FP64, no FMA.
// HBM bound: 48 B per point (24 read + 24 written), 96 B with normals.
struct Mat4 { double m[16]; };
#ifndef TR_UNR
#define TR_UNR 1
#endif

__device__ __forceinline__ double row_dot(const double *r, double x, double y, double z, double w)
{
    return __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(r[0], x), __dmul_rn(r[1], y)), __dmul_rn(r[2], z)),
                     __dmul_rn(r[3], w));
}

// point = (T [p,1]).xyz / w.  With an affine T (last row exactly 0 0 0 1) w is 1.0 for every finite point and r / 1.0 == r
// bit for bit, so the three FP64 divisions (~30 instructions each, the kernel is FP64-issue bound with them) are skipped
// unless w != 1.0 (a non-finite coordinate makes w NaN: that point takes the division and gets Open3D's NaN).
__device__ __forceinline__ double persp(double r, double w) { return w == 1.0 ? r : __ddiv_rn(r, w); }

__global__ void __launch_bounds__(256)
transform_kernel(double *__restrict__ x, double *__restrict__ y, double *__restrict__ z, int64_t n,
                 const __grid_constant__ Mat4 T, int is_normal)
{
    // two points per 16-byte access (planes are 256 B aligned) and TR_UNR accesses per plane and step: the loads of a
    // step are issued together (3 TR_UNR x 16 B in flight per thread), the kernel is bound by HBM latency x bytes in flight
    int64_t n2 = n >> 1;
    double2 *x2 = (double2 *)x, *y2 = (double2 *)y, *z2 = (double2 *)z;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n2; i0 += TR_UNR * stride) {
        double2 a[TR_UNR], b[TR_UNR], cc[TR_UNR];
#pragma unroll
        for (int u = 0; u < TR_UNR; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < n2) { a[u] = x2[i]; b[u] = y2[i]; cc[u] = z2[i]; }
        }
#pragma unroll
        for (int u = 0; u < TR_UNR; ++u) {
            const int64_t i = i0 + u * stride;
            if (i >= n2) break;
            double2 ox, oy, oz;
            if (is_normal) {
                ox.x = row_dot(T.m + 0, a[u].x, b[u].x, cc[u].x, 0.0); ox.y = row_dot(T.m + 0, a[u].y, b[u].y, cc[u].y, 0.0);
                oy.x = row_dot(T.m + 4, a[u].x, b[u].x, cc[u].x, 0.0); oy.y = row_dot(T.m + 4, a[u].y, b[u].y, cc[u].y, 0.0);
                oz.x = row_dot(T.m + 8, a[u].x, b[u].x, cc[u].x, 0.0); oz.y = row_dot(T.m + 8, a[u].y, b[u].y, cc[u].y, 0.0);
            } else {
                double w0 = row_dot(T.m + 12, a[u].x, b[u].x, cc[u].x, 1.0), w1 = row_dot(T.m + 12, a[u].y, b[u].y, cc[u].y, 1.0);
                ox.x = persp(row_dot(T.m + 0, a[u].x, b[u].x, cc[u].x, 1.0), w0);
                ox.y = persp(row_dot(T.m + 0, a[u].y, b[u].y, cc[u].y, 1.0), w1);
                oy.x = persp(row_dot(T.m + 4, a[u].x, b[u].x, cc[u].x, 1.0), w0);
                oy.y = persp(row_dot(T.m + 4, a[u].y, b[u].y, cc[u].y, 1.0), w1);
                oz.x = persp(row_dot(T.m + 8, a[u].x, b[u].x, cc[u].x, 1.0), w0);
                oz.y = persp(row_dot(T.m + 8, a[u].y, b[u].y, cc[u].y, 1.0), w1);
            }
            x2[i] = ox; y2[i] = oy; z2[i] = oz;
        }
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
        int64_t i = n - 1;
        double a = x[i], b = y[i], cc = z[i];
        if (is_normal) {
            x[i] = row_dot(T.m + 0, a, b, cc, 0.0);
            y[i] = row_dot(T.m + 4, a, b, cc, 0.0);
            z[i] = row_dot(T.m + 8, a, b, cc, 0.0);
        } else {
            double w = row_dot(T.m + 12, a, b, cc, 1.0);
            x[i] = persp(row_dot(T.m + 0, a, b, cc, 1.0), w);
            y[i] = persp(row_dot(T.m + 4, a, b, cc, 1.0), w);
            z[i] = persp(row_dot(T.m + 8, a, b, cc, 1.0), w);
        }
    }
}

static int transform_impl(o3dpp_ctx *c, const double *T16, int what = 3)      // what: 1 = points, 2 = normals, 3 = both
{
    if (!T16) return set_err(O3DPP_E_ARG, "null matrix");
    Mat4 T;
    memcpy(T.m, T16, sizeof T.m);
    int64_t n = c->n;
    if (n == 0) return 0;
    invalidate_derived(c);
    int grid = grid_for(c, (n + 1) / 2, 256, 8);
    cudaEventRecord(c->ev[0], c->stream);
    if (what & 1) LAUNCH(c, transform_kernel, grid, 256, 0, c->pts.x(), c->pts.y(), c->pts.z(), n, T, 0);
    if ((what & 2) && c->nrm.present)
        LAUNCH(c, transform_kernel, grid, 256, 0, c->nrm.x(), c->nrm.y(), c->nrm.z(), n, T, 1);
    cudaEventRecord(c->ev[1], c->stream);
    return check_launch("transform");
}

// ---- selection -----------------------------------------------------------------------------
__global__ void mark_kernel(const int32_t *__restrict__ idx, int64_t m, uint8_t *__restrict__ flags, int64_t n)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
        int32_t j = idx[i];
        if (j >= 0 && j < n) flags[j] = 1;
    }
}

struct FlagPred {
    const uint8_t *flags;
    int want;
    __device__ bool operator()(int64_t i) const { return (flags[i] != 0) == (want != 0); }
};

__global__ void gather3_kernel(const int32_t *__restrict__ idx, int64_t m, const double *__restrict__ sx,
                               const double *__restrict__ sy, const double *__restrict__ sz,
                               double *__restrict__ dx, double *__restrict__ dy, double *__restrict__ dz)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
        int32_t j = idx[i];
        dx[i] = sx[j]; dy[i] = sy[j]; dz[i] = sz[j];
    }
}

// keep the points listed in d_sel (ascending, m entries, device): gathers every present attribute
static int apply_selection(o3dpp_ctx *c, const int32_t *d_sel, int64_t m)
{
    Attr3 *as[3] = {&c->pts, &c->nrm, &c->col};
    int64_t st = plane_stride(m);
    for (int k = 0; k < 3; ++k) {
        Attr3 *a = as[k];
        if (!a->present) continue;
        DevBuf nb;
        O3_TRY(nb.ensure((size_t)st * 3 * sizeof(double)));
        double *d = nb.as<double>();
        if (m > 0) {
            int grid = grid_for(c, m, 256, 16);
            LAUNCH(c, gather3_kernel, grid, 256, 0, d_sel, m, a->x(), a->y(), a->z(), d, d + st, d + 2 * st);
            O3_TRY(check_launch("gather3"));
        }
        CU_TRY(cudaStreamSynchronize(c->stream));
        a->buf.release();
        a->buf = nb;
        a->stride = st;
    }
    c->n = m;
    invalidate_derived(c);
    return 0;
}

static int select_impl(o3dpp_ctx *c, const int32_t *d_idx, int64_t m, int invert)
{
    int64_t n = c->n;
    DevBuf &flags = c->sel_flags, &sel = c->sel_idx;
    O3_TRY(flags.ensure((size_t)n + 64));
    O3_TRY(sel.ensure(((size_t)n + 64) * sizeof(int32_t)));
    int rc = 0;
    int64_t cnt = 0;
    do {
        if (cudaMemsetAsync(flags.p, 0, (size_t)n + 64, c->stream) != cudaSuccess) { rc = set_err(O3DPP_E_CUDA, "memset failed"); break; }
        if (m > 0) {
            LAUNCH(c, mark_kernel, grid_for(c, m, 256, 16), 256, 0, d_idx, m, flags.as<uint8_t>(), n);
        }
        FlagPred pred{flags.as<uint8_t>(), invert ? 0 : 1};
        rc = compact_indices(c, pred, n, sel.as<int32_t>(), &cnt);
        if (rc) break;
        rc = apply_selection(c, sel.as<int32_t>(), cnt);
    } while (0);
    cudaStreamSynchronize(c->stream);
    return rc;
}
