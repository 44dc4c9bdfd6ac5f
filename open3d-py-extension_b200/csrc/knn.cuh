// knn.cuh — neighbour search and normal estimation.
// Replaces pcd.estimate_normals(param) (reference PointCloud.py:68-73; Open3D
// PointCloud::EstimateNormals + KDTreeFlann::Search* + ComputeCovariance + FastEigen3x3,
// SURVEY.md §3.2) and the batched form of the KD-tree queries at PointCloud.py:148-163.
//
// Search structure ("multi-resolution uniform grid"): points are keyed by the Hilbert index of
// their cell in a 2^b-per-axis uniform grid over the bounding box and radix sorted, so every
// cell of every coarser grid level is one contiguous range.  Consecutive groups of 32 sorted
// points form the finest cell-range table ("cell groups"); 32 groups form a node of the next
// table, and so on, each entry carrying the FP64 bounding box of its range.  LiDAR density
// varies by >1000x over a scan, which a single-resolution grid cannot serve; the range tables
// adapt because they are built over the sorted points, not over space.
//
// Search: one warp owns one cell group and answers its 32 queries one after another.
//   - the 32 lanes hold the sorted candidate list (rank r lives in slot r/32 of lane r%32),
//     keyed by (d2, original index): FP64 d2 = (dx*dx+dy*dy)+dz*dz unfused, as nanoflann;
//   - the list is seeded by a bitonic sort of the query's own cell group;
//   - the range tables are walked top-down: each lane tests one child box against the
//     current k-th distance (ballot), surviving groups are loaded one point per lane
//     (coalesced 256 B per plane) and merged by warp-shuffle insertion;
//   - results of the 32 queries are parked in shared memory; then every lane finishes ONE
//     query: covariance in (d2,index) order (the order Open3D sums in), closed-form
//     smallest eigenvector, sign alignment, scattered store to the original index.
#pragma once
#include "voxel.cuh"
#include "hilbert_lut.cuh"
#include <stdlib.h>

#define KNN_MAX_LEVELS 6
#ifndef KNN_NEAREST_FIRST
#define KNN_NEAREST_FIRST 1
#endif
#ifndef KNN_BOXQ
#define KNN_BOXQ 1
#endif

struct GridDesc {
    int64_t n;
    int levels;                       // number of range tables; table 0 = cell groups of 32 points
    int64_t cnt[KNN_MAX_LEVELS];      // entries per table
    int64_t off[KNN_MAX_LEVELS];      // offset (doubles) of the table's 6 box planes in `box`
    int64_t pitch[KNN_MAX_LEVELS];    // plane pitch (doubles)
    const double *sx, *sy, *sz;       // sorted points
    const double4 *sp4;               // the same points as (x, y, z, 0) records: one 32-byte sector per gather
    const uint32_t *sidx;             // sorted position -> original index
    const double *box;
    const float4 *boxq;               // the same boxes in FP32, rounded outwards (lo down, hi up): conservative culling;
                                      // two records per entry {lo.x lo.y lo.z hi.x} {hi.y hi.z - -}, entry e of table l at qoff[l] + e
    int qoff[KNN_MAX_LEVELS];
    int64_t qfirst;                   // queries = the points with original index < qfirst (n: all); O3DPP_OPT_QUERY_FIRST
};

// ---- build -------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t spread3(uint64_t v)
{
    v &= 0x1fffffull;
    v = (v | (v << 32)) & 0x1f00000000ffffull;
    v = (v | (v << 16)) & 0x1f0000ff0000ffull;
    v = (v | (v << 8)) & 0x100f00f00f00f00full;
    v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}

// Hilbert index of a cell, `bits` bits per axis (the curve of Skilling's transpose algorithm; tools/gen_hilbert_lut.py holds the
// bit-twiddling form and checks the table against it).  Unlike the Morton code, consecutive cells are always face neighbours,
// so 32 consecutive sorted points form a compact blob: measured on the 1 M-point LiDAR-like scan, the number of cell groups
// within half a box diagonal of a group drops from 124 (Morton) to 43 (tools/knn_leaf_study.py).
// The index comes from the generated state machine (hilbert_lut.cuh): Morton code, then two octants per table lookup.
// ~110 instructions per point against ~400 of the bit-twiddling form, which made curve_key_kernel issue bound (62 % of the
// issue slots, 0.21 ms at 10 M for a 0.10 ms HBM pass, ncu r02y)
__device__ __forceinline__ uint64_t hilbert3_lut(uint32_t x, uint32_t y, uint32_t z, int bits, const unsigned char *lut1,
                                                 const unsigned short *lut2)
{
    const uint64_t m = (spread3(x) << 2) | (spread3(y) << 1) | spread3(z);
    unsigned s = 0;
    uint64_t h = 0;
    int k = bits;
    if (k & 1) {
        --k;
        const unsigned e = lut1[(unsigned)(m >> (3 * k)) & 7u];
        h = e & 7u;
        s = e >> 3;
    }
    while (k > 0) {
        k -= 2;
        const unsigned e = lut2[s * 64u + ((unsigned)(m >> (3 * k)) & 63u)];
        h = (h << 6) | (e & 63u);
        s = e >> 6;
    }
    return h;
}

__global__ void __launch_bounds__(256)
curve_key_kernel(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                  int64_t n, const double *__restrict__ mnmx, int bits, int pack /* index bits below the key, 0 = none */,
                  uint64_t *__restrict__ keys, double4 *__restrict__ p4)
{
    __shared__ unsigned short lut2[HILBERT_STATES * 64];
    __shared__ unsigned char lut1[8];
    for (int t = threadIdx.x; t < HILBERT_STATES * 64; t += blockDim.x) lut2[t] = HILBERT_LUT2[t];
    if (threadIdx.x < 8) lut1[threadIdx.x] = HILBERT_LUT1[threadIdx.x];          // state 0 only: the first step
    __syncthreads();
    double m0 = mnmx[0], m1 = mnmx[1], m2 = mnmx[2];
    double ext = fmax(fmax(mnmx[3] - m0, mnmx[4] - m1), mnmx[5] - m2);
    double cells = (double)(1u << bits);
    double scale = ext > 0.0 ? cells / (ext * 1.000000001) : 0.0;
    uint32_t cmax = (1u << bits) - 1u;
    // two points per thread and step: six loads in flight per thread (HBM bound: 24 B read, 40 B written per point)
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 2 * stride) {
        const int64_t i2 = i + stride < n ? i + stride : i;
        double px[2] = {x[i], x[i2]}, py[2] = {y[i], y[i2]}, pz[2] = {z[i], z[i2]};
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            if (u == 1 && i2 == i) break;
            double fx = (px[u] - m0) * scale, fy = (py[u] - m1) * scale, fz = (pz[u] - m2) * scale;
            uint32_t cx = fx > 0.0 ? (fx < cells ? (uint32_t)fx : cmax) : 0u;   // NaN lands in cell 0
            uint32_t cy = fy > 0.0 ? (fy < cells ? (uint32_t)fy : cmax) : 0u;
            uint32_t cz = fz > 0.0 ? (fz < cells ? (uint32_t)fz : cmax) : 0u;
            if (cx > cmax) cx = cmax;
            if (cy > cmax) cy = cmax;
            if (cz > cmax) cz = cmax;
            const int64_t j = u ? i2 : i;
            const uint64_t hk = hilbert3_lut(cx, cy, cz, bits, lut1, lut2);
            keys[j] = pack ? (hk << pack) | (uint64_t)j : hk;
            p4[j] = make_double4(px[u], py[u], pz[u], 0.0);  // one 32-byte record per point: the gather below reads one sector, not three
        }
    }
}

__device__ __forceinline__ double warp_min_d(double v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { double o = shfl_xor_d(v, d); v = o < v ? o : v; }
    return v;
}
__device__ __forceinline__ double warp_max_d(double v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { double o = shfl_xor_d(v, d); v = o > v ? o : v; }
    return v;
}

// sorted copies of the points AND table 0 of the range tables: a warp holds the 32 points of one cell group, so their box
// costs twelve shuffle reductions here instead of a kernel of its own that reads the sorted planes back (0.11 ms at 10 M)
__global__ void __launch_bounds__(256)
gather_sorted_kernel(const uint32_t *__restrict__ sidx, int64_t n, const double4 *__restrict__ p4, double *__restrict__ sx,
                     double *__restrict__ sy, double *__restrict__ sz, double4 *__restrict__ sp4, const GridDesc g,
                     double *__restrict__ box, float4 *__restrict__ boxq)
{
    const int lane = lane_id();
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const double INF = pos_inf();
    for (int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; grp < g.cnt[0]; grp += nw) {
        const int64_t i = grp * 32 + lane;
        double lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
        if (i < n) {
            const double4 q = p4[sidx[i]];
            sx[i] = q.x; sy[i] = q.y; sz[i] = q.z;
            sp4[i] = make_double4(q.x, q.y, q.z, 0.0);
            lo[0] = hi[0] = q.x; lo[1] = hi[1] = q.y; lo[2] = hi[2] = q.z;
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) { lo[k] = warp_min_d(lo[k]); hi[k] = warp_max_d(hi[k]); }
        if (lane == 0) {
            double *b = box + g.off[0];
            const int64_t p = g.pitch[0];
            b[grp] = lo[0]; b[p + grp] = lo[1]; b[2 * p + grp] = lo[2];
            b[3 * p + grp] = hi[0]; b[4 * p + grp] = hi[1]; b[5 * p + grp] = hi[2];
            float4 *bq = boxq + 2 * ((int64_t)g.qoff[0] + grp);
            bq[0] = make_float4(__double2float_rd(lo[0]), __double2float_rd(lo[1]), __double2float_rd(lo[2]), __double2float_ru(hi[0]));
            bq[1] = make_float4(__double2float_ru(hi[1]), __double2float_ru(hi[2]), 0.f, 0.f);
        }
    }
}

// one warp per entry of table `lvl` >= 1 (table 0 is built by gather_sorted_kernel); children are entries of table lvl-1
__global__ void __launch_bounds__(256)
box_level_kernel(GridDesc g, int lvl, double *__restrict__ box, float4 *__restrict__ boxq)
{
    int64_t node = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (node >= g.cnt[lvl]) return;
    int lane = lane_id();
    const double INF = pos_inf();
    double lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
    int64_t ch = node * 32 + lane;
    if (ch < g.cnt[lvl - 1]) {
        const double *b = box + g.off[lvl - 1];
        int64_t p = g.pitch[lvl - 1];
        lo[0] = b[ch]; lo[1] = b[p + ch]; lo[2] = b[2 * p + ch];
        hi[0] = b[3 * p + ch]; hi[1] = b[4 * p + ch]; hi[2] = b[5 * p + ch];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) { lo[k] = warp_min_d(lo[k]); hi[k] = warp_max_d(hi[k]); }
    if (lane == 0) {
        double *b = box + g.off[lvl];
        int64_t p = g.pitch[lvl];
        b[node] = lo[0]; b[p + node] = lo[1]; b[2 * p + node] = lo[2];
        b[3 * p + node] = hi[0]; b[4 * p + node] = hi[1]; b[5 * p + node] = hi[2];
        float4 *bq = boxq + 2 * ((int64_t)g.qoff[lvl] + node);
        bq[0] = make_float4(__double2float_rd(lo[0]), __double2float_rd(lo[1]), __double2float_rd(lo[2]), __double2float_ru(hi[0]));
        bq[1] = make_float4(__double2float_ru(hi[1]), __double2float_ru(hi[2]), 0.f, 0.f);
    }
}

// ---- closed-form smallest eigenvector (Open3D FastEigen3x3) ---------------
// C = [c00 c01 c02 c11 c12 c22].  Written with plain operators: the TU is compiled -fmad=false.
__device__ __forceinline__ void cross3d(const double *a, const double *b, double *o)
{
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ double dot3d(const double *a, const double *b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }

// FP64 division and square root out of line (IEEE results, same bits as the inline expansion): the normal solve has
// ~30 of them at ~25 instructions each, and the kernels that finish queries are bound by instruction fetch
#ifndef O3_NIDIV
#define O3_NIDIV 1
#endif
#if O3_NIDIV
__device__ __noinline__ double o3_div(double a, double b) { return a / b; }
__device__ __noinline__ double o3_sqrt(double a) { return sqrt(a); }
__device__ __noinline__ double o3_cos(double a) { return cos(a); }
#else
__device__ __forceinline__ double o3_div(double a, double b) { return a / b; }
__device__ __forceinline__ double o3_sqrt(double a) { return sqrt(a); }
__device__ __forceinline__ double o3_cos(double a) { return cos(a); }
#endif

__device__ __forceinline__ void eig_vec0(const double *A, double ev, double *out)
{
    double r0[3] = {A[0] - ev, A[1], A[2]};
    double r1[3] = {A[1], A[3] - ev, A[4]};
    double r2[3] = {A[2], A[4], A[5] - ev};
    double x01[3], x02[3], x12[3];
    cross3d(r0, r1, x01); cross3d(r0, r2, x02); cross3d(r1, r2, x12);
    double d0 = dot3d(x01, x01), d1 = dot3d(x02, x02), d2 = dot3d(x12, x12);
    double dmax = d0;
    int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) { imax = 2; }
    double sx = imax == 0 ? x01[0] : (imax == 1 ? x02[0] : x12[0]);
    double sy = imax == 0 ? x01[1] : (imax == 1 ? x02[1] : x12[1]);
    double sz = imax == 0 ? x01[2] : (imax == 1 ? x02[2] : x12[2]);
    double dd = imax == 0 ? d0 : (imax == 1 ? d1 : d2);
    double s = o3_sqrt(dd);
    out[0] = o3_div(sx, s); out[1] = o3_div(sy, s); out[2] = o3_div(sz, s);
}

__device__ __forceinline__ void eig_vec1(const double *A, const double *e0, double ev1, double *out)
{
    double U[3], V[3];
    if (fabs(e0[0]) > fabs(e0[1])) {
        double il = o3_div(1.0, o3_sqrt(e0[0] * e0[0] + e0[2] * e0[2]));
        U[0] = -e0[2] * il; U[1] = 0.0; U[2] = e0[0] * il;
    } else {
        double il = o3_div(1.0, o3_sqrt(e0[1] * e0[1] + e0[2] * e0[2]));
        U[0] = 0.0; U[1] = e0[2] * il; U[2] = -e0[1] * il;
    }
    cross3d(e0, U, V);
    double AU[3] = {(A[0] * U[0] + A[1] * U[1]) + A[2] * U[2], (A[1] * U[0] + A[3] * U[1]) + A[4] * U[2],
                    (A[2] * U[0] + A[4] * U[1]) + A[5] * U[2]};
    double AV[3] = {(A[0] * V[0] + A[1] * V[1]) + A[2] * V[2], (A[1] * V[0] + A[3] * V[1]) + A[4] * V[2],
                    (A[2] * V[0] + A[4] * V[1]) + A[5] * V[2]};
    double m00 = dot3d(U, AU) - ev1;
    double m01 = dot3d(U, AV);
    double m11 = dot3d(V, AV) - ev1;
    double a00 = fabs(m00), a01 = fabs(m01), a11 = fabs(m11);
    if (a00 >= a11) {
        double mx = a00 > a01 ? a00 : a01;
        if (mx > 0) {
            if (a00 >= a01) { m01 = o3_div(m01, m00); m00 = o3_div(1.0, o3_sqrt(1.0 + m01 * m01)); m01 *= m00; }
            else            { m00 = o3_div(m00, m01); m01 = o3_div(1.0, o3_sqrt(1.0 + m00 * m00)); m00 *= m01; }
            for (int a = 0; a < 3; ++a) out[a] = m01 * U[a] - m00 * V[a];
        } else { for (int a = 0; a < 3; ++a) out[a] = U[a]; }
    } else {
        double mx = a11 > a01 ? a11 : a01;
        if (mx > 0) {
            if (a11 >= a01) { m01 = o3_div(m01, m11); m11 = o3_div(1.0, o3_sqrt(1.0 + m01 * m01)); m01 *= m11; }
            else            { m11 = o3_div(m11, m01); m01 = o3_div(1.0, o3_sqrt(1.0 + m11 * m11)); m11 *= m01; }
            for (int a = 0; a < 3; ++a) out[a] = m11 * U[a] - m01 * V[a];
        } else { for (int a = 0; a < 3; ++a) out[a] = U[a]; }
    }
}

// sep (optional): distance between the two smallest eigenvalues in units of the largest coefficient of C, i.e. how
// well the returned direction is determined (0 when C is zero)
__device__ __forceinline__ void fast_eigen3x3(const double *C, double *out, double *sep = nullptr)
{
    if (sep) *sep = 0.0;
    double mc = C[0];
#pragma unroll
    for (int a = 1; a < 6; ++a) if (C[a] > mc) mc = C[a];
    if (mc == 0.0) { out[0] = out[1] = out[2] = 0.0; return; }
    double A[6];
#pragma unroll
    for (int a = 0; a < 6; ++a) A[a] = o3_div(C[a], mc);
    double norm = (A[1] * A[1] + A[2] * A[2]) + A[4] * A[4];
    if (norm > 0.0) {
        double q = o3_div((A[0] + A[3]) + A[5], 3.0);
        double b00 = A[0] - q, b11 = A[3] - q, b22 = A[5] - q;
        double p = o3_sqrt(o3_div(((b00 * b00 + b11 * b11) + b22 * b22) + norm * 2.0, 6.0));
        double c00 = b11 * b22 - A[4] * A[4];
        double c01 = A[1] * b22 - A[4] * A[2];
        double c02 = A[1] * A[4] - b11 * A[2];
        double det = o3_div((b00 * c00 - A[1] * c01) + A[2] * c02, p * p * p);
        double half_det = det * 0.5;
        if (half_det < -1.0) half_det = -1.0;
        if (half_det > 1.0) half_det = 1.0;
        double angle = o3_div(acos(half_det), 3.0);
        const double two_thirds_pi = 2.09439510239319549;
        double beta2 = o3_cos(angle) * 2.0;
        double beta0 = o3_cos(angle + two_thirds_pi) * 2.0;
        double beta1 = -(beta0 + beta2);
        double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;
        if (sep) *sep = p * (beta1 - beta0);      // angle in [0, pi/3]: beta0 <= beta1 <= beta2, so e0 <= e1 <= e2
        // one copy of each solve (the kernels that inline this are instruction-cache bound): the extreme eigenvalue on the
        // well-conditioned side first (e2 when half_det >= 0, else e0), then e1, the third by the cross product
        const bool top = half_det >= 0.0;
        const double ea = top ? e2 : e0, eo = top ? e0 : e2;
        double va[3], v1[3];
        eig_vec0(A, ea, va);
        if (ea < eo && ea < e1) { out[0] = va[0]; out[1] = va[1]; out[2] = va[2]; return; }
        eig_vec1(A, va, e1, v1);
        if (e1 < e0 && e1 < e2) { out[0] = v1[0]; out[1] = v1[1]; out[2] = v1[2]; return; }
        double l[3], r[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) { l[a] = top ? v1[a] : va[a]; r[a] = top ? va[a] : v1[a]; }
        cross3d(l, r, out);
    } else {
        if (sep) {
            const double lo = fmin(A[0], fmin(A[3], A[5])), hi = fmax(A[0], fmax(A[3], A[5]));
            *sep = ((A[0] + A[3] + A[5]) - lo - hi) - lo;
        }
        if (C[0] < C[3] && C[0] < C[5]) { out[0] = 1; out[1] = 0; out[2] = 0; }
        else if (C[3] < C[0] && C[3] < C[5]) { out[0] = 0; out[1] = 1; out[2] = 0; }
        else { out[0] = 0; out[1] = 0; out[2] = 1; }
    }
}

// single-pass cumulants -> covariance (Open3D utility::ComputeCovariance)
struct Cumul {
    double c[9];
    int n;
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int a = 0; a < 9; ++a) c[a] = 0.0;
        n = 0;
    }
    __device__ __forceinline__ void add(double x, double y, double z)
    {
        c[0] += x; c[1] += y; c[2] += z;
        c[3] += x * x; c[4] += x * y; c[5] += x * z;
        c[6] += y * y; c[7] += y * z; c[8] += z * z;
        n++;
    }
    __device__ __forceinline__ void cov(double *C) const
    {
        if (n < 3) { C[0] = 1; C[1] = 0; C[2] = 0; C[3] = 1; C[4] = 0; C[5] = 1; return; }
        double m[9], dn = (double)n;
#pragma unroll
        for (int a = 0; a < 9; ++a) m[a] = o3_div(c[a], dn);
        C[0] = m[3] - m[0] * m[0];
        C[1] = m[4] - m[0] * m[1];
        C[2] = m[5] - m[0] * m[2];
        C[3] = m[6] - m[1] * m[1];
        C[4] = m[7] - m[1] * m[2];
        C[5] = m[8] - m[2] * m[2];
    }
};

// ---- the search kernel ---------------------------------------------------------------------
struct KnnOut {
    double *nx, *ny, *nz;       // normals by original index (nullable: search only)
    int has_old_normals;        // align with the previous content of nx/ny/nz
    int32_t *idx;               // n*K original indices (nullable)
    double *d2;                 // n*K (nullable)
    int32_t *cnt;               // n (nullable)
    unsigned long long *stats;  // debug counters (nullable): visits, groups loaded, inserts, candidates
    // tier chain of estimate_normals (knn_lane.cuh): the selected neighbours as sorted positions, [j * nbr_pitch + query
    // position], KNN_NONE beyond the count.  Tier 1 always writes them (knn_finish_kernel reads them back); with keep != 0
    // tiers 2 and 3 record theirs too and `tier` says which tier answered (o3dpp_fetch_neighbors)
    uint32_t *nbr;
    int64_t nbr_pitch;
    unsigned char *tier;
    int keep;
};

template <int S>
struct TopK {
    double d2[S];
    int pos[S];
};

__device__ __forceinline__ bool key_less(double a, int ap, double b, int bp, const uint32_t *__restrict__ sidx)
{
    if (a < b) return true;
    if (a > b) return false;
    if (ap < 0) return false;   // invalid entries (d2 = +inf) sort last
    if (bp < 0) return true;
    return sidx[ap] < sidx[bp];
}

template <int MODE, int S>
struct Search {
    const GridDesc &g;
    int lane;
    int K;
    double r2;
    double qx, qy, qz;
    TopK<S> L;
    double worst_d2;
    int worst_pos;
    double bound;        // prune boxes with lower bound > bound
    double cap;          // caller-known upper bound of the k-th d2 (+inf if none)
    int home;
    Cumul cum;           // MODE 1 only: lane-private partial sums
    unsigned st_visit, st_leaf, st_ins, st_cand;

    __device__ __forceinline__ Search(const GridDesc &g_) : g(g_) {}

    __device__ __forceinline__ void refresh_worst()
    {
        int wl = (K - 1) & 31;
        worst_d2 = shfl_d(L.d2[S - 1], wl);
        worst_pos = __shfl_sync(FULL, L.pos[S - 1], wl);
        bound = worst_d2;
        if (cap < bound) bound = cap;
        if (MODE == 2 && r2 < bound) bound = r2;
    }

    // insert (nd, np) — identical on all lanes — into the sorted list
    __device__ __forceinline__ void insert(double nd, int np)
    {
        int ins = 0;
#pragma unroll
        for (int s = 0; s < S; ++s) ins += __popc(__ballot_sync(FULL, key_less(L.d2[s], L.pos[s], nd, np, g.sidx)));
        double rd[S];
        int rp[S];
        int src = (lane + 31) & 31;
#pragma unroll
        for (int s = 0; s < S; ++s) { rd[s] = shfl_d(L.d2[s], src); rp[s] = __shfl_sync(FULL, L.pos[s], src); }
#pragma unroll
        for (int s = 0; s < S; ++s) {
            int r = s * 32 + lane;
            double sd = rd[s];
            int sp = rp[s];
            if (lane == 0 && s > 0) { sd = rd[s - 1]; sp = rp[s - 1]; }
            if (r == ins) { L.d2[s] = nd; L.pos[s] = np; }
            else if (r > ins) { L.d2[s] = sd; L.pos[s] = sp; }
        }
    }

    __device__ __forceinline__ void leaf(int64_t leaf_id)
    {
        int64_t p = leaf_id * 32 + lane;
        bool v = p < g.n;
        double d = pos_inf();
        double cx = 0, cy = 0, cz = 0;
        if (v) { const double4 c4 = ld_rec(g.sp4 + p); cx = c4.x; cy = c4.y; cz = c4.z; d = dist2_rn(qx, qy, qz, cx, cy, cz); }
        if (MODE == 1) {
            if (v && d < r2) cum.add(cx, cy, cz);
            return;
        }
        unsigned m = __ballot_sync(FULL, v && d <= bound && (MODE != 2 || d < r2));
        st_leaf++; st_cand += __popc(m);
        while (m) {
            int s = __ffs(m) - 1;
            m &= m - 1;
            double nd = shfl_d(d, s);
            int np = (int)(leaf_id * 32 + s);
            if (!key_less(nd, np, worst_d2, worst_pos, g.sidx)) continue;
            insert(nd, np);
            st_ins++;
            refresh_worst();
            m &= __ballot_sync(FULL, d <= bound);
        }
    }

    // lower bound of the squared distance from the query to box c of table lvl.  KNN_BOXQ: from the FP32 record (two loads,
    // FP32 arithmetic rounded down, boxes rounded outwards), never above the FP64 value, so nothing that matters is pruned
    __device__ __forceinline__ double box_lb(int lvl, int64_t c)
    {
#if KNN_BOXQ
        const float4 *e = g.boxq + 2 * (g.qoff[lvl] + (int)c);
        const float4 a = e[0];
        const float2 b = *(const float2 *)(e + 1);
        const float qxl = __double2float_rd(qx), qxh = __double2float_ru(qx), qyl = __double2float_rd(qy), qyh = __double2float_ru(qy),
                    qzl = __double2float_rd(qz), qzh = __double2float_ru(qz);
        float gx = fmaxf(fmaxf(__fsub_rd(a.x, qxh), __fsub_rd(qxl, a.w)), 0.f);
        float gy = fmaxf(fmaxf(__fsub_rd(a.y, qyh), __fsub_rd(qyl, b.x)), 0.f);
        float gz = fmaxf(fmaxf(__fsub_rd(a.z, qzh), __fsub_rd(qzl, b.y)), 0.f);
        return (double)__fadd_rd(__fadd_rd(__fmul_rd(gx, gx), __fmul_rd(gy, gy)), __fmul_rd(gz, gz));
#else
        const double *b = g.box + g.off[lvl];
        int64_t p = g.pitch[lvl];
        double gx = fmax(fmax(__dsub_rn(b[c], qx), __dsub_rn(qx, b[3 * p + c])), 0.0);
        double gy = fmax(fmax(__dsub_rn(b[p + c], qy), __dsub_rn(qy, b[4 * p + c])), 0.0);
        double gz = fmax(fmax(__dsub_rn(b[2 * p + c], qz), __dsub_rn(qz, b[5 * p + c])), 0.0);
        return __dadd_rn(__dadd_rn(__dmul_rn(gx, gx), __dmul_rn(gy, gy)), __dmul_rn(gz, gz));
#endif
    }

    template <int LVL>
    __device__ __forceinline__ void visit(int64_t parent)
    {
        int64_t c = parent * 32 + lane;
        bool ex = c < g.cnt[LVL];
        double lb = ex ? box_lb(LVL, c) : pos_inf();
        unsigned m = __ballot_sync(FULL, ex && lb <= bound);
        st_visit++;
#if KNN_NEAREST_FIRST
        const unsigned nkey = __float_as_uint(__double2float_rd(lb));      // non-negative: the bit patterns order like the values
#endif
        while (m) {
#if KNN_NEAREST_FIRST
            // nearest child first (MODE 1 sums everything inside the radius: the order is irrelevant there): the bound tightens
            // early and the far children, which a walk in index order would open and insert from, are never touched
            int b;
            if (MODE == 1) b = __ffs(m) - 1;
            else {
                const bool mine = (m >> lane) & 1u;
                const unsigned kmin = __reduce_min_sync(FULL, mine ? nkey : 0xffffffffu);
                b = __ffs(__ballot_sync(FULL, mine && nkey == kmin)) - 1;
            }
            m &= ~(1u << b);
#else
            int b = __ffs(m) - 1;
            m &= m - 1;
#endif
            int64_t child = parent * 32 + b;
            if constexpr (LVL == 0) {
                if (MODE == 1 || child != home) leaf(child);
            } else {
                visit<LVL - 1>(child);
            }
            if (MODE != 1) m &= __ballot_sync(FULL, lb <= bound);
        }
    }

    __device__ __forceinline__ void walk()
    {
        switch (g.levels - 1) {
        case 0: visit<0>(0); break;
        case 1: visit<1>(0); break;
        case 2: visit<2>(0); break;
        case 3: visit<3>(0); break;
        case 4: visit<4>(0); break;
        default: visit<5>(0); break;
        }
    }
};

template <int S>
__device__ __forceinline__ void bitonic_sort32(double &d, int &p, int lane, const uint32_t *__restrict__ sidx)
{
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            double od = shfl_xor_d(d, j);
            int op = __shfl_xor_sync(FULL, p, j);
            bool up = (lane & k) == 0 || k == 32;   // final merge ascending
            bool lower = (lane & j) == 0;
            bool mine_less = key_less(d, p, od, op, sidx);
            bool other_less = key_less(od, op, d, p, sidx);
            bool take_other = (lower == up) ? other_less : mine_less;
            if (take_other) { d = od; p = op; }
        }
    }
}

template <int MODE, int S, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
knn_normals_kernel(const GridDesc g, int K, double radius, KnnOut out)
{
    constexpr int ROW = 32 * S + 1;
    extern __shared__ int nb_all[];                 // [WARPS][32][ROW] neighbour positions
    __shared__ int cnt_all[WARPS][32];
    const int lane = lane_id(), w = threadIdx.x >> 5;
    const int64_t leaf_id = (int64_t)blockIdx.x * WARPS + w;
    if (leaf_id >= g.cnt[0]) return;
    int *nb = nb_all + (size_t)w * 32 * ROW;
    int *cnts = cnt_all[w];

    const int64_t base = leaf_id * 32;
    const int64_t mypos = base + lane;
    const bool valid = mypos < g.n;
    double px = 0, py = 0, pz = 0;
    if (valid) { px = g.sx[mypos]; py = g.sy[mypos]; pz = g.sz[mypos]; }
    const int nq = (int)(g.n - base < 32 ? g.n - base : 32);

    Search<MODE, S> st(g);
    st.lane = lane; st.K = K; st.r2 = __dmul_rn(radius, radius); st.home = (int)leaf_id; st.cap = pos_inf();
    Cumul mine;          // MODE 1: totals of the query this lane will finish
    mine.clear();
    st.st_visit = st.st_leaf = st.st_ins = st.st_cand = 0;

    for (int i = 0; i < nq; ++i) {
        st.qx = shfl_d(px, i); st.qy = shfl_d(py, i); st.qz = shfl_d(pz, i);
        if (MODE == 1) {
            st.cum.clear();
            st.bound = st.r2;
            st.walk();
            // warp totals (tree order; Open3D sums in d2 order: same set, rounding differs)
#pragma unroll
            for (int a = 0; a < 9; ++a) {
                double v = st.cum.c[a];
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) v += shfl_xor_d(v, d);
                if (lane == i) mine.c[a] = v;
            }
            int tn = st.cum.n;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) tn += __shfl_xor_sync(FULL, tn, d);
            if (lane == i) mine.n = tn;
            continue;
        }
        // seed: own cell group, bitonic sorted by (d2, index)
        double d = valid ? dist2_rn(st.qx, st.qy, st.qz, px, py, pz) : pos_inf();
        int p = valid ? (int)mypos : -1;
        bitonic_sort32<S>(d, p, lane, g.sidx);
        st.L.d2[0] = d; st.L.pos[0] = p;
#pragma unroll
        for (int s = 1; s < S; ++s) { st.L.d2[s] = pos_inf(); st.L.pos[s] = -1; }
        st.refresh_worst();
        st.walk();
        // park the result
        int c = 0;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            int r = s * 32 + lane;
            bool ok = r < K && st.L.pos[s] >= 0 && (MODE != 2 || st.L.d2[s] < st.r2);
            c += __popc(__ballot_sync(FULL, ok));
            nb[i * ROW + r] = st.L.pos[s];
        }
        if (lane == 0) cnts[i] = c;
        if (out.idx) {
            int64_t qo = g.sidx[base + i];
#pragma unroll
            for (int s = 0; s < S; ++s) {
                int r = s * 32 + lane;
                if (r < K) {
                    bool ok = r < c;
                    out.idx[qo * K + r] = ok ? (int32_t)g.sidx[st.L.pos[s]] : -1;
                    if (out.d2) out.d2[qo * K + r] = ok ? st.L.d2[s] : -1.0;
                }
            }
            if (out.cnt && lane == 0) out.cnt[qo] = c;
        }
    }
    __syncwarp();
    if (out.stats && lane == 0) {
        atomicAdd(out.stats + 0, (unsigned long long)st.st_visit);
        atomicAdd(out.stats + 1, (unsigned long long)st.st_leaf);
        atomicAdd(out.stats + 2, (unsigned long long)st.st_ins);
        atomicAdd(out.stats + 3, (unsigned long long)st.st_cand);
    }
    if (!valid || !out.nx) return;
    // every lane finishes one query
    double C[6];
    if (MODE == 1) {
        mine.cov(C);
    } else {
        Cumul cu;
        cu.clear();
        int c = cnts[lane];
        for (int j = 0; j < c; ++j) {
            int p = nb[lane * ROW + j];
            cu.add(g.sx[p], g.sy[p], g.sz[p]);
        }
        cu.cov(C);
    }
    double nv[3];
    fast_eigen3x3(C, nv);
    uint32_t qo = g.sidx[mypos];
    double nn = sqrt((nv[0] * nv[0] + nv[1] * nv[1]) + nv[2] * nv[2]);
    double ox = 0, oy = 0, oz = 0;
    if (out.has_old_normals) { ox = out.nx[qo]; oy = out.ny[qo]; oz = out.nz[qo]; }
    if (nn == 0.0) {
        if (out.has_old_normals) { nv[0] = ox; nv[1] = oy; nv[2] = oz; }
        else { nv[0] = 0; nv[1] = 0; nv[2] = 1; }
    }
    if (out.has_old_normals && ((nv[0] * ox + nv[1] * oy) + nv[2] * oz) < 0.0) { nv[0] = -nv[0]; nv[1] = -nv[1]; nv[2] = -nv[2]; }
    out.nx[qo] = nv[0]; out.ny[qo] = nv[1]; out.nz[qo] = nv[2];
}

// ---- host side -------------------------------------------------------------------------------
static int grid_build(o3dpp_ctx *c, GridDesc *gd)
{
    int64_t n = c->n;
    memset(gd, 0, sizeof *gd);
    gd->n = n;
    gd->qfirst = (c->tune.query_first > 0 && c->tune.query_first < n) ? c->tune.query_first : n;
    O3_TRY(bounds_impl(c, nullptr, nullptr));   // d_small[0..5] = min, max
    int bits = 4;
    { int64_t m = 1; int lg = 0; while (m < n) { m <<= 1; ++lg; } bits = (lg + 1) / 2 + 1; }   // ~2 n^(1/2) cells per axis: measured as good as finer grids, one radix pass fewer at 10 M
    if (bits < 8) bits = 8;
    if (bits > 21) bits = 21;
    if (c->tune.bits >= 4 && c->tune.bits <= 21) bits = c->tune.bits;   // tuning (O3DPP_OPT_KNN_BITS): curve cells per axis = 2^bits
    size_t nn = (size_t)n + 64;
    O3_TRY(c->s_keys0.ensure(nn * 8)); O3_TRY(c->s_keys1.ensure(nn * 8));
    O3_TRY(c->s_vals0.ensure(nn * 4)); O3_TRY(c->s_vals1.ensure(nn * 4));
    O3_TRY(c->g_sx.ensure(nn * 8)); O3_TRY(c->g_sy.ensure(nn * 8)); O3_TRY(c->g_sz.ensure(nn * 8));
    O3_TRY(c->g_sidx.ensure(nn * 4));
    O3_TRY(c->g_sp4.ensure(nn * 32));
    O3_TRY(c->g_p4.ensure(nn * 32));
    int grid = grid_for(c, n, 256, 16);
    // up to 16.7 M points with a key of at most 40 bits: the original index rides in the low 24 bits of the key and the sort
    // moves 8 bytes per element and pass instead of 12
    const int pack = (n <= ((int64_t)1 << RS_PACK_BITS) && 3 * bits + RS_PACK_BITS <= 64) ? RS_PACK_BITS : 0;
    LAUNCH(c, curve_key_kernel, grid, 256, 0, c->pts.x(), c->pts.y(), c->pts.z(), n, c->d_small.as<double>(), bits, pack,
           c->s_keys0.as<uint64_t>(), c->g_p4.as<double4>());
    uint64_t *ks; uint32_t *vs;
    if (pack) {
        O3_TRY(radix_sort_packed(c, c->s_keys0.as<uint64_t>(), c->s_keys1.as<uint64_t>(), n, pack, 3 * bits, &ks));
        LAUNCH(c, unpack_index_kernel, grid, 256, 0, ks, n, c->g_sidx.as<uint32_t>());
    } else {
        O3_TRY(radix_sort_pairs(c, c->s_keys0.as<uint64_t>(), c->s_keys1.as<uint64_t>(), c->s_vals0.as<uint32_t>(),
                                c->s_vals1.as<uint32_t>(), n, 3 * bits, &ks, &vs));
        CU_TRY(cudaMemcpyAsync(c->g_sidx.p, vs, (size_t)n * 4, cudaMemcpyDeviceToDevice, c->stream));
    }
    // range tables
    int L = 0;
    int64_t cnt = (n + 31) / 32, off = 0;
    for (;;) {
        gd->cnt[L] = cnt;
        gd->pitch[L] = ((cnt + 31) / 32) * 32;
        gd->off[L] = off;
        gd->qoff[L] = (int)(off / 6);
        off += gd->pitch[L] * 6;
        ++L;
        if (cnt <= 32 || L == KNN_MAX_LEVELS) break;
        cnt = (cnt + 31) / 32;
    }
    if (gd->cnt[L - 1] > 32) return set_err(O3DPP_E_UNSUPPORTED, "cloud too large for the range-table depth");
    gd->levels = L;
    O3_TRY(c->g_box.ensure((size_t)off * 8 + 256));
    if (off / 6 > 0x3fffffff) return set_err(O3DPP_E_UNSUPPORTED, "cloud too large for 32-bit range-table indices");
    O3_TRY(c->g_boxf.ensure((size_t)(off / 6) * 32 + 256));
    gd->sx = c->g_sx.as<double>(); gd->sy = c->g_sy.as<double>(); gd->sz = c->g_sz.as<double>();
    gd->sidx = c->g_sidx.as<uint32_t>();
    gd->sp4 = c->g_sp4.as<double4>();
    gd->box = c->g_box.as<double>();
    gd->boxq = c->g_boxf.as<float4>();
    // sorted copies + table 0 in one pass, then the upper tables
    LAUNCH(c, gather_sorted_kernel, grid, 256, 0, c->g_sidx.as<uint32_t>(), n, c->g_p4.as<double4>(),
           c->g_sx.as<double>(), c->g_sy.as<double>(), c->g_sz.as<double>(), c->g_sp4.as<double4>(), *gd,
           c->g_box.as<double>(), c->g_boxf.as<float4>());
    for (int l = 1; l < L; ++l) {
        int64_t warps = gd->cnt[l];
        int blocks = (int)((warps * 32 + 255) / 256);
        LAUNCH(c, box_level_kernel, blocks, 256, 0, *gd, l, c->g_box.as<double>(), c->g_boxf.as<float4>());
    }
    return check_launch("grid_build");
}
